// Device-side data layout shared by the kernels of the path (see DESIGN.md "Data layout in HBM").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "orfx.h"

namespace orfx {

// Flat view of one resident batch.  Arrays marked [in] are uploaded as given by the caller
// (orfx_batch); the others are derived on the device by prep kernels.
struct DevBatch {
    int32_t n_bundles, n_tx;
    int64_t n_segs, seq_bytes;   // seq_bytes counts bases (== bytes unless seq_format == ORFX_SEQ_NT4)
    int32_t seq_format;
    int32_t hb_max_segs, hb_max_pairs;   // <= HB_MAX_SEGS / HB_MAX_PAIRS: bundles above them take the heavy kernels' global-memory branches
    // [in]
    const uint8_t* seq;
    const int64_t* seq_off;
    const int32_t* bundle_start;
    const int32_t* bundle_tx_off;
    const uint8_t* tx_kind;
    const uint8_t* tx_strand;
    const int32_t* tx_seg_off;
    const int32_t* tx_cds_start;
    const int32_t* tx_cds_end;
    const int32_t* tx_tid_class;
    const orfx_seg* segs;
    const int8_t* seg_phase;
    const int32_t* seg_seq_off;   // islands layout (else nullptr)
    const int32_t* tx_ext_lo;     // --overhang: usable flank below the first / above the last exon of each query (else nullptr)
    const int32_t* tx_ext_hi;
    const int64_t* seq_exc_pos;   // ORFX_SEQ_NT2: bases that are none of A C G T (position, ORFX_NT4_* code)
    const uint8_t* seq_exc_code;
    int64_t n_seq_exc;
    const uint8_t* tx_ext_flags;  // ORFX_EXT_*_VIRTUAL per query (nullptr when no flank of the batch is a re-read edge base)
    // derived
    int32_t* seg_cum;        // [n_segs] bases of the transcript before this segment (genomic order)
    int32_t* tx_len;         // [n_tx] spliced length
    int32_t* tx_bundle;      // [n_tx]
    int64_t* tx_mask_off;    // [n_tx+1] word offset of the transcript's masks (queries only; 4-word aligned)
    int64_t* tx_tile_off;    // [n_tx+1] first codon-scan tile of the transcript (queries only)
    int64_t* tx_q_off;       // [n_tx+1] query ordinal (exclusive scan of is_query)
    int32_t* q_list;         // [n_tx capacity] transcript index of each query
    int64_t* q_row_off;      // [n_tx+1] first candidate row of each query (entries past n_queries = total)
    int32_t* tile_tx;        // [n_tiles] transcript of each codon-scan tile
    uint8_t* bundle_need;    // [n_bundles] bit 0: holds a '+' query, bit 1: holds a query on another strand
    uint32_t* planes;        // base planes of the window bytes, one uint4 per 32-byte chunk c and strand class:
                             //   ((uint4*)planes)[c] = {T,A,G,R} ('+'), [plane_words + c] = {A,T|U,C,Y} (reverse strand: T A G R)
    int64_t plane_words;     // chunks per strand class (bit i of a plane word = byte 32 c + i of `seq`)
    // codon-scan output: bit p of a transcript's region = codon starting at spliced position p (5'->3')
    uint32_t* stop_mask;
    uint32_t* start_mask;
    // alignment stage (params.align_enabled, translate.cu; else nullptr): residue code of the codon starting at every
    // position of every QUERY (flanked chain, the masks' coordinates) and TEMPLATE (spliced CDS chain) transcript
    int64_t* tx_aa_off;      // [n_tx+1] first byte of the transcript's region in aa_pos (16-byte aligned)
    uint8_t* aa_pos;
};

// Lookup tables of translate.cu: letter class of a sequence byte on the forward / on any other strand (upper-cased, and
// complemented first on the latter), residue code of a triplet of letter classes.
struct TranslateTables { uint8_t cls_fwd[256], cls_rev[256], aa3[17 * 17 * 17]; };

// Candidate rows, SoA (orfx_results row_* arrays mirror these)
struct DevRows {
    int64_t n_rows;
    int32_t *q_tx, *t_tx, *seg_start, *seg_end, *cds_start, *cds_end, *n_dup;
    int32_t *qlen, *tlen, *ulen, *query_start, *extra, *missing, *match, *outframe, *inframe;
    float *lpi, *ilpi, *mlpi;
    uint8_t *status, *pass, *start_match;
    int64_t* print_order;
    int32_t *ext_lo, *ext_hi;   // --overhang: exon-chain growth of the row's query copy (else nullptr)
    // alignment stage (else nullptr): the row's ORF as SEQ holds it -- first codon (mask coordinate) and codons, 0 when the
    // row never reached TX::score -- and the four ksw_extz_t fields of its alignment against the template's protein
    int32_t *aln_q0, *aln_qn, *aln_max, *aln_score, *aln_mqe, *aln_mte;
};

static constexpr uint8_t BUNDLE_HEAVY = 4;     // bundle_need bit: a transcript of the bundle has more than HEAVY_TX_SEGS segments (pipeline.cu)
static constexpr int HB_MAX_SEGS = 9216;      // heavy-bundle kernels (pipeline.cu): capacity of the shared-memory stage
static constexpr int HB_MAX_PAIRS = 8192;
static constexpr int SCAN_TILE_WORDS = 16;     // mask words (32 spliced positions each) per codon-scan run (one thread)
static constexpr int CSCAN_THREADS = 256;

}  // namespace orfx
