/*
 * orfx.h — C-ABI of the B200-native ORFanage hot path.
 *
 * The reference (alevar/ORFanage v1.2.0) has no plugin / FFI interface: its hot path is a code
 * region, the body of the bundle loop in orfanage.cpp:376-817 and the SEGTP/CHAIN/TX/SEQ methods
 * it calls (src/Transcriptome.{h,cpp}, src/common.hpp).  This header is the seam a maintainer
 * would cut there: everything upstream (GTF/FASTA loading, template preparation, bundling —
 * orfanage.cpp:313-367) packs an orfx_batch; everything downstream (attribute maps, GTF / stats
 * formatting — orfanage.cpp:662-815) consumes orfx_results.  INTEGRATION.md shows the binding.
 *
 * Conventions: plain C, plain pointers and sizes, no exceptions across the boundary.  Every call
 * returns 0 on success or a negative ORFX_E_* code; orfx_last_error() gives the message.  There is
 * no CPU fallback: orfx_create fails when no sm_100 device / CUDA runtime is usable.
 */
#ifndef ORFX_H
#define ORFX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORFX_ABI_VERSION 7

/* error codes */
#define ORFX_OK            0
#define ORFX_E_INVALID    -1   /* bad argument / malformed batch */
#define ORFX_E_CUDA       -2   /* CUDA runtime or kernel failure */
#define ORFX_E_NODEVICE   -3   /* no usable GPU: the product path never computes on the CPU */
#define ORFX_E_UNSUPPORTED -4  /* option outside this build's scope (e.g. gap costs with gapo + gape > 64) */
#define ORFX_E_STATE      -5   /* call sequence error (wait without submit, ...) */

/* --mode selector entries; replaces enum FindMode, orfanage.cpp:29-39 */
#define ORFX_MODE_START_MATCH   0
#define ORFX_MODE_ALL           1
#define ORFX_MODE_LONGEST       2
#define ORFX_MODE_LONGEST_MATCH 3
#define ORFX_MODE_BEST          4
#define ORFX_MODE_FIRST         5

/* candidate-row status; replaces the `notes` strings written at orfanage.cpp:510-627 */
#define ORFX_ST_OK             0   /* "-" (reached scoring; see `pass`)        */
#define ORFX_ST_NO_OVERLAP     1   /* "no-overlap"      orfanage.cpp:509-514   */
#define ORFX_ST_SEGMENT_LT3    2   /* "segment_len<3"   orfanage.cpp:538-542   */
#define ORFX_ST_CDS_LT3        3   /* "cds_len<3"       orfanage.cpp:549-561   */
#define ORFX_ST_NOT_RESCUED    4   /* "not_rescued"     orfanage.cpp:578-582   */
#define ORFX_ST_MISSING_STOP   5   /* "missing_stop"    orfanage.cpp:584-598   */
#define ORFX_ST_MISSING_START  6   /* "missing_start"   orfanage.cpp:584-598   */
#define ORFX_ST_MINLEN         7   /* "minlen"          orfanage.cpp:607-611   */
#define ORFX_ST_DUP            8   /* "dup"             orfanage.cpp:622-627   */

/* transcript kinds inside a batch */
#define ORFX_TX_QUERY     0
#define ORFX_TX_TEMPLATE  1
#define ORFX_TX_SKIP      2   /* present in the bundle but not evaluated (e.g. --keep_all_cds) */

/* Hot-path parameters; replaces the fields of GlobalParameters (orfanage.cpp:99-142) that the
 * bundle loop reads. */
typedef struct orfx_params {
    int32_t n_modes;
    int32_t mode_array[8];      /* ORFX_MODE_*, priority order (orfanage.cpp:120-122)          */
    int32_t lpi, ilpi, mlpi;    /* int thresholds, -1 = off (orfanage.cpp:615-617)             */
    int32_t cds_minlen;         /* -1 = off (orfanage.cpp:607)                                 */
    int32_t non_aug;            /* --non_aug                                                   */
    int32_t overhang;           /* --overhang N: the stop search may run up to N bases past the 3' end of the
                                   transcript, the unguided start search up to N bases before its 5' end
                                   (TX::extend_seq, Transcriptome.cpp:342-399); needs orfx_batch::tx_ext_lo/hi */
    int32_t spliced_overhang;   /* informational here: the cut at the first GT / last AG is applied by whoever
                                   fills tx_ext_lo / tx_ext_hi                                               */
    int32_t has_reference;      /* 0: no FASTA, candidates are scored as seeded (orfanage.cpp:564) */
    int32_t want_rows;          /* 1: return every candidate row (needed by --stats and mode ALL) */
    /* Protein alignment of every scored candidate (Finder::align, src/aln.hpp:27-41, with the build definition of
     * m = 24, mat50, aa_table, unbanded, no z-drop, score only).  Needs has_reference and a batch whose
     * `seq` also covers the template CDS segments (seg_seq_off set for TEMPLATE segments, raw bytes: ORFX_SEQ_RAW).
     * Query protein = SEQ::translate (common.hpp:477-495) of the row's rescued CDS on the spliced query, template protein =
     * the same over the template's spliced CDS chain from its first base; both keep their stop codon; a codon that is
     * not a codon_map key translates to '\0' there and to residue 23 here, like every byte aa_table does not know. */
    int32_t align_enabled;
    int32_t align_gapo;         /* gap open, 0 = the reference's default 12 (aln.hpp:146)                       */
    int32_t align_gape;         /* gap extend, 0 = the reference's default 2                                    */
    int32_t host_results_pageable;   /* 1: the library's result buffers are ordinary memory instead of page-locked memory.  For a
                                        caller that evaluates one batch per process (the command-line tool) page-locking tens of
                                        megabytes costs more than the one download gains                                        */
} orfx_params;

/* 1-based inclusive genomic interval; replaces SEGTP (Transcriptome.h:23-169) */
typedef struct orfx_seg { int32_t start, end; } orfx_seg;

/* A packed set of locus bundles; replaces the bundled Transcriptome handed to the loop at
 * orfanage.cpp:376 (Bundle: Transcriptome.h:840-891).  All arrays are caller-owned host memory
 * (pinned memory makes the upload asynchronous) and must stay valid until orfx_wait returns.
 * Only bundles that hold at least one template and one query belong here (orfanage.cpp:383). */
typedef struct orfx_batch {
    int32_t n_bundles;
    int32_t n_tx;
    int64_t n_segs;
    int64_t seq_bytes;
    /* per bundle */
    const uint8_t* seq;            /* per bundle, raw FASTA bytes, case preserved; may be NULL when
                                      params.has_reference == 0.  Two layouts:
                                      windows (seg_seq_off == NULL): the whole locus window
                                        [bundle_start, bundle end], as Bundle::bundle_seq;
                                      islands (seg_seq_off != NULL): only the bases under the query
                                        exons of the bundle (overlapping exons merged), which is all
                                        the path reads without --overhang; 4-15x fewer bytes to upload */
    const int64_t* seq_off;        /* [n_bundles+1] byte offsets into seq                          */
    const int32_t* bundle_start;   /* [n_bundles] genomic coordinate of the first window byte      */
    const int32_t* bundle_tx_off;  /* [n_bundles+1] transcripts of a bundle are contiguous, in the
                                      reference's sorted bundle order                              */
    /* per transcript */
    const uint8_t* tx_kind;        /* [n_tx] ORFX_TX_*                                             */
    const uint8_t* tx_strand;      /* [n_tx] '+', '-' or '.'                                       */
    const int32_t* tx_seg_off;     /* [n_tx+1] CSR into segs                                       */
    const int32_t* tx_cds_start;   /* [n_tx] templates: TX::cds_start member; others 0             */
    const int32_t* tx_cds_end;     /* [n_tx] templates: TX::cds_end member; others 0               */
    const int32_t* tx_tid_class;   /* [n_tx] equal transcript_id strings <=> equal value           */
    /* per segment: query -> exon chain; template -> CDS chain (== its exon chain after
     * set_cds_as_exons, orfanage.cpp:351) with the phase column of each CDS segment */
    const orfx_seg* segs;          /* [n_segs] ascending, disjoint                                 */
    const int8_t*   seg_phase;     /* [n_segs] templates: SEGTP::phase of the CDS chain; else 0    */
    const int32_t*  seg_seq_off;   /* [n_segs] or NULL.  Islands layout: for every segment of a QUERY
                                      transcript, the offset of its first base relative to
                                      seq_off[bundle]; ignored for other transcripts               */
    int32_t seq_format;            /* ORFX_SEQ_RAW: `seq` holds one FASTA byte per base.
                                      ORFX_SEQ_NT4: two bases per byte (base 2k in the low nibble of byte k),
                                      ORFX_NT4_* codes; seq_bytes / seq_off / seg_seq_off still count BASES and
                                      the buffer holds (seq_bytes + 1) / 2 bytes.  Halves the largest upload.
                                      The codes keep exactly what the path can see of a base: the reference
                                      upper-cases it and only the codons of codon_map (common.hpp:279) that
                                      translate to '.' or 'M' steer the path -- TAA TAG TGA TAR TRA and ATG -- so
                                      A C G T U R Y are told apart and every other byte is "other"            */
    int32_t reserved0;
    /* --overhang (params.overhang > 0; NULL otherwise): for every QUERY transcript the number of genomic bases
     * that TX::extend_seq (Transcriptome.cpp:342-399) would add below the first exon / above the last exon: at most
     * params.overhang, clamped to the bundle window, and with --spliced_overhang cut at the last AG (5' side) /
     * first GT (3' side) of the strand-corrected flank.  `seq` must hold these bases: in the islands layout they
     * sit right before the first / right after the last exon's bases (the island is widened). */
    const int32_t* tx_ext_lo;      /* [n_tx] */
    const int32_t* tx_ext_hi;      /* [n_tx] */
    /* Which side the bases come from is the packer's business (TX::extend_seq picks it, cpp:348-364): the path reads the
     * 3' flank next to the 3' terminal exon and the 5' flank next to the 5' one, where "3'" is above the last exon for
     * strand '+' and below the first exon for every other strand byte.  A strand '.' transcript takes BOTH its flanks from
     * the low genomic side in the reference; its packer gives it private island bases with the far end of the low-side
     * flank copied above its last exon (host/runner.cpp: query_flanks) -- islands layout only.
     * tx_ext_flags ([n_tx] or NULL): ORFX_EXT_LO_VIRTUAL / ORFX_EXT_HI_VIRTUAL mark a one-base flank that is a re-read of the
     * transcript's own edge base: the transcript touches the end of its chromosome and the clamp arithmetic of
     * cpp:352-364 yields that base again.  The position does not exist: a stop codon completed by it grows the exon by
     * one, and the re-loaded sequence then ends one base short, so the row ends as "missing_stop" (cpp:418-442, 155-171). */
    const uint8_t* tx_ext_flags;
    /* ORFX_SEQ_NT2 only: the bases that are none of A C G T (case-insensitive) -- their position (offset in bases into `seq`,
     * ascending) and their ORFX_NT4_* code; `seq` holds 0 in their place.  Genomes have few of them under exons. */
    const int64_t* seq_exc_pos;    /* [n_seq_exc] */
    const uint8_t* seq_exc_code;   /* [n_seq_exc] */
    int64_t n_seq_exc;
    /* Segment dictionary (optional; seg_ref == NULL: not used).  The transcripts of a locus share most of their exons
     * (alternative isoforms), so the three per-segment arrays above repeat the same few records many times.  With
     * seg_ref != NULL every segment k of a transcript of bundle b is the dictionary record
     *     d = bundle_dict_off[b] + seg_ref[k]        (seg_ref[k] < bundle_dict_off[b+1] - bundle_dict_off[b] <= 65536)
     * i.e. segs[k] = dict_segs[d], seg_phase[k] = dict_phase[d], seg_seq_off[k] = dict_seq_off[d] (islands layout iff
     * dict_seq_off != NULL).  The dictionary is then the ONLY source of the chains: segs / seg_phase / seg_seq_off are not
     * read and may be NULL; the library expands the records on the device before it validates the batch.  2 bytes per
     * segment and 13 per distinct record instead of 13 per segment: 28 MB less to upload per 235 k queries. */
    const uint16_t* seg_ref;          /* [n_segs] */
    const int32_t*  bundle_dict_off;  /* [n_bundles+1] */
    const orfx_seg* dict_segs;        /* [n_dict] */
    const int8_t*   dict_phase;       /* [n_dict] */
    const int32_t*  dict_seq_off;     /* [n_dict] or NULL (windows layout) */
    int64_t n_dict;
} orfx_batch;

#define ORFX_EXT_LO_VIRTUAL 1
#define ORFX_EXT_HI_VIRTUAL 2

#define ORFX_SEQ_RAW 0
#define ORFX_SEQ_NT4 1
#define ORFX_SEQ_NT2 2   /* islands layout only: four bases per byte (base 4k in bits 0-1 of byte k), A = 0, C = 1, G = 2, T = 3; every
                            other base is listed in seq_exc_pos / seq_exc_code.  seq_bytes / seq_off / seg_seq_off count BASES, seq_off
                            entries are multiples of 4 and the buffer holds (seq_bytes + 3) / 4 bytes.  A quarter of the raw upload. */
/* 4-bit base codes of ORFX_SEQ_NT4 (case-insensitive); U is kept apart from T because the reference only
 * equates them through ntComplement on the reverse strand */
#define ORFX_NT4_OTHER 0
#define ORFX_NT4_A 1
#define ORFX_NT4_C 2
#define ORFX_NT4_U 3
#define ORFX_NT4_G 4
#define ORFX_NT4_R 5
#define ORFX_NT4_T 8
#define ORFX_NT4_Y 10

/* One chosen ORF per query transcript; what orfanage.cpp:753-791 needs to write the GTF. */
typedef struct orfx_best {
    int32_t q_tx;          /* transcript index in the batch                                   */
    int32_t t_tx;          /* template index of the winning row, -1 = no ORF found            */
    int32_t cds_start;     /* genomic, 1-based inclusive                                      */
    int32_t cds_end;
    int32_t duplicity;     /* TX::num_dups() of the emitted transcript (orfanage_duplicity)   */
    int32_t n_rows;        /* candidate rows of this query                                    */
    int64_t row_off;       /* first row of this query in the row arrays                       */
    int64_t best_row;      /* absolute row index of the winner, -1 = none                     */
    float   lpi, ilpi, mlpi;
    int32_t start_match;
    int32_t ext_lo, ext_hi; /* --overhang: bases by which the emitted transcript's first exon starts earlier /
                               its last exon ends later than the query's (Transcriptome.cpp:418-429, 461-474) */
} orfx_best;

/* Results of one batch.  Library-owned pinned host buffers, valid until orfx_release / the next
 * orfx_submit on the same context.  Row arrays are filled only when params.want_rows != 0. */
typedef struct orfx_results {
    int32_t n_queries;
    int64_t n_rows;
    const orfx_best* best;         /* [n_queries] in bundle order, then query order inside the bundle */
    /* candidate rows, SoA, in the reference's (query, template, segment) order */
    const int32_t* row_q_tx;
    const int32_t* row_t_tx;
    const int32_t* row_seg_start;  /* original intersection segment (orfanage.cpp:522)           */
    const int32_t* row_seg_end;
    const int32_t* row_cds_start;  /* flatten() key (orfanage.cpp:265-267)                       */
    const int32_t* row_cds_end;
    const uint8_t* row_status;     /* ORFX_ST_*                                                  */
    const uint8_t* row_pass;
    const uint8_t* row_start_match;
    const int32_t* row_n_dup;      /* TX::num_dups() of the row's query copy (stats column 5 - 1) */
    const int32_t* row_qlen;       /* Score (common.hpp:32-65)                                   */
    const int32_t* row_tlen;
    const int32_t* row_ulen;
    const int32_t* row_query_start;
    const int32_t* row_extra;
    const int32_t* row_missing;
    const int32_t* row_match;
    const int32_t* row_outframe;
    const int32_t* row_inframe;
    const float*   row_lpi;
    const float*   row_ilpi;
    const float*   row_mlpi;
    const int32_t* row_ext_lo;     /* --overhang only (else NULL): exon-chain growth of the row's query copy, */
    const int32_t* row_ext_hi;     /* genomic low / high side, as orfx_best::ext_lo / ext_hi                  */
    const int64_t* print_order;    /* [n_rows] absolute row indices in the stats print order:
                                      per query, keys ascending, inside a key the std::sort by
                                      score_gt (orfanage.cpp:250-283)                            */
    /* device-side timing of the last batch, milliseconds (CUDA events).  orfx_submit uploads the sequence bytes
     * on a copy stream while the chain kernels already run: ms_h2d (submit -> last byte on the device) and
     * ms_kernels (chains on the device -> last kernel) overlap, their sum exceeds the batch time. */
    float ms_h2d, ms_kernels, ms_d2h;
    /* params.align_enabled && params.want_rows (else NULL): ksw_extz_t fields of the alignment of the row's ORF protein
     * (query) against its template's protein (target), for every row that reached TX::score (status "-" or "dup");
     * 0 / INT32_MIN-like sentinels never appear: rows that were not scored hold 0 in all four. */
    const int32_t* row_aln_max;    /* ez.max   */
    const int32_t* row_aln_score;  /* ez.score */
    const int32_t* row_aln_mqe;    /* ez.mqe   */
    const int32_t* row_aln_mte;    /* ez.mte   */
    float ms_align;                /* device time of the alignment stage of the last batch (inside ms_kernels) */
    float reserved1;
} orfx_results;

typedef struct orfx_ctx orfx_ctx;

const char* orfx_last_error(void);
int  orfx_abi_version(void);

/* device = CUDA ordinal. */
int  orfx_create(orfx_ctx** out, int device, const orfx_params* params);
int  orfx_set_params(orfx_ctx* ctx, const orfx_params* params);
/* Asynchronous: uploads the batch and enqueues every kernel of the path on the context's stream. */
int  orfx_submit(orfx_ctx* ctx, const orfx_batch* batch);
/* Blocks until the batch is done and the results are in host memory. */
int  orfx_wait(orfx_ctx* ctx, orfx_results* out);
int  orfx_release(orfx_ctx* ctx);
void orfx_destroy(orfx_ctx* ctx);

/* Pinned (page-locked, portable) host memory for batch arrays: makes the uploads of orfx_submit
 * asynchronous and full-speed.  Optional — any host memory is accepted by orfx_submit. */
void* orfx_host_alloc(size_t bytes);
void  orfx_host_free(void* p);

/* ---- device-resident variant used by the benchmark's kernel-only timing ------------------- */
/* Uploads a batch once and keeps it in HBM. */
int  orfx_upload(orfx_ctx* ctx, const orfx_batch* batch);
/* Runs the whole path on the resident batch; no host<->device copies.  Returns the number of
 * kernels launched in *n_launches (may be NULL). */
int  orfx_run_resident(orfx_ctx* ctx, int* n_launches);
/* Copies the results of the last resident run to the host (same buffers as orfx_wait). */
int  orfx_fetch(orfx_ctx* ctx, orfx_results* out);
/* Per-kernel device times of the last run, in launch order.  names[i] are static strings. */
int  orfx_kernel_times(orfx_ctx* ctx, int max_n, const char** names, float* ms, int* n);
/* Device time of the last run of the path (first to last kernel, CUDA events on the launch stream;
 * includes the one host round trip that sizes the row buffers), and its number of kernel launches.
 * Valid after orfx_wait or orfx_run_resident. */
float orfx_last_run_ms(orfx_ctx* ctx);
int   orfx_last_run_launches(orfx_ctx* ctx);
/* Bytes the last orfx_submit / orfx_upload copied host -> device (every array of the batch it actually sent). */
int64_t orfx_last_upload_bytes(orfx_ctx* ctx);
/* Sum of spliced query nucleotides in the resident batch (codon-scan roofline numerator). */
int64_t orfx_resident_query_nt(orfx_ctx* ctx);

/* ---- per-stage access for differential tests (SURVEY.md §4 iii) --------------------------- */
/* Codon-scan masks of the resident batch after a run: for query transcript tx, bit p of
 * stop/start says whether the codon starting at spliced position p (5'->3') is a stop / ATG.
 * Buffers must hold ceil(L/32) words, L = spliced length.  Returns L or a negative error. */
int64_t orfx_get_masks(orfx_ctx* ctx, int32_t tx, uint32_t* stop_bits, uint32_t* start_bits, int64_t max_words);

/* Test hook: lowers an internal capacity so that small inputs reach the code paths of very large ones.
 * "hb_max_segs" / "hb_max_pairs": segments of a titin-class bundle / (query, transcript) pairs of a bundle slice that
 * the heavy-bundle kernels keep in shared memory (defaults 9216 / 8192); bundles above take their global-memory branches.
 * "serialize": every kernel on one stream (clean per-kernel times).  "blocking_sync": while the host waits for the stream
 * (twice inside orfx_submit, once in orfx_wait) the thread sleeps on a cudaEventBlockingSync event instead of spinning in
 * cudaStreamSynchronize -- for processes in which more threads wait for GPUs than the host has cores.
 * Takes effect at the next orfx_submit / orfx_upload.  Results never depend on these values. */
int  orfx_debug_set(orfx_ctx* ctx, const char* key, int64_t value);

/* ---- template cleaning: --cleant [--rescue] -------------------------- */
/* Replaces the per-transcript body of Transcriptome::clean_cds (Transcriptome.cpp:1063-1100): load_seq,
 * optionally TX::rescue_cds(non_aug, 0, false, nullptr) (cpp:641-677), then "first codon is M (unless
 * --non_aug), last codon is a stop, CDS length % 3 == 0".  The batch is a *cleaning batch*: every coding
 * template of the bundles made by the first bundleup (cpp:1066) is packed as ORFX_TX_QUERY with segs = its
 * exon chain and tx_cds_start / tx_cds_end = its CDS bounds after build_cds + correct_chain_len; anything
 * else is ORFX_TX_SKIP.  Outputs are caller-owned arrays with one entry per ORFX_TX_QUERY transcript, in
 * batch order.  Uses params.non_aug and params.overhang of the context (with --overhang the unguided searches of the
 * rescue run into the flanks, tx_ext_lo / tx_ext_hi as for orfx_submit). */
typedef struct orfx_clean_out {
    uint8_t* keep;        /* 1: the CDS survives cleaning; 0: TX::remove_cds()                       */
    int32_t* cds_start;   /* genomic CDS bounds after the (optional) rescue; 0 when keep == 0          */
    int32_t* cds_end;
} orfx_clean_out;
int  orfx_clean_templates(orfx_ctx* ctx, const orfx_batch* batch, int rescue, orfx_clean_out* out, int64_t* n_out);

/* ---- orfcompare: pair scoring only ----------------------------------- */
/* Replaces the `q->score(*t)` loop of orfcompare.cpp:76-160: TX::score (Transcriptome.cpp:699-762) for every
 * (query, template) pair of every bundle, no bounding-box test.  The batch is a *compare batch*: queries
 * (ORFX_TX_QUERY) and templates (ORFX_TX_TEMPLATE) both carry their CDS chain as `segs`; no sequence is read.
 * Synchronous.  On return out->n_rows pairs are described by row_q_tx, row_t_tx, row_qlen, row_tlen, row_ulen,
 * row_match, row_inframe, row_outframe, row_extra, row_missing, row_lpi, row_ilpi, row_mlpi (query-major, templates
 * in bundle order); every other member of *out is NULL / 0.  Buffers are library-owned, valid until the next call. */
int  orfx_score_pairs(orfx_ctx* ctx, const orfx_batch* batch, orfx_results* out);

/* ---- protein alignment (src/aln.hpp:27-41 with the build definition) ---- */
typedef struct orfx_aln_out {
    int32_t* max;      /* ez.max  : max(0, best H over all cells) */
    int32_t* score;    /* ez.score: H(tlen-1, qlen-1)             */
    int32_t* mqe;      /* ez.mqe  : best H in the last query column */
    int32_t* mte;      /* ez.mte  : best H in the last target row   */
} orfx_aln_out;

/* aa: concatenated residues already encoded 0..23 through aa_table (common.hpp:268-277);
 * off[n_seq+1]; pair_q / pair_t index sequences.  Substitution matrix mat50 (common.hpp:239-265),
 * gap open gapo, gap extend gape (aln.hpp:146: 12 / 2).  Output arrays are caller-owned host memory
 * of n_pairs entries.  cells (optional) receives sum of qlen*tlen; ms (optional) the device time. */
int  orfx_align_batch(orfx_ctx* ctx, const uint8_t* aa, const int64_t* off, int64_t n_seq,
                      const int32_t* pair_q, const int32_t* pair_t, int64_t n_pairs,
                      int gapo, int gape, orfx_aln_out* out, double* cells, float* ms);

/* orfx_align_batch with NO allocation inside the call.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The call makes no cudaMalloc / cudaFree / cudaMallocHost and creates no events; its scratch (pair descriptors, the
 * longest-first schedule, the per-warp stripe edges) is one file-scope `static __device__` array of ORFX_STATIC_SCRATCH_MIB
 * MiB that every GPU which loads the library holds until the process exits.
 *   - aa, off[n_seq+1], pair_q, pair_t and the four arrays of *out are DEVICE memory of the context's device (the caller's
 *     own allocations); meaning and results are those of orfx_align_batch, bit for bit.
 *   - limits: n_pairs <= ORFX_STATIC_MAX_PAIRS per call, every protein <= ORFX_STATIC_MAX_LEN residues; beyond either the call
 *     answers ORFX_E_UNSUPPORTED and names the limit in orfx_last_error() (use orfx_align_batch, or split the call).  A pair
 *     index outside [0, n_seq) or descending offsets: ORFX_E_INVALID.  After an error the output arrays are unspecified.
 *   - the call returns after the context's stream has drained (it synchronises it twice).
 *   - all calls share the one array, which keeps the previous call's contents (what has to start at zero is cleared by every
 *     call): calls that overlap on the GPU -- two contexts of one device driven from two threads -- race.  One at a time per
 *     device. */
#define ORFX_STATIC_MAX_PAIRS 65536
#define ORFX_STATIC_MAX_LEN 1024
#define ORFX_STATIC_SCRATCH_MIB 21
int  orfx_align_pairs_static(orfx_ctx* ctx, const uint8_t* aa, const int64_t* off, int64_t n_seq,
                             const int32_t* pair_q, const int32_t* pair_t, int64_t n_pairs,
                             int gapo, int gape, const orfx_aln_out* out);

/* Integer-ALU issue-rate microbenchmark on the context's device: the denominator of the alignment roofline
 * (SURVEY.md §8d).  ops4[0] = IADD3 instructions/s; ops4[1] = IADD + VIMNMX3 pairs (2 operations each)/s;
 * ops4[2] = VIADDMNMX (add + max, 2 operations each)/s; ops4[3] = IDP.4A instructions/s; all SMs, 8
 * independent chains per thread. */
int  orfx_int_alu_peak(orfx_ctx* ctx, double* ops4);

#ifdef __cplusplus
}
#endif
#endif /* ORFX_H */
