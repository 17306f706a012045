// Translation of every transcript of a batch, all three frames at once, for the alignment stage
// (orfx_params::align_enabled).
//
//   aa_pos[tx_aa_off[tx] + p] = residue code (aa_table, common.hpp:268-277) of the codon that starts at position p of
//                               the transcript's sequence, 5'->3'
// for QUERY transcripts over the flanked exon chain (the masks' coordinates: [5' flank][transcript][3' flank]) and for
// TEMPLATE transcripts over their spliced CDS chain.  A protein is then every third byte from its first codon on: the
// alignment kernel reads the row's ORF (SEQ::translate of the rescued CDS, common.hpp:477-495) and the template's
// protein (the same from the first base of its CDS chain) straight out of this array with stride 3 -- nothing is
// materialised per candidate row.
//
// Replaces, for the build definition of SURVEY.md §8a-A1: TX::_get_nt_seq (Transcriptome.cpp:155-171: per-base
// ntComplement on every strand but '+'), the toupper of SEQ::SEQ (common.hpp:303-308), the codon_map lookups of
// SEQ::translate (a triplet that is no key of the map -- and any incomplete one -- gives '\0') and the aa_table encoding
// of Finder::align (aln.hpp:33-36; '\0', '.' and every byte the table does not know are residue 23).
#include "orfx_device.cuh"
#include "orfx_buffers.cuh"
#include "orfx_codon_table.h"
#include "orfx_mat50.h"
#include <cctype>
#include <climits>
#include <cstring>

namespace orfx {

// 17 letter classes: the 16 letters codon_map's keys are written in, and "anything else"
static const char CODON_LETTERS[] = "ACGTBDHKMNRSVWYX";
constexpr int NCLS = 17;

void build_translate_tables(TranslateTables& T) {
    auto cls_of = [](int c) { const char* at = c ? strchr(CODON_LETTERS, c) : nullptr; return at ? (int)(at - CODON_LETTERS) : NCLS - 1; };
    // ntComplement (gclib), IUPAC, case preserving -- as oracle/shim provides it to the reference's sources
    unsigned char comp[256];
    for (int i = 0; i < 256; i++) comp[i] = (unsigned char)i;
    const char* defs = "AaCcTtGgUuMmRrWwSsYyKkVvHhDdBbNnXx-*";
    const char* cmpl = "TtGgAaCcAaKkYyWwSsRrMmBbDdHhVvNnXx-*";
    for (int i = 0; defs[i]; i++) comp[(unsigned char)defs[i]] = (unsigned char)cmpl[i];
    for (int b = 0; b < 256; b++) {
        T.cls_fwd[b] = (uint8_t)cls_of(b < 128 ? toupper(b) : b);
        const int c = comp[b];
        T.cls_rev[b] = (uint8_t)cls_of(c < 128 ? toupper(c) : c);
    }
    memset(T.aa3, 23, sizeof T.aa3);
    const char* order = ORFX_AA_ORDER;
    for (int i = 0; i < ORFX_CODON_TABLE_N; i++) {
        const char* e = ORFX_CODON_TABLE[i];
        const char* at = strchr(order, e[3]);   // '.' (stop) is not a letter of aa_table: residue 23
        const int code = at && e[3] ? (int)(at - order) : 23;
        T.aa3[(cls_of(e[0]) * NCLS + cls_of(e[1])) * NCLS + cls_of(e[2])] = (uint8_t)code;
    }
}

namespace {
constexpr int TR_THREADS = 256;
constexpr int TR_CHUNK = 16;   // positions per thread: one 128-bit store

__global__ void __launch_bounds__(256) aa_len_kernel(DevBatch B, int64_t* cnt) {
    const int tx = blockIdx.x * blockDim.x + threadIdx.x;
    if (tx >= B.n_tx) return;
    const int kind = B.tx_kind[tx];
    int64_t len = 0;
    if (kind == ORFX_TX_QUERY) len = (int64_t)B.tx_len[tx] + (B.tx_ext_lo ? B.tx_ext_lo[tx] + B.tx_ext_hi[tx] : 0);
    else if (kind == ORFX_TX_TEMPLATE) len = B.tx_len[tx];
    cnt[tx] = (len + TR_CHUNK - 1) / TR_CHUNK * TR_CHUNK;   // 16-byte aligned regions
}

__global__ void __launch_bounds__(TR_THREADS) translate_kernel(DevBatch B, const TranslateTables* __restrict__ T, int64_t n_chunks) {
    __shared__ uint8_t s_cls[2][256];
    __shared__ uint8_t s_aa3[NCLS * NCLS * NCLS];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) { s_cls[0][i] = T->cls_fwd[i]; s_cls[1][i] = T->cls_rev[i]; }
    for (int i = threadIdx.x; i < NCLS * NCLS * NCLS; i += blockDim.x) s_aa3[i] = T->aa3[i];
    __syncthreads();
    const int64_t chunk = (int64_t)blockIdx.x * TR_THREADS + threadIdx.x;
    if (chunk >= n_chunks) return;
    const int64_t pos0 = chunk * TR_CHUNK;
    int tx;
    {   // last transcript whose region starts at or before pos0 (empty regions share an offset with their successor)
        int lo = 0, hi = B.n_tx - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (B.tx_aa_off[mid] <= pos0) lo = mid; else hi = mid - 1; }
        tx = lo;
    }
    const int p_lo = (int)(pos0 - B.tx_aa_off[tx]);
    const bool isq = B.tx_kind[tx] == ORFX_TX_QUERY;
    const int xlo = isq && B.tx_ext_lo ? B.tx_ext_lo[tx] : 0, xhi = isq && B.tx_ext_hi ? B.tx_ext_hi[tx] : 0;
    const int L = B.tx_len[tx] + xlo + xhi;
    uint4 out = make_uint4(0x17171717u, 0x17171717u, 0x17171717u, 0x17171717u);   // residue 23 everywhere
    if (p_lo < L) {
        const bool plus = B.tx_strand[tx] == '+';
        const int s0 = B.tx_seg_off[tx], nseg = B.tx_seg_off[tx + 1] - s0;
        const int bundle = B.tx_bundle[tx];
        const bool isl = B.seg_seq_off != nullptr;
        const uint8_t* __restrict__ seq = B.seq + B.seq_off[bundle];
        const int bstart = isl ? 0 : B.bundle_start[bundle];
        const int32_t* __restrict__ cum0 = B.seg_cum + s0;
        const orfx_seg* __restrict__ segs = B.segs + s0;
        const int32_t* __restrict__ sso = isl ? B.seg_seq_off + s0 : nullptr;
        auto cum = [&](int k) { return cum0[k] + (k > 0 ? xlo : 0); };
        auto cum_at = [&](int k) { return k < nseg ? cum(k) : L; };
        auto pos_at = [&](int k) { return (isl ? sso[k] : segs[k].start - bstart) - (k == 0 ? xlo : 0); };
        const int nb = min(TR_CHUNK + 2, L - p_lo);   // letters this thread reads, the two look-ahead bases included
        const int u = plus ? p_lo : L - 1 - p_lo;     // genomic-order spliced offset of the first one
        int k;
        { int lo = 0, hi = nseg - 1; while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (cum(mid) <= u) lo = mid; else hi = mid - 1; } k = lo; }
        int rem = plus ? cum_at(k + 1) - u : u - cum(k) + 1;
        int off = pos_at(k) + (u - cum(k));
        const uint8_t* __restrict__ cls = s_cls[plus ? 0 : 1];
        uint32_t c[TR_CHUNK + 2];
#pragma unroll
        for (int i = 0; i < TR_CHUNK + 2; i++) {
            c[i] = NCLS - 1;
            if (i < nb) {
                c[i] = cls[seq[off]];
                off += plus ? 1 : -1;
                if (--rem == 0 && i + 1 < nb) {   // next exon in reading direction
                    k += plus ? 1 : -1;
                    rem = cum_at(k + 1) - cum(k);
                    off = pos_at(k) + (plus ? 0 : rem - 1);
                }
            }
        }
        uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < TR_CHUNK; i++) {
            const uint32_t a = i + 2 < nb ? s_aa3[(c[i] * NCLS + c[i + 1]) * NCLS + c[i + 2]] : 23u;
            w[i >> 2] |= a << (8 * (i & 3));
        }
        out = make_uint4(w[0], w[1], w[2], w[3]);
    }
    *reinterpret_cast<uint4*>(B.aa_pos + pos0) = out;
}

// One alignment pair per candidate row that reached TX::score: query = the row's ORF as SEQ holds it at the end of the
// rescue (process_row: first codon at mask coordinate aln_q0, aln_qn codons, stop codon included), target = the
// template's protein (floor(L / 3) codons from the first base of its CDS chain).  Other rows: not aligned.
__global__ void __launch_bounds__(256) row_pairs_kernel(DevBatch B, DevRows R, AlnPair* __restrict__ pairs) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= R.n_rows) return;
    AlnPair p;
    p.q_off = 0; p.t_off = 0; p.qlen = -1; p.tlen = 0;
    const int n = R.aln_qn[row];
    if (n > 0) {
        const int qx = R.q_tx[row], t = R.t_tx[row];
        p.q_off = B.tx_aa_off[qx] + R.aln_q0[row];
        p.t_off = B.tx_aa_off[t];
        p.qlen = n;
        p.tlen = B.tx_len[t] / 3;
    }
    pairs[row] = p;
}
}  // namespace

int launch_aa_len(const DevBatch& B, int64_t* cnt, cudaStream_t s) {
    if (B.n_tx == 0) return 0;
    aa_len_kernel<<<(B.n_tx + 255) / 256, 256, 0, s>>>(B, cnt);
    return 1;
}
int launch_translate(const DevBatch& B, const TranslateTables* d_tables, int64_t total_positions, cudaStream_t s) {
    const int64_t n_chunks = total_positions / TR_CHUNK;
    if (n_chunks <= 0) return 0;
    translate_kernel<<<(unsigned)((n_chunks + TR_THREADS - 1) / TR_THREADS), TR_THREADS, 0, s>>>(B, d_tables, n_chunks);
    return 1;
}
int launch_row_pairs(const DevBatch& B, const DevRows& R, AlnPair* pairs, cudaStream_t s) {
    if (R.n_rows == 0) return 0;
    row_pairs_kernel<<<(unsigned)((R.n_rows + 255) / 256), 256, 0, s>>>(B, R, pairs);
    return 1;
}

}  // namespace orfx
