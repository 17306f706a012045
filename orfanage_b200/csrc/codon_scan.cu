// Codon scan: stop / ATG bit masks of every query transcript, all three frames at once.
//
//   bit p of `stop`  = the codon starting at spliced position p (5'->3') is a stop codon
//                      (TAA TAG TGA TAR TRA — the '.' entries of codon_map, common.hpp:279)
//   bit p of `start` = it is ATG
// Every later question of the rescue step ("first in-frame stop at or after a", "all in-frame ATG
// between the previous stop and a") becomes a bit scan over these masks (chain_ops.cuh).
//
// Replaces, for the whole batch at once, what the reference recomputes ~2.4x per candidate:
// TX::_get_nt_seq (Transcriptome.cpp:155-171), SEQ::SEQ toupper (common.hpp:303-308), SEQ::translate
// and SEQ::find_inframe_codon's codon_map lookups (common.hpp:420-450, 477-495).
//
// Two kernels, both HBM/L2-streaming, no shared memory needed:
//
//  1. base_planes_kernel — once per WINDOW byte (not per spliced base: a locus window is shared by all
//     its isoforms).  Each thread reads 32 bytes with two 128-bit loads, classifies them with SWAR byte
//     compares (case folded, exact for all 256 byte values) and writes one 32-bit word into each of
//     each of the base planes the window's strand needs:  '+' : T A G R;  other strands (transcript
//     base = complement of the genomic byte, gclib ntComplement): A (->T)  T|U (->A)  C (->G)  Y (->R);
//     the four planes of a 32-byte chunk are stored together as one uint4.
//  2. codon_scan_kernel — one warp per work item (up to 256 mask words of one transcript), exon table
//     staged in shared memory.  A lane assembles the four plane words of its 32 spliced positions by
//     funnel-shifting bit runs out of the planes exon piece by exon piece (reversed with brev for the
//     reverse strand), takes the next word's low bits from its neighbour lane (shuffle), evaluates
//     the codon logic on whole words and stores the two mask words, coalesced.
//
// Roofline accounting: algorithmic bytes = 1 B per spliced nucleotide read + 2 bits
// written; the planes make the real DRAM traffic smaller than that (window bytes once + bits).
#include "orfx_device.cuh"

namespace orfx {

namespace {

__device__ __forceinline__ uint32_t eq_flags(uint32_t w, uint32_t fold, uint32_t lit) {
    // 0x80 in every byte of w that equals lit after masking with fold; exact for all 256 byte values
    const uint32_t y = (w & fold) ^ lit;
    const uint32_t t = ((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y;
    return ~t & 0x80808080u;
}
__device__ __forceinline__ uint32_t gather4(uint32_t flags) {
    // bits 7,15,23,31 -> bits 0..3
    return (flags * 0x00204081u) >> 28;
}

constexpr uint32_t FOLD = 0xDFDFDFDFu;      // clear the ASCII case bit
constexpr uint32_t FOLD_TU = 0xDEDEDEDEu;   // ... and bit 0: 'T' (0x54) and 'U' (0x55) coincide

template <uint32_t FOLDM, uint32_t LIT>
__device__ __forceinline__ uint32_t plane_word(const uint32_t (&w)[8]) {
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) r |= gather4(eq_flags(w[k], FOLDM, LIT)) << (4 * k);
    return r;
}

}  // namespace

// ---- 1. window bytes -> base planes -------------------------------------------------------------------
// Plane layout: for every 32-byte chunk c of `seq` one uint4 per strand class, so that a scan lane gets
// the four planes of 32 bases with ONE 128-bit load:
//   planes4[c]               = {T, A, G, R}                       ('+' transcripts)
//   planes4[plane_words + c] = {A (->T), T|U (->A), C (->G), Y (->R)}   (other strands: complement classes)
__global__ void __launch_bounds__(256) base_planes_kernel(DevBatch B, int64_t n_chunks) {
    const int64_t chunk = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // 32 bytes each
    // which strands do the windows under this warp's 1 KB need?  (warp-uniform)
    const int64_t warp_lo = (chunk & ~(int64_t)31) * 32;
    if (warp_lo >= B.seq_bytes) return;
    const int64_t warp_hi = min(warp_lo + 1024, B.seq_bytes);
    int lo = 0, hi = B.n_bundles - 1;
    while (lo < hi) {   // last bundle with seq_off <= warp_lo
        const int mid = (lo + hi + 1) >> 1;
        if (B.seq_off[mid] <= warp_lo) lo = mid; else hi = mid - 1;
    }
    uint32_t need = 0;
    for (int b = lo; b < B.n_bundles && B.seq_off[b] < warp_hi; b++) need |= B.bundle_need[b];
    if (chunk >= n_chunks || need == 0) return;
    const uint4* src = reinterpret_cast<const uint4*>(B.seq) + 2 * chunk;
    const uint4 v0 = __ldg(src), v1 = __ldg(src + 1);
    const uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    uint4* out = reinterpret_cast<uint4*>(B.planes) + chunk;
    const uint32_t a = plane_word<FOLD, 0x41414141u>(w);   // A: '+' A plane and reverse-strand T plane
    if (need & 1u)
        out[0] = make_uint4(plane_word<FOLD, 0x54545454u>(w), a, plane_word<FOLD, 0x47474747u>(w), plane_word<FOLD, 0x52525252u>(w));
    if (need & 2u)
        out[B.plane_words] = make_uint4(a, plane_word<FOLD_TU, 0x54545454u>(w), plane_word<FOLD, 0x43434343u>(w), plane_word<FOLD, 0x59595959u>(w));
}

// ---- 2. planes -> per-transcript stop / start masks ---------------------------------------------------
// One warp per work item = up to SCAN_ITEM_WORDS mask words (8192 spliced positions) of one transcript,
// walked 32 words at a time.  The warp reads the transcript's header once and stages its exon table
// (spliced offset + bit position in the planes of every exon) in shared memory; a lane then finds the
// exon piece(s) under its 32 positions with a short walk at shared-memory latency and fetches each
// piece with two 128-bit plane loads.  The next 32 words are gathered before the current ones are
// finished (software pipeline), which also supplies lane 31's look-ahead bits without a halo lane.
namespace {
constexpr int SCAN_SMEM_EXONS = 128;   // exon tables up to this size are staged per warp; longer chains stay in global memory

struct P4 { uint32_t T, A, G, R; };

struct ExonTable {
    const int32_t* cum;     // [n+1] spliced offset of each exon (genomic order), cum[n] = L   (shared)   or seg_cum + s0 (global, no sentinel)
    const int32_t* pos;     // [n]   bit position of the exon's first base relative to the warp's plane base (shared only)
    const orfx_seg* segs;   // global fallback
    const int32_t* sso;     // global fallback (islands) or nullptr
    int n, L, bstart, base_sh;
    bool staged;
    __device__ __forceinline__ int cum_at(int k) const { return (staged || k < n) ? cum[k] : L; }
    __device__ __forceinline__ int pos_at(int k) const {
        if (staged) return pos[k];
        return (sso ? sso[k] : segs[k].start - bstart) + base_sh;
    }
};

// the four plane words of spliced offsets [u0, u0+n) in genomic order; k = exon holding u0
__device__ __forceinline__ P4 gather_word(const ExonTable& E, const uint4* __restrict__ pl, int u0, int n, int k) {
    P4 r{0, 0, 0, 0};
    int filled = 0;
    while (filled < n && k < E.n) {
        const int u = u0 + filled;
        const int take = min(n - filled, E.cum_at(k + 1) - u);
        const int bit = E.pos_at(k) + (u - E.cum_at(k));
        const uint4 lo = __ldg(pl + (bit >> 5)), hi = __ldg(pl + (bit >> 5) + 1);
        const uint32_t sh = (uint32_t)bit & 31u, m = 0xFFFFFFFFu >> (32 - take);
        r.T |= (__funnelshift_r(lo.x, hi.x, sh) & m) << filled;
        r.A |= (__funnelshift_r(lo.y, hi.y, sh) & m) << filled;
        r.G |= (__funnelshift_r(lo.z, hi.z, sh) & m) << filled;
        r.R |= (__funnelshift_r(lo.w, hi.w, sh) & m) << filled;
        filled += take;
        k++;
    }
    return r;
}
}  // namespace

__global__ void __launch_bounds__(SCAN_WARPS_PER_CTA * 32) codon_scan_kernel(DevBatch B, int64_t n_items) {
    __shared__ int32_t s_cum[SCAN_WARPS_PER_CTA][SCAN_SMEM_EXONS + 1];
    __shared__ int32_t s_pos[SCAN_WARPS_PER_CTA][SCAN_SMEM_EXONS];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t item = (int64_t)blockIdx.x * SCAN_WARPS_PER_CTA + wib;
    if (item >= n_items) return;
    const int tx = B.tile_tx[item];
    const int L = B.tx_len[tx];
    const bool plus = B.tx_strand[tx] == '+';
    const int s0 = B.tx_seg_off[tx], nseg = B.tx_seg_off[tx + 1] - s0;
    const int bundle = B.tx_bundle[tx];
    const int nwords = (L + 31) >> 5;
    const int w_begin = (int)(item - B.tx_tile_off[tx]) * SCAN_TILE_WORDS;
    const int w_end = min(nwords, w_begin + SCAN_TILE_WORDS);
    const bool isl = B.seg_seq_off != nullptr;
    const int64_t base = B.seq_off[bundle];   // byte (= plane bit) offset of the bundle's first base
    const uint4* pl = reinterpret_cast<const uint4*>(B.planes) + (plus ? 0 : B.plane_words) + (base >> 5);
    ExonTable E;
    E.n = nseg; E.L = L; E.bstart = isl ? 0 : B.bundle_start[bundle]; E.base_sh = (int)(base & 31);
    E.segs = B.segs + s0; E.sso = isl ? B.seg_seq_off + s0 : nullptr;
    E.staged = nseg <= SCAN_SMEM_EXONS;
    if (E.staged) {
        for (int i = lane; i < nseg; i += 32) {
            s_cum[wib][i] = B.seg_cum[s0 + i];
            s_pos[wib][i] = (isl ? B.seg_seq_off[s0 + i] : B.segs[s0 + i].start - E.bstart) + E.base_sh;
        }
        if (lane == 0) s_cum[wib][nseg] = L;
        __syncwarp();
        E.cum = s_cum[wib]; E.pos = s_pos[wib];
    } else {
        E.cum = B.seg_cum + s0; E.pos = nullptr;
    }
    int k = 0;          // exon under this lane's current offset
    bool first = true;
    // planes of mask word w of this transcript, in transcript order (bit j = spliced position 32 w + j)
    auto word_planes = [&](int w) -> P4 {
        const int p0 = w << 5, n = min(32, L - p0);
        const int u0 = plus ? p0 : L - p0 - n;
        if (first) {   // exon holding u0: last k with cum[k] <= u0
            int lo = 0, hi = nseg - 1;
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (E.cum_at(mid) <= u0) lo = mid; else hi = mid - 1; }
            k = lo; first = false;
        } else {
            while (k + 1 < nseg && E.cum_at(k + 1) <= u0) k++;
            while (k > 0 && E.cum_at(k) > u0) k--;
        }
        P4 r = gather_word(E, pl, u0, n, k);
        if (!plus) {   // genomic order -> transcript order: bit j <- bit n-1-j
            const int sh = 32 - n;
            r.T = __brev(r.T) >> sh; r.A = __brev(r.A) >> sh; r.G = __brev(r.G) >> sh; r.R = __brev(r.R) >> sh;
        }
        return r;
    };
    const int64_t mbase = B.tx_mask_off[tx];
    P4 cur{0, 0, 0, 0};
    if (w_begin + lane < w_end) cur = word_planes(w_begin + lane);
    for (int wb = w_begin; wb < w_end; wb += 32) {
        const int wn = wb + 32 + lane;
        P4 nxt{0, 0, 0, 0};
        if (wn < nwords && (wn < w_end || lane == 0)) nxt = word_planes(wn);   // past the item only lane 0: the look-ahead word
        // the next word's planes live in the neighbour lane; lane 31 takes lane 0's of the next round
        uint32_t T1w = __shfl_down_sync(0xffffffffu, cur.T, 1), A1w = __shfl_down_sync(0xffffffffu, cur.A, 1);
        uint32_t G1w = __shfl_down_sync(0xffffffffu, cur.G, 1), R1w = __shfl_down_sync(0xffffffffu, cur.R, 1);
        const uint32_t T1n = __shfl_sync(0xffffffffu, nxt.T, 0), A1n = __shfl_sync(0xffffffffu, nxt.A, 0);
        const uint32_t G1n = __shfl_sync(0xffffffffu, nxt.G, 0), R1n = __shfl_sync(0xffffffffu, nxt.R, 0);
        if (lane == 31) { T1w = T1n; A1w = A1n; G1w = G1n; R1w = R1n; }
        const int w = wb + lane;
        if (w < w_end) {
            const uint32_t A1 = __funnelshift_r(cur.A, A1w, 1), A2 = __funnelshift_r(cur.A, A1w, 2);
            const uint32_t G1 = __funnelshift_r(cur.G, G1w, 1), G2 = __funnelshift_r(cur.G, G1w, 2);
            const uint32_t R1 = __funnelshift_r(cur.R, R1w, 1), R2 = __funnelshift_r(cur.R, R1w, 2);
            const uint32_t T1 = __funnelshift_r(cur.T, T1w, 1);
            // TAA TAG TAR | TGA TRA ; ATG
            B.stop_mask[mbase + w] = cur.T & ((A1 & (A2 | G2 | R2)) | ((G1 | R1) & A2));
            B.start_mask[mbase + w] = cur.A & T1 & G2;
        }
        cur = nxt;
    }
}

int launch_base_planes(const DevBatch& B, cudaStream_t stream) {
    if (B.seq_bytes <= 0 || B.n_bundles <= 0) return 0;
    const int64_t n_chunks = (B.seq_bytes + 31) / 32;
    const int64_t threads = (n_chunks + 31) / 32 * 32;
    base_planes_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(B, n_chunks);
    return 1;
}

void launch_codon_scan(const DevBatch& B, int64_t n_items, int n_sm, cudaStream_t stream) {
    (void)n_sm;
    if (n_items <= 0) return;
    const int64_t grid = (n_items + SCAN_WARPS_PER_CTA - 1) / SCAN_WARPS_PER_CTA;
    codon_scan_kernel<<<(unsigned)grid, SCAN_WARPS_PER_CTA * 32, 0, stream>>>(B, n_items);
}

}  // namespace orfx
