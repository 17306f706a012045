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
//     the base planes the window's strand needs:  '+' : T A G R;  other strands (transcript base =
//     complement of the genomic byte, gclib ntComplement): A (->T)  T|U (->A)  C (->G)  Y (->R).
//  2. codon_scan_kernel — one warp per tile of 31 mask words of one transcript.  A lane assembles the
//     four plane words of its 32 spliced positions by funnel-shifting bit runs out of the window
//     planes exon piece by exon piece (reversed with brev for the reverse strand), takes the next
//     word's low bits from its neighbour lane (shuffle; lane 31 only supplies the halo), evaluates the
//     codon logic on whole words and stores the two mask words, coalesced.
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
    uint32_t* out = B.planes + chunk;
    const int64_t ps = B.plane_words;
    const uint32_t a = plane_word<FOLD, 0x41414141u>(w);   // A: '+' A plane and reverse-strand T plane
    out[1 * ps] = a;
    if (need & 1u) {
        out[0 * ps] = plane_word<FOLD, 0x54545454u>(w);    // T
        out[2 * ps] = plane_word<FOLD, 0x47474747u>(w);    // G
        out[3 * ps] = plane_word<FOLD, 0x52525252u>(w);    // R
    }
    if (need & 2u) {
        out[4 * ps] = plane_word<FOLD_TU, 0x54545454u>(w); // T or U -> A
        out[5 * ps] = plane_word<FOLD, 0x43434343u>(w);    // C -> G
        out[6 * ps] = plane_word<FOLD, 0x59595959u>(w);    // Y -> R
    }
}

// ---- 2. planes -> per-transcript stop / start masks ---------------------------------------------------
namespace {
// n (1..32) bits of a plane starting at absolute bit position `pos`
__device__ __forceinline__ uint32_t plane_bits(const uint32_t* __restrict__ plane, int64_t pos, int n) {
    const int64_t wi = pos >> 5;
    const uint32_t lo = __ldg(plane + wi), hi = __ldg(plane + wi + 1);
    const uint32_t v = __funnelshift_r(lo, hi, (uint32_t)(pos & 31));
    return n >= 32 ? v : (v & ((1u << n) - 1u));
}
}  // namespace

__global__ void __launch_bounds__(SCAN_WARPS_PER_CTA * 32) codon_scan_kernel(DevBatch B, int64_t n_tiles) {
    const int lane = threadIdx.x & 31;
    const int64_t tile = (int64_t)blockIdx.x * SCAN_WARPS_PER_CTA + (threadIdx.x >> 5);
    if (tile >= n_tiles) return;
    const int tx = B.tile_tx[tile];
    const int tidx = (int)(tile - B.tx_tile_off[tx]);
    const int L = B.tx_len[tx];
    const bool plus = B.tx_strand[tx] == '+';
    const int w = tidx * SCAN_TILE_WORDS + lane;   // mask word of this lane (lane 31: halo for lane 30)
    const int p0 = w * 32;
    uint32_t T0 = 0, A0 = 0, G0 = 0, R0 = 0;
    if (p0 < L) {
        const int n = min(32, L - p0);
        const int s0 = B.tx_seg_off[tx], s1 = B.tx_seg_off[tx + 1];
        const int bundle = B.tx_bundle[tx];
        const bool isl = B.seg_seq_off != nullptr;
        // windows: byte offset of genomic coordinate 0; islands: byte offset of the bundle's first island
        const int64_t win = isl ? B.seq_off[bundle] : B.seq_off[bundle] - (int64_t)B.bundle_start[bundle];
        const int64_t ps = B.plane_words;
        const uint32_t* pT = B.planes + (plus ? 0 : 1) * ps;
        const uint32_t* pA = B.planes + (plus ? 1 : 4) * ps;
        const uint32_t* pG = B.planes + (plus ? 2 : 5) * ps;
        const uint32_t* pR = B.planes + (plus ? 3 : 6) * ps;
        // spliced offsets in genomic order covered by this word
        const int u0 = plus ? p0 : L - p0 - n;
        int lo = s0, hi = s1 - 1;
        while (lo < hi) {   // exon holding u0: last k with cum[k] <= u0
            const int mid = (lo + hi + 1) >> 1;
            if (B.seg_cum[mid] <= u0) lo = mid; else hi = mid - 1;
        }
        int filled = 0;
        for (int k = lo; filled < n && k < s1; k++) {
            const orfx_seg sg = B.segs[k];
            const int offs = u0 + filled - B.seg_cum[k];
            const int take = min(n - filled, sg.end - sg.start + 1 - offs);
            const int64_t pos = win + (isl ? B.seg_seq_off[k] : sg.start) + offs;
            T0 |= plane_bits(pT, pos, take) << filled;
            A0 |= plane_bits(pA, pos, take) << filled;
            G0 |= plane_bits(pG, pos, take) << filled;
            R0 |= plane_bits(pR, pos, take) << filled;
            filled += take;
        }
        if (!plus) {   // genomic order -> transcript order: bit j <- bit n-1-j
            const int sh = 32 - n;
            T0 = __brev(T0) >> sh; A0 = __brev(A0) >> sh; G0 = __brev(G0) >> sh; R0 = __brev(R0) >> sh;
        }
    }
    // the next word's planes live in the neighbour lane
    const uint32_t T1w = __shfl_down_sync(0xffffffffu, T0, 1), A1w = __shfl_down_sync(0xffffffffu, A0, 1);
    const uint32_t G1w = __shfl_down_sync(0xffffffffu, G0, 1), R1w = __shfl_down_sync(0xffffffffu, R0, 1);
    if (lane == 31 || p0 >= L) return;
    const uint32_t A1 = __funnelshift_r(A0, A1w, 1), A2 = __funnelshift_r(A0, A1w, 2);
    const uint32_t G1 = __funnelshift_r(G0, G1w, 1), G2 = __funnelshift_r(G0, G1w, 2);
    const uint32_t R1 = __funnelshift_r(R0, R1w, 1), R2 = __funnelshift_r(R0, R1w, 2);
    const uint32_t T1 = __funnelshift_r(T0, T1w, 1);
    // TAA TAG TAR | TGA TRA ; ATG
    const uint32_t stop = T0 & ((A1 & (A2 | G2 | R2)) | ((G1 | R1) & A2));
    const uint32_t start = A0 & T1 & G2;
    const int64_t base = B.tx_mask_off[tx] + w;
    B.stop_mask[base] = stop;
    B.start_mask[base] = start;
}

int launch_base_planes(const DevBatch& B, cudaStream_t stream) {
    if (B.seq_bytes <= 0 || B.n_bundles <= 0) return 0;
    const int64_t n_chunks = (B.seq_bytes + 31) / 32;
    const int64_t threads = (n_chunks + 31) / 32 * 32;
    base_planes_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(B, n_chunks);
    return 1;
}

void launch_codon_scan(const DevBatch& B, int64_t n_tiles, int n_sm, cudaStream_t stream) {
    (void)n_sm;
    if (n_tiles <= 0) return;
    const int64_t grid = (n_tiles + SCAN_WARPS_PER_CTA - 1) / SCAN_WARPS_PER_CTA;
    codon_scan_kernel<<<(unsigned)grid, SCAN_WARPS_PER_CTA * 32, 0, stream>>>(B, n_tiles);
}

}  // namespace orfx
