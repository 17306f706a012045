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
#include <climits>

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

// Same planes from 4-bit base codes (ORFX_SEQ_NT4: two bases per byte): one 128-bit load per 32 bases.
namespace {
template <uint32_t CODE>
__device__ __forceinline__ uint32_t nibble_eq8(uint32_t w) {   // bit k = (nibble k of w == CODE), k = 0..7
    const uint32_t y = w ^ (CODE * 0x11111111u);
    const uint32_t t = ((y & 0x77777777u) + 0x77777777u) | y;
    uint32_t f = (~t & 0x88888888u) >> 3;           // bit 4k
    f = (f | (f >> 3)) & 0x03030303u;                // two bits per byte
    f = (f | (f >> 6)) & 0x000F000Fu;                // four bits per half
    return (f | (f >> 12)) & 0xFFu;
}
template <uint32_t CODE>
__device__ __forceinline__ uint32_t plane_word4(const uint4& v) {
    return nibble_eq8<CODE>(v.x) | (nibble_eq8<CODE>(v.y) << 8) | (nibble_eq8<CODE>(v.z) << 16) | (nibble_eq8<CODE>(v.w) << 24);
}
}  // namespace
__global__ void __launch_bounds__(256) base_planes_nt4_kernel(DevBatch B, int64_t n_chunks) {
    const int64_t chunk = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // 32 bases = 16 bytes each
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
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(B.seq) + chunk);
    uint4* out = reinterpret_cast<uint4*>(B.planes) + chunk;
    const uint32_t a = plane_word4<ORFX_NT4_A>(v);
    if (need & 1u) out[0] = make_uint4(plane_word4<ORFX_NT4_T>(v), a, plane_word4<ORFX_NT4_G>(v), plane_word4<ORFX_NT4_R>(v));
    if (need & 2u)
        out[B.plane_words] = make_uint4(a, plane_word4<ORFX_NT4_T>(v) | plane_word4<ORFX_NT4_U>(v), plane_word4<ORFX_NT4_C>(v), plane_word4<ORFX_NT4_Y>(v));
}

// Same planes from 2-bit base codes (ORFX_SEQ_NT2: four bases per byte, A C G T only): one 64-bit load per 32 bases.  The few
// bases that are none of the four are listed apart (seq_exc_*) and patched into the planes by plane_exceptions_kernel.
namespace {
__device__ __forceinline__ uint32_t compact_even16(uint32_t x) {   // bits 0, 2, 4, ... 30 -> bits 0 .. 15
    x = (x | (x >> 1)) & 0x33333333u;
    x = (x | (x >> 2)) & 0x0F0F0F0Fu;
    x = (x | (x >> 4)) & 0x00FF00FFu;
    return (x | (x >> 8)) & 0x0000FFFFu;
}
}  // namespace
__global__ void __launch_bounds__(256) base_planes_nt2_kernel(DevBatch B, int64_t n_chunks) {
    const int64_t chunk = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // 32 bases = 8 bytes each
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
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(B.seq) + chunk);
    const uint32_t M = 0x55555555u;
    const uint32_t l0 = v.x & M, h0 = (v.x >> 1) & M, l1 = v.y & M, h1 = (v.y >> 1) & M;
    const uint32_t a = compact_even16(~h0 & ~l0 & M) | (compact_even16(~h1 & ~l1 & M) << 16);
    const uint32_t c = compact_even16(~h0 & l0) | (compact_even16(~h1 & l1) << 16);
    const uint32_t g = compact_even16(h0 & ~l0) | (compact_even16(h1 & ~l1) << 16);
    const uint32_t t = compact_even16(h0 & l0) | (compact_even16(h1 & l1) << 16);
    uint4* out = reinterpret_cast<uint4*>(B.planes) + chunk;
    if (need & 1u) out[0] = make_uint4(t, a, g, 0u);
    if (need & 2u) out[B.plane_words] = make_uint4(a, t, c, 0u);
}
// one thread per listed base: its bit in the eight plane words of its chunk is rewritten from its ORFX_NT4_* code
__global__ void __launch_bounds__(256) plane_exceptions_kernel(DevBatch B) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B.n_seq_exc) return;
    const int64_t pos = B.seq_exc_pos[i];
    if (pos < 0 || pos >= B.seq_bytes) return;
    const uint32_t code = B.seq_exc_code[i], bit = 1u << (pos & 31);
    uint32_t* fwd = B.planes + (pos >> 5) * 4;                     // {T, A, G, R}
    uint32_t* rev = B.planes + (B.plane_words + (pos >> 5)) * 4;   // {A, T|U, C, Y}
    const bool f[4] = {code == ORFX_NT4_T, code == ORFX_NT4_A, code == ORFX_NT4_G, code == ORFX_NT4_R};
    const bool r[4] = {code == ORFX_NT4_A, code == ORFX_NT4_T || code == ORFX_NT4_U, code == ORFX_NT4_C, code == ORFX_NT4_Y};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (f[k]) atomicOr(fwd + k, bit); else atomicAnd(fwd + k, ~bit);
        if (r[k]) atomicOr(rev + k, bit); else atomicAnd(rev + k, ~bit);
    }
}

// ---- 2. planes -> per-transcript stop / start masks ---------------------------------------------------
// One THREAD per run of SCAN_TILE_WORDS consecutive mask words (256 spliced positions) of one transcript.
// The thread locates the exon under the run's first position once, then streams: every round appends up
// to 32 plane bits of the current exon piece to a 64-bit window per plane (one new 128-bit plane load
// per round while it stays inside an exon) and, once 34 bits are there (32 positions + the two
// look-ahead bases of the last codons), evaluates the codon logic on whole words and emits one word of
// each mask.  Per-word costs of a gather formulation (exon search, junction masks, neighbour shuffles)
// are paid once per run or once per exon junction instead.  The run's words are parked in shared memory
// (dynamic index) and leave as 128-bit stores.
namespace {
struct Win { uint64_t T, A, G, R; };

__device__ __forceinline__ void win_append(Win& w, const uint4& lo, const uint4& hi, uint32_t sh, int take, int fill, bool reverse) {
    const uint32_t m = 0xFFFFFFFFu >> (32 - take);
    uint32_t t = __funnelshift_r(lo.x, hi.x, sh) & m, a = __funnelshift_r(lo.y, hi.y, sh) & m;
    uint32_t g = __funnelshift_r(lo.z, hi.z, sh) & m, r = __funnelshift_r(lo.w, hi.w, sh) & m;
    if (reverse) {   // genomic order -> transcript order
        const int rs = 32 - take;
        t = __brev(t) >> rs; a = __brev(a) >> rs; g = __brev(g) >> rs; r = __brev(r) >> rs;
    }
    w.T |= (uint64_t)t << fill; w.A |= (uint64_t)a << fill; w.G |= (uint64_t)g << fill; w.R |= (uint64_t)r << fill;
}
}  // namespace

__global__ void __launch_bounds__(CSCAN_THREADS) codon_scan_kernel(DevBatch B, int64_t n_tiles) {
    __shared__ uint32_t s_out[2][SCAN_TILE_WORDS][CSCAN_THREADS];
    const int tid = threadIdx.x;
    const int64_t tile = (int64_t)blockIdx.x * CSCAN_THREADS + tid;
    if (tile >= n_tiles) return;
    const int tx = B.tile_tx[tile];
    // --overhang: the scanned chain is the exon chain with its first segment started xlo bases earlier and its last
    // one ended xhi bases later (the usable flanks, orfx_batch::tx_ext_lo / tx_ext_hi)
    const int xlo = B.tx_ext_lo ? B.tx_ext_lo[tx] : 0, xhi = B.tx_ext_hi ? B.tx_ext_hi[tx] : 0;
    const int L = B.tx_len[tx] + xlo + xhi;
    const int nwords = (L + 31) >> 5;
    const int w0 = (int)(tile - B.tx_tile_off[tx]) * SCAN_TILE_WORDS;
    const int nw = min(SCAN_TILE_WORDS, nwords - w0);
    if (nw <= 0) return;
    const bool plus = B.tx_strand[tx] == '+';
    const int s0 = B.tx_seg_off[tx], nseg = B.tx_seg_off[tx + 1] - s0;
    const int bundle = B.tx_bundle[tx];
    const bool isl = B.seg_seq_off != nullptr;
    const int64_t base = B.seq_off[bundle];   // byte (= plane bit) offset of the bundle's first base
    const int base_sh = (int)(base & 31);
    const int bstart = isl ? 0 : B.bundle_start[bundle];
    const uint4* __restrict__ pl = reinterpret_cast<const uint4*>(B.planes) + (plus ? 0 : B.plane_words) + (base >> 5);
    const int32_t* __restrict__ cum0 = B.seg_cum + s0;
    const orfx_seg* __restrict__ segs = B.segs + s0;
    const int32_t* __restrict__ sso = isl ? B.seg_seq_off + s0 : nullptr;
    auto cum = [&](int k) { return cum0[k] + (k > 0 ? xlo : 0); };
    auto cum_at = [&](int k) { return k < nseg ? cum(k) : L; };
    auto pos_at = [&](int k) { return (isl ? sso[k] : segs[k].start - bstart) + base_sh - (k == 0 ? xlo : 0); };

    const int p_lo = w0 << 5;
    const int nbits = min(L, ((w0 + nw) << 5) + 2) - p_lo;   // spliced positions this run reads, look-ahead included
    // exon under the first position (genomic-order spliced offset u): last k with cum[k] <= u
    const int u = plus ? p_lo : L - 1 - p_lo;
    int k;
    {
        int lo = 0, hi = nseg - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (cum(mid) <= u) lo = mid; else hi = mid - 1; }
        k = lo;
    }
    int c0 = cum(k), c1 = cum_at(k + 1);
    int rem = plus ? c1 - u : u - c0 + 1;     // positions left in the current exon, in reading direction
    int bit = pos_at(k) + (u - c0);           // plane bit of the next position to read
    Win w{0, 0, 0, 0};
    uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
    int ci = INT_MIN;                         // plane chunk held in lo (hi = chunk ci + 1)
    int fill = 0, produced = 0, emitted = 0;
    // Tried and measured slower on configs[1] (0.176 ms for this loop): requesting the next chunk a round ahead and the
    // next exon's table entries an exon ahead (+27 % instructions on a kernel whose ALU pipe is 70 % busy: 0.204 ms); a
    // per-word formulation with a 3-funnel-shift fast path for words inside one exon and piecewise assembly for the rest,
    // in one loop (0.330 ms: lanes work on different transcripts, so the warp pays both paths at every word) or with the
    // junction words deferred to a second pass (0.289 ms).
    while (emitted < nw) {
        if (fill < 34 && produced < nbits) {
            const int take = min(min(32, rem), min(nbits - produced, 64 - fill));
            const int lowbit = plus ? bit : bit - take + 1;
            const int ni = lowbit >> 5;
            if (ni != ci) {   // steady state inside an exon: one new chunk per round, the other is carried over
                const bool fwd = ni == ci + 1, bwd = ni == ci - 1;
                const uint4 keep_lo = lo;
                if (fwd) lo = hi; else lo = __ldg(pl + ni);
                if (bwd) hi = keep_lo; else hi = __ldg(pl + ni + 1);
                ci = ni;
            }
            win_append(w, lo, hi, (uint32_t)lowbit & 31u, take, fill, !plus);
            fill += take; produced += take; rem -= take;
            bit += plus ? take : -take;
            if (rem == 0 && produced < nbits) {   // next exon in reading direction
                k += plus ? 1 : -1;
                c0 = cum(k); c1 = cum_at(k + 1);
                rem = c1 - c0;
                bit = pos_at(k) + (plus ? 0 : rem - 1);
            }
        }
        if (fill >= 34 || produced >= nbits) {
            const uint32_t T0 = (uint32_t)w.T, A0 = (uint32_t)w.A;
            const uint32_t T1 = (uint32_t)(w.T >> 1);
            const uint32_t A1 = (uint32_t)(w.A >> 1), A2 = (uint32_t)(w.A >> 2);
            const uint32_t G1 = (uint32_t)(w.G >> 1), G2 = (uint32_t)(w.G >> 2);
            const uint32_t R1 = (uint32_t)(w.R >> 1), R2 = (uint32_t)(w.R >> 2);
            // TAA TAG TAR | TGA TRA ; ATG
            s_out[0][emitted][tid] = T0 & ((A1 & (A2 | G2 | R2)) | ((G1 | R1) & A2));
            s_out[1][emitted][tid] = A0 & T1 & G2;
            emitted++;
            w.T >>= 32; w.A >>= 32; w.G >>= 32; w.R >>= 32;
            fill = max(fill - 32, 0);
        }
    }
    // mask regions are 16-byte aligned and padded to 4 words; a run starts on a 4-word boundary
    const int64_t at = B.tx_mask_off[tx] + w0;
#pragma unroll
    for (int q = 0; q < SCAN_TILE_WORDS / 4; q++) {
        if (q * 4 >= nw) break;
        uint4 so, sa;
        so.x = s_out[0][q * 4][tid]; sa.x = s_out[1][q * 4][tid];
        so.y = nw > q * 4 + 1 ? s_out[0][q * 4 + 1][tid] : 0u; sa.y = nw > q * 4 + 1 ? s_out[1][q * 4 + 1][tid] : 0u;
        so.z = nw > q * 4 + 2 ? s_out[0][q * 4 + 2][tid] : 0u; sa.z = nw > q * 4 + 2 ? s_out[1][q * 4 + 2][tid] : 0u;
        so.w = nw > q * 4 + 3 ? s_out[0][q * 4 + 3][tid] : 0u; sa.w = nw > q * 4 + 3 ? s_out[1][q * 4 + 3][tid] : 0u;
        *reinterpret_cast<uint4*>(B.stop_mask + at + q * 4) = so;
        *reinterpret_cast<uint4*>(B.start_mask + at + q * 4) = sa;
    }
}

int launch_base_planes(const DevBatch& B, cudaStream_t stream) {
    if (B.seq_bytes <= 0 || B.n_bundles <= 0) return 0;
    const int64_t n_chunks = (B.seq_bytes + 31) / 32;
    const int64_t threads = (n_chunks + 31) / 32 * 32;
    if (B.seq_format == ORFX_SEQ_NT2) {
        base_planes_nt2_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(B, n_chunks);
        if (B.n_seq_exc <= 0) return 1;
        plane_exceptions_kernel<<<(unsigned)((B.n_seq_exc + 255) / 256), 256, 0, stream>>>(B);
        return 2;
    }
    if (B.seq_format == ORFX_SEQ_NT4) base_planes_nt4_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(B, n_chunks);
    else base_planes_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(B, n_chunks);
    return 1;
}

void launch_codon_scan(const DevBatch& B, int64_t n_items, int, cudaStream_t stream) {
    if (n_items <= 0) return;
    const int64_t grid = (n_items + CSCAN_THREADS - 1) / CSCAN_THREADS;
    codon_scan_kernel<<<(unsigned)grid, CSCAN_THREADS, 0, stream>>>(B, n_items);
}

}  // namespace orfx
