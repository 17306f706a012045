// Device helpers on CSR interval chains and on the codon-scan bit masks.
#pragma once
#include "orfx_device.cuh"

namespace orfx {

// One transcript's chain: segments ascending by genomic coordinate, `cum` = bases before each segment.
struct TxView {
    const orfx_seg* segs;
    const int32_t* cum;
    int n;        // segments
    int L;        // spliced length
    bool plus;    // strand == '+'; every other strand byte is handled as the reverse strand, as the
                  // reference does (`strand == '+' ? ... : ...` throughout Transcriptome.cpp)
    // --overhang: the row's query copy may have grown (Transcriptome.cpp:418-429, 461-474): its first segment
    // starts xlo bases earlier, its last one ends xhi bases later; L includes both.  0 everywhere else.
    int xlo, xhi;
};
__device__ __forceinline__ orfx_seg vseg(const TxView& v, int k) {
    orfx_seg s = v.segs[k];
    if (k == 0) s.start -= v.xlo;
    if (k == v.n - 1) s.end += v.xhi;
    return s;
}
__device__ __forceinline__ int vcum(const TxView& v, int k) { return v.cum[k] + (k > 0 ? v.xlo : 0); }
__device__ __forceinline__ void tx_grow(TxView& v, int dlo, int dhi) { v.xlo += dlo; v.xhi += dhi; v.L += dlo + dhi; }

__device__ __forceinline__ TxView tx_view(const DevBatch& B, int tx) {
    TxView v;
    int s0 = B.tx_seg_off[tx];
    v.segs = B.segs + s0;
    v.cum = B.seg_cum + s0;
    v.n = B.tx_seg_off[tx + 1] - s0;
    v.L = B.tx_len[tx];
    v.plus = B.tx_strand[tx] == '+';
    v.xlo = v.xhi = 0;
    return v;
}

// index of the segment containing genomic position g, or -1
__device__ __forceinline__ int seg_of_genomic(const TxView& v, int g) {
    int lo = 0, hi = v.n - 1;
    if (hi < 0 || g < v.segs[0].start - v.xlo) return -1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (v.segs[mid].start <= g) lo = mid; else hi = mid - 1;   // mid >= 1: never the grown start
    }
    return g <= vseg(v, lo).end ? lo : -1;
}
// genomic -> transcript-order offset (CHAIN::genome2nt, Transcriptome.h:447-462); -1 when not exonic
__device__ __forceinline__ int g2t(const TxView& v, int g) {
    int k = seg_of_genomic(v, g);
    if (k < 0) return -1;
    int u = vcum(v, k) + (g - vseg(v, k).start);
    return v.plus ? u : v.L - 1 - u;
}
// index of the segment holding genomic-order spliced offset u (0 <= u < L)
__device__ __forceinline__ int seg_of_offset(const TxView& v, int u) {
    int lo = 0, hi = v.n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (vcum(v, mid) <= u) lo = mid; else hi = mid - 1;
    }
    return lo;
}
// transcript-order offset -> genomic (CHAIN::nt2genome, Transcriptome.h:463-481); p must be in [0,L)
__device__ __forceinline__ int t2g(const TxView& v, int p) {
    int u = v.plus ? p : v.L - 1 - p;
    int k = seg_of_offset(v, u);
    return vseg(v, k).start + (u - vcum(v, k));
}

// ---- in-frame scans over a transcript's bit mask (bit p = codon starting at spliced position p) ----
__device__ __forceinline__ uint32_t frame_pattern(int word, int residue) {
    // bits b of word `word` with (32*word + b) % 3 == residue
    int s = (residue - (word * 2) % 3 + 3) % 3;   // 32 == 2 (mod 3)
    return s == 0 ? 0x49249249u : (s == 1 ? 0x92492492u : 0x24924924u);
}
__device__ __forceinline__ bool mask_bit(const uint32_t* m, int p) { return (m[p >> 5] >> (p & 31)) & 1u; }
// The mask region of a transcript starts 16-byte aligned and is padded to a multiple of 4 words
// (tx_prep_kernel), so long scans move 4 words per load with the next load already in flight.
// smallest q >= p with q == p (mod 3), q < L and bit set; -1 when none
__device__ __forceinline__ int next_inframe(const uint32_t* m, int p, int L) {
    if (p < 0 || p >= L) return -1;
    const int r = p % 3, nw = (L + 31) >> 5;
    int w = p >> 5;
    uint32_t x = m[w] & frame_pattern(w, r) & (0xFFFFFFFFu << (p & 31));
    while (true) {   // head: up to the next 4-word boundary
        if (x) { const int q = (w << 5) + __ffs(x) - 1; return q < L ? q : -1; }
        if (++w >= nw) return -1;
        if ((w & 3) == 0) break;
        x = m[w] & frame_pattern(w, r);
    }
    // 16 words per round, four independent 128-bit loads in flight plus the next round's four: an ORF's
    // stop is typically kilobases away, and one dependent load per 128 positions would make that a
    // latency chain
    uint32_t pa = frame_pattern(w, r), pb = frame_pattern(w + 1, r), pc = frame_pattern(w + 2, r);
    const uint4 zero = make_uint4(0, 0, 0, 0);
    auto load4 = [&](int at) { return at < nw ? *reinterpret_cast<const uint4*>(m + at) : zero; };
    auto hit4 = [&](const uint4& v, int at, uint32_t a, uint32_t b, uint32_t c) {   // first set in-frame bit of words [at, at+4), or -1
        const uint32_t x0 = v.x & a, x1 = v.y & b, x2 = v.z & c, x3 = v.w & a;
        if (!(x0 | x1 | x2 | x3)) return -1;
        if (x0) return (at << 5) + __ffs(x0) - 1;
        if (x1) return ((at + 1) << 5) + __ffs(x1) - 1;
        if (x2) return ((at + 2) << 5) + __ffs(x2) - 1;
        return ((at + 3) << 5) + __ffs(x3) - 1;
    };
    uint4 v0 = load4(w), v1 = load4(w + 4), v2 = load4(w + 8), v3 = load4(w + 12);
    while (true) {
        const int wn = w + 16;
        const uint4 n0 = load4(wn), n1 = load4(wn + 4), n2 = load4(wn + 8), n3 = load4(wn + 12);
        int q = hit4(v0, w, pa, pb, pc);                 // 4 words on, the frame pattern advances by one
        if (q < 0) q = hit4(v1, w + 4, pb, pc, pa);
        if (q < 0) q = hit4(v2, w + 8, pc, pa, pb);
        if (q < 0) q = hit4(v3, w + 12, pa, pb, pc);
        if (q >= 0) return q < L ? q : -1;               // words past the last real one are padding: any bit there is >= L
        if (wn >= nw) return -1;
        w = wn; v0 = n0; v1 = n1; v2 = n2; v3 = n3;
        const uint32_t t = pa; pa = pb; pb = pc; pc = t;  // 16 words on: one step as well
    }
}
// The same for a whole warp (heavy bundles: all 32 lanes work on one row and call this with the same arguments).  A
// titin-class ORF is tens of kilobases long; the scalar scan above walks it 16 words per round, one lane's loads at a
// time.  Here every lane takes its own 4-word block of the round (128 words = 4 096 positions per round), the lowest
// lane with a hit wins.
__device__ __forceinline__ int next_inframe_warp(const uint32_t* m, int p, int L, int lane) {
    if (p < 0 || p >= L) return -1;
    const int r = p % 3, nw = (L + 31) >> 5;
    const int w_first = p >> 5;
    for (int base = w_first & ~3; base < nw; base += 128) {
        const int w = base + 4 * lane;                        // this lane's block: words [w, w + 4)
        int q = -1;
        if (w < nw) {                                         // regions are padded to 4 words: the load stays inside
            const uint4 v = *reinterpret_cast<const uint4*>(m + w);
            uint32_t x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 3; k >= 0; k--) {
                uint32_t bits = x[k] & frame_pattern(w + k, r);
                if (w + k < w_first) bits = 0;
                else if (w + k == w_first) bits &= 0xFFFFFFFFu << (p & 31);
                if (bits) q = ((w + k) << 5) + __ffs(bits) - 1;
            }
        }
        const unsigned hit = __ballot_sync(0xffffffffu, q >= 0);
        if (hit) {
            q = __shfl_sync(0xffffffffu, q, __ffs(hit) - 1);
            return q < L ? q : -1;                            // bits past the last real position are padding
        }
    }
    return -1;
}
// largest q <= p with q == p (mod 3), q >= 0 and bit set; -1 when none
__device__ __forceinline__ int prev_inframe(const uint32_t* m, int p) {
    if (p < 0) return -1;
    const int r = p % 3;
    int w = p >> 5;
    uint32_t x = m[w] & frame_pattern(w, r) & (0xFFFFFFFFu >> (31 - (p & 31)));
    while (true) {   // head: down to a 4-word boundary
        if (x) return (w << 5) + 31 - __clz(x);
        if (w == 0) return -1;
        if ((w & 3) == 0) break;
        --w;
        x = m[w] & frame_pattern(w, r);
    }
    w -= 4;   // block [w, w+4)
    uint32_t pa = frame_pattern(w, r), pb = frame_pattern(w + 1, r), pc = frame_pattern(w + 2, r);
    uint4 v = *reinterpret_cast<const uint4*>(m + w);
    while (true) {
        uint4 nx = make_uint4(0, 0, 0, 0);
        if (w >= 4) nx = *reinterpret_cast<const uint4*>(m + w - 4);
        const uint32_t x0 = v.x & pa, x1 = v.y & pb, x2 = v.z & pc, x3 = v.w & pa;
        if (x0 | x1 | x2 | x3) {
            if (x3) return ((w + 3) << 5) + 31 - __clz(x3);
            if (x2) return ((w + 2) << 5) + 31 - __clz(x2);
            if (x1) return ((w + 1) << 5) + 31 - __clz(x1);
            return (w << 5) + 31 - __clz(x0);
        }
        if (w < 4) return -1;
        w -= 4; v = nx;
        const uint32_t t = pc; pc = pb; pb = pa; pa = t;   // 4 words back: the pattern steps the other way
    }
}

}  // namespace orfx
