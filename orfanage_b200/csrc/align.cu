// Protein pairwise alignment: affine-gap extension DP (the ksw2 "extz" recurrence, score only),
// 24-letter alphabet, BLOSUM50-style matrix mat50 (common.hpp:239-265), int32 cells, no tensor cores.
//
//   H(i,j) = max{ H(i-1,j-1) + s(t_i,q_j), E(i,j), F(i,j) }
//   E(i+1,j) = max{ H(i,j) - gapo, E(i,j) } - gape          (gap along the target)
//   F(i,j+1) = max{ H(i,j) - gapo, F(i,j) } - gape          (gap along the query)
//   borders  H(-1,-1) = 0, H(-1,j) = -(gapo + (j+1) gape), H(i,-1) = -(gapo + (i+1) gape)
//   outputs  max  = max(0, max H), score = H(tlen-1,qlen-1), mqe = max_i H(i,qlen-1), mte = max_j H(tlen-1,j)
// (src/aln.hpp:27-41 calls ksw_extz2_sse with w=-1, zdrop=-1, end_bonus=0; build definition SURVEY.md §8a A1.)
//
// Shape: one warp per pair, pairs handed out longest-first from an atomic counter (persistent warps).
// Lanes own CPL consecutive query columns (H of the previous row and E live in registers); lane l works
// on target row (step - l), so the anti-diagonal wavefront moves through the warp and only the lane's
// right edge (H, F) and the target residue travel to the next lane by shuffle.  Queries wider than
// 32*CPL columns are processed in stripes; the right edge of a stripe (2 int32 per target row) goes
// through a per-warp scratch in global memory, read and written 32 rows at a time, coalesced.
#include "orfx_device.cuh"
#include "orfx_buffers.cuh"
#include "orfx_mat50.h"
#include <algorithm>
#include <mutex>
#include <numeric>
#include <string>
#include <type_traits>
#include <vector>

namespace orfx {

namespace {

// mat50, the reference's protein substitution matrix (src/common.hpp:239-265): include/orfx_mat50.h
__constant__ int8_t c_mat50[24 * 24] = ORFX_MAT50_LITERAL;

constexpr int NEG_INF = -0x40000000;
constexpr int MAT_STRIDE = 32;
#ifndef ORFX_ALN_WARPS
#define ORFX_ALN_WARPS 16
#endif
constexpr int ALN_WARPS = ORFX_ALN_WARPS;       // one CTA per SM; its warps work alone on short pairs and as a pipeline on long ones

struct AlnAcc { int gmax, mqe, mte, score; };

// What a stripe reads on its left and writes on its right: H(i, c_base-1) / F(i, c_base) + (q+e) per target row.
// PIPE: the stripes of one pair run concurrently on the warps of a CTA; a stripe may read row chunk
// [s, s+32) of its left edge once the warp of the previous stripe has posted that many rows.
struct StripeIO {
    const int32_t* h_in; const int32_t* f_in;      // nullptr for the first stripe
    int32_t* h_out; int32_t* f_out;                // nullptr for the last stripe
    unsigned long long* wait_flag; unsigned long long wait_base;   // shared-memory progress counters, read and written with atomics
    unsigned long long* post_flag; unsigned long long post_base;
};

// Per-warp shared memory of the DP (dynamic shared memory, one WarpSmem per warp).
struct WarpSmem {
    uint2 prof[24][32];     // query profile: prof[r][lane] = the lane's CPL scores against target residue r, one int8 each
    int   tl[96];           // target residues of rows blk - 32 .. blk + 63 (blk = first step of the current 32-step block), as byte
                            // offsets into a lane's profile column (residue * 256): a step's lookup is base + immediate, no index math
    int2  in_ring[32];      // left edge {H(i, c_base-1), F(i, c_base) + (q+e)} of the current chunk (borders for the first stripe)
    int2  out_ring[64];     // right edge of the rows lane 31 finished, until their chunk is stored
    int   q_ring[32][8];    // last stripe: H of the CPL columns of the lane that owns column qlen-1, one row per step of the block
};
constexpr int PAD_SCORE8 = -128;    // score of the virtual columns right of the query's end: padded cells stay below
                                    // H(i, qlen-1) - (gapo+gape) per column as long as gapo+gape <= 64

// One stripe of 32*CPL query columns over all target rows.  Lanes own CPL consecutive columns; lane l works on
// target row (step - l), so only the lane's right edge (H, F) moves to the next lane (two shuffles per step).
// The substitution scores of a step come as ONE shared-memory load of the lane's profile entry (fetched one step
// ahead) and are added to the diagonal with dp4a (one-hot selector) on the FMA pipe, which leaves the ALU pipe
// to the max / add-max chain:
//   x = d + s;  u = max(x, E);  h = max(u, F);  F' = max(F - e, u - (q+e));  E' = max(E - e, h - (q+e))
// F' may take u instead of h because F - (q+e) <= F - e; with E and F kept shifted up by (q+e) all four are add-max
// instructions (VIADDMNMX), no plain add is left, and the chain carried from cell to cell is ONE instruction deep (F' needs
// only F and u, and u does not depend on F):
//   u = max(Eh - (q+e), x);  h = max(Fh - (q+e), u);  Fh' = max(Fh - e, u);  Eh' = max(Eh - e, h)      (4.5 ALU + 1 FMA per cell)
// Steps go in blocks of 32; blocks in which every lane has a row run without per-step range checks.
// LAST: the stripe holding column qlen-1 (tracks mqe / score, writes no right edge).
// A protein as the kernels read it: residue i is p[i * stride] (stride 1: caller-supplied proteins; stride 3: the codons of one
// reading frame in the per-position translation of a transcript, translate.cu).
struct Res {
    const uint8_t* p; int stride;
    __device__ __forceinline__ int operator[](int i) const { return min((int)p[(long long)i * stride], 23); }
};
template <int CPL, bool PIPE, bool LAST>
__device__ __forceinline__ void dp_stripe(const int* __restrict__ s_mat, WarpSmem& ws, const Res q, const Res t,
                                          int qlen, int tlen, int c_base, int gapoe, int gape, const StripeIO& io, AlnAcc& acc, int lane) {
    static_assert(CPL <= 8, "a profile entry holds 8 scores");
    const int c0 = c_base + lane * CPL;
    const bool first_stripe = io.h_in == nullptr;
    int hp[CPL], ee[CPL];
    {
        int qcode[CPL];
#pragma unroll
        for (int k = 0; k < CPL; k++) {
            const int c = c0 + k;
            qcode[k] = c < qlen ? q[c] : -1;
            hp[k] = -(gapoe + gape * c);                 // H(-1, c)
            ee[k] = -(gapoe + gape * c);                 // E(0, c) + (q+e)
        }
        __syncwarp();   // the previous stripe's profile and rings are no longer read
        for (int r = 0; r < 24; r++) {
            uint32_t lo = 0, hi = 0;
#pragma unroll
            for (int k = 0; k < CPL; k++) {
                const uint32_t b = (uint32_t)(qcode[k] >= 0 ? s_mat[r * MAT_STRIDE + qcode[k]] : PAD_SCORE8) & 0xFFu;
                if (k < 4) lo |= b << (8 * k); else hi |= b << (8 * (k - 4));
            }
            ws.prof[r][lane] = make_uint2(lo, hi);
        }
        ws.tl[lane] = 0;                                      // rows -32 .. -1
        ws.tl[32 + lane] = lane < tlen ? t[lane] << 8 : 0;     // rows 0 .. 31; every block loads the chunk after its own
    }
    // lane-local column of qlen-1 (last stripe only)
    const int jq_lane = (qlen - 1 - c_base) / CPL, jq = (qlen - 1 - c_base) % CPL;
    int diag = c0 == 0 ? 0 : -(gapoe + gape * (c0 - 1));   // H(i-1, c0-1), starts at row -1
    int h_out = 0, f_out = 0;                              // what this lane hands to lane+1
    int gmax = acc.gmax, mqe = acc.mqe, mte = acc.mte, score = acc.score;
    uint2 sc = make_uint2(0, 0);                           // profile entry of this lane's next row
    const int n_steps = tlen + 31;
    int stored = 0;                                        // rows of the right edge already in global memory
    int blk = 0;                                           // first step of the current block
    const char* const prof_lane = reinterpret_cast<const char*>(&ws.prof[0][lane]);
    const int* const tl_lane = ws.tl + 33 - lane;          // tl_lane[s - blk] = profile offset of the lane's row at step s + 1

    auto step = [&](int s, auto full_tag) {
        constexpr bool FULL = decltype(full_tag)::value;
        // values arriving from the left neighbour (computed in the previous step); lane 0 reads the edge ring.  (Measured and
        // dropped: lane 31 loading the ring value into its own registers and a rotating shuffle instead of the two selects --
        // the divergent block costs what the selects cost: 2 935 against 2 965 GCUPS.)
        int h_left = __shfl_up_sync(0xffffffffu, h_out, 1);
        int f_in = __shfl_up_sync(0xffffffffu, f_out, 1);
        const int2 edge = ws.in_ring[s & 31];
        if (lane == 0) { h_left = edge.x; f_in = edge.y; }
        const int i = s - lane;
        if (FULL || (i >= 0 && i < tlen)) {
            const uint2 cur = sc;
            sc = *reinterpret_cast<const uint2*>(prof_lane + tl_lane[s - blk]);   // next step's scores: off the critical path
            int f = f_in, d = diag, h = 0, m0 = NEG_INF, m1 = NEG_INF;
#pragma unroll
            for (int k = 0; k < CPL; k++) {
                const int x = __dp4a((int)(k < 4 ? cur.x : cur.y), 1 << (8 * (k & 3)), d);
                const int u = __viaddmax_s32(ee[k], -gapoe, x);
                h = __viaddmax_s32(f, -gapoe, u);
                f = __viaddmax_s32(f, -gape, u);
                ee[k] = __viaddmax_s32(ee[k], -gape, h);
                d = hp[k];
                hp[k] = h;
                if (k & 1) m1 = __vimax3_s32(m1, m0, h); else m0 = h;
            }
            const int rowmax = (CPL & 1) ? max(m1, m0) : m1;
            gmax = max(gmax, rowmax);
            diag = h_left;
            h_out = h; f_out = f;
            if (!FULL && i == tlen - 1) mte = max(mte, rowmax);
            // mqe / score need H(i, qlen-1): column jq of lane jq_lane, a runtime register index.  That lane parks its CPL values
            // (vector stores, no select chain on the ALU pipe); the block's 32 rows are picked out of shared memory afterwards.
            if (LAST && lane == jq_lane) {
                int* qr = ws.q_ring[(s - blk) & 31];
                if (CPL == 8) { *reinterpret_cast<int4*>(qr) = make_int4(hp[0], hp[1], hp[2], hp[3]); *reinterpret_cast<int4*>(qr + 4) = make_int4(hp[4], hp[5], hp[6], hp[7]); }
                else if (CPL == 4) *reinterpret_cast<int4*>(qr) = make_int4(hp[0], hp[1], hp[2], hp[3]);
                else {
#pragma unroll
                    for (int k = 0; k < CPL; k++) qr[k] = hp[k];
                }
            }
        } else if (i == -1) {
            sc = *reinterpret_cast<const uint2*>(prof_lane + ws.tl[32]);   // the step before the lane's first row (first block: row 0 is tl[32])
        }
        if (!LAST && lane == 31 && (FULL || s >= 31)) ws.out_ring[(s - 31) & 63] = make_int2(h_out, f_out);
    };

    for (int s0 = 0; s0 < n_steps; s0 += 32) {
        {   // next 32 target rows' left edge into the ring, the residues of the chunk after it
            const int r = s0 + lane;
            int2 e = make_int2(-(gapoe + gape * r), -(gapoe + gape * r));   // first stripe: H(r, -1), F(r, 0) + (q+e)
            if (!first_stripe) {
                if (PIPE && s0 < tlen) {
                    const unsigned long long need = io.wait_base + (unsigned long long)min(s0 + 32, tlen);
                    if (lane == 0) while (atomicAdd(io.wait_flag, 0ULL) < need) __nanosleep(20);
                    __syncwarp();
                }
                e.x = r < tlen ? __ldcg(io.h_in + r) : 0;
                e.y = r < tlen ? __ldcg(io.f_in + r) : 0;
            }
            __syncwarp();   // every lane is done with the rings' previous contents
            blk = s0;
            if (s0 > 0) { const int a = ws.tl[32 + lane], b = ws.tl[64 + lane]; ws.tl[lane] = a; ws.tl[32 + lane] = b; }   // the window moves on by 32 rows
            ws.tl[64 + lane] = r + 32 < tlen ? t[r + 32] << 8 : 0;
            ws.in_ring[lane] = e;
            __syncwarp();
            if (s0 == 0 && lane == 0) sc = *reinterpret_cast<const uint2*>(prof_lane + ws.tl[32]);
        }
        if (s0 >= 32 && s0 + 33 <= tlen) {   // rows s - lane of the whole block lie in [1, tlen - 2]
#pragma unroll 4
            for (int u = 0; u < 32; u++) step(s0 + u, std::true_type{});
        } else {
            const int ue = min(32, n_steps - s0);
            for (int u = 0; u < ue; u++) step(s0 + u, std::false_type{});
        }
        if (LAST) {   // column qlen-1 of the rows lane jq_lane worked on in this block: lane u looks at step s0 + u
            __syncwarp();
            const int i = s0 + lane - jq_lane;
            if (s0 + lane < n_steps && i >= 0 && i < tlen) {
                const int hq = ws.q_ring[lane][jq];
                mqe = max(mqe, hq);
                if (i == tlen - 1) score = hq;
            }
        }
        if (!LAST) {
            // rows finished by lane 31 so far; store them 32 at a time, coalesced
            const int done = min(tlen, s0 + 32 - 31);
            while (stored + 32 <= done || (done == tlen && stored < tlen)) {
                __syncwarp();
                const int r = stored + lane;
                if (r < min(stored + 32, tlen)) { const int2 o = ws.out_ring[r & 63]; __stcg(io.h_out + r, o.x); __stcg(io.f_out + r, o.y); }
                stored += 32;
                if (PIPE) {
                    __threadfence_block();   // reader and writer are warps of one CTA; the edges go through L2 (.cg) on both sides
                    __syncwarp();
                    if (lane == 0) atomicExch(io.post_flag, io.post_base + (unsigned long long)min(stored, tlen));
                }
            }
        }
    }
    acc.gmax = gmax; acc.mqe = mqe; acc.mte = mte; acc.score = score;
}

__device__ __forceinline__ void acc_reduce_warp(AlnAcc& a) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        a.gmax = max(a.gmax, __shfl_xor_sync(0xffffffffu, a.gmax, d));
        a.mte = max(a.mte, __shfl_xor_sync(0xffffffffu, a.mte, d));
        a.mqe = max(a.mqe, __shfl_xor_sync(0xffffffffu, a.mqe, d));
        a.score = max(a.score, __shfl_xor_sync(0xffffffffu, a.score, d));
    }
}

// a whole pair on one warp, stripe after stripe (edges through the warp's global scratch): full stripes of 256
// columns, then the remainder with the narrowest lane width that covers it (64 / 128 / 256 columns), so that a
// 300-residue query costs 320 columns of work, not 512
__device__ __forceinline__ void pair_on_warp(const int* s_mat, WarpSmem& ws, const Res q, const Res t, int qlen, int tlen, int gapoe, int gape,
                                             int32_t* hb, int32_t* fb, AlnAcc& acc, int lane) {
    for (int c_base = 0; c_base < qlen; c_base += 256) {
        const int rem = qlen - c_base;
        StripeIO io;
        io.h_in = c_base > 0 ? hb : nullptr; io.f_in = c_base > 0 ? fb : nullptr;
        io.h_out = rem > 256 ? hb : nullptr; io.f_out = rem > 256 ? fb : nullptr;
        io.wait_flag = io.post_flag = nullptr; io.wait_base = io.post_base = 0;
        if (rem > 256) dp_stripe<8, false, false>(s_mat, ws, q, t, qlen, tlen, c_base, gapoe, gape, io, acc, lane);
        else if (rem <= 64) dp_stripe<2, false, true>(s_mat, ws, q, t, qlen, tlen, c_base, gapoe, gape, io, acc, lane);
        else if (rem <= 128) dp_stripe<4, false, true>(s_mat, ws, q, t, qlen, tlen, c_base, gapoe, gape, io, acc, lane);
        else dp_stripe<8, false, true>(s_mat, ws, q, t, qlen, tlen, c_base, gapoe, gape, io, acc, lane);
        __syncwarp();
    }
}

constexpr int LONG_CPL = 8;

// Persistent CTAs.  Phase A: the n_long longest pairs (order[0 .. n_long)) are taken one per CTA; the stripes of
// a pair go round-robin to the CTA's warps, which run them as a software pipeline (block-cooperative tiles):
// stripe k starts as soon as stripe k-1 has posted its first 32 rows.  Edges of neighbouring stripes live in two
// global buffers per CTA (stripe parity), handed over with shared-memory progress counters.  Phase B: every warp
// takes the remaining pairs on its own from a second counter.
__global__ void __launch_bounds__(ALN_WARPS * 32, 1)
align_kernel(const uint8_t* __restrict__ aa, int q_stride, int t_stride, const AlnPair* __restrict__ pairs,
             const int32_t* __restrict__ order, long long n_pairs, long long n_long, int gapo, int gape,
             int32_t* __restrict__ out_max, int32_t* __restrict__ out_score, int32_t* __restrict__ out_mqe, int32_t* __restrict__ out_mte,
             int32_t* __restrict__ scratch, long long scratch_stride, int32_t* __restrict__ cta_scratch, long long cta_stride,
             unsigned long long* counters) {
    __shared__ int s_mat[25 * MAT_STRIDE];
    __shared__ unsigned long long s_progress[ALN_WARPS];      // per warp: rows posted, monotone over (pair, stripe, row)
    __shared__ unsigned long long s_slot_ring[8];             // long pairs taken by this CTA, published by warp 0
    __shared__ unsigned s_fetched;                            // how many of them are published
    __shared__ unsigned s_pair_done[4];                       // slot k & 3: k + 1 once pair k's results are written
    __shared__ int s_acc[4][4];
    __shared__ unsigned s_cnt[4];
    for (int i = threadIdx.x; i < 25 * MAT_STRIDE; i += blockDim.x) {
        int r = i / MAT_STRIDE, c = i % MAT_STRIDE;
        s_mat[i] = (r < 24 && c < 24) ? (int)c_mat50[r * 24 + c] : PAD_SCORE8;
    }
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int gapoe = gapo + gape;
    extern __shared__ __align__(16) unsigned char aln_smem[];
    WarpSmem& ws = reinterpret_cast<WarpSmem*>(aln_smem)[wib];
    if (threadIdx.x < ALN_WARPS) s_progress[threadIdx.x] = 0;
    if (threadIdx.x < 4) { s_pair_done[threadIdx.x] = 0; s_cnt[threadIdx.x] = 0; s_acc[threadIdx.x][0] = 0; s_acc[threadIdx.x][1] = s_acc[threadIdx.x][2] = s_acc[threadIdx.x][3] = NEG_INF; }
    if (threadIdx.x == 0) s_fetched = 0;
    __syncthreads();

    // ---- phase A: long pairs, taken by the CTA one after the other, no CTA-wide barrier ----
    // Warp w runs stripes w, w + 16, ... of every long pair of the CTA as a software pipeline (stripe k starts as soon
    // as stripe k-1 has posted its first 32 rows), and goes on to the next pair when its own stripes are done: the fill
    // of one pair overlaps the drain of the one before.  Edges live in global buffers indexed by (pair parity, stripe
    // parity); a warp enters pair k only after pair k-2 is complete, which frees those buffers and the result slot.
    int32_t* edge = cta_scratch + (long long)blockIdx.x * 8 * cta_stride;   // [pair parity][stripe parity][h|f][row]
    unsigned long long base = 0;
    for (unsigned k = 0; n_long > 0; k++) {
        if (k >= 2) { if (lane == 0) while (atomicAdd(&s_pair_done[(k - 2) & 3], 0u) < k - 1) __nanosleep(40); __syncwarp(); }
        if (wib == 0 && lane == 0) {
            atomicExch(&s_slot_ring[k & 7], atomicAdd(counters, 1ULL));
            __threadfence_block();
            atomicExch(&s_fetched, k + 1);
        }
        if (lane == 0) while (atomicAdd(&s_fetched, 0u) < k + 1) __nanosleep(40);
        __syncwarp();
        unsigned long long slot = 0;
        if (lane == 0) slot = atomicAdd(&s_slot_ring[k & 7], 0ULL);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= (unsigned long long)n_long) break;
        const int pair = order[slot];
        const AlnPair pr = pairs[pair];
        const Res q{aa + pr.q_off, q_stride}, t{aa + pr.t_off, t_stride};
        const int qlen = pr.qlen, tlen = pr.tlen;
        const int n_stripes = (qlen + 32 * LONG_CPL - 1) / (32 * LONG_CPL);
        AlnAcc acc{0, NEG_INF, NEG_INF, NEG_INF};
        const unsigned long long per = (unsigned long long)tlen + 1ULL;
        int32_t* edge_k = edge + (long long)(k & 1) * 4 * cta_stride;
        for (int st = wib; st < n_stripes; st += ALN_WARPS) {
            StripeIO io;
            int32_t* in = edge_k + (long long)((st + 1) & 1) * 2 * cta_stride;    // written by stripe st-1
            int32_t* outb = edge_k + (long long)(st & 1) * 2 * cta_stride;
            io.h_in = st > 0 ? in : nullptr; io.f_in = st > 0 ? in + cta_stride : nullptr;
            io.h_out = st < n_stripes - 1 ? outb : nullptr; io.f_out = st < n_stripes - 1 ? outb + cta_stride : nullptr;
            io.wait_flag = &s_progress[(wib + ALN_WARPS - 1) % ALN_WARPS]; io.wait_base = base + (unsigned long long)st * per;
            io.post_flag = &s_progress[wib]; io.post_base = base + (unsigned long long)(st + 1) * per;
            if (st == n_stripes - 1) dp_stripe<LONG_CPL, true, true>(s_mat, ws, q, t, qlen, tlen, st * 32 * LONG_CPL, gapoe, gape, io, acc, lane);
            else dp_stripe<LONG_CPL, true, false>(s_mat, ws, q, t, qlen, tlen, st * 32 * LONG_CPL, gapoe, gape, io, acc, lane);
            __syncwarp();
        }
        if (wib < n_stripes) {   // this warp had stripes of the pair
            acc_reduce_warp(acc);
            if (lane == 0) {
                int* a = s_acc[k & 3];
                atomicMax(&a[0], acc.gmax); atomicMax(&a[1], acc.score); atomicMax(&a[2], acc.mqe); atomicMax(&a[3], acc.mte);
                __threadfence_block();
                const unsigned old = atomicAdd(&s_cnt[k & 3], 1u);
                if (old + 1 == (unsigned)min(n_stripes, ALN_WARPS)) {   // last warp of the pair: publish and free the slot
                    __threadfence_block();
                    out_max[pair] = atomicExch(&a[0], 0); out_score[pair] = atomicExch(&a[1], NEG_INF);
                    out_mqe[pair] = atomicExch(&a[2], NEG_INF); out_mte[pair] = atomicExch(&a[3], NEG_INF);
                    atomicExch(&s_cnt[k & 3], 0u);
                    __threadfence_block();
                    atomicExch(&s_pair_done[k & 3], k + 1);
                }
            }
        }
        base += (unsigned long long)(n_stripes + 1) * per;
        __syncwarp();
    }

    // ---- phase B: one warp per pair ----
    const long long warp_id = (long long)blockIdx.x * ALN_WARPS + wib;
    int32_t* hb = scratch + warp_id * 2 * scratch_stride;   // right edge of the previous stripe: H(i, c_base-1)
    int32_t* fb = hb + scratch_stride;                      //                                    F(i, c_base)
    while (true) {
        unsigned long long slot = 0;
        if (lane == 0) slot = atomicAdd(counters + 1, 1ULL);
        slot = (unsigned long long)n_long + __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= (unsigned long long)n_pairs) break;
        const int pair = order[slot];
        const AlnPair pr = pairs[pair];
        const Res q{aa + pr.q_off, q_stride}, t{aa + pr.t_off, t_stride};
        const int qlen = pr.qlen, tlen = pr.tlen;
        if (qlen <= 0 || tlen <= 0) {
            if (lane == 0) { out_max[pair] = 0; out_score[pair] = NEG_INF; out_mqe[pair] = NEG_INF; out_mte[pair] = NEG_INF; }
            continue;
        }
        AlnAcc acc{0, NEG_INF, NEG_INF, NEG_INF};
        pair_on_warp(s_mat, ws, q, t, qlen, tlen, gapoe, gape, hb, fb, acc, lane);
        acc_reduce_warp(acc);
        if (lane == 0) { out_max[pair] = acc.gmax; out_score[pair] = acc.score; out_mqe[pair] = acc.mqe; out_mte[pair] = acc.mte; }
    }
}

// Issue-rate microbenchmark behind the alignment roofline: every thread runs 8 independent
// dependent chains of one integer instruction; ops/s = threads x ops / time.
template <int KIND>
__global__ void __launch_bounds__(256) int_alu_peak_kernel(int iters, int seed, int* sink) {
    int a[8];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = seed + threadIdx.x * 8 + k;
    const int b = seed * 3 + 1, c = seed - 7;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (KIND == 0) a[k] = a[k] + a[(k + 1) & 7] + b;     // IADD3 (two live chains per add: not foldable)
                else if (KIND == 1) a[k] = __vimax3_s32(a[k] - 1, b + u, c);   // IADD + VIMNMX3
                else if (KIND == 2) a[k] = __viaddmax_s32(a[k], -2, b + u);           // VIADDMNMX
                else a[k] = __dp4a(a[(k + 1) & 7], 0x00000100, a[k]);   // IDP.4A
            }
        }
    }
    int r = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) r ^= a[k];
    if (r == 0x7fffffff) *sink = r;
}

}  // namespace

// ops[0] IADD3/s, ops[1] (IADD + VIMNMX3 pairs counted as 2 ops)/s, ops[2] VIADDMNMX/s (counted as 2 ops: add + max),
// ops[3] IDP.4A instructions/s
int int_alu_peak_device(cudaStream_t stream, double* ops, std::string& err) {
    cudaDeviceProp prop;
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) { err = "cudaGetDeviceProperties"; return ORFX_E_CUDA; }
    int* sink = nullptr;
    if (cudaMalloc(&sink, 4) != cudaSuccess) { err = "cudaMalloc"; return ORFX_E_CUDA; }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = prop.multiProcessorCount * 8, iters = 4096;
    const double per_thread = (double)iters * 64;
    for (int kind = 0; kind < 4; kind++) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; rep++) {
            cudaEventRecord(e0, stream);
            if (kind == 0) int_alu_peak_kernel<0><<<grid, 256, 0, stream>>>(iters, rep + 1, sink);
            else if (kind == 1) int_alu_peak_kernel<1><<<grid, 256, 0, stream>>>(iters, rep + 1, sink);
            else if (kind == 2) int_alu_peak_kernel<2><<<grid, 256, 0, stream>>>(iters, rep + 1, sink);
            else int_alu_peak_kernel<3><<<grid, 256, 0, stream>>>(iters, rep + 1, sink);
            cudaEventRecord(e1, stream);
            if (cudaStreamSynchronize(stream) != cudaSuccess) { err = "int_alu_peak kernel failed"; cudaFree(sink); return ORFX_E_CUDA; }
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        const double insts = per_thread * (double)grid * 256.0;
        ops[kind] = insts * (kind == 0 || kind == 3 ? 1.0 : 2.0) / (best * 1e-3);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(sink);
    return ORFX_OK;
}

// ---- ordering and scheduling of the pairs, on the device ------------------------------------------------
// Long pairs (>= 1024 x 256, >= 4 M cells: whole CTAs, phase A) first, then the rest; both by descending cost in
// quarter-octave classes of qlen * tlen (a counting sort: the long tail starts early, no host-side std::sort).
namespace {
constexpr int ALN_NCLS = 128;                 // quarter octaves of the cell count, 2^0 .. 2^32
constexpr int ALN_NBINS = 2 * ALN_NCLS;       // [long, descending cost][short, descending cost]
struct AlnMeta { unsigned long long n_long, n_active, max_tlen, max_tlen_long; double cells; };

__device__ __forceinline__ int cost_class(long long cells) {
    if (cells <= 0) return 0;
    const int lg = 63 - __clzll(cells);
    const int frac = lg >= 2 ? (int)((cells >> (lg - 2)) & 3) : 0;
    return min(ALN_NCLS - 1, lg * 4 + frac);
}
__global__ void __launch_bounds__(256) aln_pairs_from_index_kernel(const int64_t* __restrict__ off, const int32_t* __restrict__ pq, const int32_t* __restrict__ pt,
                                                                  long long n, AlnPair* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int a = pq[i], b = pt[i];
    AlnPair p;
    p.q_off = off[a]; p.t_off = off[b]; p.qlen = (int)(off[a + 1] - off[a]); p.tlen = (int)(off[b + 1] - off[b]);
    out[i] = p;
}
__global__ void __launch_bounds__(256) aln_classify_kernel(const AlnPair* __restrict__ pairs, long long n, uint16_t* __restrict__ cls, unsigned* __restrict__ hist,
                                                          AlnMeta* meta) {
    __shared__ unsigned s_hist[ALN_NBINS];
    __shared__ double s_cells[8];
    for (int i = threadIdx.x; i < ALN_NBINS; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double cells = 0;
    if (i < n) {
        const AlnPair p = pairs[i];
        int bin = 0xFFFF;
        if (p.qlen >= 0) {
            const long long c = (long long)p.qlen * (long long)max(p.tlen, 0);
            cells = (double)c;
            // a whole CTA (16 warps, one stripe of 256 columns each at a time) pays off from ~12 stripes on, or when the pair is
            // so large that a single warp would be the tail of the launch
            const bool is_long = p.qlen >= 1024 && p.tlen >= 256 && (p.qlen >= 3072 ? c >= 4194304LL : c >= 33554432LL);
            bin = (is_long ? 0 : ALN_NCLS) + (ALN_NCLS - 1 - cost_class(c));
            atomicAdd(&s_hist[bin], 1u);
            if (is_long) atomicMax(&meta->max_tlen_long, (unsigned long long)p.tlen);
            else if (p.qlen > 256) atomicMax(&meta->max_tlen, (unsigned long long)max(p.tlen, 0));   // only a multi-stripe pair parks edges in the warp's scratch
        }
        cls[i] = (uint16_t)bin;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, d);
    if ((threadIdx.x & 31) == 0) s_cells[threadIdx.x >> 5] = cells;
    __syncthreads();
    if (threadIdx.x == 0) { double c = 0; for (int w = 0; w < 8; w++) c += s_cells[w]; if (c != 0) atomicAdd(&meta->cells, c); }
    for (int b = threadIdx.x; b < ALN_NBINS; b += blockDim.x) if (s_hist[b]) atomicAdd(&hist[b], s_hist[b]);
}
// exclusive prefix of the bins (in place: hist becomes the scatter cursors); totals into meta
__global__ void aln_scan_kernel(unsigned* hist, AlnMeta* meta) {
    if (threadIdx.x != 0) return;
    unsigned long long at = 0, n_long = 0;
    for (int b = 0; b < ALN_NBINS; b++) { const unsigned v = hist[b]; hist[b] = (unsigned)at; at += v; if (b == ALN_NCLS - 1) n_long = at; }
    meta->n_long = n_long; meta->n_active = at;
}
__global__ void __launch_bounds__(256) aln_scatter_kernel(const uint16_t* __restrict__ cls, long long n, unsigned* cursor, int32_t* __restrict__ order) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int bin = cls[i];
    if (bin == 0xFFFF) return;
    order[atomicAdd(&cursor[bin], 1u)] = (int32_t)i;
}
}  // namespace

// Aligns the pairs described by d_pairs (device memory, n entries) over the residues at d_aa; results into the four device
// arrays d_out[0..3] (max, score, mqe, mte), indexed like d_pairs.  Buffers come from `w` (grow-only).  One host round trip
// (five numbers that size the stripe-edge scratch).  Returns the number of kernels launched in *n_launches.
int align_pairs_device(AlnWork& w, cudaStream_t stream, int n_sm, const uint8_t* d_aa, int q_stride, int t_stride, const AlnPair* d_pairs, long long n,
                       int gapo, int gape, int32_t* const d_out[4], double* cells, float* ms, int* n_launches, std::string& err) {
    auto ck = [&](cudaError_t e, const char* what) { if (e != cudaSuccess) { err = std::string(what) + ": " + cudaGetErrorString(e); return false; } return true; };
    if (n_launches) *n_launches = 0;
    if (cells) *cells = 0;
    if (ms) *ms = 0;
    if (n <= 0) return ORFX_OK;
    if (gapo < 0 || gape < 0 || gapo + gape > 64) { err = "gap costs outside this build's range (0 <= gapo + gape <= 64)"; return ORFX_E_UNSUPPORTED; }
    if (n > 0x7FFFFFFFLL) { err = "more than 2^31 pairs in one call"; return ORFX_E_INVALID; }
    if (!ck(w.cls.ensure((size_t)n * 2 + 16), "cudaMalloc cls") || !ck(w.order.ensure((size_t)n * 4 + 16), "cudaMalloc order") ||
        !ck(w.hist.ensure(ALN_NBINS * 4), "cudaMalloc hist") || !ck(w.meta.ensure(sizeof(AlnMeta)), "cudaMalloc meta") ||
        !ck(w.counters.ensure(16), "cudaMalloc counters") || !ck(w.h_meta.ensure(sizeof(AlnMeta)), "cudaMallocHost meta"))
        return ORFX_E_CUDA;
    cudaMemsetAsync(w.hist.p, 0, ALN_NBINS * 4, stream);
    cudaMemsetAsync(w.meta.p, 0, sizeof(AlnMeta), stream);
    cudaMemsetAsync(w.counters.p, 0, 16, stream);
    const unsigned nblk = (unsigned)((n + 255) / 256);
    aln_classify_kernel<<<nblk, 256, 0, stream>>>(d_pairs, n, w.cls.as<uint16_t>(), w.hist.as<unsigned>(), w.meta.as<AlnMeta>());
    aln_scan_kernel<<<1, 32, 0, stream>>>(w.hist.as<unsigned>(), w.meta.as<AlnMeta>());
    cudaMemcpyAsync(w.h_meta.p, w.meta.p, sizeof(AlnMeta), cudaMemcpyDeviceToHost, stream);
    aln_scatter_kernel<<<nblk, 256, 0, stream>>>(w.cls.as<uint16_t>(), n, w.hist.as<unsigned>(), w.order.as<int32_t>());
    if (!ck(cudaStreamSynchronize(stream), "alignment scheduling")) return ORFX_E_CUDA;
    const AlnMeta m = *w.h_meta.as<AlnMeta>();
    if (n_launches) *n_launches = 3;
    if (cells) *cells = m.cells;
    if (m.n_active == 0) return ORFX_OK;
    const long long n_long = (long long)m.n_long, n_active = (long long)m.n_active;
    int grid = n_sm;
    const long long max_grid = std::max<long long>(n_long, (n_active - n_long + ALN_WARPS - 1) / ALN_WARPS);
    if (grid > max_grid) grid = (int)std::max<long long>(1, max_grid);
    const long long n_warps = (long long)grid * ALN_WARPS;
    const long long stride = ((long long)std::max<unsigned long long>(m.max_tlen, 1) + 31) / 32 * 32;
    const long long cta_stride = ((long long)std::max<unsigned long long>(m.max_tlen_long, 1) + 31) / 32 * 32;
    if (!ck(w.scratch.ensure((size_t)(n_warps * 2 * stride) * 4 + 16), "cudaMalloc stripe scratch") ||
        !ck(w.cta.ensure((size_t)((long long)grid * 8 * cta_stride) * 4 + 16), "cudaMalloc CTA scratch"))
        return ORFX_E_CUDA;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ms) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, stream); }
    cudaFuncSetAttribute(align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ALN_WARPS * sizeof(WarpSmem)));
    align_kernel<<<grid, ALN_WARPS * 32, ALN_WARPS * sizeof(WarpSmem), stream>>>(d_aa, q_stride, t_stride, d_pairs, w.order.as<int32_t>(), n_active, n_long, gapo, gape,
                                                                               d_out[0], d_out[1], d_out[2], d_out[3], w.scratch.as<int32_t>(), stride,
                                                                               w.cta.as<int32_t>(), cta_stride, w.counters.as<unsigned long long>());
    if (n_launches) *n_launches = 4;
    if (ms) {
        cudaEventRecord(e1, stream);
        const bool ok = ck(cudaStreamSynchronize(stream), "alignment kernel");
        if (ok) cudaEventElapsedTime(ms, e0, e1);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (!ok) return ORFX_E_CUDA;
    }
    if (!ck(cudaGetLastError(), "alignment kernel launch")) return ORFX_E_CUDA;
    return ORFX_OK;
}

// orfx_align_batch: caller-supplied proteins in host memory (BASELINE.json configs[4])
int align_batch_device(AlnWork& w, cudaStream_t stream, int n_sm, const uint8_t* aa, const int64_t* off, int64_t n_seq, const int32_t* pair_q,
                       const int32_t* pair_t, int64_t n_pairs, int gapo, int gape, orfx_aln_out* out, double* cells, float* ms, std::string& err) {
    auto ck = [&](cudaError_t e, const char* what) { if (e != cudaSuccess) { err = std::string(what) + ": " + cudaGetErrorString(e); return false; } return true; };
    if (n_pairs == 0) { if (cells) *cells = 0; if (ms) *ms = 0; return ORFX_OK; }
    for (int64_t p = 0; p < n_pairs; p++)
        if (pair_q[p] < 0 || pair_q[p] >= n_seq || pair_t[p] < 0 || pair_t[p] >= n_seq) { err = "pair index out of range"; return ORFX_E_INVALID; }
    const int64_t total = off[n_seq];
    if (!ck(w.aa.ensure((size_t)total + 16), "cudaMalloc aa") || !ck(w.off.ensure((size_t)(n_seq + 1) * 8), "cudaMalloc off") ||
        !ck(w.pq.ensure((size_t)n_pairs * 4), "cudaMalloc pair_q") || !ck(w.pt.ensure((size_t)n_pairs * 4), "cudaMalloc pair_t") ||
        !ck(w.pairs.ensure((size_t)n_pairs * sizeof(AlnPair)), "cudaMalloc pairs") || !ck(w.out.ensure((size_t)n_pairs * 16), "cudaMalloc out"))
        return ORFX_E_CUDA;
    if (!ck(cudaMemcpyAsync(w.aa.p, aa, (size_t)total, cudaMemcpyHostToDevice, stream), "H2D aa")) return ORFX_E_CUDA;
    cudaMemcpyAsync(w.off.p, off, (size_t)(n_seq + 1) * 8, cudaMemcpyHostToDevice, stream);
    cudaMemcpyAsync(w.pq.p, pair_q, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, stream);
    cudaMemcpyAsync(w.pt.p, pair_t, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, stream);
    aln_pairs_from_index_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, stream>>>(w.off.as<int64_t>(), w.pq.as<int32_t>(), w.pt.as<int32_t>(), n_pairs, w.pairs.as<AlnPair>());
    int32_t* d_out[4] = {w.out.as<int32_t>(), w.out.as<int32_t>() + n_pairs, w.out.as<int32_t>() + 2 * n_pairs, w.out.as<int32_t>() + 3 * n_pairs};
    float kms = 0;
    const int rc = align_pairs_device(w, stream, n_sm, w.aa.as<uint8_t>(), 1, 1, w.pairs.as<AlnPair>(), n_pairs, gapo, gape, d_out, cells, &kms, nullptr, err);
    if (rc) return rc;
    if (ms) *ms = kms;
    cudaMemcpyAsync(out->max, d_out[0], (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
    cudaMemcpyAsync(out->score, d_out[1], (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
    cudaMemcpyAsync(out->mqe, d_out[2], (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
    cudaMemcpyAsync(out->mte, d_out[3], (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
    if (!ck(cudaStreamSynchronize(stream), "alignment results")) return ORFX_E_CUDA;
    return ORFX_OK;
}

}  // namespace orfx

// ---- orfx_align_pairs_static: the alignment with no allocation inside the call --------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Everything orfx_align_batch keeps in
// context-owned, grow-on-demand buffers (pair descriptors, cost classes, the longest-first order, the histogram, the
// scheduling counters, the per-warp stripe edges) lives in ONE file-scope device array of a fixed size: every GPU that
// loads the library holds these bytes until the process exits.  All calls share the array and it keeps the last call's
// contents: the histogram, totals, counters and the error flag are cleared at the start of every call, everything else is
// written before it is read.  Two calls that overlap on the GPU (different contexts / streams of one device) race.
namespace {
constexpr size_t st_up(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr long long ST_MAX_PAIRS = ORFX_STATIC_MAX_PAIRS;   // pairs per call
constexpr long long ST_MAX_LEN = ORFX_STATIC_MAX_LEN;       // residues per protein: no pair reaches the block-cooperative class
constexpr int ST_MAX_CTAS = 148;                            // one CTA per SM of a B200
constexpr size_t ST_PAIRS = 0;
constexpr size_t ST_CLS = ST_PAIRS + st_up((size_t)ST_MAX_PAIRS * 24);
constexpr size_t ST_ORDER = ST_CLS + st_up((size_t)ST_MAX_PAIRS * 2 + 16);
constexpr size_t ST_HIST = ST_ORDER + st_up((size_t)ST_MAX_PAIRS * 4 + 16);
constexpr size_t ST_META = ST_HIST + st_up(2 * 128 * 4);    // AlnMeta (40 bytes), the error flag at +64
constexpr size_t ST_COUNTERS = ST_META + 256;
constexpr size_t ST_STRIPES = ST_COUNTERS + 256;            // [CTA][warp][H | F][ST_MAX_LEN] int32
constexpr size_t ST_BYTES = ST_STRIPES + (size_t)ST_MAX_CTAS * ORFX_ALN_WARPS * 2 * (size_t)ST_MAX_LEN * 4;
static_assert(ST_BYTES <= ((size_t)ORFX_STATIC_SCRATCH_MIB << 20), "the static scratch outgrew the size include/orfx.h states");
}  // namespace
static __device__ __align__(256) unsigned char g_static_scratch[ST_BYTES];

namespace orfx {
namespace {
static_assert(sizeof(AlnPair) == 24 && ALN_NBINS == 2 * 128, "static scratch layout");
// In host code `g_static_scratch` is a host-side stand-in, not a device address: ask the runtime once per device (with that
// device current) and hand the answer to the kernels.
std::mutex g_static_mu;
unsigned char* g_static_ptr[64] = {};
int static_scratch(int device, unsigned char** out, std::string& err) {
    if (device < 0 || device >= 64) { err = "device ordinal outside the static scratch table"; return ORFX_E_UNSUPPORTED; }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) { err = std::string("cudaGetSymbolAddress(static scratch): ") + cudaGetErrorString(e); return ORFX_E_CUDA; }
        g_static_ptr[device] = (unsigned char*)p;
    }
    *out = g_static_ptr[device];
    return ORFX_OK;
}
// pair descriptors from the caller's index arrays (all in device memory), with the checks orfx_align_batch makes on the
// host: flag bit 0 = index out of range, bit 1 = a protein longer than the static limit, bit 2 = offsets not ascending
__global__ void __launch_bounds__(256) aln_pairs_static_kernel(const int64_t* __restrict__ off, long long n_seq, const int32_t* __restrict__ pq,
                                                              const int32_t* __restrict__ pt, long long n, int max_len, AlnPair* __restrict__ out, unsigned* flag) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int a = pq[i], b = pt[i];
    AlnPair p{0, 0, -1, 0};
    unsigned bad = 0;
    if (a < 0 || a >= n_seq || b < 0 || b >= n_seq) bad = 1;
    else {
        const long long qa = off[a], qb = off[a + 1], ta = off[b], tb = off[b + 1];
        if (qa < 0 || ta < 0 || qb < qa || tb < ta) bad = 4;
        else if (qb - qa > max_len || tb - ta > max_len) bad = 2;
        else { p.q_off = qa; p.t_off = ta; p.qlen = (int)(qb - qa); p.tlen = (int)(tb - ta); }
    }
    if (bad) atomicOr(flag, bad);
    out[i] = p;
}
}  // namespace

// d_aa / d_off / d_pair_q / d_pair_t and the four output arrays are DEVICE memory of the context's device.  Synchronises the
// stream twice (after the scheduling kernels: the totals come back through a plain cudaMemcpy into a stack variable; and at
// the end).  No cudaMalloc / cudaFree / cudaMallocHost / event creation in here.
int align_pairs_static_device(int device, cudaStream_t stream, int n_sm, const uint8_t* d_aa, const int64_t* d_off, int64_t n_seq, const int32_t* d_pq,
                              const int32_t* d_pt, int64_t n_pairs, int gapo, int gape, const orfx_aln_out* d_out, std::string& err) {
    auto ck = [&](cudaError_t e, const char* what) { if (e != cudaSuccess) { err = std::string(what) + ": " + cudaGetErrorString(e); return false; } return true; };
    if (n_pairs > ST_MAX_PAIRS) {
        err = "orfx_align_pairs_static: " + std::to_string((long long)n_pairs) + " pairs in one call, the static scratch holds " + std::to_string(ST_MAX_PAIRS) +
              " (ORFX_STATIC_MAX_PAIRS); split the call or use orfx_align_batch";
        return ORFX_E_UNSUPPORTED;
    }
    if (gapo < 0 || gape < 0 || gapo + gape > 64) { err = "gap costs outside this build's range (0 <= gapo + gape <= 64)"; return ORFX_E_UNSUPPORTED; }
    if (n_pairs == 0) return ORFX_OK;
    unsigned char* base = nullptr;
    const int rc = static_scratch(device, &base, err);
    if (rc) return rc;
    AlnPair* pairs = reinterpret_cast<AlnPair*>(base + ST_PAIRS);
    uint16_t* cls = reinterpret_cast<uint16_t*>(base + ST_CLS);
    int32_t* order = reinterpret_cast<int32_t*>(base + ST_ORDER);
    unsigned* hist = reinterpret_cast<unsigned*>(base + ST_HIST);
    AlnMeta* meta = reinterpret_cast<AlnMeta*>(base + ST_META);
    unsigned* flag = reinterpret_cast<unsigned*>(base + ST_META + 64);
    unsigned long long* counters = reinterpret_cast<unsigned long long*>(base + ST_COUNTERS);
    int32_t* stripes = reinterpret_cast<int32_t*>(base + ST_STRIPES);
    // what must start at zero: histogram, totals + error flag, scheduling counters (one contiguous range)
    if (!ck(cudaMemsetAsync(base + ST_HIST, 0, ST_STRIPES - ST_HIST, stream), "clearing the static scratch")) return ORFX_E_CUDA;
    const unsigned nblk = (unsigned)((n_pairs + 255) / 256);
    aln_pairs_static_kernel<<<nblk, 256, 0, stream>>>(d_off, n_seq, d_pq, d_pt, n_pairs, (int)ST_MAX_LEN, pairs, flag);
    aln_classify_kernel<<<nblk, 256, 0, stream>>>(pairs, n_pairs, cls, hist, meta);
    aln_scan_kernel<<<1, 32, 0, stream>>>(hist, meta);
    aln_scatter_kernel<<<nblk, 256, 0, stream>>>(cls, n_pairs, hist, order);
    if (!ck(cudaStreamSynchronize(stream), "alignment scheduling")) return ORFX_E_CUDA;
    struct { AlnMeta m; unsigned char pad[64 - sizeof(AlnMeta)]; unsigned flag; } h;
    static_assert(sizeof(AlnMeta) <= 64, "flag offset");
    if (!ck(cudaMemcpy(&h, meta, 68, cudaMemcpyDeviceToHost), "reading the alignment totals")) return ORFX_E_CUDA;
    if (h.flag) {
        err = std::string("orfx_align_pairs_static: ") + ((h.flag & 1) ? "pair index out of range" : (h.flag & 4) ? "sequence offsets are not ascending"
              : "a protein is longer than " + std::to_string(ST_MAX_LEN) + " residues (ORFX_STATIC_MAX_LEN); use orfx_align_batch");
        return (h.flag & 5) ? ORFX_E_INVALID : ORFX_E_UNSUPPORTED;
    }
    if (h.m.n_active == 0) return ORFX_OK;
    if (h.m.n_long != 0 || h.m.max_tlen > (unsigned long long)ST_MAX_LEN) { err = "orfx_align_pairs_static: internal: a pair outside the static limits was scheduled"; return ORFX_E_INVALID; }
    const long long n_active = (long long)h.m.n_active;
    const int grid = (int)std::max<long long>(1, std::min<long long>(std::min(n_sm, ST_MAX_CTAS), (n_active + ALN_WARPS - 1) / ALN_WARPS));
    cudaFuncSetAttribute(align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ALN_WARPS * sizeof(WarpSmem)));
    align_kernel<<<grid, ALN_WARPS * 32, ALN_WARPS * sizeof(WarpSmem), stream>>>(d_aa, 1, 1, pairs, order, n_active, 0, gapo, gape, d_out->max, d_out->score, d_out->mqe,
                                                                               d_out->mte, stripes, ST_MAX_LEN, stripes, 32, counters);
    if (!ck(cudaGetLastError(), "alignment kernel launch")) return ORFX_E_CUDA;
    if (!ck(cudaStreamSynchronize(stream), "alignment kernel")) return ORFX_E_CUDA;
    return ORFX_OK;
}

}  // namespace orfx
