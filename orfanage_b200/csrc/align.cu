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
#include <algorithm>
#include <numeric>
#include <string>
#include <vector>

namespace orfx {

namespace {

// mat50, the reference's protein substitution matrix (src/common.hpp:239-265), order ARNDCQEGHILKMFPSTWYVBZX*
__constant__ int8_t c_mat50[24 * 24] = {
     5,-2,-1,-2,-1,-1,-1, 0,-2,-1,-2,-1,-1,-3,-1, 1, 0,-3,-2, 0,-2,-1,-1,-5,
    -2, 7,-1,-2,-4, 1, 0,-3, 0,-4,-3, 3,-2,-3,-3,-1,-1,-3,-1,-3,-1, 0,-1,-5,
    -1,-1, 7, 2,-2, 0, 0, 0, 1,-3,-4, 0,-2,-4,-2, 1, 0,-4,-2,-3, 5, 0,-1,-5,
    -2,-2, 2, 8,-4, 0, 2,-1,-1,-4,-4,-1,-4,-5,-1, 0,-1,-5,-3,-4, 6, 1,-1,-5,
    -1,-4,-2,-4,13,-3,-3,-3,-3,-2,-2,-3,-2,-2,-4,-1,-1,-5,-3,-1,-3,-3,-1,-5,
    -1, 1, 0, 0,-3, 7, 2,-2, 1,-3,-2, 2, 0,-4,-1, 0,-1,-1,-1,-3, 0, 4,-1,-5,
    -1, 0, 0, 2,-3, 2, 6,-3, 0,-4,-3, 1,-2,-3,-1,-1,-1,-3,-2,-3, 1, 5,-1,-5,
     0,-3, 0,-1,-3,-2,-3, 8,-2,-4,-4,-2,-3,-4,-2, 0,-2,-3,-3,-4,-1,-2,-1,-5,
    -2, 0, 1,-1,-3, 1, 0,-2,10,-4,-3, 0,-1,-1,-2,-1,-2,-3, 2,-4, 0, 0,-1,-5,
    -1,-4,-3,-4,-2,-3,-4,-4,-4, 5, 2,-3, 2, 0,-3,-3,-1,-3,-1, 4,-4,-3,-1,-5,
    -2,-3,-4,-4,-2,-2,-3,-4,-3, 2, 5,-3, 3, 1,-4,-3,-1,-2,-1, 1,-4,-3,-1,-5,
    -1, 3, 0,-1,-3, 2, 1,-2, 0,-3,-3, 6,-2,-4,-1, 0,-1,-3,-2,-3, 0, 1,-1,-5,
    -1,-2,-2,-4,-2, 0,-2,-3,-1, 2, 3,-2, 7, 0,-3,-2,-1,-1, 0, 1,-3,-1,-1,-5,
    -3,-3,-4,-5,-2,-4,-3,-4,-1, 0, 1,-4, 0, 8,-4,-3,-2, 1, 4,-1,-4,-4,-1,-5,
    -1,-3,-2,-1,-4,-1,-1,-2,-2,-3,-4,-1,-3,-4,10,-1,-1,-4,-3,-3,-2,-1,-1,-5,
     1,-1, 1, 0,-1, 0,-1, 0,-1,-3,-3, 0,-2,-3,-1, 5, 2,-4,-2,-2, 0, 0,-1,-5,
     0,-1, 0,-1,-1,-1,-1,-2,-2,-1,-1,-1,-1,-2,-1, 2, 5,-3,-2, 0, 0,-1,-1,-5,
    -3,-3,-4,-5,-5,-1,-3,-3,-3,-3,-2,-3,-1, 1,-4,-4,-3,15, 2,-3,-5,-2,-1,-5,
    -2,-1,-2,-3,-3,-1,-2,-3, 2,-1,-1,-2, 0, 4,-3,-2,-2, 2, 8,-1,-3,-2,-1,-5,
     0,-3,-3,-4,-1,-3,-3,-4,-4, 4, 1,-3, 1,-1,-3,-2, 0,-3,-1, 5,-3,-3,-1,-5,
    -2,-1, 5, 6,-3, 0, 1,-1, 0,-4,-4, 0,-3,-4,-2, 0, 0,-5,-3,-3, 6, 1,-1,-5,
    -1, 0, 0, 1,-3, 4, 5,-2, 0,-3,-3, 1,-1,-4,-1, 0,-1,-2,-2,-3, 1, 5,-1,-5,
    -1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-1,-5,
    -5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5,-5, 1};

constexpr int NEG_INF = -0x40000000;
constexpr int PAD_CODE = 24;        // virtual query column right of the query's end
constexpr int PAD_SCORE = -4096;    // never wins: padded cells are dominated by H(i,qlen-1)
constexpr int MAT_STRIDE = 32;
constexpr int ALN_WARPS = 4;

template <int CPL>
__global__ void __launch_bounds__(ALN_WARPS * 32)
align_kernel(const uint8_t* __restrict__ aa, const int64_t* __restrict__ off, const int32_t* __restrict__ pair_q,
             const int32_t* __restrict__ pair_t, const int32_t* __restrict__ order, long long n_pairs, int gapo, int gape,
             int32_t* __restrict__ out_max, int32_t* __restrict__ out_score, int32_t* __restrict__ out_mqe, int32_t* __restrict__ out_mte,
             int32_t* __restrict__ scratch, long long scratch_stride, unsigned long long* counter) {
    __shared__ int s_mat[25 * MAT_STRIDE];
    for (int i = threadIdx.x; i < 25 * MAT_STRIDE; i += blockDim.x) {
        int r = i / MAT_STRIDE, c = i % MAT_STRIDE;
        s_mat[i] = (r < 24 && c < 24) ? (int)c_mat50[r * 24 + c] : PAD_SCORE;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long warp_id = (long long)blockIdx.x * ALN_WARPS + (threadIdx.x >> 5);
    int32_t* hb = scratch + warp_id * 2 * scratch_stride;   // right edge of the previous stripe: H(i, c_base-1)
    int32_t* fb = hb + scratch_stride;                      //                                    F(i, c_base)
    const int gapoe = gapo + gape;

    while (true) {
        unsigned long long slot = 0;
        if (lane == 0) slot = atomicAdd(counter, 1ULL);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= (unsigned long long)n_pairs) break;
        const int pair = order[slot];
        const int qi = pair_q[pair], ti = pair_t[pair];
        const uint8_t* q = aa + off[qi];
        const uint8_t* t = aa + off[ti];
        const int qlen = (int)(off[qi + 1] - off[qi]), tlen = (int)(off[ti + 1] - off[ti]);
        if (qlen <= 0 || tlen <= 0) {
            if (lane == 0) { out_max[pair] = 0; out_score[pair] = NEG_INF; out_mqe[pair] = NEG_INF; out_mte[pair] = NEG_INF; }
            continue;
        }
        int gmax = 0, mqe = NEG_INF, mte = NEG_INF, score = NEG_INF;
        const int n_stripes = (qlen + 32 * CPL - 1) / (32 * CPL);
        for (int st = 0; st < n_stripes; st++) {
            const int c_base = st * 32 * CPL;
            const int c0 = c_base + lane * CPL;
            const bool last_stripe = st == n_stripes - 1;
            const bool have_next = !last_stripe;
            int qoff[CPL];   // byte offset of the query residue's column in s_mat
            int hp[CPL], ee[CPL];
#pragma unroll
            for (int k = 0; k < CPL; k++) {
                const int c = c0 + k;
                qoff[k] = c < qlen ? (int)q[c] : PAD_CODE;
                hp[k] = -(gapoe + gape * c);                 // H(-1, c)
                ee[k] = -(gapoe + gapoe + gape * c);         // E(0, c)
            }
            // lane-local column of qlen-1 (last stripe only)
            const int jq_lane = (qlen - 1 - c_base) / CPL, jq = (qlen - 1 - c_base) % CPL;
            int diag = c0 == 0 ? 0 : -(gapoe + gape * (c0 - 1));   // H(i-1, c0-1), starts at row -1
            int h_out = 0, f_out = 0, t_out = 0;                   // what this lane hands to lane+1
            int tcur = 0;                                          // 32 target residues, one per lane
            int hb_in = 0, fb_in = 0, hb_keep = 0, fb_keep = 0;
            const int n_steps = tlen + 31;
            for (int s = 0; s < n_steps; s++) {
                if ((s & 31) == 0) {
                    const int r = s + lane;
                    tcur = r < tlen ? (int)t[r] : 0;
                    if (st > 0) { hb_in = r < tlen ? hb[r] : 0; fb_in = r < tlen ? fb[r] : 0; }
                }
                // values arriving from the left neighbour (computed in the previous step)
                int h_left = __shfl_up_sync(0xffffffffu, h_out, 1);
                int f_in = __shfl_up_sync(0xffffffffu, f_out, 1);
                int t_res = __shfl_up_sync(0xffffffffu, t_out, 1);
                const int t0 = __shfl_sync(0xffffffffu, tcur, s & 31);
                const int hb0 = __shfl_sync(0xffffffffu, hb_in, s & 31);
                const int fb0 = __shfl_sync(0xffffffffu, fb_in, s & 31);
                if (lane == 0) {
                    t_res = t0;
                    if (st == 0) { h_left = -(gapoe + gape * s); f_in = -(gapoe + gapoe + gape * s); }
                    else { h_left = hb0; f_in = fb0; }
                }
                const int i = s - lane;
                const bool active = i >= 0 && i < tlen;
                if (active) {
                    const int* mrow = s_mat + t_res * MAT_STRIDE;
                    int f = f_in, d = diag, h = 0, rowmax = NEG_INF;
#pragma unroll
                    for (int k = 0; k < CPL; k++) {
                        h = d + mrow[qoff[k]];
                        h = __vimax3_s32(h, ee[k], f);
                        d = hp[k];
                        hp[k] = h;
                        rowmax = max(rowmax, h);
                        const int tt = h - gapoe;
                        ee[k] = __viaddmax_s32(ee[k], -gape, tt);
                        f = __viaddmax_s32(f, -gape, tt);
                    }
                    gmax = max(gmax, rowmax);
                    diag = h_left;
                    h_out = h; f_out = f; t_out = t_res;
                    if (i == tlen - 1) mte = max(mte, rowmax);
                    if (last_stripe && lane == jq_lane) {
                        int hq = hp[0];
#pragma unroll
                        for (int k = 1; k < CPL; k++) hq = jq == k ? hp[k] : hq;
                        mqe = max(mqe, hq);
                        if (i == tlen - 1) score = hq;
                    }
                }
                if (have_next) {
                    // lane 31 finished row s-31: collect 32 rows, then store them coalesced
                    const int hr = __shfl_sync(0xffffffffu, h_out, 31);
                    const int fr = __shfl_sync(0xffffffffu, f_out, 31);
                    const int row = s - 31;
                    if (row >= 0 && (row & 31) == lane) { hb_keep = hr; fb_keep = fr; }
                    if (row >= 0 && ((row & 31) == 31 || row == tlen - 1)) {
                        const int r = (row & ~31) + lane;
                        if (r <= row) { hb[r] = hb_keep; fb[r] = fb_keep; }
                    }
                }
            }
            __syncwarp();
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            gmax = max(gmax, __shfl_xor_sync(0xffffffffu, gmax, d));
            mte = max(mte, __shfl_xor_sync(0xffffffffu, mte, d));
            mqe = max(mqe, __shfl_xor_sync(0xffffffffu, mqe, d));
            score = max(score, __shfl_xor_sync(0xffffffffu, score, d));
        }
        if (lane == 0) { out_max[pair] = gmax; out_score[pair] = score; out_mqe[pair] = mqe; out_mte[pair] = mte; }
    }
}

}  // namespace

int align_batch_device(int device, cudaStream_t stream, const uint8_t* aa, const int64_t* off, int64_t n_seq, const int32_t* pair_q,
                       const int32_t* pair_t, int64_t n_pairs, int gapo, int gape, orfx_aln_out* out, double* cells, float* ms, std::string& err) {
    (void)device;
    constexpr int CPL = 8;
    auto ck = [&](cudaError_t e, const char* what) { if (e != cudaSuccess) { err = std::string(what) + ": " + cudaGetErrorString(e); return false; } return true; };
    if (n_pairs == 0) { if (cells) *cells = 0; if (ms) *ms = 0; return ORFX_OK; }
    const int64_t total = off[n_seq];
    // longest-first order (cells = qlen * tlen), so the long tail starts early
    std::vector<int32_t> order((size_t)n_pairs);
    std::iota(order.begin(), order.end(), 0);
    std::vector<double> cost((size_t)n_pairs);
    double sum_cells = 0;
    int64_t max_tlen = 1;
    for (int64_t p = 0; p < n_pairs; p++) {
        if (pair_q[p] < 0 || pair_q[p] >= n_seq || pair_t[p] < 0 || pair_t[p] >= n_seq) { err = "pair index out of range"; return ORFX_E_INVALID; }
        const int64_t ql = off[pair_q[p] + 1] - off[pair_q[p]], tl = off[pair_t[p] + 1] - off[pair_t[p]];
        cost[(size_t)p] = (double)ql * (double)tl;
        sum_cells += cost[(size_t)p];
        if (ql > 32 * CPL) max_tlen = std::max(max_tlen, tl);
    }
    std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return cost[(size_t)a] != cost[(size_t)b] ? cost[(size_t)a] > cost[(size_t)b] : a < b; });
    cudaDeviceProp prop;
    int dev = 0;
    cudaGetDevice(&dev);
    if (!ck(cudaGetDeviceProperties(&prop, dev), "cudaGetDeviceProperties")) return ORFX_E_CUDA;
    const int ctas_per_sm = 4;
    int grid = prop.multiProcessorCount * ctas_per_sm;
    const int64_t max_grid = (n_pairs + ALN_WARPS - 1) / ALN_WARPS;
    if (grid > max_grid) grid = (int)max_grid;
    const int64_t n_warps = (int64_t)grid * ALN_WARPS;
    const int64_t stride = (max_tlen + 31) / 32 * 32;

    uint8_t* d_aa = nullptr; int64_t* d_off = nullptr; int32_t *d_pq = nullptr, *d_pt = nullptr, *d_order = nullptr, *d_out = nullptr, *d_scratch = nullptr;
    unsigned long long* d_counter = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = ORFX_OK;
    do {
        if (!ck(cudaMalloc(&d_aa, (size_t)total + 16), "cudaMalloc aa")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaMalloc(&d_off, (size_t)(n_seq + 1) * 8), "cudaMalloc off")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaMalloc(&d_pq, (size_t)n_pairs * 4), "cudaMalloc pair_q")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaMalloc(&d_pt, (size_t)n_pairs * 4), "cudaMalloc pair_t")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaMalloc(&d_order, (size_t)n_pairs * 4), "cudaMalloc order")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaMalloc(&d_out, (size_t)n_pairs * 16), "cudaMalloc out")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaMalloc(&d_scratch, (size_t)(n_warps * 2 * stride) * 4 + 16), "cudaMalloc scratch")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaMalloc(&d_counter, 8), "cudaMalloc counter")) { rc = ORFX_E_CUDA; break; }
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        if (!ck(cudaMemcpyAsync(d_aa, aa, (size_t)total, cudaMemcpyHostToDevice, stream), "H2D aa")) { rc = ORFX_E_CUDA; break; }
        cudaMemcpyAsync(d_off, off, (size_t)(n_seq + 1) * 8, cudaMemcpyHostToDevice, stream);
        cudaMemcpyAsync(d_pq, pair_q, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, stream);
        cudaMemcpyAsync(d_pt, pair_t, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, stream);
        cudaMemcpyAsync(d_order, order.data(), (size_t)n_pairs * 4, cudaMemcpyHostToDevice, stream);
        cudaMemsetAsync(d_counter, 0, 8, stream);
        cudaEventRecord(e0, stream);
        align_kernel<CPL><<<grid, ALN_WARPS * 32, 0, stream>>>(d_aa, d_off, d_pq, d_pt, d_order, (long long)n_pairs, gapo, gape, d_out, d_out + n_pairs,
                                                             d_out + 2 * n_pairs, d_out + 3 * n_pairs, d_scratch, (long long)stride, d_counter);
        cudaEventRecord(e1, stream);
        cudaMemcpyAsync(out->max, d_out, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
        cudaMemcpyAsync(out->score, d_out + n_pairs, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
        cudaMemcpyAsync(out->mqe, d_out + 2 * n_pairs, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
        cudaMemcpyAsync(out->mte, d_out + 3 * n_pairs, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, stream);
        if (!ck(cudaStreamSynchronize(stream), "alignment kernel")) { rc = ORFX_E_CUDA; break; }
        if (!ck(cudaGetLastError(), "alignment kernel launch")) { rc = ORFX_E_CUDA; break; }
        float t_ms = 0;
        cudaEventElapsedTime(&t_ms, e0, e1);
        if (ms) *ms = t_ms;
        if (cells) *cells = sum_cells;
    } while (false);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(d_aa); cudaFree(d_off); cudaFree(d_pq); cudaFree(d_pt); cudaFree(d_order); cudaFree(d_out); cudaFree(d_scratch); cudaFree(d_counter);
    return rc;
}

}  // namespace orfx
