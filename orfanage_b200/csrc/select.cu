// Per-query reduction: duplicate-ORF check, flatten() by ORF key, the std::sort of every key's rows
// by score_gt, find_best_stat, and the duplicity count of the emitted transcript
// (orfanage.cpp:204-305, 622-627, 700-709, 755-766).  One thread per query transcript.
#include "chain_ops.cuh"
#include "stdsort.cuh"

namespace orfx {

namespace {

// Sort key of one candidate row, written once per row and read by the comparators as three 128-bit loads (the
// fields live in ~10 separate SoA arrays; a comparison on those is a dozen dependent scattered loads, and the rows of
// a titin-class query are compared hundreds of times).
//   k[0] pass; k[1 .. n_modes] one value per --mode entry (at most 6 distinct ones), larger = better; k[7] the row's
//   status (read by the duplicate check only); k[8] transcript_id class of the row's template; k[9] tlen; k[10], k[11]
//   the ORF key
// The records of a block's queries are contiguous (rows are laid out query by query): they are kept in shared memory
// while they fit, so that the comparisons of the std::sort replay cost a shared-memory load, not an L2 round trip.
struct __align__(16) RowKey { int k[12]; };
constexpr int SEL_THREADS = 128;
constexpr int SEL_SMEM_ROWS = 768;     // 36 KB of keys + 6 KB of order / template per block (a block's queries hold ~540 rows on average)

__device__ __forceinline__ RowKey make_key(const DevBatch& B, const DevRows& R, const orfx_params& P, int64_t row) {
    RowKey K;
#pragma unroll
    for (int i = 0; i < 12; i++) K.k[i] = 0;
    K.k[8] = B.tx_tid_class[R.t_tx[row]];
    K.k[0] = R.pass[row];
    K.k[7] = R.status[row];
    for (int i = 0; i < P.n_modes && i < 6; i++) {
        int v = 0;
        switch (P.mode_array[i]) {
            case ORFX_MODE_START_MATCH: v = R.start_match[row]; break;
            case ORFX_MODE_LONGEST: v = R.qlen[row]; break;
            case ORFX_MODE_LONGEST_MATCH: v = R.inframe[row]; break;
            case ORFX_MODE_BEST: v = __float_as_int(R.ilpi[row]); break;   // ilpi >= +0: the bit pattern orders like the float
            case ORFX_MODE_FIRST: v = -R.query_start[row]; break;          // smaller query_start is better
            default: break;                                                 // ALL: only relevant at emission
        }
        K.k[1 + i] = v;
    }
    K.k[9] = R.tlen[row]; K.k[10] = R.cds_start[row]; K.k[11] = R.cds_end[row];
    return K;
}

struct ScoreGt {   // score_gt, orfanage.cpp:204-245, on row indices relative to the query's first row
    const RowKey* keys;
    int n_modes;
    __device__ bool operator()(int a, int b) const {
        const int4* pa = reinterpret_cast<const int4*>(keys + a);
        const int4* pb = reinterpret_cast<const int4*>(keys + b);
        const int4 a0 = pa[0], b0 = pb[0], a1 = pa[1], b1 = pb[1], a2 = pa[2], b2 = pb[2];
        const int l[10] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y};
        const int r[10] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y};
        if (l[0] != r[0]) return l[0] > r[0];
#pragma unroll
        for (int i = 1; i <= 6; i++) if (i <= n_modes && l[i] != r[i]) return l[i] > r[i];
        return l[9] > r[9];
    }
};
struct KeyLess {   // std::map<std::pair<int,int>,...> order, rows of one key kept in candidate order
    const RowKey* keys;
    __device__ bool operator()(int a, int b) const {
        const int2 l = *reinterpret_cast<const int2*>(&keys[a].k[10]), r = *reinterpret_cast<const int2*>(&keys[b].k[10]);
        if (l.x != r.x) return l.x < r.x;
        if (l.y != r.y) return l.y < r.y;
        return a < b;
    }
};

// distinct template transcript_id strings among rows ord[g0..g1)
__device__ int count_distinct_tids(const RowKey* keys, const int32_t* ord, int g0, int g1) {
    int n = 0;
    for (int i = g0; i < g1; i++) {
        const int c = keys[ord[i]].k[8];
        bool seen = false;
        for (int j = g0; j < i && !seen; j++) seen = keys[ord[j]].k[8] == c;
        if (!seen) n++;
    }
    return n;
}

}  // namespace

__global__ void __launch_bounds__(SEL_THREADS) select_kernel(DevBatch B, DevRows R, orfx_params P, int n_queries, orfx_best* best_out, int32_t* ord_scratch,
                                                      int4* key_scratch) {
    __shared__ RowKey s_keys[SEL_SMEM_ROWS];
    __shared__ int32_t s_ord[SEL_SMEM_ROWS];
    __shared__ int32_t s_t[SEL_SMEM_ROWS];      // template of each staged row (the duplicate check walks runs of equal templates)
    const int q_first = blockIdx.x * blockDim.x;
    const int q_last = min(q_first + (int)blockDim.x, n_queries);
    const int64_t block_r0 = B.q_row_off[q_first], block_r1 = B.q_row_off[q_last];
    // Stage the block's rows: the records of a block's queries are contiguous, so the whole block reads them with
    // coalesced loads (a thread walking its own query's rows reads a dozen SoA arrays row after row instead).
    const int staged = (int)min(block_r1 - block_r0, (int64_t)SEL_SMEM_ROWS);
    for (int i = threadIdx.x; i < staged; i += blockDim.x) {
        s_keys[i] = make_key(B, R, P, block_r0 + i);
        s_t[i] = R.t_tx[block_r0 + i];
    }
    // (Measured and dropped: dealing the block's queries to its threads in descending order of their row count, so that a
    // warp holds queries of one size: 0.175 -> 0.21 ms.)
    __syncthreads();
    const int qi = q_first + threadIdx.x;
    if (qi >= n_queries) return;
    const int64_t r0 = B.q_row_off[qi], r1 = B.q_row_off[qi + 1];
    const int n = (int)(r1 - r0);
    const bool in_smem = r1 - block_r0 <= SEL_SMEM_ROWS;      // this query's rows fit the block's stage
    int32_t* ord = in_smem ? s_ord + (r0 - block_r0) : ord_scratch + r0;
    RowKey* keys = in_smem ? s_keys + (r0 - block_r0) : reinterpret_cast<RowKey*>(key_scratch) + r0;

    // duplicate ORFs inside one (query, template) pair: std::unordered_set<int> orf_starts keyed on the
    // genomic cds_start of every row that reaches the check (orfanage.cpp:516, 622-627)
    if (in_smem) {
        const int32_t* tt = s_t + (r0 - block_r0);
        for (int i = 0; i < n; i++) {
            if (keys[i].k[7] == ORFX_ST_OK && keys[i].k[0]) {
                const int t = tt[i], cs = keys[i].k[10];
                for (int j = i - 1; j >= 0 && tt[j] == t; j--) {
                    const bool reached = (keys[j].k[7] == ORFX_ST_OK && keys[j].k[0]) || keys[j].k[7] == ORFX_ST_DUP;
                    if (reached && keys[j].k[10] == cs) {
                        keys[i].k[7] = ORFX_ST_DUP; keys[i].k[0] = 0;
                        R.status[r0 + i] = ORFX_ST_DUP; R.pass[r0 + i] = 0;
                        break;
                    }
                }
            }
        }
        if (!P.has_reference) {   // a scored row's query copy keeps (0,0) unless the row survives (orfanage.cpp:634)
            for (int i = 0; i < n; i++) {
                const int st = keys[i].k[7];
                if (!keys[i].k[0] && (st == ORFX_ST_OK || st == ORFX_ST_DUP)) {
                    keys[i].k[10] = 0; keys[i].k[11] = 0;
                    R.cds_start[r0 + i] = 0; R.cds_end[r0 + i] = 0;
                }
            }
        }
        for (int i = 0; i < n; i++) ord[i] = i;
    } else {   // a query whose rows do not fit the stage (titin-class loci): the same on the global arrays
        for (int i = 0; i < n; i++) {
            const int64_t ri = r0 + i;
            if (R.status[ri] == ORFX_ST_OK && R.pass[ri]) {
                const int t = R.t_tx[ri], cs = R.cds_start[ri];
                for (int j = i - 1; j >= 0 && R.t_tx[r0 + j] == t; j--) {
                    const int64_t rj = r0 + j;
                    const bool reached = (R.status[rj] == ORFX_ST_OK && R.pass[rj]) || R.status[rj] == ORFX_ST_DUP;
                    if (reached && R.cds_start[rj] == cs) { R.status[ri] = ORFX_ST_DUP; R.pass[ri] = 0; break; }
                }
            }
        }
        if (!P.has_reference) {
            for (int i = 0; i < n; i++) {
                const int st = R.status[r0 + i];
                if (!R.pass[r0 + i] && (st == ORFX_ST_OK || st == ORFX_ST_DUP)) { R.cds_start[r0 + i] = 0; R.cds_end[r0 + i] = 0; }
            }
        }
        for (int i = 0; i < n; i++) { ord[i] = i; keys[i] = make_key(B, R, P, r0 + i); }
    }
    stdsort::sort(ord, ord + n, KeyLess{keys});

    const ScoreGt gt{keys, min(P.n_modes, 6)};
    const bool mode_all = P.n_modes > 0 && P.mode_array[0] == ORFX_MODE_ALL;
    int best_front = -1, best_g0 = 0, best_g1 = 0;
    for (int g0 = 0; g0 < n;) {
        int g1 = g0 + 1;
        const int ks = keys[ord[g0]].k[10], ke = keys[ord[g0]].k[11];
        while (g1 < n && keys[ord[g1]].k[10] == ks && keys[ord[g1]].k[11] == ke) g1++;
        stdsort::sort(ord + g0, ord + g1, gt);
        const int front = ord[g0];
        if (best_front < 0 || gt(front, best_front)) { best_front = front; best_g0 = g0; best_g1 = g1; }   // find_best_stat
        if (mode_all && keys[front].k[0] && g1 - g0 > 1) R.n_dup[r0 + front] = count_distinct_tids(keys, ord, g0, g1);
        g0 = g1;
    }
    orfx_best b;
    b.q_tx = B.q_list[qi];
    b.t_tx = -1; b.cds_start = 0; b.cds_end = 0; b.duplicity = 0; b.n_rows = n; b.row_off = r0; b.best_row = -1;
    b.lpi = b.ilpi = b.mlpi = 0.f; b.start_match = 0; b.ext_lo = b.ext_hi = 0;
    if (best_front >= 0 && keys[best_front].k[0]) {
        const int64_t rb = r0 + best_front;
        if (!mode_all) R.n_dup[rb] = count_distinct_tids(keys, ord, best_g0 + 1, best_g1);
        b.t_tx = R.t_tx[rb]; b.cds_start = R.cds_start[rb]; b.cds_end = R.cds_end[rb];
        b.duplicity = R.n_dup[rb]; b.best_row = rb;
        b.lpi = R.lpi[rb]; b.ilpi = R.ilpi[rb]; b.mlpi = R.mlpi[rb]; b.start_match = R.start_match[rb];
        if (R.ext_lo) { b.ext_lo = R.ext_lo[rb]; b.ext_hi = R.ext_hi[rb]; }
    }
    best_out[qi] = b;
    for (int i = 0; i < n; i++) R.print_order[r0 + i] = r0 + ord[i];
}

int launch_select(const DevBatch& B, const DevRows& R, const orfx_params& P, int n_queries, orfx_best* best_out, int32_t* ord_scratch, int4* key_scratch,
                  cudaStream_t s) {
    if (n_queries == 0) return 0;
    select_kernel<<<(n_queries + SEL_THREADS - 1) / SEL_THREADS, SEL_THREADS, 0, s>>>(B, R, P, n_queries, best_out, ord_scratch, key_scratch);
    return 1;
}

}  // namespace orfx
