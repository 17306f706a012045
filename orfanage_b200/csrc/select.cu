// Per-query reduction: duplicate-ORF check, flatten() by ORF key, the std::sort of every key's rows
// by score_gt, find_best_stat, and the duplicity count of the emitted transcript
// (orfanage.cpp:204-305, 622-627, 700-709, 755-766).  One thread per query transcript.
#include "chain_ops.cuh"
#include "stdsort.cuh"

namespace orfx {

namespace {

struct ScoreGt {   // score_gt, orfanage.cpp:204-245, on row indices relative to `base`
    const DevRows* R;
    const orfx_params* P;
    int64_t base;
    __device__ bool operator()(int a, int b) const {
        const int64_t l = base + a, r = base + b;
        if (R->pass[l] != R->pass[r]) return R->pass[l] > R->pass[r];
        for (int i = 0; i < P->n_modes; i++) {
            switch (P->mode_array[i]) {
                case ORFX_MODE_START_MATCH: if (R->start_match[l] != R->start_match[r]) return R->start_match[l] > R->start_match[r]; break;
                case ORFX_MODE_LONGEST: if (R->qlen[l] != R->qlen[r]) return R->qlen[l] > R->qlen[r]; break;
                case ORFX_MODE_LONGEST_MATCH: if (R->inframe[l] != R->inframe[r]) return R->inframe[l] > R->inframe[r]; break;
                case ORFX_MODE_BEST: if (R->ilpi[l] != R->ilpi[r]) return R->ilpi[l] > R->ilpi[r]; break;
                case ORFX_MODE_FIRST: if (R->query_start[l] != R->query_start[r]) return R->query_start[l] < R->query_start[r]; break;
                default: break;   // ALL: only relevant at emission
            }
        }
        return R->tlen[l] > R->tlen[r];
    }
};
struct KeyLess {   // std::map<std::pair<int,int>,...> order, rows of one key kept in candidate order
    const DevRows* R;
    int64_t base;
    __device__ bool operator()(int a, int b) const {
        const int64_t l = base + a, r = base + b;
        if (R->cds_start[l] != R->cds_start[r]) return R->cds_start[l] < R->cds_start[r];
        if (R->cds_end[l] != R->cds_end[r]) return R->cds_end[l] < R->cds_end[r];
        return a < b;
    }
};

// distinct template transcript_id strings among rows ord[g0..g1)
__device__ int count_distinct_tids(const DevBatch& B, const DevRows& R, int64_t base, const int32_t* ord, int g0, int g1) {
    int n = 0;
    for (int i = g0; i < g1; i++) {
        const int c = B.tx_tid_class[R.t_tx[base + ord[i]]];
        bool seen = false;
        for (int j = g0; j < i && !seen; j++) seen = B.tx_tid_class[R.t_tx[base + ord[j]]] == c;
        if (!seen) n++;
    }
    return n;
}

}  // namespace

__global__ void __launch_bounds__(128) select_kernel(DevBatch B, DevRows R, orfx_params P, int n_queries, orfx_best* best_out, int32_t* ord_scratch) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= n_queries) return;
    const int64_t r0 = B.q_row_off[qi], r1 = B.q_row_off[qi + 1];
    const int n = (int)(r1 - r0);
    int32_t* ord = ord_scratch + r0;

    // duplicate ORFs inside one (query, template) pair: std::unordered_set<int> orf_starts keyed on the
    // genomic cds_start of every row that reaches the check (orfanage.cpp:516, 622-627)
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
    if (!P.has_reference) {   // a scored row's query copy keeps (0,0) unless the row survives (orfanage.cpp:634)
        for (int i = 0; i < n; i++) {
            const int st = R.status[r0 + i];
            if (!R.pass[r0 + i] && (st == ORFX_ST_OK || st == ORFX_ST_DUP)) { R.cds_start[r0 + i] = 0; R.cds_end[r0 + i] = 0; }
        }
    }
    for (int i = 0; i < n; i++) ord[i] = i;
    stdsort::sort(ord, ord + n, KeyLess{&R, r0});

    const ScoreGt gt{&R, &P, r0};
    const bool mode_all = P.n_modes > 0 && P.mode_array[0] == ORFX_MODE_ALL;
    int best_front = -1, best_g0 = 0, best_g1 = 0;
    for (int g0 = 0; g0 < n;) {
        int g1 = g0 + 1;
        const int ks = R.cds_start[r0 + ord[g0]], ke = R.cds_end[r0 + ord[g0]];
        while (g1 < n && R.cds_start[r0 + ord[g1]] == ks && R.cds_end[r0 + ord[g1]] == ke) g1++;
        stdsort::sort(ord + g0, ord + g1, gt);
        const int front = ord[g0];
        if (best_front < 0 || gt(front, best_front)) { best_front = front; best_g0 = g0; best_g1 = g1; }   // find_best_stat
        if (mode_all && R.pass[r0 + front] && g1 - g0 > 1) R.n_dup[r0 + front] = count_distinct_tids(B, R, r0, ord, g0, g1);
        g0 = g1;
    }
    orfx_best b;
    b.q_tx = B.q_list[qi];
    b.t_tx = -1; b.cds_start = 0; b.cds_end = 0; b.duplicity = 0; b.n_rows = n; b.row_off = r0; b.best_row = -1;
    b.lpi = b.ilpi = b.mlpi = 0.f; b.start_match = 0;
    if (best_front >= 0 && R.pass[r0 + best_front]) {
        const int64_t rb = r0 + best_front;
        if (!mode_all) R.n_dup[rb] = count_distinct_tids(B, R, r0, ord, best_g0 + 1, best_g1);
        b.t_tx = R.t_tx[rb]; b.cds_start = R.cds_start[rb]; b.cds_end = R.cds_end[rb];
        b.duplicity = R.n_dup[rb]; b.best_row = rb;
        b.lpi = R.lpi[rb]; b.ilpi = R.ilpi[rb]; b.mlpi = R.mlpi[rb]; b.start_match = R.start_match[rb];
    }
    best_out[qi] = b;
    for (int i = 0; i < n; i++) R.print_order[r0 + i] = r0 + ord[i];
}

int launch_select(const DevBatch& B, const DevRows& R, const orfx_params& P, int n_queries, orfx_best* best_out, int32_t* ord_scratch, cudaStream_t s) {
    if (n_queries == 0) return 0;
    select_kernel<<<(n_queries + 127) / 128, 128, 0, s>>>(B, R, P, n_queries, best_out, ord_scratch);
    return 1;
}

}  // namespace orfx
