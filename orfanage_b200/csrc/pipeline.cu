// Batch preparation, candidate enumeration, ORF rescue and pair scoring kernels.
//
//   prep        per-transcript spliced length / per-segment prefix, query list, codon-scan tile list
//   rows        for every query, every template of its bundle that passes the bounding-box test
//               (orfanage.cpp:500-503) -> CHAIN::intersection(..., segments=true) (Transcriptome.h:321-361)
//               -> one candidate row per intersection segment (or one "no-overlap" row)
//   rescue      per row: template phase at the segment's 5' end, seed CDS, stop trim / stop extension,
//               template-guided start choice, start trim (orfanage.cpp:532-599; TX::rescue_cds and
//               callees, Transcriptome.cpp:182-227, 400-677) restated in transcript coordinates as
//               in-frame bit scans over the codon-scan masks
//   score       TX::score (Transcriptome.cpp:699-762): lossy union, sweep classification, lpi/ilpi/mlpi
#include "chain_ops.cuh"
#include <climits>
#include <algorithm>

namespace orfx {

// ------------------------------------------------------------------------------------------------
// exclusive scan of int64 (small hand-written three-phase scan; sizes here are n_tx / n_queries)
// ------------------------------------------------------------------------------------------------
namespace {
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_BLOCK = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* total) {
    __shared__ int64_t warp_sums[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int64_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int64_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int64_t s = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0;
#pragma unroll
        for (int d = 1; d < SCAN_THREADS / 32; d <<= 1) { int64_t t = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += t; }
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = s;
    }
    __syncthreads();
    int64_t base = wid > 0 ? warp_sums[wid - 1] : 0;
    *total = warp_sums[SCAN_THREADS / 32 - 1];
    __syncthreads();
    return base + incl - v;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_local_kernel(const int64_t* in, int64_t* out, int64_t* block_sums, int64_t n) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_BLOCK + (int64_t)threadIdx.x * SCAN_ITEMS;
    int64_t v[SCAN_ITEMS], sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = base + i < n ? in[base + i] : 0; sum += v[i]; }
    int64_t total;
    int64_t ex = block_exclusive_scan(sum, &total);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { if (base + i < n) out[base + i] = ex; ex += v[i]; }
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_THREADS) scan_tops_kernel(int64_t* block_sums, int64_t n_blocks, int64_t* grand_total) {
    int64_t carry = 0;
    for (int64_t c = 0; c < n_blocks; c += SCAN_THREADS) {
        int64_t i = c + threadIdx.x;
        int64_t v = i < n_blocks ? block_sums[i] : 0, total;
        int64_t ex = block_exclusive_scan(v, &total);
        if (i < n_blocks) block_sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) *grand_total = carry;
}
__global__ void __launch_bounds__(SCAN_THREADS) scan_add_kernel(int64_t* out, const int64_t* block_sums, int64_t n) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_BLOCK + (int64_t)threadIdx.x * SCAN_ITEMS;
    const int64_t add = block_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) if (base + i < n) out[base + i] += add;
}
}  // namespace

// out[0..n) = exclusive prefix of in[0..n); out[n] = total.  tmp must hold ceil(n / SCAN_BLOCK) int64.
int launch_exclusive_scan(const int64_t* in, int64_t* out, int64_t* tmp, int64_t n, cudaStream_t s) {
    if (n <= 0) { cudaMemsetAsync(out, 0, sizeof(int64_t), s); return 0; }
    int64_t nb = (n + SCAN_BLOCK - 1) / SCAN_BLOCK;
    scan_local_kernel<<<(unsigned)nb, SCAN_THREADS, 0, s>>>(in, out, tmp, n);
    scan_tops_kernel<<<1, SCAN_THREADS, 0, s>>>(tmp, nb, out + n);
    scan_add_kernel<<<(unsigned)nb, SCAN_THREADS, 0, s>>>(out, tmp, n);
    return 3;
}
int64_t scan_tmp_elems(int64_t n) { return 3 * ((n + SCAN_BLOCK - 1) / SCAN_BLOCK + 1); }   // room for scan3

// Three independent exclusive scans of the same length in three launches instead of nine (the prep stage is a chain of
// small kernels: launch latency is what it costs).  blockIdx.y picks the array.
namespace {
struct Scan3 { const int64_t* in[3]; int64_t* out[3]; };
__global__ void __launch_bounds__(SCAN_THREADS) scan3_local_kernel(Scan3 a, int64_t* block_sums, int64_t nb, int64_t n) {
    const int w = blockIdx.y;
    const int64_t base = (int64_t)blockIdx.x * SCAN_BLOCK + (int64_t)threadIdx.x * SCAN_ITEMS;
    const int64_t* in = a.in[w];
    int64_t* out = a.out[w];
    int64_t v[SCAN_ITEMS], sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = base + i < n ? in[base + i] : 0; sum += v[i]; }
    int64_t total;
    int64_t ex = block_exclusive_scan(sum, &total);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) { if (base + i < n) out[base + i] = ex; ex += v[i]; }
    if (threadIdx.x == 0) block_sums[(int64_t)w * (nb + 1) + blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_THREADS) scan3_tops_kernel(Scan3 a, int64_t* block_sums, int64_t nb, int64_t n) {
    const int w = blockIdx.x;
    int64_t* bs = block_sums + (int64_t)w * (nb + 1);
    int64_t carry = 0;
    for (int64_t c = 0; c < nb; c += SCAN_THREADS) {
        int64_t i = c + threadIdx.x;
        int64_t v = i < nb ? bs[i] : 0, total;
        int64_t ex = block_exclusive_scan(v, &total);
        if (i < nb) bs[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) a.out[w][n] = carry;
}
__global__ void __launch_bounds__(SCAN_THREADS) scan3_add_kernel(Scan3 a, const int64_t* block_sums, int64_t nb, int64_t n) {
    const int w = blockIdx.y;
    const int64_t base = (int64_t)blockIdx.x * SCAN_BLOCK + (int64_t)threadIdx.x * SCAN_ITEMS;
    const int64_t add = block_sums[(int64_t)w * (nb + 1) + blockIdx.x];
    int64_t* out = a.out[w];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) if (base + i < n) out[base + i] += add;
}
}  // namespace
int launch_exclusive_scan3(const int64_t* in0, int64_t* out0, const int64_t* in1, int64_t* out1, const int64_t* in2, int64_t* out2, int64_t* tmp, int64_t n,
                           cudaStream_t s) {
    if (n <= 0) { cudaMemsetAsync(out0, 0, 8, s); cudaMemsetAsync(out1, 0, 8, s); cudaMemsetAsync(out2, 0, 8, s); return 0; }
    const int64_t nb = (n + SCAN_BLOCK - 1) / SCAN_BLOCK;
    Scan3 a{{in0, in1, in2}, {out0, out1, out2}};
    scan3_local_kernel<<<dim3((unsigned)nb, 3), SCAN_THREADS, 0, s>>>(a, tmp, nb, n);
    scan3_tops_kernel<<<3, SCAN_THREADS, 0, s>>>(a, tmp, nb, n);
    scan3_add_kernel<<<dim3((unsigned)nb, 3), SCAN_THREADS, 0, s>>>(a, tmp, nb, n);
    return 3;
}

// ------------------------------------------------------------------------------------------------
// validation of a caller-supplied batch (include/orfx.h is a public seam: the host packer is not the only caller)
// ------------------------------------------------------------------------------------------------
// Bit set in *flag: 1 offsets of the bundle / transcript / sequence CSR not monotone or out of range, 2 unknown
// transcript kind, 4 QUERY / TEMPLATE transcript without a segment, 8 segment with end < start or start < 0 or a chain
// that is not ascending and disjoint, 16 negative --overhang flank, 32 a query segment (with its flank) outside its
// bundle's bases.  Every later kernel trusts these properties.
// eight lanes per transcript (as tx_prep_kernel): the segment loads of a chain are coalesced
__global__ void __launch_bounds__(256) validate_batch_kernel(DevBatch B, int has_seq, int allow_empty, int* flag) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = gt >> 3;
    const int sub = (int)(gt & 7);
    int bad = 0;
    if (gt == 0) {
        if (B.bundle_tx_off[0] != 0 || B.bundle_tx_off[B.n_bundles] != B.n_tx || B.tx_seg_off[0] != 0 || (int64_t)B.tx_seg_off[B.n_tx] != B.n_segs) bad |= 1;
        if (has_seq && (B.seq_off[0] != 0 || B.seq_off[B.n_bundles] > B.seq_bytes)) bad |= 1;
    }
    if (gt < B.n_bundles) {
        const int x0 = B.bundle_tx_off[gt], x1 = B.bundle_tx_off[gt + 1];
        if (x0 < 0 || x1 < x0 || x1 > B.n_tx) bad |= 1;
        if (has_seq) { const int64_t a = B.seq_off[gt], b = B.seq_off[gt + 1]; if (a < 0 || b < a || b > B.seq_bytes) bad |= 1; }
    }
    if (i < B.n_tx) {
        const int s0 = B.tx_seg_off[i], s1 = B.tx_seg_off[i + 1];
        const int kind = B.tx_kind[i];
        if (kind > ORFX_TX_SKIP) bad |= 2;
        if (s0 < 0 || s1 < s0 || (int64_t)s1 > B.n_segs) bad |= 1;
        else {
            if (kind != ORFX_TX_SKIP && s1 == s0 && !allow_empty) bad |= 4;
            int xlo = 0, xhi = 0;
            if (B.tx_ext_lo && kind == ORFX_TX_QUERY) { xlo = B.tx_ext_lo[i]; xhi = B.tx_ext_hi[i]; if (xlo < 0 || xhi < 0) { bad |= 16; xlo = xhi = 0; } }
            // the transcript's bundle (a broken bundle_tx_off is reported by its own threads); only the bases of a query are looked up
            const bool check_seq = has_seq && kind == ORFX_TX_QUERY && B.n_bundles > 0;
            const int lo = check_seq ? B.tx_bundle[i] : 0;   // tx_bundle_kernel: a valid index whatever the batch holds
            const int64_t nbases = check_seq ? B.seq_off[lo + 1] - B.seq_off[lo] : 0;
            const int bstart = check_seq && !B.seg_seq_off ? B.bundle_start[lo] : 0;
            for (int k = s0 + sub; k < s1; k += 8) {
                const orfx_seg sg = B.segs[k];
                if (sg.end < sg.start || sg.start < 0 || (k > s0 && sg.start <= B.segs[k - 1].end)) bad |= 8;
                if (check_seq) {
                    const int64_t at = B.seg_seq_off ? (int64_t)B.seg_seq_off[k] : (int64_t)sg.start - bstart;
                    const int64_t first = at - (k == s0 ? xlo : 0);
                    const int64_t last = at + (sg.end - sg.start) + (k == s1 - 1 ? xhi : 0);
                    if (first < 0 || last >= nbases) bad |= 32;
                }
            }
        }
    }
    if (bad) atomicOr(flag, bad);
}
// Segment dictionary (orfx_batch::seg_ref, include/orfx.h): every segment of a transcript of bundle b is record
// bundle_dict_off[b] + seg_ref[k] of the dictionary.  Writes the three per-segment arrays the path reads; runs before the
// validation, which then checks the expanded chains like uploaded ones.  *flag != 0: a reference outside its bundle's
// records or dictionary offsets that are not monotone (the segment is written as (0, 0)).
// Eight lanes per transcript; the transcript's bundle comes from tx_bundle (tx_bundle_kernel below, launched first).  Offsets
// are only trusted as far as staying inside the arrays: the validation that follows reports what is wrong with them.
// (Measured: one warp per bundle, lanes striding over the bundle's contiguous segments, has the titin-class loci as its
// tail -- 80 us; a bundle search per lane group 73 us.)
__global__ void __launch_bounds__(256) expand_dict_kernel(DevBatch B, const uint16_t* __restrict__ ref, const int32_t* __restrict__ bundle_dict_off,
                                                         const orfx_seg* __restrict__ dict_segs, const int8_t* __restrict__ dict_phase,
                                                         const int32_t* __restrict__ dict_seq_off, int64_t n_dict, orfx_seg* __restrict__ segs,
                                                         int8_t* __restrict__ phase, int32_t* __restrict__ sso, int* flag) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t tx = gt >> 3;
    const int sub = (int)(gt & 7);
    if (tx >= B.n_tx) return;
    const int s0 = B.tx_seg_off[tx], s1 = B.tx_seg_off[tx + 1];
    if (s0 < 0 || s1 < s0 || (int64_t)s1 > B.n_segs) return;   // reported by validate_batch_kernel
    const int b = B.tx_bundle[tx];                              // in [0, n_bundles) whatever the batch holds
    int64_t d0 = bundle_dict_off[b], d1 = bundle_dict_off[b + 1];
    bool bad = false;
    if (d0 < 0 || d1 < d0 || d1 > n_dict || d1 - d0 > 65536) { bad = true; d1 = d0 = 0; }
    for (int k = s0 + sub; k < s1; k += 8) {
        const int64_t d = d0 + (int64_t)ref[k];
        orfx_seg sg{0, 0};
        int ph = 0, so = 0;
        if (d < d1) { sg = dict_segs[d]; ph = dict_phase[d]; if (dict_seq_off) so = dict_seq_off[d]; }
        else bad = true;
        segs[k] = sg; phase[k] = (int8_t)ph;
        if (sso) sso[k] = so;
    }
    if (bad) atomicOr(flag, 1);
}
// tx_bundle[x] = the bundle whose [bundle_tx_off[b], bundle_tx_off[b+1]) holds x, for the upload-time kernels (expansion,
// validation).  Written over a zeroed array and clamped to it, so every entry is a valid bundle index even when the offsets
// are garbage (which the validation then reports); exact when they are a partition of the transcripts.
__global__ void tx_bundle_kernel(DevBatch B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B.n_bundles) return;
    const int x0 = max(B.bundle_tx_off[b], 0), x1 = min(B.bundle_tx_off[b + 1], B.n_tx);
    for (int x = x0; x < x1; x++) B.tx_bundle[x] = b;
}
int launch_tx_bundle(const DevBatch& B, cudaStream_t s) {
    if (B.n_tx <= 0 || B.n_bundles <= 0) return 0;
    cudaMemsetAsync(B.tx_bundle, 0, (size_t)B.n_tx * 4, s);
    tx_bundle_kernel<<<(B.n_bundles + 127) / 128, 128, 0, s>>>(B);
    return 1;
}
int launch_expand_dict(const DevBatch& B, const uint16_t* ref, const int32_t* bundle_dict_off, const orfx_seg* dict_segs, const int8_t* dict_phase,
                       const int32_t* dict_seq_off, int64_t n_dict, orfx_seg* segs, int8_t* phase, int32_t* sso, int* flag, cudaStream_t s) {
    cudaMemsetAsync(flag, 0, sizeof(int), s);
    if (B.n_tx <= 0 || B.n_bundles <= 0 || B.n_segs <= 0) return 0;
    const int64_t n = (int64_t)B.n_tx * 8;
    expand_dict_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(B, ref, bundle_dict_off, dict_segs, dict_phase, dict_seq_off, n_dict, segs, phase, sso, flag);
    return 1;
}
int launch_validate(const DevBatch& B, int has_seq, int allow_empty, int* flag, cudaStream_t s) {
    cudaMemsetAsync(flag, 0, sizeof(int), s);
    const int64_t n = std::max<int64_t>(std::max<int64_t>((int64_t)B.n_tx * 8, B.n_bundles), 1);
    validate_batch_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(B, has_seq, allow_empty, flag);
    return 1;
}

// ------------------------------------------------------------------------------------------------
// prep
// ------------------------------------------------------------------------------------------------
// eight lanes per transcript (a chain has ~10 segments): coalesced segment loads, group scan of the segment lengths
constexpr int PREP_LANES = 8;
__global__ void __launch_bounds__(256) tx_prep_kernel(DevBatch B, int64_t* cnt_q, int64_t* cnt_words, int64_t* cnt_tiles) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int tx = (int)(gt / PREP_LANES), sub = (int)(gt % PREP_LANES);
    if (tx >= B.n_tx) return;   // whole groups leave together
    const unsigned gmask = ((1u << PREP_LANES) - 1u) << ((threadIdx.x & 31) & ~(PREP_LANES - 1));
    const int s0 = B.tx_seg_off[tx], s1 = B.tx_seg_off[tx + 1];
    int acc = 0;
    for (int base = s0; base < s1; base += PREP_LANES) {
        const int k = base + sub;
        int len = 0;
        if (k < s1) { const orfx_seg sg = B.segs[k]; len = sg.end - sg.start + 1; }
        int incl = len;
#pragma unroll
        for (int d = 1; d < PREP_LANES; d <<= 1) { const int v = __shfl_up_sync(gmask, incl, d, PREP_LANES); if (sub >= d) incl += v; }
        if (k < s1) B.seg_cum[k] = acc + incl - len;
        acc += __shfl_sync(gmask, incl, PREP_LANES - 1, PREP_LANES);
    }
    if (sub != 0) return;
    B.tx_len[tx] = acc;
    bool isq = B.tx_kind[tx] == ORFX_TX_QUERY;
    if (isq && B.tx_ext_lo) acc += B.tx_ext_lo[tx] + B.tx_ext_hi[tx];   // the scanned chain covers the --overhang flanks
    const int64_t words = acc > 0 ? (acc + 31) / 32 : 1;
    cnt_q[tx] = isq ? 1 : 0;
    cnt_tiles[tx] = isq ? (words + SCAN_TILE_WORDS - 1) / SCAN_TILE_WORDS : 0;
    cnt_words[tx] = isq ? ((words + 3) & ~(int64_t)3) : 0;   // 16-byte aligned mask regions
}
// A bundle is "heavy" when one of its transcripts has more than HEAVY_TX_SEGS segments (titin-class loci):
// chain walks over such transcripts are long dependent-load chains that would make single threads the
// tail of every row kernel.  Heavy bundles are listed here and handled by the CTA-per-bundle kernels
// below, which stage the bundle's chains in shared memory; the thread-per-query / thread-per-row
// kernels skip them.
constexpr int HEAVY_TX_SEGS = 32;

__global__ void bundle_prep_kernel(DevBatch B, int32_t* heavy_list, unsigned* heavy_count) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B.n_bundles) return;
    uint32_t need = 0;
    int max_segs = 0;
    bool has_q = false, has_t = false;
    for (int x = B.bundle_tx_off[b]; x < B.bundle_tx_off[b + 1]; x++) {
        B.tx_bundle[x] = b;
        const int kind = B.tx_kind[x];
        if (kind == ORFX_TX_QUERY) { need |= B.tx_strand[x] == '+' ? 1u : 2u; has_q = true; }
        if (kind == ORFX_TX_TEMPLATE) has_t = true;
        if (kind != ORFX_TX_SKIP) max_segs = max(max_segs, B.tx_seg_off[x + 1] - B.tx_seg_off[x]);
    }
    if (max_segs > HEAVY_TX_SEGS && has_q && has_t) {
        need |= BUNDLE_HEAVY;
        heavy_list[atomicAdd(heavy_count, 1u)] = b;
    }
    B.bundle_need[b] = (uint8_t)need;
}
__global__ void fill_q_list_kernel(DevBatch B) {
    int tx = blockIdx.x * blockDim.x + threadIdx.x;
    if (tx >= B.n_tx) return;
    if (B.tx_kind[tx] == ORFX_TX_QUERY) B.q_list[B.tx_q_off[tx]] = tx;
}
__global__ void __launch_bounds__(256) fill_tile_list_kernel(DevBatch B) {   // 8 lanes per transcript
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int tx = (int)(gt >> 3), sub = (int)(gt & 7);
    if (tx >= B.n_tx || B.tx_kind[tx] != ORFX_TX_QUERY) return;
    const int64_t t1 = B.tx_tile_off[tx + 1];
    for (int64_t t = B.tx_tile_off[tx] + sub; t < t1; t += 8) B.tile_tx[t] = tx;
}
// Everything that needs no host-side size: lengths, prefixes, the three per-transcript scans, the
// query list (capacity n_tx).  The totals land in tx_q_off[n_tx], tx_mask_off[n_tx], tx_tile_off[n_tx].
int launch_prep(const DevBatch& B, int64_t* cnt_q, int64_t* cnt_words, int64_t* cnt_tiles, int64_t* scan_tmp, int32_t* heavy_list, unsigned* heavy_count,
                cudaStream_t s) {
    int n = 0;
    if (B.n_tx == 0) return 0;
    tx_prep_kernel<<<(unsigned)(((int64_t)B.n_tx * PREP_LANES + 255) / 256), 256, 0, s>>>(B, cnt_q, cnt_words, cnt_tiles); n++;
    cudaMemsetAsync(heavy_count, 0, sizeof(unsigned), s);
    bundle_prep_kernel<<<(B.n_bundles + 127) / 128, 128, 0, s>>>(B, heavy_list, heavy_count); n++;
    n += launch_exclusive_scan3(cnt_q, B.tx_q_off, cnt_words, B.tx_mask_off, cnt_tiles, B.tx_tile_off, scan_tmp, B.n_tx, s);
    fill_q_list_kernel<<<(B.n_tx + 255) / 256, 256, 0, s>>>(B); n++;
    return n;
}
int launch_fill_tiles(const DevBatch& B, cudaStream_t s) {
    if (B.n_tx == 0) return 0;
    fill_tile_list_kernel<<<(unsigned)(((int64_t)B.n_tx * 8 + 255) / 256), 256, 0, s>>>(B);
    return 1;
}

// ------------------------------------------------------------------------------------------------
// rows: bounding box + intersection segments
// ------------------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ void put_row(const DevRows& R, int64_t at, int qx, int t, int s, int e, int status) {
    R.q_tx[at] = qx; R.t_tx[at] = t; R.seg_start[at] = s; R.seg_end[at] = e; R.status[at] = (uint8_t)status;
}

// Candidate rows of one (query, template) pair in the reference's order: CHAIN::intersection(...,
// segments = true) (Transcriptome.h:321-361) merges consecutive pieces into one genomic span while they
// share intron boundaries; one row per span, or one "no-overlap" row.  FILL = false only counts.
// `segs` is B.segs or a shared-memory copy of the bundle's segments biased to accept the same absolute
// indices.  Same visiting order and state as the reference's double loop at O(exons + segments) per pair.
template <class Emit>
__device__ __forceinline__ int pair_rows_core(const DevBatch& B, const orfx_seg* segs, int qx, int t, Emit&& emit) {
    const int q0 = B.tx_seg_off[qx], q1 = B.tx_seg_off[qx + 1];
    const int t0 = B.tx_seg_off[t], t1 = B.tx_seg_off[t + 1];
    int nseg = 0, cur_s = 0, cur_e = 0, n = 0;
    if (q0 < q1 && t0 < t1) {
        // The reference's double loop, flattened into one merge step per exon or template segment (its
        // `start` pointer is j here); the next element of either chain is already loaded when it is needed.
        int i = q0, j = t0;
        orfx_seg s1 = segs[i], c2 = segs[j];
        orfx_seg s1n = segs[min(i + 1, q1 - 1)], c2n = segs[min(j + 1, t1 - 1)];
        bool continue_prev = false, found_yet = false;
        while (true) {
            bool next_seg;
            if (c2.end < s1.start) {            // segment wholly left of the exon: walked over (low > start)
                continue_prev = false;
                next_seg = true;
            } else if (c2.start > s1.end) {     // nothing (more) under this exon
                if (!found_yet) continue_prev = false;
                next_seg = false;
            } else {
                const int is = max(s1.start, c2.start), ie = min(s1.end, c2.end);
                found_yet = true;
                const bool ends_match = ie == s1.end && s1.end == c2.end;
                const bool starts_match = is == s1.start && s1.start == c2.start;
                if (!continue_prev || !starts_match) {
                    if (nseg > 0) { emit(n, cur_s, cur_e, ORFX_ST_OK); n++; }
                    nseg++;
                    cur_s = is; cur_e = ie;
                } else {
                    cur_s = min(cur_s, is); cur_e = max(cur_e, ie);
                }
                continue_prev = ends_match;
                next_seg = ie == c2.end;        // else the segment reaches past the exon: next exon, same segment
            }
            if (next_seg) {
                if (++j >= t1) break;
                c2 = c2n; c2n = segs[min(j + 1, t1 - 1)];
            } else {
                if (++i >= q1) break;
                s1 = s1n; s1n = segs[min(i + 1, q1 - 1)];
                found_yet = false;
            }
        }
    }
    if (nseg > 0) { emit(n, cur_s, cur_e, ORFX_ST_OK); n++; }
    else { emit(n, -1, -1, ORFX_ST_NO_OVERLAP); n++; }
    return n;
}
template <bool FILL>
__device__ int pair_rows(const DevBatch& B, const orfx_seg* segs, const DevRows& R, int qx, int t, int64_t out) {
    if (FILL) return pair_rows_core(B, segs, qx, t, [&](int i, int s, int e, int st) { put_row(R, out + i, qx, t, s, e, st); });
    return pair_rows_core(B, segs, qx, t, [](int, int, int, int) {});
}
// one walk: the first ROWS_BUF rows of the pair are kept in registers (nearly every pair has one or two)
constexpr int ROWS_BUF = 3;
struct RowBuf { int s[ROWS_BUF], e[ROWS_BUF]; int st0; };
__device__ __forceinline__ int pair_rows_buffered(const DevBatch& B, const orfx_seg* segs, int qx, int t, RowBuf& rb) {
    return pair_rows_core(B, segs, qx, t, [&](int i, int s, int e, int st) {
#pragma unroll
        for (int k = 0; k < ROWS_BUF; k++) if (i == k) { rb.s[k] = s; rb.e[k] = e; }
        if (i == 0) rb.st0 = st;
    });
}

// does template slot t of the query's bundle take part?  (bounding box, orfanage.cpp:500-503)
__device__ __forceinline__ bool pair_live(const DevBatch& B, int t, int x1, int qstart, int qend) {
    return t < x1 && B.tx_kind[t] == ORFX_TX_TEMPLATE && !(qstart > B.tx_cds_end[t] || qend < B.tx_cds_start[t]);
}
}  // namespace

// ROWS_GROUP lanes per query slot (8 queries per warp); the lanes of a group take the transcripts of
// the query's bundle ROWS_GROUP at a time.  The query count is still on the device here: slots past
// the last query count zero, so that q_row_off[n_queries .. n_tx] all hold the total.
constexpr int ROWS_GROUP = 4;

__global__ void __launch_bounds__(256) count_rows_kernel(DevBatch B, int64_t* cnt) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int qi = (int)(gt / ROWS_GROUP), sub = (int)(gt % ROWS_GROUP);
    const int64_t nq = B.tx_q_off[B.n_tx];
    int total = 0;
    bool heavy = false;
    if (qi < nq) {
        const int qx = B.q_list[qi];
        const int b = B.tx_bundle[qx];
        heavy = (B.bundle_need[b] & BUNDLE_HEAVY) != 0;   // counted by heavy_count_kernel, which may run at the same time
        const int x0 = B.bundle_tx_off[b], x1 = heavy ? x0 : B.bundle_tx_off[b + 1];
        const int q0 = B.tx_seg_off[qx], q1 = B.tx_seg_off[qx + 1];
        const int qstart = B.segs[q0].start, qend = B.segs[q1 - 1].end;
        DevRows dummy{};
        for (int t = x0 + sub; t < x1; t += ROWS_GROUP)
            if (pair_live(B, t, x1, qstart, qend)) total += pair_rows<false>(B, B.segs, dummy, qx, t, 0);
    }
#pragma unroll
    for (int d = ROWS_GROUP / 2; d > 0; d >>= 1) total += __shfl_xor_sync(0xffffffffu, total, d);
    if (sub == 0 && qi < B.n_tx && !heavy) cnt[qi] = total;
}
__global__ void __launch_bounds__(256) fill_rows_kernel(DevBatch B, DevRows R, int n_queries) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int qi = (int)(gt / ROWS_GROUP), sub = (int)(gt % ROWS_GROUP);
    bool valid = qi < n_queries;
    const int qx = valid ? B.q_list[qi] : 0;
    int x0 = 0, x1 = 0, qstart = 0, qend = 0;
    int64_t out = 0;
    if (valid && (B.bundle_need[B.tx_bundle[qx]] & BUNDLE_HEAVY)) valid = false;   // filled by heavy_fill_rescue_kernel
    if (valid) {
        const int b = B.tx_bundle[qx];
        x0 = B.bundle_tx_off[b]; x1 = B.bundle_tx_off[b + 1];
        qstart = B.segs[B.tx_seg_off[qx]].start; qend = B.segs[B.tx_seg_off[qx + 1] - 1].end;
        out = B.q_row_off[qi];
    }
    // the groups of a warp iterate together (shuffles are warp-wide): as many rounds as the largest bundle
    int rounds = (x1 - x0 + ROWS_GROUP - 1) / ROWS_GROUP;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) rounds = max(rounds, __shfl_xor_sync(0xffffffffu, rounds, d));
    for (int it = 0; it < rounds; it++) {
        const int t = x0 + it * ROWS_GROUP + sub;
        const bool live = valid && pair_live(B, t, x1, qstart, qend);
        RowBuf rb;
        const int n = live ? pair_rows_buffered(B, B.segs, qx, t, rb) : 0;
        int incl = n;
#pragma unroll
        for (int d = 1; d < ROWS_GROUP; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d, ROWS_GROUP); if (sub >= d) incl += v; }
        if (live) {
            if (n <= ROWS_BUF) {   // a "no-overlap" pair has exactly one row; every other row is ORFX_ST_OK
#pragma unroll
                for (int k = 0; k < ROWS_BUF; k++) if (k < n) put_row(R, out + incl - n + k, qx, t, rb.s[k], rb.e[k], k == 0 ? rb.st0 : ORFX_ST_OK);
            } else pair_rows<true>(B, B.segs, R, qx, t, out + incl - n);
        }
        out += __shfl_sync(0xffffffffu, incl, ROWS_GROUP - 1, ROWS_GROUP);
    }
}
int launch_heavy_count(const DevBatch& B, int64_t* cnt, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm, cudaStream_t s);
// The heavy bundles' kernel runs beside the thread-per-query kernel on `side` (disjoint queries; both are latency-bound
// and neither fills the GPU): fork after the zeroing, join before the scan.
int launch_count_rows(const DevBatch& B, int64_t* cnt, int64_t* scan_tmp, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm, cudaStream_t s,
                      cudaStream_t side, cudaEvent_t ev_fork, cudaEvent_t ev_join) {
    if (B.n_tx == 0) { cudaMemsetAsync(B.q_row_off, 0, sizeof(int64_t), s); return 0; }
    cudaMemsetAsync(cnt, 0, (size_t)B.n_tx * sizeof(int64_t), s);   // heavy_count_kernel adds into its queries' counters
    cudaEventRecord(ev_fork, s);
    cudaStreamWaitEvent(side, ev_fork, 0);
    count_rows_kernel<<<(unsigned)(((int64_t)B.n_tx * ROWS_GROUP + 255) / 256), 256, 0, s>>>(B, cnt);
    int n = 1 + launch_heavy_count(B, cnt, heavy_list, heavy_count, n_sm, side);   // (plans the heavy work items first)
    cudaEventRecord(ev_join, side);
    cudaStreamWaitEvent(s, ev_join, 0);
    return n + launch_exclusive_scan(cnt, B.q_row_off, scan_tmp, B.n_tx, s);
}
int launch_fill_rows(const DevBatch& B, const DevRows& R, int n_queries, cudaStream_t s) {
    if (n_queries == 0) return 0;
    fill_rows_kernel<<<(unsigned)(((int64_t)n_queries * ROWS_GROUP + 255) / 256), 256, 0, s>>>(B, R, n_queries);
    return 1;
}

// ------------------------------------------------------------------------------------------------
// rescue + score, one thread per candidate row
// ------------------------------------------------------------------------------------------------
namespace {

// Template-guided choice among the in-frame ATGs upstream of `a` (TX::extend_to_start(TX*,bool),
// Transcriptome.cpp:564-619): the pieces of (query CDS extended to the farthest ATG) x (template CDS)
// are walked 5'->3' until the first piece where both CDSs are in the same frame; the start is the
// nearest ATG whose transcript position does not exceed that piece's CDS-relative position (mixed
// coordinate systems exactly as written, cpp:581-590), else the farthest ATG.
// `base`: 5' end of the query CDS that is intersected and the origin of its CDS-relative positions -- the farthest ATG
// (`far`) everywhere except on strand '.', where the reference extends the chain at the wrong end (see process_row).
__device__ int guided_start(const TxView& q, const TxView& t, const uint32_t* start_mask, int far, int base, int a, int e, int lb, int o) {
    const int g_far = t2g(q, base), g_e = t2g(q, e);
    const int glo = min(g_far, g_e), ghi = max(g_far, g_e);
    const bool plus = q.plus;
    bool seen = false, done = false;
    uint32_t Q = 0;
    int qi = seg_of_genomic(q, plus ? glo : ghi);
    for (; qi >= 0 && qi < q.n && !done; qi += plus ? 1 : -1) {
        const orfx_seg qs = vseg(q, qi);
        if (qs.end < glo || qs.start > ghi) break;
        const int s = max(qs.start, glo), en = min(qs.end, ghi);
        if (plus) {
            int lo = 0, hi = t.n;   // first template segment with end >= s
            while (lo < hi) { int mid = (lo + hi) >> 1; if (t.segs[mid].end < s) lo = mid + 1; else hi = mid; }
            for (int tj = lo; tj < t.n; tj++) {
                const orfx_seg ts = t.segs[tj];
                if (ts.start > en) break;
                const int pos5 = max(s, ts.start);
                const uint32_t q_tpos = (uint32_t)(vcum(q, qi) + (pos5 - qs.start) - base);
                const uint32_t t_tpos = (uint32_t)(t.cum[tj] + (pos5 - ts.start));
                seen = true; Q = q_tpos;
                if (t_tpos % 3 == q_tpos % 3) { done = true; break; }
            }
        } else {
            int lo = -1, hi = t.n - 1;   // last template segment with start <= en
            while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (t.segs[mid].start <= en) lo = mid; else hi = mid - 1; }
            for (int tj = lo; tj >= 0; tj--) {
                const orfx_seg ts = t.segs[tj];
                if (ts.end < s) break;
                const int pos5 = min(en, ts.end);
                const uint32_t q_tpos = (uint32_t)(q.L - 1 - (vcum(q, qi) + (pos5 - qs.start)) - base);
                const uint32_t t_tpos = (uint32_t)(t.L - 1 - (t.cum[tj] + (pos5 - ts.start)));
                seen = true; Q = q_tpos;
                if (t_tpos % 3 == q_tpos % 3) { done = true; break; }
            }
        }
    }
    if (!seen) return far;
    const uint32_t lim = Q < (uint32_t)a ? Q : (uint32_t)a;
    const uint32_t r = (uint32_t)(a % 3);
    if (lim < r) return far;
    const int p = (int)(lim - ((lim - r) % 3));
    const int m = prev_inframe(start_mask, p + o) - o;   // o: mask index of transcript position 0 (--overhang flank)
    return m >= lb ? m : far;
}

// ---- warp-cooperative chain walks (heavy bundles: chains of hundreds of segments) ------------------
// The per-thread walks above are dependent-load chains of O(exons + segments) steps; on titin-class
// transcripts a warp does them 32 exons at a time instead.  Every lane of the warp calls these with the
// same arguments; results are warp-uniform.

// first k in [lo, hi) with segs[k].end >= g, else hi
__device__ __forceinline__ int first_end_ge(const orfx_seg* segs, int lo, int hi, int g) {
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (segs[mid].end < g) lo = mid + 1; else hi = mid; }
    return lo;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}
__device__ __forceinline__ int warp_excl_scan(int v, int lane) {
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
    return incl - v;
}

// pair_rows for a whole warp.  The reference's merge rule (CHAIN::intersection with segments = true,
// Transcriptome.h:321-361) joins piece (exon i, segment j) to the piece before it exactly when that piece
// is (i-1, j-1) with equal ends and this one has equal starts (a shared intron); any other predecessor
// resets `continue_prev`.  So every piece knows locally whether it opens a row (head) and whether it closes
// one (tail); row r runs from the r-th head's start to the r-th tail's end.
template <bool FILL>
__device__ int pair_rows_warp(const DevBatch& B, const orfx_seg* segs, const DevRows& R, int qx, int t, int64_t out, int lane) {
    const int q0 = B.tx_seg_off[qx], q1 = B.tx_seg_off[qx + 1];
    const int t0 = B.tx_seg_off[t], t1 = B.tx_seg_off[t + 1];
    int heads = 0, tails = 0;
    if (t0 < t1) {
        for (int base = q0; base < q1; base += 32) {
            const int i = base + lane;
            int nh = 0, nt = 0, j0 = t1;
            orfx_seg s1 = {0, -1};
            int prev_end = INT_MIN, next_start = INT_MIN;
            if (i < q1) {
                s1 = segs[i];
                if (i > q0) prev_end = segs[i - 1].end;
                if (i + 1 < q1) next_start = segs[i + 1].start;
                j0 = first_end_ge(segs, t0, t1, s1.start);
                for (int j = j0; j < t1; j++) {
                    const orfx_seg c2 = segs[j];
                    if (c2.start > s1.end) break;
                    const bool lp = j > t0 && prev_end != INT_MIN && s1.start == c2.start && prev_end == segs[j - 1].end;
                    const bool ln = j + 1 < t1 && next_start != INT_MIN && s1.end == c2.end && next_start == segs[j + 1].start;
                    nh += !lp; nt += !ln;
                }
            }
            if (FILL) {
                int64_t hp = out + heads + warp_excl_scan(nh, lane), tp = out + tails + warp_excl_scan(nt, lane);
                if (i < q1) {
                    for (int j = j0; j < t1; j++) {
                        const orfx_seg c2 = segs[j];
                        if (c2.start > s1.end) break;
                        const bool lp = j > t0 && prev_end != INT_MIN && s1.start == c2.start && prev_end == segs[j - 1].end;
                        const bool ln = j + 1 < t1 && next_start != INT_MIN && s1.end == c2.end && next_start == segs[j + 1].start;
                        if (!lp) { R.q_tx[hp] = qx; R.t_tx[hp] = t; R.seg_start[hp] = max(s1.start, c2.start); R.status[hp] = ORFX_ST_OK; hp++; }
                        if (!ln) { R.seg_end[tp] = min(s1.end, c2.end); tp++; }
                    }
                }
            }
            heads += warp_sum(nh); tails += warp_sum(nt);
        }
    }
    if (heads == 0) {
        if (FILL && lane == 0) put_row(R, out, qx, t, -1, -1, ORFX_ST_NO_OVERLAP);
        return 1;
    }
    return heads;
}

// CHAIN::compare + the frame walk of TX::score (Transcriptome.h:514-558, cpp:719-745) for a whole warp.
// The walk bins every match piece by (extra - missing) mod 3 seen so far on the genomic-left side; that
// difference is (query CDS bases left of the piece) - (template bases left of it), which the prefix arrays
// give directly, so the pieces are independent.  q segments are clipped to [cs, ce] (indices qk0..qk1).
struct SweepSums { int match, bin0, bin1, bin2; };
// CHAIN::get_union (Transcriptome.h:291-319; lossy: both chains advance on an overlap, the tail of the
// longer chain is dropped) for a whole warp.  When the walk reaches query segment i its template index is
// F(i) or F(i) + 1, F(i) = first template segment ending at or after the segment's start (skipped segments
// are added on the way), so a query segment is a map {0, 1} -> (next state, length added); state 2 = template
// exhausted.  The maps compose associatively: one ordered warp scan per 32 query segments.
struct UMap { int n0, c0, n1, c1; };
__device__ __forceinline__ UMap umap_then(const UMap& a, const UMap& b) {
    UMap r;
    r.n0 = a.n0 == 2 ? 2 : (a.n0 == 0 ? b.n0 : b.n1);
    r.c0 = a.c0 + (a.n0 == 2 ? 0 : (a.n0 == 0 ? b.c0 : b.c1));
    r.n1 = a.n1 == 2 ? 2 : (a.n1 == 0 ? b.n0 : b.n1);
    r.c1 = a.c1 + (a.n1 == 2 ? 0 : (a.n1 == 0 ? b.c0 : b.c1));
    return r;
}
// Union and sweep in one pass: both need F(i), found once per query segment (the next segment's F comes from
// the neighbouring lane).
__device__ int union_sweep_warp(const TxView& q, const TxView& t, int qk0, int qk1, int cs, int ce, int lane, SweepSums& S) {
    auto tcum = [&](int k) { return k < t.n ? t.cum[k] : t.L; };
    S.match = S.bin0 = S.bin1 = S.bin2 = 0;
    const int qbase = vcum(q, qk0) + (cs - vseg(q, qk0).start);
    int state = 0;
    int total = tcum(first_end_ge(t.segs, 0, t.n, max(vseg(q, qk0).start, cs)));
    for (int base = qk0; base <= qk1; base += 32) {
        const int i = base + lane;
        const bool valid = i <= qk1;
        orfx_seg e = {0, -1};
        int c1s = 0, c1e = -1, F = t.n;
        if (valid) {
            e = vseg(q, i);
            c1s = max(e.start, cs); c1e = min(e.end, ce);
            F = first_end_ge(t.segs, 0, t.n, c1s);
        }
        int Fn = __shfl_down_sync(0xffffffffu, F, 1);
        if (lane == 31 && i < qk1) Fn = first_end_ge(t.segs, F, t.n, vseg(q, i + 1).start);
        UMap m = {0, 0, 1, 0};   // identity
        if (valid) {
            // sweep: the match pieces under this query segment
            for (int j = F; j < t.n; j++) {
                const orfx_seg c = t.segs[j];
                if (c.start > c1e) break;
                const int lo = max(c1s, c.start), hi = min(c1e, c.end);
                const int qb = vcum(q, i) + (lo - e.start) - qbase, tb = t.cum[j] + (lo - c.start);
                int d = (qb - tb) % 3; if (d < 0) d += 3;
                const int len = hi - lo + 1;
                S.match += len;
                if (d == 0) S.bin0 += len; else if (d == 1) S.bin1 += len; else S.bin2 += len;
            }
            // union: this segment's map
#pragma unroll
            for (int st = 0; st < 2; st++) {
                const int jp = F + st;
                int ns, cost;
                if (jp >= t.n) { ns = 2; cost = 0; }
                else {
                    const orfx_seg c2 = t.segs[jp];
                    int jo;
                    if (c2.start > c1e) { cost = c1e - c1s + 1; jo = jp; }
                    else { cost = max(c1e, c2.end) - min(c1s, c2.start) + 1; jo = jp + 1; }
                    if (i == qk1) ns = 0;
                    else if (jo < Fn) { cost += tcum(Fn) - tcum(jo); ns = 0; }
                    else ns = jo - Fn;
                }
                if (st == 0) { m.n0 = ns; m.c0 = cost; } else { m.n1 = ns; m.c1 = cost; }
            }
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            UMap o;
            o.n0 = __shfl_up_sync(0xffffffffu, m.n0, d); o.c0 = __shfl_up_sync(0xffffffffu, m.c0, d);
            o.n1 = __shfl_up_sync(0xffffffffu, m.n1, d); o.c1 = __shfl_up_sync(0xffffffffu, m.c1, d);
            if (lane >= d) m = umap_then(o, m);
        }
        const int n0 = __shfl_sync(0xffffffffu, m.n0, 31), c0 = __shfl_sync(0xffffffffu, m.c0, 31);
        const int n1 = __shfl_sync(0xffffffffu, m.n1, 31), c1 = __shfl_sync(0xffffffffu, m.c1, 31);
        if (state != 2) {
            total += state == 0 ? c0 : c1;
            state = state == 0 ? n0 : n1;
        }
    }
    S.match = warp_sum(S.match); S.bin0 = warp_sum(S.bin0); S.bin1 = warp_sum(S.bin1); S.bin2 = warp_sum(S.bin2);
    return total;
}

__device__ __forceinline__ float pct(int num, int den) {
    // 100.0 * ((float)num / (float)den) narrowed to float (Transcriptome.cpp:748-750)
    return __double2float_rn(__dmul_rn(100.0, (double)__fdiv_rn((float)num, (float)den)));
}

}  // namespace

// One candidate row: rescue (when a genome is loaded) + TX::score.  q / t may view the chains in
// global or in shared memory.
// TX::score's chain arithmetic (Transcriptome.cpp:699-750) for a query CDS = q's segments clipped to the
// genomic span [cs, ce] (qlen bases) against the template chain t: lossy union, sweep classification, frame walk.
struct ChainScore { int ulen, extra, missing, match, inframe, outframe; };
template <bool WARP>
__device__ __forceinline__ ChainScore score_chains(const TxView& q, const TxView& t, int cs, int ce, int qlen, bool plus, int lane) {
    int ulen = 0, extra = 0, missing = 0, match = 0, inframe = 0, outframe = 0;
    const int tlen = t.L;
    // query CDS segments = exons clipped to [cs,ce]
    const int qk0 = seg_of_genomic(q, cs), qk1 = seg_of_genomic(q, ce);
    if (WARP) {
        SweepSums S;
        ulen = union_sweep_warp(q, t, qk0, qk1, cs, ce, lane, S);
        match = S.match; extra = qlen - match; missing = tlen - match;
        int want = 0;
        if (!plus) { want = (extra - missing) % 3; if (want < 0) want += 3; }
        inframe = want == 0 ? S.bin0 : (want == 1 ? S.bin1 : S.bin2);
        outframe = match - inframe;
    }
    // lossy union (CHAIN::get_union, Transcriptome.h:291-319)
    if (!WARP) {
        int i = qk0, j = 0;
        while (i <= qk1 && j < t.n) {
            const int c1s = max(vseg(q, i).start, cs), c1e = min(vseg(q, i).end, ce);
            const int c2s = t.segs[j].start, c2e = t.segs[j].end;
            if (c1e < c2s) { ulen += c1e - c1s + 1; i++; }
            else if (c2e < c1s) { ulen += c2e - c2s + 1; j++; }
            else { ulen += max(c1e, c2e) - min(c1s, c2s) + 1; i++; j++; }
        }
    }
    // sweep classification (CHAIN::compare + the frame walk, Transcriptome.h:514-558, cpp:719-745):
    // match pieces are binned by (extra - missing) mod 3 seen so far on the genomic-left side
    if (!WARP) {
        int bin[3] = {0, 0, 0};
        int i = qk0, j = 0;
        int qs = max(vseg(q, i).start, cs), qe = min(vseg(q, i).end, ce);
        int ts = 0, te = -1;
        if (t.n > 0) { ts = t.segs[0].start; te = t.segs[0].end; }
        bool hq = true, ht = t.n > 0;
        while (hq || ht) {
            if (hq && (!ht || qe < ts)) {          // query-only remainder
                extra += qe - qs + 1;
                i++; hq = i <= qk1;
                if (hq) { qs = max(vseg(q, i).start, cs); qe = min(vseg(q, i).end, ce); }
            } else if (ht && (!hq || te < qs)) {   // template-only remainder
                missing += te - ts + 1;
                j++; ht = j < t.n;
                if (ht) { ts = t.segs[j].start; te = t.segs[j].end; }
            } else {
                const int lo = max(qs, ts), hi = min(qe, te);
                if (qs < lo) extra += lo - qs;
                if (ts < lo) missing += lo - ts;
                int d = (extra - missing) % 3; if (d < 0) d += 3;
                bin[d] += hi - lo + 1;
                match += hi - lo + 1;
                qs = hi + 1; ts = hi + 1;
                if (qs > qe) { i++; hq = i <= qk1; if (hq) { qs = max(vseg(q, i).start, cs); qe = min(vseg(q, i).end, ce); } }
                if (ts > te) { j++; ht = j < t.n; if (ht) { ts = t.segs[j].start; te = t.segs[j].end; } }
            }
        }
        // 5'->3' on '+' is genomic-left to right: in frame when (extra-missing) so far == 0 (mod 3);
        // on the reverse strand the 5' side is the genomic-right side: totals minus the left part
        int want = 0;
        if (!plus) { want = (extra - missing) % 3; if (want < 0) want += 3; }
        inframe = bin[want];
        outframe = match - inframe;
    }
    ChainScore r = {ulen, extra, missing, match, inframe, outframe};
    return r;
}

// WARP = true: all 32 lanes of a warp work on the same row (heavy bundles); the two O(chain) walks of
// TX::score are done cooperatively, everything else redundantly, lane 0 stores.
template <bool WARP>
__device__ void process_row(const DevBatch& B, const DevRows& R, const orfx_params& P, const int64_t row, const TxView& q_in, const TxView& t) {
    const int lane = WARP ? (int)(threadIdx.x & 31) : 0;
    TxView q = q_in;   // the row's query copy: --overhang may grow its terminal exons
    // defaults: an all-zero Score that does not pass
    int status = R.status[row];
    int key_s = 0, key_e = 0;
    int qlen = 0, tlen = 0, ulen = 0, query_start = 0, extra = 0, missing = 0, match = 0, outframe = 0, inframe = 0;
    float lpi = 0.f, ilpi = 0.f, mlpi = 0.f;
    int pass = 0, start_match = 0;
    int aln_0 = 0, aln_n = 0, seq_0 = 0, seq_n = 0;   // the ORF as SEQ holds it (alignment stage)

    const int qx = R.q_tx[row], tx = R.t_tx[row];
    const int seg_s = R.seg_start[row], seg_e = R.seg_end[row];
    const bool plus = q.plus;
    const int t_cds_start = B.tx_cds_start[tx], t_cds_end = B.tx_cds_end[tx];
    const int t5 = plus ? t_cds_start : t_cds_end;
    int start = 0, e = 0;   // final ORF, transcript coordinates
    // Strand '.': the reference reads such a transcript as reverse-strand sequence, but TX::extend_cds_chain is told
    // `strand == '-'` (Transcriptome.cpp:487, 566), so every 5' extension of the CDS chain lands on its 3' end instead:
    // the chain keeps its start and grows downstream by the same number of bases (CHAIN::extend, h:365-397, stops at the
    // end of the exon chain).  Reproduced as written; `sq_*` is what SEQ (get_aa) holds when that differs from the chain.
    const bool dot = !plus && B.tx_strand[qx] != '-';
    int sq_start = -1, sq_e = -1;
    // --overhang: the masks cover [5' flank][transcript][3' flank]; transcript position p is mask bit p + o
    int E5 = 0, E3 = 0;
    bool virt3 = false;   // the one-base 3' flank re-reads the edge base of the chromosome (ORFX_EXT_*_VIRTUAL)
    if (B.tx_ext_lo) {
        const int lo = B.tx_ext_lo[qx], hi = B.tx_ext_hi[qx]; E5 = plus ? lo : hi; E3 = plus ? hi : lo;
        if (B.tx_ext_flags) virt3 = (B.tx_ext_flags[qx] & (plus ? ORFX_EXT_HI_VIRTUAL : ORFX_EXT_LO_VIRTUAL)) != 0;
    }
    bool stop_in_virtual = false;
    int o = E5;
    const uint32_t* stopm = nullptr;
    const uint32_t* startm = nullptr;
    auto mbit = [&](const uint32_t* m, int p) { return mask_bit(m, p + o); };
    auto nxt = [&](const uint32_t* m, int p, int lim) {   // first hit in [p, lim)
        const int r = WARP ? next_inframe_warp(m, p + o, lim + o, lane) : next_inframe(m, p + o, lim + o);
        return r < 0 ? -1 : r - o;
    };
    auto prv = [&](const uint32_t* m, int p) { const int r = p < 0 ? -1 : prev_inframe(m, p + o); return r < o ? -1 : r - o; };       // last hit in [0, p]

    do {
        if (status != ORFX_ST_OK) break;   // no-overlap
        // template phase at the segment's 5' end (orfanage.cpp:532-534; SEGTP::get_phase, Transcriptome.h:80-89)
        const int pos5_t = t.plus ? seg_s : seg_e;
        int tk = seg_of_genomic(t, pos5_t);
        uint32_t phase;
        if (tk >= 0) {
            const uint32_t off = t.plus ? (uint32_t)(pos5_t - t.segs[tk].start) : (uint32_t)(t.segs[tk].end - pos5_t);
            phase = ((uint32_t)(int)B.seg_phase[B.tx_seg_off[tx] + tk] + (3u - off % 3u) % 3u) % 3u;
        } else {
            phase = (uint32_t)(int)B.seg_phase[B.tx_seg_off[tx] + (t.plus ? 0 : t.n - 1)];
        }
        if ((uint32_t)(seg_e + 1 - seg_s) < 3u) { status = ORFX_ST_SEGMENT_LT3; break; }
        // seed CDS: TX::build_cds + correct_chain_len on the query copy (orfanage.cpp:543-561)
        const int a0 = g2t(q, plus ? seg_s : seg_e);
        const int b0 = g2t(q, plus ? seg_e : seg_s);
        int a = a0;
        if (phase != 0) { a = a0 + (int)phase; if (a > q.L - 1) a = q.L - 1; }
        const int g_a = t2g(q, a);
        if (a > b0 || b0 - a + 1 < 3) {
            status = ORFX_ST_CDS_LT3;
            key_s = plus ? g_a : seg_s;
            key_e = plus ? seg_e : g_a;
            break;
        }
        const int b = a + ((b0 - a + 1) / 3) * 3 - 1;
        start = a; e = b;
        if (P.has_reference) {
            stopm = B.stop_mask + B.tx_mask_off[qx];
            startm = B.start_mask + B.tx_mask_off[qx];
            // adjust_stop + extend_to_stop (Transcriptome.cpp:182-201, 400-442): first in-frame stop at or after a,
            // searched through the 3' flank too (TX::extend_seq, cpp:342-399)
            const int p0 = nxt(stopm, a, q.L + E3);
            if (p0 == a) { status = ORFX_ST_NOT_RESCUED; break; }
            const bool has_stop = p0 >= 0;
            e = has_stop ? p0 + 2 : b;
            if (has_stop && P.overhang > 0) {   // stop inside the flank: the terminal exon grows up to it (cpp:418-429)
                const int g3 = max(p0 + 2 - q.L + 1, 0);
                if (g3 > 0) tx_grow(q, plus ? 0 : g3, plus ? g3 : 0);
                // ... but a base past the end of the chromosome is not there when the sequence is loaded again (cpp:437,
                // Bundle::get_nts): the last codon comes back incomplete, '\0', and the row ends as "missing_stop"
                stop_in_virtual = g3 > 0 && virt3;
            }
            // extend_to_start(TX*, non_aug) (Transcriptome.cpp:496-620)
            const bool m_at_a = mbit(startm, a);
            bool found_start = false;
            bool searched = false;   // the walk over the upstream in-frame ATGs ran and found none
            int prev_stop = -1, lb = 0;
            if ((m_at_a || P.non_aug) && g_a == t5) found_start = true;
            if (!found_start && P.non_aug) {
                const int ref = g2t(q, t5);
                if (ref >= 0 && ref % 3 == a % 3) {
                    bool stop_between = false;
                    for (int i = a;; i -= 3) {   // SEQ::find_inframe_codon('.',';',...,backward,first), common.hpp:420-450
                        if (mbit(stopm, i)) { stop_between = true; break; }
                        if (i < ref) break;
                        if (i < 3) break;
                    }
                    if (!stop_between) {
                        if (ref < a) { if (dot) e = min(e + (a - ref), q.L - 1); else start = ref; }
                        found_start = true;
                    }
                }
            }
            if (!found_start) {
                // all in-frame ATG walking back from a until a stop codon or the transcript start
                prev_stop = a >= 3 ? prv(stopm, a - 3) : -1;
                lb = prev_stop >= 0 ? prev_stop + 3 : a % 3;
                int far = nxt(startm, lb, a + 1);
                if (far >= 0) {
                    const int ref = g2t(q, t5);
                    if (ref >= lb && ref <= a && ref % 3 == a % 3 && mbit(startm, ref)) {
                        if (dot) e = min(e + (a - ref), q.L - 1); else start = ref;
                    } else if (!dot) start = guided_start(q, t, startm, far, far, a, e, lb, o);
                    else {
                        // strand '.': the chain [a, e] grows by a - far bases at its 3' end, is intersected with the template
                        // from `a`, and loses tb - far bases at its 5' end (trim_to_pos, h:483-494: a start past the chain's
                        // end lands on its last base); SEQ takes the chosen ATG as its start and keeps its end (cpp:611-613)
                        const int e2 = min(e + (a - far), q.L - 1);
                        const int tb = guided_start(q, t, startm, far, a, a, e2, lb, o);
                        sq_start = tb; sq_e = e;
                        start = min(a + (tb - far), e2); e = e2;
                    }
                } else searched = true;
            }
            // no upstream ATG: extend_to_start(-1, overhang, spliced) (cpp:446-492) walks on into the 5' flank — nearest
            // in-frame ATG before a stop.  Without a flank it finds nothing either.
            if (searched && P.overhang > 0 && prev_stop < 0 && lb - 3 + o >= 0) {
                const int top = lb - 3 + o;   // mask coordinates from here
                const int m5 = prev_inframe(startm, top), s5 = prev_inframe(stopm, top);
                if (m5 >= 0 && m5 > s5) {
                    // as written: the exon grows by overhang - position in the flanked sequence, not by the distance
                    // to the old 5' end (cpp:461-474); the flank may be shorter than --overhang
                    const int g5 = max(P.overhang - m5, 0);
                    tx_grow(q, plus ? g5 : 0, plus ? 0 : g5);
                    if (dot) { start += g5; e = min(e + g5 + (start - g5 + o - m5), q.L - 1); o -= g5; }   // a + E5 - m5 bases, at the 3' end
                    else { start = m5 - o + g5; e += g5; o -= g5; }
                }
            }
            // adjust_start (Transcriptome.cpp:202-227) unless --non_aug kept the template's start
            const bool at_t5 = t2g(q, start) == t5;
            if (sq_start < 0 && !(P.non_aug && at_t5) && !mbit(startm, start)) {   // (sq_start >= 0: SEQ starts with the chosen ATG)
                const int m = nxt(startm, start, e - 1);
                if (m >= 0 && m + 2 <= e) start = m;
            }
        }
        const int gs = t2g(q, start), ge = t2g(q, e);
        const int cs = min(gs, ge), ce = max(gs, ge);
        if (P.has_reference) {
            key_s = cs; key_e = ce;
            // get_aa(): SEQ::translate of SEQ's own bounds -- the chain's, except after the strand '.' guided choice; the
            // last amino acid is the last whole codon (a length that is no multiple of 3 only arises on strand '.')
            const int s0 = sq_start >= 0 ? sq_start : start, e0 = sq_start >= 0 ? sq_e : e;
            const int ncod = (e0 - s0 + 1) / 3;
            const bool ends_in_stop = ncod >= 1 && mbit(stopm, s0 + 3 * (ncod - 1)) && !stop_in_virtual;
            const bool front_m = mbit(startm, s0);
            seq_0 = s0 + o; seq_n = ncod;
            const bool at_t5 = (plus ? cs : ce) == t5;
            if (!ends_in_stop) { status = ORFX_ST_MISSING_STOP; break; }
            if ((!P.non_aug && !front_m) || (P.non_aug && !front_m && !at_t5)) { status = ORFX_ST_MISSING_START; break; }
        }
        qlen = e - start + 1;
        if (qlen < P.cds_minlen) { status = ORFX_ST_MINLEN; qlen = 0; break; }

        // ---- TX::score (Transcriptome.cpp:699-762) ----
        aln_0 = seq_0; aln_n = seq_n;
        tlen = t.L;
        {
            const ChainScore cs_ = score_chains<WARP>(q, t, cs, ce, qlen, plus, lane);
            ulen = cs_.ulen; extra = cs_.extra; missing = cs_.missing; match = cs_.match; inframe = cs_.inframe; outframe = cs_.outframe;
        }
        lpi = pct(qlen, ulen); ilpi = pct(inframe, ulen); mlpi = pct(match, ulen);
        query_start = plus ? cs - vseg(q, 0).start : vseg(q, q.n - 1).end - ce;
        start_match = (B.tx_strand[qx] == B.tx_strand[tx]) && ((plus ? cs : ce) == t5);
        pass = 1;
        if (lpi < (float)P.lpi || ilpi < (float)P.ilpi || mlpi < (float)P.mlpi) pass = 0;
        // with a reference the row's query copy already carries the rescued coordinates; without one
        // they are stored only for rows that survive the duplicate check (orfanage.cpp:577 vs 634)
        key_s = cs; key_e = ce;
    } while (false);

    if (WARP && lane != 0) return;
    R.status[row] = (uint8_t)status;
    R.cds_start[row] = key_s; R.cds_end[row] = key_e;
    R.qlen[row] = qlen; R.tlen[row] = tlen; R.ulen[row] = ulen; R.query_start[row] = query_start;
    R.extra[row] = extra; R.missing[row] = missing; R.match[row] = match; R.outframe[row] = outframe; R.inframe[row] = inframe;
    R.lpi[row] = lpi; R.ilpi[row] = ilpi; R.mlpi[row] = mlpi;
    R.pass[row] = (uint8_t)pass; R.start_match[row] = (uint8_t)start_match;
    R.n_dup[row] = 0;
    if (R.ext_lo) { R.ext_lo[row] = q.xlo; R.ext_hi[row] = q.xhi; }
    if (R.aln_qn) { R.aln_q0[row] = aln_0; R.aln_qn[row] = aln_n; }
}

// 8 CTAs per SM = 64 registers (16 bytes of spill): 0.237 ms against 0.245 ms at the 72 the compiler takes unasked, 0.263 ms at 78
__global__ void __launch_bounds__(128, 8) rescue_score_kernel(DevBatch B, DevRows R, orfx_params P) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= R.n_rows) return;
    const int qx = R.q_tx[row];
    if (B.bundle_need[B.tx_bundle[qx]] & BUNDLE_HEAVY) return;   // done by heavy_fill_rescue_kernel
    const TxView q = tx_view(B, qx), t = tx_view(B, R.t_tx[row]);
    process_row<false>(B, R, P, row, q, t);
}

// ------------------------------------------------------------------------------------------------
// heavy bundles: one CTA per bundle, the bundle's chains staged in shared memory
// ------------------------------------------------------------------------------------------------
constexpr int HB_THREADS = 512;   // x 2 CTAs per SM
// HB_MAX_SEGS / HB_MAX_PAIRS (orfx_device.cuh): segments of one bundle that fit the stage (8 B + 4 B each: 108 KB, two CTAs
// per SM) and (query, transcript slot) pairs whose row counts fit it

namespace {
struct HeavyBundle {
    int b, x0, x1, ntx, qbase, nq, seg_lo, ns;
    bool staged;
    const orfx_seg* segs;    // accepts absolute segment indices
    const int32_t* cum;
};
// loads bundle h of the heavy list; stages its segments (and prefixes when `with_cum`) when they fit
__device__ HeavyBundle hb_load(const DevBatch& B, int b, orfx_seg* s_seg, int32_t* s_cum, bool with_cum) {
    HeavyBundle H;
    H.b = b;
    H.x0 = B.bundle_tx_off[b]; H.x1 = B.bundle_tx_off[b + 1]; H.ntx = H.x1 - H.x0;
    H.qbase = (int)B.tx_q_off[H.x0]; H.nq = (int)B.tx_q_off[H.x1] - H.qbase;
    H.seg_lo = B.tx_seg_off[H.x0]; H.ns = B.tx_seg_off[H.x1] - H.seg_lo;
    H.staged = H.ns <= B.hb_max_segs;   // HB_MAX_SEGS unless a test lowered it (orfx_debug_set)
    if (H.staged) {
        for (int i = threadIdx.x; i < H.ns; i += blockDim.x) {
            s_seg[i] = B.segs[H.seg_lo + i];
            if (with_cum) s_cum[i] = B.seg_cum[H.seg_lo + i];
        }
        H.segs = s_seg - H.seg_lo; H.cum = s_cum - H.seg_lo;
    } else {
        H.segs = B.segs; H.cum = B.seg_cum;
    }
    __syncthreads();
    return H;
}
__device__ __forceinline__ TxView hb_view(const DevBatch& B, const HeavyBundle& H, int tx) {
    TxView v;
    const int s0 = B.tx_seg_off[tx];
    v.segs = H.segs + s0; v.cum = H.cum + s0;
    v.n = B.tx_seg_off[tx + 1] - s0; v.L = B.tx_len[tx]; v.plus = B.tx_strand[tx] == '+'; v.xlo = v.xhi = 0;
    return v;
}
}  // namespace

// Work item of a heavy kernel = (heavy bundle, slice k of S): a slice takes every
// S-th query of the bundle (pair kernels) or every S-th group of rows (row kernel), so that a
// titin-class bundle is spread over several SMs.  Every CTA stages the whole bundle's chains.
constexpr int HB_MAX_SLICES = 16;

// Work plan of the heavy kernels.  The buffer behind heavy_list holds [list: n_bundles ints][scratch: n_bundles int4]
// [work: HB_MAX_SLICES * n_bundles int4]; heavy_count[0] = heavy bundles, heavy_count[1] = work items.
// A work item is (bundle, slice k, slices S of that bundle): slice k takes every S-th query (pair kernels) or every S-th
// group of rows (row kernel).  S grows with the bundle's queries (a third of them, 1..8: fewer queries per slice than
// that leaves warps idle, measured with 16 slices for all; the largest loci get one slice per 16 K of cost, up to 16),
// and the items are listed by descending cost (queries x
// segments) so that the static round-robin over the CTAs starts with the largest loci: without it one 8 500-segment locus
// made one rank's step 0.24 ms longer than another's.
__device__ __forceinline__ int4* heavy_scratch(const DevBatch& B, const int32_t* heavy_list) {
    return reinterpret_cast<int4*>(const_cast<int32_t*>(heavy_list) + ((B.n_bundles + 3) & ~3));
}
__device__ __forceinline__ int4* heavy_work(const DevBatch& B, const int32_t* heavy_list) { return heavy_scratch(B, heavy_list) + B.n_bundles; }

__global__ void __launch_bounds__(1024) heavy_plan_kernel(DevBatch B, int32_t* heavy_list, unsigned* heavy_count) {
    int4* scr = heavy_scratch(B, heavy_list);
    int4* work = heavy_work(B, heavy_list);
    const int n = (int)heavy_count[0];
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int b = heavy_list[i];
        const int x0 = B.bundle_tx_off[b], x1 = B.bundle_tx_off[b + 1];
        const int nq = (int)(B.tx_q_off[x1] - B.tx_q_off[x0]);
        const int ns = B.tx_seg_off[x1] - B.tx_seg_off[x0];
        const int64_t cost = (int64_t)nq * ns;
        const int by_cost = (int)min(cost >> 14, (int64_t)HB_MAX_SLICES);   // the largest loci: more slices than a third of their queries
        scr[i] = make_int4((int)(cost & 0x7FFFFFFF), (int)(cost >> 31), max(1, min(max((nq + 2) / 3, by_cost), min(HB_MAX_SLICES, nq))), b);
    }
    __syncthreads();
    const bool sorted = n <= 4096;   // rank sort, quadratic: beyond this the list keeps its order
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        int rank = i;
        if (sorted) {
            const int4 me = scr[i];
            rank = 0;
            for (int j = 0; j < n; j++) {
                const int4 o = scr[j];
                const bool before = o.y != me.y ? o.y > me.y : (o.x != me.x ? o.x > me.x : o.w < me.w);
                rank += before ? 1 : 0;
            }
        }
        work[i].x = rank;   // parked in the (still unused) work area: the scratch records are being read by other threads
    }
    __syncthreads();
    // first work item of every bundle: exclusive prefix of the slice counts in rank order (in shared memory while it fits)
    __shared__ int s_off[4096];
    if (sorted) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) s_off[work[i].x] = scr[i].z;
        __syncthreads();
        if (threadIdx.x == 0) {
            int off = 0;
            for (int r = 0; r < n; r++) { const int v = s_off[r]; s_off[r] = off; off += v; }
            heavy_count[1] = (unsigned)off;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) scr[i].y = s_off[work[i].x];
    } else {
        if (threadIdx.x == 0) {   // list order
            int off = 0;
            for (int i = 0; i < n; i++) { scr[i].y = off; off += scr[i].z; }
            heavy_count[1] = (unsigned)off;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int4 me = scr[i];
        for (int k = 0; k < me.z; k++) work[me.y + k] = make_int4(me.w, k, me.z, 0);
    }
}

// rows of every query of the heavy bundles, one warp per (query, transcript slot) pair
__global__ void __launch_bounds__(HB_THREADS, 2) heavy_count_kernel(DevBatch B, int64_t* cnt, const int32_t* heavy_list, const unsigned* heavy_count) {
    extern __shared__ __align__(16) unsigned char hb_smem[];
    orfx_seg* s_seg = reinterpret_cast<orfx_seg*>(hb_smem);
    const unsigned n = heavy_count[1];
    const int4* work = heavy_work(B, heavy_list);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    DevRows dummy{};
    for (unsigned w = blockIdx.x; w < n; w += gridDim.x) {
        const int4 item = work[w];
        const int k = item.y, HB_SLICES = item.z;
        const HeavyBundle H = hb_load(B, item.x, s_seg, nullptr, false);
        const int nq_slice = (H.nq - k + HB_SLICES - 1) / HB_SLICES;
        const int npairs = nq_slice * H.ntx;
        for (int p = warp; p < npairs; p += nwarps) {
            const int ql = k + HB_SLICES * (p / H.ntx), t = H.x0 + p % H.ntx;
            const int qx = B.q_list[H.qbase + ql];
            const int qstart = H.segs[B.tx_seg_off[qx]].start, qend = H.segs[B.tx_seg_off[qx + 1] - 1].end;
            if (!pair_live(B, t, H.x1, qstart, qend)) continue;
            const int r = pair_rows_warp<false>(B, H.segs, dummy, qx, t, 0, lane);
            if (lane == 0) atomicAdd(reinterpret_cast<unsigned long long*>(cnt + H.qbase + ql), (unsigned long long)r);   // zeroed by launch_count_rows
        }
        __syncthreads();
    }
}

// candidate rows of the heavy bundles, in the reference's order; one warp per pair
__global__ void __launch_bounds__(HB_THREADS, 2) heavy_fill_kernel(DevBatch B, DevRows R, const int32_t* heavy_list, const unsigned* heavy_count) {
    extern __shared__ __align__(16) unsigned char hb_smem[];
    orfx_seg* s_seg = reinterpret_cast<orfx_seg*>(hb_smem);
    int32_t* s_n = reinterpret_cast<int32_t*>(hb_smem + (size_t)HB_MAX_SEGS * sizeof(orfx_seg));   // rows of a pair: not bounded by 65 535 when the chains are walked in global memory
    const unsigned n = heavy_count[1];
    const int4* work = heavy_work(B, heavy_list);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (unsigned w = blockIdx.x; w < n; w += gridDim.x) {
        const int4 item = work[w];
        const int k = item.y, HB_SLICES = item.z;
        const HeavyBundle H = hb_load(B, item.x, s_seg, nullptr, false);
        const int nq_slice = (H.nq - k + HB_SLICES - 1) / HB_SLICES;
        const int64_t npairs = (int64_t)nq_slice * H.ntx;
        if (npairs <= B.hb_max_pairs) {
            // rows of every pair, then each pair fills at (rows of the query's earlier pairs)
            for (int p = warp; p < (int)npairs; p += nwarps) {
                const int ql = k + HB_SLICES * (p / H.ntx), t = H.x0 + p % H.ntx;
                const int qx = B.q_list[H.qbase + ql];
                const int qstart = H.segs[B.tx_seg_off[qx]].start, qend = H.segs[B.tx_seg_off[qx + 1] - 1].end;
                const int r = pair_live(B, t, H.x1, qstart, qend) ? pair_rows_warp<false>(B, H.segs, R, qx, t, 0, lane) : 0;
                if (lane == 0) s_n[p] = r;
            }
            __syncthreads();
            for (int p = warp; p < (int)npairs; p += nwarps) {
                if (s_n[p] == 0) continue;   // a live pair has at least its "no-overlap" row
                const int ql = k + HB_SLICES * (p / H.ntx), tl = p % H.ntx;
                int before = 0;
                for (int j = p - tl + lane; j < p; j += 32) before += s_n[j];
                const int64_t out = B.q_row_off[H.qbase + ql] + warp_sum(before);
                pair_rows_warp<true>(B, H.segs, R, B.q_list[H.qbase + ql], H.x0 + tl, out, lane);
            }
        } else {
            for (int qs = warp; qs < nq_slice; qs += nwarps) {
                const int ql = k + HB_SLICES * qs;
                const int qx = B.q_list[H.qbase + ql];
                const int qstart = H.segs[B.tx_seg_off[qx]].start, qend = H.segs[B.tx_seg_off[qx + 1] - 1].end;
                int64_t out = B.q_row_off[H.qbase + ql];
                for (int t = H.x0; t < H.x1; t++)
                    if (pair_live(B, t, H.x1, qstart, qend)) out += pair_rows_warp<true>(B, H.segs, R, qx, t, out, lane);
            }
        }
        __syncthreads();
    }
}

// rescue + score of the heavy bundles' rows against the staged chains; one warp per row
__global__ void __launch_bounds__(HB_THREADS, 2) heavy_rescue_kernel(DevBatch B, DevRows R, orfx_params P, const int32_t* heavy_list,
                                                                    const unsigned* heavy_count) {
    extern __shared__ __align__(16) unsigned char hb_smem[];
    orfx_seg* s_seg = reinterpret_cast<orfx_seg*>(hb_smem);
    int32_t* s_cum = reinterpret_cast<int32_t*>(hb_smem + (size_t)HB_MAX_SEGS * sizeof(orfx_seg));
    const unsigned n = heavy_count[1];
    const int4* work = heavy_work(B, heavy_list);
    const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (unsigned w = blockIdx.x; w < n; w += gridDim.x) {
        const int4 item = work[w];
        const int k = item.y, HB_SLICES = item.z;
        const HeavyBundle H = hb_load(B, item.x, s_seg, s_cum, true);
        const int64_t r0 = B.q_row_off[H.qbase], r1 = B.q_row_off[H.qbase + H.nq];
        for (int64_t row = r0 + (int64_t)k * nwarps + warp; row < r1; row += (int64_t)HB_SLICES * nwarps)
            process_row<true>(B, R, P, row, hb_view(B, H, R.q_tx[row]), hb_view(B, H, R.t_tx[row]));
        __syncthreads();
    }
}

constexpr size_t HB_SMEM_PAIRS = (size_t)HB_MAX_SEGS * sizeof(orfx_seg) + (size_t)HB_MAX_PAIRS * sizeof(int32_t);    // segments + row counts
constexpr size_t HB_SMEM_ROWS = (size_t)HB_MAX_SEGS * (sizeof(orfx_seg) + sizeof(int32_t));                          // segments + prefixes

int launch_heavy_count(const DevBatch& B, int64_t* cnt, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm, cudaStream_t s) {
    if (B.n_bundles == 0) return 0;
    heavy_plan_kernel<<<1, 1024, 0, s>>>(B, const_cast<int32_t*>(heavy_list), const_cast<unsigned*>(heavy_count));
    cudaFuncSetAttribute(heavy_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HB_SMEM_PAIRS);   // per device
    heavy_count_kernel<<<(unsigned)std::min<int64_t>(4 * n_sm, (int64_t)B.n_bundles * HB_MAX_SLICES), HB_THREADS, HB_SMEM_PAIRS, s>>>(B, cnt, heavy_list, heavy_count);
    return 2;
}
// candidate rows of the heavy bundles: needs the chains only (runs while the sequence is still being uploaded)
int launch_heavy_fill(const DevBatch& B, const DevRows& R, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm, cudaStream_t s) {
    if (B.n_bundles == 0 || R.n_rows == 0) return 0;
    const unsigned grid = (unsigned)std::min<int64_t>(4 * n_sm, (int64_t)B.n_bundles * HB_MAX_SLICES);
    cudaFuncSetAttribute(heavy_fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HB_SMEM_PAIRS);
    heavy_fill_kernel<<<grid, HB_THREADS, HB_SMEM_PAIRS, s>>>(B, R, heavy_list, heavy_count);
    return 1;
}
// rescue + score of those rows: needs the codon-scan masks
int launch_heavy_rescue(const DevBatch& B, const DevRows& R, const orfx_params& P, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm,
                        cudaStream_t s) {
    if (B.n_bundles == 0 || R.n_rows == 0) return 0;
    // 2 CTAs per SM take every register of the SM: beside rescue_score_kernel the two kernels end up one after the other
    // (0.51 ms together).  Smaller grids that leave room for the other kernel were measured and are slower (1 / 2 / 3 CTA
    // waves per SM: 0.81 / 0.60 / 0.54 ms for the pair).
    const unsigned grid = (unsigned)std::min<int64_t>(4 * n_sm, (int64_t)B.n_bundles * HB_MAX_SLICES);
    cudaFuncSetAttribute(heavy_rescue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HB_SMEM_ROWS);
    heavy_rescue_kernel<<<grid, HB_THREADS, HB_SMEM_ROWS, s>>>(B, R, P, heavy_list, heavy_count);
    return 1;
}

// ------------------------------------------------------------------------------------------------
// template cleaning (--cleant [--rescue]): Transcriptome::clean_cds, Transcriptome.cpp:1063-1100
// ------------------------------------------------------------------------------------------------
// One thread per transcript of a cleaning batch (every transcript there is a coding template packed as
// ORFX_TX_QUERY: segs = its exon chain, tx_cds_start / tx_cds_end = its CDS bounds after build_cds +
// correct_chain_len).  The checks of clean_cds and, with `rescue`, TX::rescue_cds(non_aug, 0, false,
// nullptr) (cpp:641-677: adjust_stop, extend_to_stop, unguided extend_to_start, adjust_start) are bit
// scans over the transcript's codon-scan masks, as in process_row.
__global__ void __launch_bounds__(128) clean_templates_kernel(DevBatch B, int n_queries, int rescue, int non_aug, uint8_t* keep, int32_t* out_start,
                                                              int32_t* out_end) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= n_queries) return;
    const int qx = B.q_list[qi];
    const TxView q = tx_view(B, qx);
    const int gs = B.tx_cds_start[qx], ge = B.tx_cds_end[qx];
    int ok = 0, cs = 0, ce = 0;
    do {
        if (q.n == 0 || (gs == 0 && ge == 0)) break;
        const int a = g2t(q, q.plus ? gs : ge), b = g2t(q, q.plus ? ge : gs);
        if (a < 0 || b < 0 || b < a) break;                    // CDS bounds outside the exon chain: not a usable template
        const uint32_t* stopm = B.stop_mask + B.tx_mask_off[qx];
        const uint32_t* startm = B.start_mask + B.tx_mask_off[qx];
        // --overhang: the masks cover [5' flank][transcript][3' flank]; transcript position p is mask bit p + o.  Only the
        // CDS bounds leave this kernel, and t2g extrapolates past the terminal exons, so the chain itself is not grown.
        int o = 0, E3 = 0;
        bool virt3 = false;
        if (B.tx_ext_lo) {
            const int lo = B.tx_ext_lo[qx], hi = B.tx_ext_hi[qx]; o = q.plus ? lo : hi; E3 = q.plus ? hi : lo;
            if (B.tx_ext_flags) virt3 = (B.tx_ext_flags[qx] & (q.plus ? ORFX_EXT_HI_VIRTUAL : ORFX_EXT_LO_VIRTUAL)) != 0;
        }
        const bool dot = !q.plus && B.tx_strand[qx] != '-';      // see process_row: 5' extensions of the chain land on its 3' end
        int start = a, e = b;
        if (b - a + 1 < 3) break;                              // rescue_cds: cds_nt_len < 3; no rescue: no codon to test
        if (rescue) {
            int p0 = next_inframe(stopm, a + o, q.L + E3 + o);  // adjust_stop + extend_to_stop (3' flank included)
            if (p0 >= 0) p0 -= o;
            if (p0 == a) break;                                // nothing left before the first stop: CDS removed
            if (p0 >= 0) e = p0 + 2;
            if (virt3 && e >= q.L) break;                      // the stop needs a base past the chromosome's end: it is gone after the reload
            int Lg = max(q.L, e + 1);                          // exon chain after the growth towards the stop (cpp:418-429)
            if (!non_aug && !mask_bit(startm, a + o) && a - 3 + o >= 0) {  // extend_to_start(-1, ...): nearest upstream ATG before a stop, 5' flank included
                const int prev_stop = prev_inframe(stopm, a - 3 + o);
                const int m = prev_inframe(startm, a - 3 + o);
                if (m >= 0 && m > prev_stop) {
                    if (dot) e = min(e + (a - (m - o)), Lg - 1); else start = m - o;
                }
            }
            if (!mask_bit(startm, start + o)) {                // adjust_start: first ATG inside the CDS
                int m = next_inframe(startm, start + o, e - 1 + o);
                if (m >= 0 && m - o + 2 <= e) start = m - o;
            }
        }
        const int naa = (e - start + 1) / 3;
        const int last = start + 3 * (naa - 1);
        if (!non_aug && !mask_bit(startm, start + o)) break;   // start codon must be M unless --non_aug
        if (!mask_bit(stopm, last + o)) break;                 // last codon must be a stop
        if ((e - start + 1) % 3 != 0) break;
        const int g0 = t2g(q, start), g1 = t2g(q, e);
        cs = min(g0, g1); ce = max(g0, g1);
        ok = 1;
    } while (false);
    keep[qi] = (uint8_t)ok; out_start[qi] = cs; out_end[qi] = ce;
}
int launch_clean_templates(const DevBatch& B, int n_queries, int rescue, int non_aug, uint8_t* keep, int32_t* out_start, int32_t* out_end, cudaStream_t s) {
    if (n_queries == 0) return 0;
    clean_templates_kernel<<<(n_queries + 127) / 128, 128, 0, s>>>(B, n_queries, rescue, non_aug, keep, out_start, out_end);
    return 1;
}

// ------------------------------------------------------------------------------------------------
// orfcompare: TX::score of every (query CDS, template CDS) pair of a bundle (orfcompare.cpp:76-160)
// ------------------------------------------------------------------------------------------------
// compare batch: queries and templates both carry their CDS chain as `segs`.  Pairs are laid out per query
// (q_row_off, filled by count_pairs_kernel + scan), templates in bundle order; no bounding-box test.
__global__ void __launch_bounds__(256) count_pairs_kernel(DevBatch B, int64_t* cnt) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= B.n_tx) return;
    const int64_t nq = B.tx_q_off[B.n_tx];
    int n = 0;
    if (qi < nq) {
        const int b = B.tx_bundle[B.q_list[qi]];
        for (int x = B.bundle_tx_off[b]; x < B.bundle_tx_off[b + 1]; x++) n += B.tx_kind[x] == ORFX_TX_TEMPLATE;
    }
    cnt[qi] = n;
}
__global__ void __launch_bounds__(128) score_pairs_kernel(DevBatch B, DevRows R, int n_queries) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= n_queries) return;
    const int qx = B.q_list[qi];
    const TxView q = tx_view(B, qx);
    const int b = B.tx_bundle[qx];
    int64_t row = B.q_row_off[qi];
    const int qlen = q.L;
    const int cs = q.n ? vseg(q, 0).start : 0, ce = q.n ? vseg(q, q.n - 1).end : 0;
    for (int x = B.bundle_tx_off[b]; x < B.bundle_tx_off[b + 1]; x++) {
        if (B.tx_kind[x] != ORFX_TX_TEMPLATE) continue;
        const TxView t = tx_view(B, x);
        ChainScore c = {0, 0, t.L, 0, 0, 0};
        if (q.n > 0) c = score_chains<false>(q, t, cs, ce, qlen, q.plus, 0);
        R.q_tx[row] = qx; R.t_tx[row] = x;
        R.qlen[row] = qlen; R.tlen[row] = t.L; R.ulen[row] = c.ulen;
        R.extra[row] = c.extra; R.missing[row] = c.missing; R.match[row] = c.match; R.inframe[row] = c.inframe; R.outframe[row] = c.outframe;
        R.lpi[row] = pct(qlen, c.ulen); R.ilpi[row] = pct(c.inframe, c.ulen); R.mlpi[row] = pct(c.match, c.ulen);
        row++;
    }
}
int launch_count_pairs(const DevBatch& B, int64_t* cnt, int64_t* scan_tmp, cudaStream_t s) {
    if (B.n_tx == 0) { cudaMemsetAsync(B.q_row_off, 0, sizeof(int64_t), s); return 0; }
    count_pairs_kernel<<<(B.n_tx + 255) / 256, 256, 0, s>>>(B, cnt);
    return 1 + launch_exclusive_scan(cnt, B.q_row_off, scan_tmp, B.n_tx, s);
}
int launch_score_pairs(const DevBatch& B, const DevRows& R, int n_queries, cudaStream_t s) {
    if (n_queries == 0) return 0;
    score_pairs_kernel<<<(n_queries + 127) / 128, 128, 0, s>>>(B, R, n_queries);
    return 1;
}

int launch_rescue_score(const DevBatch& B, const DevRows& R, const orfx_params& P, cudaStream_t s) {
    if (R.n_rows == 0) return 0;
    rescue_score_kernel<<<(unsigned)((R.n_rows + 127) / 128), 128, 0, s>>>(B, R, P);
    return 1;
}

}  // namespace orfx
