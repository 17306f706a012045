// C-ABI of the device library (include/orfx.h): context, buffers, stream orchestration.
// One context = one GPU, one stream.  No CPU fallback: every entry point that computes needs a
// CUDA device, and orfx_create fails without one.
#include "orfx_device.cuh"
#include "orfx_buffers.cuh"

#include <climits>
#include <nvtx3/nvToolsExt.h>   // header-only; ranges show up in Nsight timelines, cost nothing without a tool attached
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace orfx {
int launch_exclusive_scan(const int64_t* in, int64_t* out, int64_t* tmp, int64_t n, cudaStream_t s);
int64_t scan_tmp_elems(int64_t n);
int launch_prep(const DevBatch& B, int64_t* cnt_q, int64_t* cnt_words, int64_t* cnt_tiles, int64_t* scan_tmp, int32_t* heavy_list, unsigned* heavy_count,
                cudaStream_t s);
int launch_fill_tiles(const DevBatch& B, cudaStream_t s);
int launch_validate(const DevBatch& B, int has_seq, int allow_empty, int* flag, cudaStream_t s);
int launch_tx_bundle(const DevBatch& B, cudaStream_t s);
int launch_expand_dict(const DevBatch& B, const uint16_t* ref, const int32_t* bundle_dict_off, const orfx_seg* dict_segs, const int8_t* dict_phase,
                       const int32_t* dict_seq_off, int64_t n_dict, orfx_seg* segs, int8_t* phase, int32_t* sso, int* flag, cudaStream_t s);
int launch_count_rows(const DevBatch& B, int64_t* cnt, int64_t* scan_tmp, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm, cudaStream_t s,
                      cudaStream_t side, cudaEvent_t ev_fork, cudaEvent_t ev_join);
int launch_fill_rows(const DevBatch& B, const DevRows& R, int n_queries, cudaStream_t s);
int launch_rescue_score(const DevBatch& B, const DevRows& R, const orfx_params& P, cudaStream_t s);
int launch_heavy_fill(const DevBatch& B, const DevRows& R, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm, cudaStream_t s);
int launch_heavy_rescue(const DevBatch& B, const DevRows& R, const orfx_params& P, const int32_t* heavy_list, const unsigned* heavy_count, int n_sm,
                        cudaStream_t s);
int launch_select(const DevBatch& B, const DevRows& R, const orfx_params& P, int n_queries, orfx_best* best_out, int32_t* ord_scratch, int4* key_scratch,
                  cudaStream_t s);
void launch_codon_scan(const DevBatch& B, int64_t n_tiles, int which, cudaStream_t stream);
int launch_base_planes(const DevBatch& B, cudaStream_t stream);
int launch_clean_templates(const DevBatch& B, int n_queries, int rescue, int non_aug, uint8_t* keep, int32_t* out_start, int32_t* out_end, cudaStream_t s);
int launch_count_pairs(const DevBatch& B, int64_t* cnt, int64_t* scan_tmp, cudaStream_t s);
int launch_score_pairs(const DevBatch& B, const DevRows& R, int n_queries, cudaStream_t s);
int align_batch_device(AlnWork& w, cudaStream_t stream, int n_sm, const uint8_t* aa, const int64_t* off, int64_t n_seq, const int32_t* pair_q,
                       const int32_t* pair_t, int64_t n_pairs, int gapo, int gape, orfx_aln_out* out, double* cells, float* ms, std::string& err);
int align_pairs_static_device(int device, cudaStream_t stream, int n_sm, const uint8_t* d_aa, const int64_t* d_off, int64_t n_seq, const int32_t* d_pq,
                              const int32_t* d_pt, int64_t n_pairs, int gapo, int gape, const orfx_aln_out* d_out, std::string& err);
int align_pairs_device(AlnWork& w, cudaStream_t stream, int n_sm, const uint8_t* d_aa, int q_stride, int t_stride, const AlnPair* d_pairs, long long n,
                       int gapo, int gape, int32_t* const d_out[4], double* cells, float* ms, int* n_launches, std::string& err);
void build_translate_tables(TranslateTables& T);
int launch_aa_len(const DevBatch& B, int64_t* cnt, cudaStream_t s);
int launch_translate(const DevBatch& B, const TranslateTables* d_tables, int64_t total_positions, cudaStream_t s);
int launch_row_pairs(const DevBatch& B, const DevRows& R, AlnPair* pairs, cudaStream_t s);
int int_alu_peak_device(cudaStream_t stream, double* ops, std::string& err);
}  // namespace orfx

using namespace orfx;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(ORFX_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

enum { ST_PREP = 0, ST_COUNT, ST_PLANES, ST_SCAN, ST_FILL, ST_HEAVY, ST_RESCUE, ST_SELECT, ST_TRANSLATE, ST_ALIGN, ST_N };
const char* STAGE_NAME[ST_N] = {"prep", "count_rows", "base_planes", "codon_scan", "fill_rows", "heavy_bundles", "rescue_score", "select", "translate", "align"};

}  // namespace

struct orfx_ctx {
    int device = 0;
    int n_sm = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // sequence bytes go up here while the chain kernels already run on `stream`
    cudaStream_t side_stream = nullptr;   // the heavy bundles' kernels run here, beside the per-query / per-row kernels on `stream`
    cudaEvent_t ev_fork[3] = {nullptr, nullptr, nullptr}, ev_join[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t ev_chains = nullptr, ev_seq = nullptr;
    bool seq_in_flight = false;
    orfx_params P;
    // inputs
    DBuf d_sso;
    DBuf d_seq, d_seq_off, d_bstart, d_btx_off, d_kind, d_strand, d_seg_off, d_cds_s, d_cds_e, d_tid, d_segs, d_phase;
    // derived
    DBuf d_seg_cum, d_tx_len, d_tx_bundle, d_mask_off, d_tile_off, d_q_off, d_q_list, d_row_off, d_tile_tx, d_stop, d_start;
    DBuf d_need, d_planes, d_heavy, d_heavy_count;
    DBuf d_cnt_q, d_cnt_words, d_cnt_tiles, d_cnt_rows, d_scan_tmp, d_ord;
    DBuf d_clean_keep, d_clean_s, d_clean_e, d_keys;
    DBuf d_xlo, d_xhi, d_xfl;   // --overhang flanks
    DBuf d_flag, d_exc_pos, d_exc_code;
    DBuf d_ref, d_bdo, d_dsegs, d_dphase, d_dsso;   // segment dictionary (orfx_batch::seg_ref)
    bool blocking_sync = false;                      // orfx_debug_set("blocking_sync"): sleep, not spin, while waiting for the stream
    cudaEvent_t ev_block = nullptr;
    size_t upload_bytes = 0;
    // alignment stage (params.align_enabled)
    AlnWork aln;
    DBuf d_aa_off, d_aa_pos, d_cnt_aa, d_tables, d_row_pairs, r_aln[6];
    HBuf h_aln[4];
    double aln_cells = 0;
    float ms_align = 0;
    int64_t n_aa = 0;
    int hb_max_segs = HB_MAX_SEGS, hb_max_pairs = HB_MAX_PAIRS;   // orfx_debug_set
    bool serialize = false;            // orfx_debug_set "serialize": every kernel on the one stream (per-kernel times with nothing beside them)
    bool allow_empty_chains = false;   // compare batches (orfx_score_pairs): a query without a CDS is legal there
    // rows
    DBuf r_i32[16], r_f32[3], r_u8[3], r_order, d_best, r_ext[2];
    HBuf h_totals, h_best, h_i32[16], h_f32[3], h_u8[3], h_order, h_ext[2];
    DevBatch B;
    DevRows R;
    bool resident = false, pending = false, have_results = false;
    int64_t n_queries = 0, n_rows = 0, n_tiles = 0, n_words = 0;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t st_ev[ST_N][2];
    cudaEvent_t run_ev[2] = {nullptr, nullptr};
    float run_ms = 0;
    float st_ms[ST_N];
    int n_launches = 0;
};

namespace {

// Waiting for the stream from the host.  Default: cudaStreamSynchronize (the driver spins, lowest latency).  With
// blocking_sync the thread sleeps on an event created with cudaEventBlockingSync: what a process wants when more threads wait
// for GPUs than the host has cores (eight ranks with three contexts each on a 16-core box), where spinning waiters keep the
// threads that have work to enqueue off the cores.
cudaError_t sync_stream(orfx_ctx* c, cudaStream_t s) {
    if (!c->blocking_sync || !c->ev_block) return cudaStreamSynchronize(s);
    cudaError_t e = cudaEventRecord(c->ev_block, s);
    if (e != cudaSuccess) return e;
    return cudaEventSynchronize(c->ev_block);
}

// overlap = true (orfx_submit): the chain arrays go first on the launch stream, the sequence bytes (two thirds
// of the upload) follow on the copy stream; prep / candidate-row kernels only need the chains and start while
// the sequence is still on the bus.  run_path waits for ev_seq before the first kernel that reads `seq`.
int upload_batch(orfx_ctx* c, const orfx_batch* b, bool overlap = false) {
    if (!b) return fail(ORFX_E_INVALID, "null batch");
    if (b->n_bundles < 0 || b->n_tx < 0 || b->n_segs < 0) return fail(ORFX_E_INVALID, "negative batch size");
    if (c->P.has_reference && b->seq_bytes > 0 && !b->seq) return fail(ORFX_E_INVALID, "params.has_reference set but batch.seq is NULL");
    if ((b->n_bundles > 0 || b->n_tx > 0) && (!b->bundle_tx_off || !b->tx_seg_off || !b->tx_kind || !b->tx_strand || !b->tx_cds_start || !b->tx_cds_end || !b->tx_tid_class ||
                                              !b->bundle_start || (b->n_segs > 0 && !b->seg_ref && (!b->segs || !b->seg_phase)) || (c->P.has_reference && b->seq_bytes > 0 && !b->seq_off)))
        return fail(ORFX_E_INVALID, "null array in a non-empty batch");
    cudaStream_t s = c->stream;
    const size_t nb = (size_t)b->n_bundles, nt = (size_t)b->n_tx, ns = (size_t)b->n_segs;
    const size_t seqb = c->P.has_reference ? (size_t)b->seq_bytes : 0;                    // bases
    if (b->seq_format != ORFX_SEQ_RAW && b->seq_format != ORFX_SEQ_NT4 && b->seq_format != ORFX_SEQ_NT2) return fail(ORFX_E_INVALID, "unknown batch.seq_format");
    const bool dict = b->seg_ref != nullptr && b->n_segs > 0;
    if (dict && (!b->bundle_dict_off || b->n_dict <= 0 || !b->dict_segs || !b->dict_phase)) return fail(ORFX_E_INVALID, "batch.seg_ref without a dictionary (bundle_dict_off, dict_segs, dict_phase, n_dict)");
    const int32_t* const host_sso = dict ? b->dict_seq_off : b->seg_seq_off;   // non-NULL <=> islands layout
    if (b->seq_format == ORFX_SEQ_NT2 && seqb) {
        if (!host_sso) return fail(ORFX_E_INVALID, "ORFX_SEQ_NT2 is defined for the islands layout (seg_seq_off)");
        if (b->n_seq_exc < 0 || (b->n_seq_exc > 0 && (!b->seq_exc_pos || !b->seq_exc_code))) return fail(ORFX_E_INVALID, "ORFX_SEQ_NT2: bad exception list");
    }
    const size_t seq_store = b->seq_format == ORFX_SEQ_NT4 ? (seqb + 1) / 2 : (b->seq_format == ORFX_SEQ_NT2 ? (seqb + 3) / 4 : seqb);   // bytes behind batch.seq
    CU(c->d_seq.ensure(seq_store + 64));
    CU(c->d_seq_off.ensure((nb + 1) * 8));
    CU(c->d_bstart.ensure(nb * 4 + 4));
    CU(c->d_btx_off.ensure((nb + 1) * 4));
    CU(c->d_kind.ensure(nt + 4));
    CU(c->d_strand.ensure(nt + 4));
    CU(c->d_seg_off.ensure((nt + 1) * 4));
    CU(c->d_cds_s.ensure(nt * 4 + 4));
    CU(c->d_cds_e.ensure(nt * 4 + 4));
    CU(c->d_tid.ensure(nt * 4 + 4));
    CU(c->d_segs.ensure(ns * 8 + 8));
    CU(c->d_phase.ensure(ns + 4));
    c->upload_bytes = 0;
    auto cp = [&](DBuf& d, const void* src, size_t bytes) -> cudaError_t {
        if (bytes == 0) return cudaSuccess;
        c->upload_bytes += bytes;
        return cudaMemcpyAsync(d.p, src, bytes, cudaMemcpyHostToDevice, s);
    };
    if (seqb && !overlap) {
        CU(cp(c->d_seq, b->seq, seq_store));
        CU(cudaMemsetAsync((char*)c->d_seq.p + seq_store, 0, 64, s));   // the last aligned 16-byte chunk may read past the end
    }
    CU(cp(c->d_seq_off, b->seq_off, (nb + 1) * 8));
    CU(cp(c->d_bstart, b->bundle_start, nb * 4));
    CU(cp(c->d_btx_off, b->bundle_tx_off, (nb + 1) * 4));
    CU(cp(c->d_kind, b->tx_kind, nt));
    CU(cp(c->d_strand, b->tx_strand, nt));
    CU(cp(c->d_seg_off, b->tx_seg_off, (nt + 1) * 4));
    CU(cp(c->d_cds_s, b->tx_cds_start, nt * 4));
    CU(cp(c->d_cds_e, b->tx_cds_end, nt * 4));
    CU(cp(c->d_tid, b->tx_tid_class, nt * 4));
    const size_t nd = dict ? (size_t)b->n_dict : 0;
    if (!dict) {
        CU(cp(c->d_segs, b->segs, ns * 8));
        CU(cp(c->d_phase, b->seg_phase, ns));
    } else {   // 2 bytes per segment + the distinct records; expanded on the device below
        CU(c->d_ref.ensure(ns * 2 + 4)); CU(c->d_bdo.ensure((nb + 1) * 4)); CU(c->d_dsegs.ensure(nd * 8 + 8)); CU(c->d_dphase.ensure(nd + 4));
        CU(cp(c->d_ref, b->seg_ref, ns * 2)); CU(cp(c->d_bdo, b->bundle_dict_off, (nb + 1) * 4));
        CU(cp(c->d_dsegs, b->dict_segs, nd * 8)); CU(cp(c->d_dphase, b->dict_phase, nd));
        if (b->dict_seq_off) { CU(c->d_dsso.ensure(nd * 4 + 4)); CU(cp(c->d_dsso, b->dict_seq_off, nd * 4)); }
    }
    const size_t nexc = (seqb && b->seq_format == ORFX_SEQ_NT2) ? (size_t)b->n_seq_exc : 0;
    if (nexc) {
        CU(c->d_exc_pos.ensure(nexc * 8)); CU(c->d_exc_code.ensure(nexc + 4));
        CU(cp(c->d_exc_pos, b->seq_exc_pos, nexc * 8)); CU(cp(c->d_exc_code, b->seq_exc_code, nexc));
    }
    const bool islands = seqb && host_sso;
    if (islands) { CU(c->d_sso.ensure(ns * 4 + 4)); if (!dict) CU(cp(c->d_sso, b->seg_seq_off, ns * 4)); }
    const bool flanks = seqb && c->P.overhang > 0;
    if (flanks) {
        if (nt && (!b->tx_ext_lo || !b->tx_ext_hi)) return fail(ORFX_E_INVALID, "params.overhang > 0 needs batch.tx_ext_lo / tx_ext_hi");
        CU(c->d_xlo.ensure(nt * 4 + 4)); CU(c->d_xhi.ensure(nt * 4 + 4));
        CU(cp(c->d_xlo, b->tx_ext_lo, nt * 4)); CU(cp(c->d_xhi, b->tx_ext_hi, nt * 4));
        if (b->tx_ext_flags) { CU(c->d_xfl.ensure(nt + 4)); CU(cp(c->d_xfl, b->tx_ext_flags, nt)); }
        if (!host_sso)   // a strand '.' query takes its 5' flank from the low side: only a private copy of its bases can hold that
            for (size_t i = 0; i < nt; i++)
                if (b->tx_kind[i] == ORFX_TX_QUERY && b->tx_strand[i] != '+' && b->tx_strand[i] != '-' && b->tx_ext_hi[i] > 0)
                    return fail(ORFX_E_UNSUPPORTED, "--overhang with strand '.' query transcripts needs the islands layout (seg_seq_off)");
    }
    const bool align = c->P.align_enabled && seqb;
    if (align && b->seq_format != ORFX_SEQ_RAW)
        return fail(ORFX_E_UNSUPPORTED, "params.align_enabled needs raw sequence bytes (ORFX_SEQ_RAW): the 4-bit codes do not tell the IUPAC letters of codon_map apart");
    if (align) { CU(c->d_aa_off.ensure((nt + 1) * 8)); CU(c->d_cnt_aa.ensure(nt * 8 + 8)); }
    c->seq_in_flight = false;
    if (seqb && overlap) {
        c->upload_bytes += seq_store;
        CU(cudaEventRecord(c->ev_chains, s));
        CU(cudaStreamWaitEvent(c->copy_stream, c->ev_chains, 0));   // after the chains, and after whatever still reads d_seq
        CU(cudaMemcpyAsync(c->d_seq.p, b->seq, seq_store, cudaMemcpyHostToDevice, c->copy_stream));
        CU(cudaMemsetAsync((char*)c->d_seq.p + seq_store, 0, 64, c->copy_stream));
        CU(cudaEventRecord(c->ev_seq, c->copy_stream));
        c->seq_in_flight = true;
    }
    // derived buffers whose size is known now
    CU(c->d_seg_cum.ensure(ns * 4 + 4));
    CU(c->d_tx_len.ensure(nt * 4 + 4));
    CU(c->d_tx_bundle.ensure(nt * 4 + 4));
    CU(c->d_mask_off.ensure((nt + 1) * 8));
    CU(c->d_tile_off.ensure((nt + 1) * 8));
    CU(c->d_q_off.ensure((nt + 1) * 8));
    CU(c->d_q_list.ensure(nt * 4 + 4));
    CU(c->d_row_off.ensure((nt + 1) * 8));
    CU(c->d_cnt_q.ensure(nt * 8 + 8));
    CU(c->d_cnt_words.ensure(nt * 8 + 8));
    CU(c->d_cnt_tiles.ensure(nt * 8 + 8));
    CU(c->d_cnt_rows.ensure(nt * 8 + 8));
    CU(c->d_scan_tmp.ensure((size_t)scan_tmp_elems((int64_t)nt) * 8 + 8));
    CU(c->d_need.ensure(nb + 4));
    CU(c->d_heavy.ensure((nb + 4) * 4 + nb * 16 + nb * 16 * 16 + 64));   // heavy list, plan scratch, work items (pipeline.cu: heavy_plan_kernel)
    CU(c->d_heavy_count.ensure(16));
    const size_t plane_words = (seqb + 31) / 32 + 4;
    if (seqb) CU(c->d_planes.ensure(2 * plane_words * 16));
    CU(c->h_totals.ensure(128));
    DevBatch& B = c->B;
    memset(&B, 0, sizeof B);
    B.n_bundles = b->n_bundles; B.n_tx = b->n_tx; B.n_segs = b->n_segs; B.seq_bytes = (int64_t)seqb; B.seq_format = b->seq_format;
    B.seq = c->d_seq.as<uint8_t>(); B.seq_off = c->d_seq_off.as<int64_t>(); B.bundle_start = c->d_bstart.as<int32_t>();
    B.bundle_tx_off = c->d_btx_off.as<int32_t>(); B.tx_kind = c->d_kind.as<uint8_t>(); B.tx_strand = c->d_strand.as<uint8_t>();
    B.tx_seg_off = c->d_seg_off.as<int32_t>(); B.tx_cds_start = c->d_cds_s.as<int32_t>(); B.tx_cds_end = c->d_cds_e.as<int32_t>();
    B.tx_tid_class = c->d_tid.as<int32_t>(); B.segs = c->d_segs.as<orfx_seg>(); B.seg_phase = c->d_phase.as<int8_t>();
    B.seg_cum = c->d_seg_cum.as<int32_t>(); B.tx_len = c->d_tx_len.as<int32_t>(); B.tx_bundle = c->d_tx_bundle.as<int32_t>();
    B.tx_mask_off = c->d_mask_off.as<int64_t>(); B.tx_tile_off = c->d_tile_off.as<int64_t>(); B.tx_q_off = c->d_q_off.as<int64_t>();
    B.q_list = c->d_q_list.as<int32_t>(); B.q_row_off = c->d_row_off.as<int64_t>();
    B.seg_seq_off = islands ? c->d_sso.as<int32_t>() : nullptr;
    B.tx_ext_lo = flanks ? c->d_xlo.as<int32_t>() : nullptr;
    B.tx_ext_hi = flanks ? c->d_xhi.as<int32_t>() : nullptr;
    B.tx_ext_flags = flanks && b->tx_ext_flags ? c->d_xfl.as<uint8_t>() : nullptr;
    B.bundle_need = c->d_need.as<uint8_t>(); B.planes = c->d_planes.as<uint32_t>(); B.plane_words = (int64_t)plane_words;
    B.hb_max_segs = c->hb_max_segs; B.hb_max_pairs = c->hb_max_pairs;
    B.seq_exc_pos = nexc ? c->d_exc_pos.as<int64_t>() : nullptr; B.seq_exc_code = nexc ? c->d_exc_code.as<uint8_t>() : nullptr; B.n_seq_exc = (int64_t)nexc;
    B.tx_aa_off = align ? c->d_aa_off.as<int64_t>() : nullptr;
    c->have_results = false;
    // The batch is caller-supplied: check what every kernel relies on (monotone CSR offsets, ordered chains, bases in
    // range) before anything walks it.  One small kernel and one flag read-back; the sequence bytes are already on their way.
    if (nt || nb) {
        CU(c->d_flag.ensure(16));
        launch_tx_bundle(B, s);
        if (dict)
            launch_expand_dict(B, c->d_ref.as<uint16_t>(), c->d_bdo.as<int32_t>(), c->d_dsegs.as<orfx_seg>(), c->d_dphase.as<int8_t>(),
                               b->dict_seq_off && islands ? c->d_dsso.as<int32_t>() : nullptr, (int64_t)nd, c->d_segs.as<orfx_seg>(), c->d_phase.as<int8_t>(),
                               islands ? c->d_sso.as<int32_t>() : nullptr, c->d_flag.as<int>() + 1, s);
        else CU(cudaMemsetAsync(c->d_flag.as<int>() + 1, 0, sizeof(int), s));
        launch_validate(B, seqb ? 1 : 0, c->allow_empty_chains ? 1 : 0, c->d_flag.as<int>(), s);
        int* hflag = reinterpret_cast<int*>(c->h_totals.as<int64_t>() + 6);
        CU(cudaMemcpyAsync(hflag, c->d_flag.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, s));
        CU(sync_stream(c, s));
        if (hflag[1]) {
            c->resident = false;
            if (c->seq_in_flight) { cudaStreamSynchronize(c->copy_stream); c->seq_in_flight = false; }
            return fail(ORFX_E_INVALID, "malformed batch: a seg_ref outside its bundle's dictionary records, or bundle_dict_off not monotone / beyond n_dict");
        }
        if (*hflag) {
            c->resident = false;
            if (c->seq_in_flight) { cudaStreamSynchronize(c->copy_stream); c->seq_in_flight = false; }
            char msg[160];
            snprintf(msg, sizeof msg, "malformed batch (check bits 0x%x: 1 offsets, 2 transcript kind, 4 empty chain, 8 segment order, 16 negative flank, 32 bases out of range)", *hflag);
            return fail(ORFX_E_INVALID, msg);
        }
    }
    c->resident = true;
    return ORFX_OK;
}

// a stage = a CUDA event pair on the stream it is launched on (orfx_kernel_times) + an NVTX range around its launches
void stage_begin(orfx_ctx* c, int st, cudaStream_t s = nullptr) { nvtxRangePushA(STAGE_NAME[st]); cudaEventRecord(c->st_ev[st][0], s ? s : c->stream); }
void stage_end(orfx_ctx* c, int st, cudaStream_t s = nullptr) { cudaEventRecord(c->st_ev[st][1], s ? s : c->stream); nvtxRangePop(); }

// the whole path on the resident batch; leaves results in device memory
int run_path(orfx_ctx* c) {
    cudaStream_t s = c->stream;
    DevBatch& B = c->B;
    c->n_launches = 0;
    c->n_queries = c->n_rows = c->n_tiles = c->n_words = 0;
    cudaEventRecord(c->run_ev[0], s);
    if (B.n_tx == 0) {
        if (c->seq_in_flight) { CU(cudaStreamWaitEvent(s, c->ev_seq, 0)); c->seq_in_flight = false; }
        cudaEventRecord(c->run_ev[1], s);
        return ORFX_OK;
    }
    stage_begin(c, ST_PREP);
    c->n_launches += launch_prep(B, c->d_cnt_q.as<int64_t>(), c->d_cnt_words.as<int64_t>(), c->d_cnt_tiles.as<int64_t>(), c->d_scan_tmp.as<int64_t>(),
                                 c->d_heavy.as<int32_t>(), c->d_heavy_count.as<unsigned>(), s);
    stage_end(c, ST_PREP);
    stage_begin(c, ST_COUNT);
    c->n_launches += launch_count_rows(B, c->d_cnt_rows.as<int64_t>(), c->d_scan_tmp.as<int64_t>(), c->d_heavy.as<int32_t>(), c->d_heavy_count.as<unsigned>(),
                                       c->n_sm, s, c->serialize ? s : c->side_stream, c->ev_fork[0], c->ev_join[0]);
    stage_end(c, ST_COUNT);
    if (c->P.align_enabled && c->P.has_reference && B.seq_bytes > 0 && !B.tx_aa_off)
        return fail(ORFX_E_STATE, "params.align_enabled was set after the batch was uploaded: upload it again");
    const bool align = B.tx_aa_off != nullptr && c->P.align_enabled;
    if (align) {
        c->n_launches += launch_aa_len(B, c->d_cnt_aa.as<int64_t>(), s);
        c->n_launches += launch_exclusive_scan(c->d_cnt_aa.as<int64_t>(), B.tx_aa_off, c->d_scan_tmp.as<int64_t>(), B.n_tx, s);
    }
    // the only host round trip of the path: four totals that size the remaining buffers
    int64_t* tot = c->h_totals.as<int64_t>();
    tot[4] = 0;
    if (align) CU(cudaMemcpyAsync(tot + 4, B.tx_aa_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(tot + 0, B.tx_q_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(tot + 1, B.tx_mask_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(tot + 2, B.tx_tile_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(tot + 3, B.q_row_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(sync_stream(c, s));
    const int64_t nq = tot[0], words = tot[1], tiles = tot[2], nrows = tot[3];
    c->n_queries = nq; c->n_words = words; c->n_tiles = tiles; c->n_rows = nrows;
    if (c->P.has_reference) {
        CU(c->d_tile_tx.ensure((size_t)tiles * 4 + 4));
        CU(c->d_stop.ensure((size_t)words * 4 + 16));
        CU(c->d_start.ensure((size_t)words * 4 + 16));
        B.tile_tx = c->d_tile_tx.as<int32_t>(); B.stop_mask = c->d_stop.as<uint32_t>(); B.start_mask = c->d_start.as<uint32_t>();
    }
    if (nrows > (int64_t)INT32_MAX) return fail(ORFX_E_INVALID, "more than 2^31 candidate rows in one batch: shard it");
    for (int i = 0; i < 16; i++) CU(c->r_i32[i].ensure((size_t)nrows * 4 + 4));
    for (int i = 0; i < 3; i++) CU(c->r_f32[i].ensure((size_t)nrows * 4 + 4));
    for (int i = 0; i < 3; i++) CU(c->r_u8[i].ensure((size_t)nrows + 4));
    CU(c->r_order.ensure((size_t)nrows * 8 + 8));
    CU(c->d_ord.ensure((size_t)nrows * 4 + 4));
    CU(c->d_keys.ensure((size_t)nrows * 48 + 16));
    CU(c->d_best.ensure((size_t)nq * sizeof(orfx_best) + 8));
    DevRows& R = c->R;
    R.n_rows = nrows;
    int32_t** i32[16] = {&R.q_tx, &R.t_tx, &R.seg_start, &R.seg_end, &R.cds_start, &R.cds_end, &R.n_dup, &R.qlen, &R.tlen, &R.ulen,
                         &R.query_start, &R.extra, &R.missing, &R.match, &R.outframe, &R.inframe};
    for (int i = 0; i < 16; i++) *i32[i] = c->r_i32[i].as<int32_t>();
    R.lpi = c->r_f32[0].as<float>(); R.ilpi = c->r_f32[1].as<float>(); R.mlpi = c->r_f32[2].as<float>();
    R.status = c->r_u8[0].as<uint8_t>(); R.pass = c->r_u8[1].as<uint8_t>(); R.start_match = c->r_u8[2].as<uint8_t>();
    R.print_order = c->r_order.as<int64_t>();
    R.aln_q0 = R.aln_qn = R.aln_max = R.aln_score = R.aln_mqe = R.aln_mte = nullptr;
    c->n_aa = 0; c->aln_cells = 0; c->ms_align = 0;
    if (align) {
        c->n_aa = tot[4];
        CU(c->d_aa_pos.ensure((size_t)tot[4] + 64));
        B.aa_pos = c->d_aa_pos.as<uint8_t>();
        for (int i = 0; i < 6; i++) CU(c->r_aln[i].ensure((size_t)nrows * 4 + 4));
        CU(c->d_row_pairs.ensure((size_t)nrows * sizeof(AlnPair) + 16));
        R.aln_q0 = c->r_aln[0].as<int32_t>(); R.aln_qn = c->r_aln[1].as<int32_t>(); R.aln_max = c->r_aln[2].as<int32_t>();
        R.aln_score = c->r_aln[3].as<int32_t>(); R.aln_mqe = c->r_aln[4].as<int32_t>(); R.aln_mte = c->r_aln[5].as<int32_t>();
    }
    R.ext_lo = R.ext_hi = nullptr;
    if (B.tx_ext_lo) {
        for (int i = 0; i < 2; i++) CU(c->r_ext[i].ensure((size_t)nrows * 4 + 4));
        R.ext_lo = c->r_ext[0].as<int32_t>(); R.ext_hi = c->r_ext[1].as<int32_t>();
    }

    // The heavy bundles (titin-class loci, ~1 % of the bundles, a fifth of the chain work) have their own CTA-per-bundle
    // kernels over disjoint queries / rows; they run on the side stream next to the thread-per-query / thread-per-row
    // kernels, which leave most of the GPU's latency-hiding capacity unused.
    cudaStream_t side = c->serialize ? s : c->side_stream;
    stage_begin(c, ST_FILL);
    CU(cudaEventRecord(c->ev_fork[1], s));
    CU(cudaStreamWaitEvent(side, c->ev_fork[1], 0));
    c->n_launches += launch_fill_rows(B, R, (int)nq, s);
    c->n_launches += launch_heavy_fill(B, R, c->d_heavy.as<int32_t>(), c->d_heavy_count.as<unsigned>(), c->n_sm, side);
    stage_end(c, ST_FILL);
    if (c->seq_in_flight) { CU(cudaStreamWaitEvent(s, c->ev_seq, 0)); c->seq_in_flight = false; }
    stage_begin(c, ST_PLANES);
    if (c->P.has_reference) {
        c->n_launches += launch_fill_tiles(B, s);
        c->n_launches += launch_base_planes(B, s);
    }
    stage_end(c, ST_PLANES);
    stage_begin(c, ST_SCAN);   // the codon_scan_kernel launch alone: this is the roofline's kernel time
    if (c->P.has_reference && tiles > 0) { launch_codon_scan(B, tiles, 0, s); c->n_launches++; }
    stage_end(c, ST_SCAN);
    stage_begin(c, ST_TRANSLATE);
    if (align) c->n_launches += launch_translate(B, c->d_tables.as<TranslateTables>(), c->n_aa, s);
    stage_end(c, ST_TRANSLATE);
    // (Measured and dropped: scanning the heavy bundles' transcripts in a first launch so that heavy_rescue_kernel could start
    // beside the scan of the rest.  The kernels only trade SM time with each other: the step went from 1.34 to 1.37 ms.)
    CU(cudaEventRecord(c->ev_fork[2], s));                 // masks ready
    CU(cudaStreamWaitEvent(side, c->ev_fork[2], 0));       // (heavy_fill is already ahead of it on the side stream)
    stage_begin(c, ST_HEAVY, side);
    c->n_launches += launch_heavy_rescue(B, R, c->P, c->d_heavy.as<int32_t>(), c->d_heavy_count.as<unsigned>(), c->n_sm, side);
    stage_end(c, ST_HEAVY, side);
    CU(cudaEventRecord(c->ev_join[2], side));
    stage_begin(c, ST_RESCUE);
    c->n_launches += launch_rescue_score(B, R, c->P, s);
    stage_end(c, ST_RESCUE);
    CU(cudaStreamWaitEvent(s, c->ev_join[2], 0));
    stage_begin(c, ST_SELECT);
    c->n_launches += launch_select(B, R, c->P, (int)nq, c->d_best.as<orfx_best>(), c->d_ord.as<int32_t>(), c->d_keys.as<int4>(), s);
    stage_end(c, ST_SELECT);
    // alignment of every scored row's ORF against its template's protein (Finder::align, aln.hpp:27-41)
    stage_begin(c, ST_ALIGN);
    if (align && nrows > 0) {
        for (int i = 2; i < 6; i++) CU(cudaMemsetAsync(c->r_aln[i].p, 0, (size_t)nrows * 4, s));   // rows that were not scored hold 0
        c->n_launches += launch_row_pairs(B, R, c->d_row_pairs.as<AlnPair>(), s);
        int32_t* outp[4] = {R.aln_max, R.aln_score, R.aln_mqe, R.aln_mte};
        std::string err;
        int nl = 0;
        const int rc = align_pairs_device(c->aln, s, c->n_sm, B.aa_pos, 3, 3, c->d_row_pairs.as<AlnPair>(), nrows, c->P.align_gapo ? c->P.align_gapo : 12,
                                          c->P.align_gape ? c->P.align_gape : 2, outp, &c->aln_cells, nullptr, &nl, err);
        if (rc) return fail(rc, err);
        c->n_launches += nl;
    }
    stage_end(c, ST_ALIGN);
    cudaEventRecord(c->run_ev[1], s);
    CU(cudaGetLastError());
    return ORFX_OK;
}

void collect_times(orfx_ctx* c) {
    c->run_ms = 0;
    for (int i = 0; i < ST_N; i++) c->st_ms[i] = 0;
    if (c->B.n_tx == 0) return;
    cudaEventElapsedTime(&c->run_ms, c->run_ev[0], c->run_ev[1]);
    for (int i = 0; i < ST_N; i++) cudaEventElapsedTime(&c->st_ms[i], c->st_ev[i][0], c->st_ev[i][1]);
}

int download_results(orfx_ctx* c) {
    cudaStream_t s = c->stream;
    const size_t nq = (size_t)c->n_queries, nr = (size_t)c->n_rows;
    const bool pg = c->P.host_results_pageable != 0;
    CU(c->h_best.ensure(nq * sizeof(orfx_best) + 8, pg));
    if (nq) CU(cudaMemcpyAsync(c->h_best.p, c->d_best.p, nq * sizeof(orfx_best), cudaMemcpyDeviceToHost, s));
    if (c->P.want_rows && nr) {
        for (int i = 0; i < 16; i++) { CU(c->h_i32[i].ensure(nr * 4, pg)); CU(cudaMemcpyAsync(c->h_i32[i].p, c->r_i32[i].p, nr * 4, cudaMemcpyDeviceToHost, s)); }
        for (int i = 0; i < 3; i++) { CU(c->h_f32[i].ensure(nr * 4, pg)); CU(cudaMemcpyAsync(c->h_f32[i].p, c->r_f32[i].p, nr * 4, cudaMemcpyDeviceToHost, s)); }
        for (int i = 0; i < 3; i++) { CU(c->h_u8[i].ensure(nr, pg)); CU(cudaMemcpyAsync(c->h_u8[i].p, c->r_u8[i].p, nr, cudaMemcpyDeviceToHost, s)); }
        CU(c->h_order.ensure(nr * 8, pg));
        CU(cudaMemcpyAsync(c->h_order.p, c->r_order.p, nr * 8, cudaMemcpyDeviceToHost, s));
        if (c->R.ext_lo) for (int i = 0; i < 2; i++) { CU(c->h_ext[i].ensure(nr * 4, pg)); CU(cudaMemcpyAsync(c->h_ext[i].p, c->r_ext[i].p, nr * 4, cudaMemcpyDeviceToHost, s)); }
        if (c->R.aln_max) for (int i = 0; i < 4; i++) { CU(c->h_aln[i].ensure(nr * 4, pg)); CU(cudaMemcpyAsync(c->h_aln[i].p, c->r_aln[2 + i].p, nr * 4, cudaMemcpyDeviceToHost, s)); }
    }
    return ORFX_OK;
}

void fill_results(orfx_ctx* c, orfx_results* r) {
    memset(r, 0, sizeof *r);
    r->n_queries = (int32_t)c->n_queries;
    r->n_rows = c->n_rows;
    r->best = c->h_best.as<orfx_best>();
    if (c->P.want_rows) {
        const int32_t** i32[16] = {&r->row_q_tx, &r->row_t_tx, &r->row_seg_start, &r->row_seg_end, &r->row_cds_start, &r->row_cds_end, &r->row_n_dup,
                                   &r->row_qlen, &r->row_tlen, &r->row_ulen, &r->row_query_start, &r->row_extra, &r->row_missing, &r->row_match,
                                   &r->row_outframe, &r->row_inframe};
        for (int i = 0; i < 16; i++) *i32[i] = c->h_i32[i].as<int32_t>();
        r->row_lpi = c->h_f32[0].as<float>(); r->row_ilpi = c->h_f32[1].as<float>(); r->row_mlpi = c->h_f32[2].as<float>();
        r->row_status = c->h_u8[0].as<uint8_t>(); r->row_pass = c->h_u8[1].as<uint8_t>(); r->row_start_match = c->h_u8[2].as<uint8_t>();
        r->print_order = c->h_order.as<int64_t>();
        if (c->R.ext_lo && c->n_rows) { r->row_ext_lo = c->h_ext[0].as<int32_t>(); r->row_ext_hi = c->h_ext[1].as<int32_t>(); }
        if (c->R.aln_max && c->n_rows) {
            r->row_aln_max = c->h_aln[0].as<int32_t>(); r->row_aln_score = c->h_aln[1].as<int32_t>();
            r->row_aln_mqe = c->h_aln[2].as<int32_t>(); r->row_aln_mte = c->h_aln[3].as<int32_t>();
        }
    }
}

// A repeated --mode entry never decides a comparison (the first occurrence already did): drop the repeats, so that
// the device sees at most the six distinct modes.
orfx_params canonical_params(const orfx_params& in) {
    orfx_params p = in;
    int n = 0;
    for (int i = 0; i < in.n_modes; i++) {
        bool seen = false;
        for (int j = 0; j < n; j++) seen = seen || p.mode_array[j] == in.mode_array[i];
        if (!seen) p.mode_array[n++] = in.mode_array[i];
    }
    p.n_modes = n;
    return p;
}

int check_params(const orfx_params* p) {
    if (!p) return fail(ORFX_E_INVALID, "null params");
    if (p->n_modes < 0 || p->n_modes > 8) return fail(ORFX_E_INVALID, "n_modes out of range");
    for (int i = 0; i < p->n_modes; i++) if (p->mode_array[i] < 0 || p->mode_array[i] > ORFX_MODE_FIRST) return fail(ORFX_E_INVALID, "invalid mode");
    if (p->overhang < 0) return fail(ORFX_E_INVALID, "negative overhang");
    if (p->align_enabled) {
        if (!p->has_reference) return fail(ORFX_E_INVALID, "params.align_enabled needs params.has_reference (proteins are translated from the genome)");
        const int go = p->align_gapo ? p->align_gapo : 12, ge = p->align_gape ? p->align_gape : 2;
        if (go < 0 || ge < 0 || go + ge > 64) return fail(ORFX_E_UNSUPPORTED, "gap costs outside this build's range (0 <= gapo + gape <= 64)");
    }
    return ORFX_OK;
}

}  // namespace

extern "C" {

const char* orfx_last_error(void) { return g_err.c_str(); }
int orfx_abi_version(void) { return ORFX_ABI_VERSION; }

void* orfx_host_alloc(size_t bytes) {
    void* p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) { g_err = std::string("cudaHostAlloc: ") + cudaGetErrorString(e); return nullptr; }
    return p;
}
void orfx_host_free(void* p) { if (p) cudaFreeHost(p); }

int orfx_create(orfx_ctx** out, int device, const orfx_params* params) {
    if (!out) return fail(ORFX_E_INVALID, "null out");
    int rc = check_params(params);
    if (rc) return rc;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(ORFX_E_NODEVICE, std::string("no CUDA device available (") + (e == cudaSuccess ? "device count 0" : cudaGetErrorString(e)) +
                                         "); this library has no CPU path");
    if (device < 0 || device >= ndev) return fail(ORFX_E_NODEVICE, "device ordinal out of range");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(ORFX_E_NODEVICE, std::string("device ") + prop.name + " is not sm_100-class; this library is built for sm_100a only");
    orfx_ctx* c = new orfx_ctx();
    c->device = device;
    c->n_sm = prop.multiProcessorCount;
    c->P = canonical_params(*params);
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return fail(ORFX_E_CUDA, "cudaStreamCreate failed"); }
    cudaEventCreateWithFlags(&c->ev_chains, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_block, cudaEventDisableTiming | cudaEventBlockingSync);
    cudaEventCreate(&c->ev_seq);
    for (int i = 0; i < 3; i++) { cudaEventCreateWithFlags(&c->ev_fork[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&c->ev_join[i], cudaEventDisableTiming); }
    for (auto& ev : c->ev) cudaEventCreate(&ev);
    for (int i = 0; i < ST_N; i++) { cudaEventCreate(&c->st_ev[i][0]); cudaEventCreate(&c->st_ev[i][1]); c->st_ms[i] = 0; }
    cudaEventCreate(&c->run_ev[0]); cudaEventCreate(&c->run_ev[1]);
    {
        TranslateTables T;
        build_translate_tables(T);
        if (c->d_tables.ensure(sizeof T) != cudaSuccess || cudaMemcpy(c->d_tables.p, &T, sizeof T, cudaMemcpyHostToDevice) != cudaSuccess) {
            orfx_destroy(c);
            return fail(ORFX_E_CUDA, "cannot upload the translation tables");
        }
    }
    *out = c;
    return ORFX_OK;
}

int orfx_set_params(orfx_ctx* c, const orfx_params* params) {
    if (!c) return fail(ORFX_E_INVALID, "null ctx");
    int rc = check_params(params);
    if (rc) return rc;
    c->P = canonical_params(*params);
    return ORFX_OK;
}

int orfx_submit(orfx_ctx* c, const orfx_batch* batch) {
    if (!c) return fail(ORFX_E_INVALID, "null ctx");
    CU(cudaSetDevice(c->device));
    CU(cudaEventRecord(c->ev[0], c->stream));
    nvtxRangePushA("orfx_submit: upload");
    int rc = upload_batch(c, batch, true);
    nvtxRangePop();
    if (rc) return rc;
    CU(cudaEventRecord(c->ev[1], c->stream));
    nvtxRangePushA("orfx_submit: path");
    rc = run_path(c);
    nvtxRangePop();
    if (rc) return rc;
    CU(cudaEventRecord(c->ev[2], c->stream));
    nvtxRangePushA("orfx_submit: download");
    rc = download_results(c);
    nvtxRangePop();
    if (rc) return rc;
    CU(cudaEventRecord(c->ev[3], c->stream));
    c->pending = true;
    return ORFX_OK;
}

int orfx_wait(orfx_ctx* c, orfx_results* out) {
    if (!c || !out) return fail(ORFX_E_INVALID, "null argument");
    if (!c->pending) return fail(ORFX_E_STATE, "orfx_wait without a pending orfx_submit");
    CU(cudaSetDevice(c->device));
    CU(sync_stream(c, c->stream));
    CU(cudaGetLastError());
    c->pending = false;
    c->have_results = true;
    collect_times(c);
    fill_results(c, out);
    cudaEventElapsedTime(&out->ms_h2d, c->ev[0], c->ev[1]);   // chain arrays; the sequence upload overlaps the first kernels
    if (c->B.seq_bytes > 0) { float t = 0; if (cudaEventElapsedTime(&t, c->ev[0], c->ev_seq) == cudaSuccess && t > out->ms_h2d) out->ms_h2d = t; }
    cudaEventElapsedTime(&out->ms_kernels, c->ev[1], c->ev[2]);
    cudaEventElapsedTime(&out->ms_d2h, c->ev[2], c->ev[3]);
    out->ms_align = c->st_ms[ST_ALIGN];
    return ORFX_OK;
}

int orfx_release(orfx_ctx* c) {
    if (!c) return fail(ORFX_E_INVALID, "null ctx");
    c->have_results = false;
    return ORFX_OK;
}

void orfx_destroy(orfx_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    DBuf* d[] = {&c->d_sso, &c->d_seq, &c->d_seq_off, &c->d_bstart, &c->d_btx_off, &c->d_kind, &c->d_strand, &c->d_seg_off, &c->d_cds_s, &c->d_cds_e, &c->d_tid,
                 &c->d_segs, &c->d_phase, &c->d_seg_cum, &c->d_tx_len, &c->d_tx_bundle, &c->d_mask_off, &c->d_tile_off, &c->d_q_off, &c->d_q_list,
                 &c->d_row_off, &c->d_tile_tx, &c->d_stop, &c->d_start, &c->d_need, &c->d_planes, &c->d_heavy, &c->d_heavy_count, &c->d_cnt_q, &c->d_cnt_words, &c->d_cnt_tiles, &c->d_cnt_rows, &c->d_scan_tmp,
                 &c->d_ord, &c->r_order, &c->d_best, &c->d_clean_keep, &c->d_clean_s, &c->d_clean_e, &c->d_keys,
                 &c->d_xlo, &c->d_xhi, &c->d_xfl, &c->d_flag, &c->d_exc_pos, &c->d_exc_code, &c->r_ext[0], &c->r_ext[1],
                 &c->d_ref, &c->d_bdo, &c->d_dsegs, &c->d_dphase, &c->d_dsso};
    for (DBuf* b : d) b->free_();
    c->aln.free_();
    c->d_aa_off.free_(); c->d_aa_pos.free_(); c->d_cnt_aa.free_(); c->d_tables.free_(); c->d_row_pairs.free_();
    for (auto& b : c->r_aln) b.free_();
    for (auto& b : c->h_aln) b.free_();
    for (auto& b : c->r_i32) b.free_();
    for (auto& b : c->r_f32) b.free_();
    for (auto& b : c->r_u8) b.free_();
    c->h_totals.free_(); c->h_best.free_(); c->h_order.free_();
    for (auto& b : c->h_i32) b.free_();
    for (auto& b : c->h_f32) b.free_();
    for (auto& b : c->h_u8) b.free_();
    for (auto& b : c->h_ext) b.free_();
    for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
    for (int i = 0; i < ST_N; i++) { cudaEventDestroy(c->st_ev[i][0]); cudaEventDestroy(c->st_ev[i][1]); }
    cudaEventDestroy(c->run_ev[0]); cudaEventDestroy(c->run_ev[1]);
    cudaStreamSynchronize(c->copy_stream);
    cudaEventDestroy(c->ev_chains); cudaEventDestroy(c->ev_seq); if (c->ev_block) cudaEventDestroy(c->ev_block);
    cudaStreamDestroy(c->copy_stream);
    cudaStreamSynchronize(c->side_stream);
    for (int i = 0; i < 3; i++) { cudaEventDestroy(c->ev_fork[i]); cudaEventDestroy(c->ev_join[i]); }
    cudaStreamDestroy(c->side_stream);
    cudaStreamDestroy(c->stream);
    delete c;
}

int orfx_upload(orfx_ctx* c, const orfx_batch* batch) {
    if (!c) return fail(ORFX_E_INVALID, "null ctx");
    CU(cudaSetDevice(c->device));
    int rc = upload_batch(c, batch);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    return ORFX_OK;
}

int orfx_run_resident(orfx_ctx* c, int* n_launches) {
    if (!c) return fail(ORFX_E_INVALID, "null ctx");
    if (!c->resident) return fail(ORFX_E_STATE, "no resident batch: call orfx_upload first");
    CU(cudaSetDevice(c->device));
    int rc = run_path(c);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    collect_times(c);
    if (n_launches) *n_launches = c->n_launches;
    return ORFX_OK;
}

int orfx_fetch(orfx_ctx* c, orfx_results* out) {
    if (!c || !out) return fail(ORFX_E_INVALID, "null argument");
    if (!c->resident) return fail(ORFX_E_STATE, "nothing to fetch");
    CU(cudaSetDevice(c->device));
    int rc = download_results(c);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    fill_results(c, out);
    return ORFX_OK;
}

int orfx_kernel_times(orfx_ctx* c, int max_n, const char** names, float* ms, int* n) {
    if (!c || !n) return fail(ORFX_E_INVALID, "null argument");
    int k = 0;
    for (; k < ST_N && k < max_n; k++) { if (names) names[k] = STAGE_NAME[k]; if (ms) ms[k] = c->st_ms[k]; }
    *n = k;
    return ORFX_OK;
}

float orfx_last_run_ms(orfx_ctx* c) { return c ? c->run_ms : -1.f; }
int orfx_last_run_launches(orfx_ctx* c) { return c ? c->n_launches : -1; }
int64_t orfx_last_upload_bytes(orfx_ctx* c) { return c ? (int64_t)c->upload_bytes : -1; }

int64_t orfx_resident_query_nt(orfx_ctx* c) {
    if (!c || !c->resident) return fail(ORFX_E_STATE, "no resident batch");
    if (cudaSetDevice(c->device) != cudaSuccess) return ORFX_E_CUDA;
    const size_t nt = (size_t)c->B.n_tx;
    std::vector<int32_t> len(nt);
    std::vector<uint8_t> kind(nt);
    if (nt) {
        if (cudaMemcpy(len.data(), c->B.tx_len, nt * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(ORFX_E_CUDA, "copy tx_len");
        if (cudaMemcpy(kind.data(), c->B.tx_kind, nt, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(ORFX_E_CUDA, "copy tx_kind");
    }
    int64_t sum = 0;
    for (size_t i = 0; i < nt; i++) if (kind[i] == ORFX_TX_QUERY) sum += len[i];
    return sum;
}

int64_t orfx_get_masks(orfx_ctx* c, int32_t tx, uint32_t* stop_bits, uint32_t* start_bits, int64_t max_words) {
    if (!c || !c->resident || !c->P.has_reference) return fail(ORFX_E_STATE, "no resident masks");
    if (tx < 0 || tx >= c->B.n_tx) return fail(ORFX_E_INVALID, "tx out of range");
    if (cudaSetDevice(c->device) != cudaSuccess) return ORFX_E_CUDA;
    int64_t off[2];
    int32_t L;
    uint8_t kind;
    if (cudaMemcpy(off, c->B.tx_mask_off + tx, 16, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(ORFX_E_CUDA, "copy mask_off");
    if (cudaMemcpy(&L, c->B.tx_len + tx, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(ORFX_E_CUDA, "copy len");
    if (cudaMemcpy(&kind, c->B.tx_kind + tx, 1, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(ORFX_E_CUDA, "copy kind");
    if (kind != ORFX_TX_QUERY) return fail(ORFX_E_INVALID, "not a query transcript");
    if (c->B.tx_ext_lo) {   // --overhang: the scanned chain covers the flanks
        int32_t x[2];
        if (cudaMemcpy(&x[0], c->B.tx_ext_lo + tx, 4, cudaMemcpyDeviceToHost) != cudaSuccess || cudaMemcpy(&x[1], c->B.tx_ext_hi + tx, 4, cudaMemcpyDeviceToHost) != cudaSuccess)
            return fail(ORFX_E_CUDA, "copy flanks");
        L += x[0] + x[1];
    }
    int64_t words = ((int64_t)L + 31) / 32;
    if (words > max_words || words > off[1] - off[0]) return fail(ORFX_E_INVALID, "mask buffer too small");
    if (cudaMemcpy(stop_bits, c->B.stop_mask + off[0], (size_t)words * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(ORFX_E_CUDA, "copy stop mask");
    if (cudaMemcpy(start_bits, c->B.start_mask + off[0], (size_t)words * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return fail(ORFX_E_CUDA, "copy start mask");
    return L;
}

int orfx_debug_set(orfx_ctx* c, const char* key, int64_t value) {
    if (!c || !key) return fail(ORFX_E_INVALID, "null argument");
    const std::string k(key);
    if (k == "hb_max_segs") { if (value < 0 || value > HB_MAX_SEGS) return fail(ORFX_E_INVALID, "hb_max_segs out of range"); c->hb_max_segs = (int)value; }
    else if (k == "hb_max_pairs") { if (value < 0 || value > HB_MAX_PAIRS) return fail(ORFX_E_INVALID, "hb_max_pairs out of range"); c->hb_max_pairs = (int)value; }
    else if (k == "serialize") c->serialize = value != 0;
    else if (k == "blocking_sync") c->blocking_sync = value != 0;
    else return fail(ORFX_E_INVALID, "unknown debug key " + k);
    return ORFX_OK;
}

int orfx_clean_templates(orfx_ctx* c, const orfx_batch* batch, int rescue, orfx_clean_out* out, int64_t* n_out) {
    if (!c || !out || !out->keep || !out->cds_start || !out->cds_end) return fail(ORFX_E_INVALID, "null argument");
    if (!c->P.has_reference) return fail(ORFX_E_INVALID, "template cleaning needs a reference genome (params.has_reference)");
    CU(cudaSetDevice(c->device));
    int rc = upload_batch(c, batch);
    if (rc) return rc;
    cudaStream_t s = c->stream;
    DevBatch& B = c->B;
    c->n_launches = 0;
    if (n_out) *n_out = 0;
    if (B.n_tx == 0) return ORFX_OK;
    c->n_launches += launch_prep(B, c->d_cnt_q.as<int64_t>(), c->d_cnt_words.as<int64_t>(), c->d_cnt_tiles.as<int64_t>(), c->d_scan_tmp.as<int64_t>(),
                                 c->d_heavy.as<int32_t>(), c->d_heavy_count.as<unsigned>(), s);
    int64_t* tot = c->h_totals.as<int64_t>();
    CU(cudaMemcpyAsync(tot + 0, B.tx_q_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(tot + 1, B.tx_mask_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(tot + 2, B.tx_tile_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    const int64_t nq = tot[0], words = tot[1], tiles = tot[2];
    CU(c->d_tile_tx.ensure((size_t)tiles * 4 + 4));
    CU(c->d_stop.ensure((size_t)words * 4 + 16));
    CU(c->d_start.ensure((size_t)words * 4 + 16));
    B.tile_tx = c->d_tile_tx.as<int32_t>(); B.stop_mask = c->d_stop.as<uint32_t>(); B.start_mask = c->d_start.as<uint32_t>();
    CU(c->d_clean_keep.ensure((size_t)nq + 4)); CU(c->d_clean_s.ensure((size_t)nq * 4 + 4)); CU(c->d_clean_e.ensure((size_t)nq * 4 + 4));
    c->n_launches += launch_fill_tiles(B, s);
    c->n_launches += launch_base_planes(B, s);
    launch_codon_scan(B, tiles, 0, s);
    if (tiles > 0) c->n_launches++;
    c->n_launches += launch_clean_templates(B, (int)nq, rescue, c->P.non_aug, c->d_clean_keep.as<uint8_t>(), c->d_clean_s.as<int32_t>(),
                                            c->d_clean_e.as<int32_t>(), s);
    CU(cudaGetLastError());
    if (nq) {
        CU(cudaMemcpyAsync(out->keep, c->d_clean_keep.p, (size_t)nq, cudaMemcpyDeviceToHost, s));
        CU(cudaMemcpyAsync(out->cds_start, c->d_clean_s.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
        CU(cudaMemcpyAsync(out->cds_end, c->d_clean_e.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    }
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    c->n_queries = nq; c->n_words = words; c->n_tiles = tiles; c->n_rows = 0;
    c->have_results = false;
    if (n_out) *n_out = nq;
    return ORFX_OK;
}

int orfx_score_pairs(orfx_ctx* c, const orfx_batch* batch, orfx_results* out) {
    if (!c || !out) return fail(ORFX_E_INVALID, "null argument");
    CU(cudaSetDevice(c->device));
    orfx_batch nb;
    if (!batch) return fail(ORFX_E_INVALID, "null batch");
    nb = *batch; nb.seq = nullptr; nb.seq_bytes = 0; nb.seg_seq_off = nullptr;   // a compare batch reads no sequence
    const int32_t has_ref = c->P.has_reference;
    c->P.has_reference = 0;
    c->allow_empty_chains = true;
    int rc = upload_batch(c, &nb);
    c->allow_empty_chains = false;
    c->P.has_reference = has_ref;
    if (rc) return rc;
    cudaStream_t s = c->stream;
    DevBatch& B = c->B;
    memset(out, 0, sizeof *out);
    c->n_launches = 0; c->n_queries = c->n_rows = 0; c->have_results = false;
    if (B.n_tx == 0) return ORFX_OK;
    c->n_launches += launch_prep(B, c->d_cnt_q.as<int64_t>(), c->d_cnt_words.as<int64_t>(), c->d_cnt_tiles.as<int64_t>(), c->d_scan_tmp.as<int64_t>(),
                                 c->d_heavy.as<int32_t>(), c->d_heavy_count.as<unsigned>(), s);
    c->n_launches += launch_count_pairs(B, c->d_cnt_rows.as<int64_t>(), c->d_scan_tmp.as<int64_t>(), s);
    int64_t* tot = c->h_totals.as<int64_t>();
    CU(cudaMemcpyAsync(tot + 0, B.tx_q_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(tot + 3, B.q_row_off + B.n_tx, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    const int64_t nq = tot[0], nrows = tot[3];
    if (nrows > (int64_t)INT32_MAX) return fail(ORFX_E_INVALID, "more than 2^31 pairs in one batch: shard it");
    static const int USED_I32[] = {0, 1, 7, 8, 9, 11, 12, 13, 14, 15};   // q_tx t_tx qlen tlen ulen extra missing match outframe inframe
    for (int i : USED_I32) CU(c->r_i32[i].ensure((size_t)nrows * 4 + 4));
    for (int i = 0; i < 3; i++) CU(c->r_f32[i].ensure((size_t)nrows * 4 + 4));
    DevRows& R = c->R;
    memset(&R, 0, sizeof R);
    R.n_rows = nrows;
    R.q_tx = c->r_i32[0].as<int32_t>(); R.t_tx = c->r_i32[1].as<int32_t>(); R.qlen = c->r_i32[7].as<int32_t>(); R.tlen = c->r_i32[8].as<int32_t>();
    R.ulen = c->r_i32[9].as<int32_t>(); R.extra = c->r_i32[11].as<int32_t>(); R.missing = c->r_i32[12].as<int32_t>(); R.match = c->r_i32[13].as<int32_t>();
    R.outframe = c->r_i32[14].as<int32_t>(); R.inframe = c->r_i32[15].as<int32_t>();
    R.lpi = c->r_f32[0].as<float>(); R.ilpi = c->r_f32[1].as<float>(); R.mlpi = c->r_f32[2].as<float>();
    c->n_launches += launch_score_pairs(B, R, (int)nq, s);
    CU(cudaGetLastError());
    if (nrows) {
        for (int i : USED_I32) { CU(c->h_i32[i].ensure((size_t)nrows * 4)); CU(cudaMemcpyAsync(c->h_i32[i].p, c->r_i32[i].p, (size_t)nrows * 4, cudaMemcpyDeviceToHost, s)); }
        for (int i = 0; i < 3; i++) { CU(c->h_f32[i].ensure((size_t)nrows * 4)); CU(cudaMemcpyAsync(c->h_f32[i].p, c->r_f32[i].p, (size_t)nrows * 4, cudaMemcpyDeviceToHost, s)); }
    }
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    c->n_queries = nq; c->n_rows = nrows;
    out->n_queries = (int32_t)nq; out->n_rows = nrows;
    out->row_q_tx = c->h_i32[0].as<int32_t>(); out->row_t_tx = c->h_i32[1].as<int32_t>(); out->row_qlen = c->h_i32[7].as<int32_t>();
    out->row_tlen = c->h_i32[8].as<int32_t>(); out->row_ulen = c->h_i32[9].as<int32_t>(); out->row_extra = c->h_i32[11].as<int32_t>();
    out->row_missing = c->h_i32[12].as<int32_t>(); out->row_match = c->h_i32[13].as<int32_t>(); out->row_outframe = c->h_i32[14].as<int32_t>();
    out->row_inframe = c->h_i32[15].as<int32_t>();
    out->row_lpi = c->h_f32[0].as<float>(); out->row_ilpi = c->h_f32[1].as<float>(); out->row_mlpi = c->h_f32[2].as<float>();
    return ORFX_OK;
}

int orfx_align_batch(orfx_ctx* c, const uint8_t* aa, const int64_t* off, int64_t n_seq, const int32_t* pair_q, const int32_t* pair_t,
                     int64_t n_pairs, int gapo, int gape, orfx_aln_out* out, double* cells, float* ms) {
    if (!c) return fail(ORFX_E_INVALID, "null ctx");
    if (!aa || !off || !pair_q || !pair_t || !out || n_pairs < 0 || n_seq < 0) return fail(ORFX_E_INVALID, "null or negative argument");
    CU(cudaSetDevice(c->device));
    std::string err;
    int rc = align_batch_device(c->aln, c->stream, c->n_sm, aa, off, n_seq, pair_q, pair_t, n_pairs, gapo, gape, out, cells, ms, err);
    if (rc) return fail(rc, err);
    return ORFX_OK;
}

// no counterpart in the reference (see include/orfx.h): the alignment over caller-owned device arrays, scratch in a static
// __device__ array (align.cu), nothing allocated in the call
int orfx_align_pairs_static(orfx_ctx* c, const uint8_t* aa, const int64_t* off, int64_t n_seq, const int32_t* pair_q, const int32_t* pair_t,
                            int64_t n_pairs, int gapo, int gape, const orfx_aln_out* out) {
    if (!c) return fail(ORFX_E_INVALID, "null ctx");
    if (n_pairs < 0 || n_seq < 0) return fail(ORFX_E_INVALID, "negative argument");
    if (n_pairs > 0 && (!aa || !off || !pair_q || !pair_t || !out || !out->max || !out->score || !out->mqe || !out->mte)) return fail(ORFX_E_INVALID, "null argument");
    if (c->pending) return fail(ORFX_E_STATE, "orfx_align_pairs_static: a submitted batch is still pending on this context (orfx_wait first)");
    CU(cudaSetDevice(c->device));
    std::string err;
    const int rc = align_pairs_static_device(c->device, c->stream, c->n_sm, aa, off, n_seq, pair_q, pair_t, n_pairs, gapo, gape, out, err);
    if (rc) return fail(rc, err);
    return ORFX_OK;
}

int orfx_int_alu_peak(orfx_ctx* c, double* ops4) {
    if (!c || !ops4) return fail(ORFX_E_INVALID, "null argument");
    CU(cudaSetDevice(c->device));
    std::string err;
    int rc = int_alu_peak_device(c->stream, ops4, err);
    if (rc) return fail(rc, err);
    return ORFX_OK;
}

}  // extern "C"
