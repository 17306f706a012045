// C-ABI of the host-side batcher (include/orfx_pack.h): load + prepare + bundle + pack, for callers
// that are not the `orfanage` binary (bench.py, the parity tests, foreign-language bindings).
#include "orfx_host.h"
#include "orfx_pack.h"
#include <memory>

using namespace orfx;

struct orfxh_workload {
    Options opt;
    Fasta fasta;
    Transcriptome tr;
    std::vector<int> active;
    TidClasses tid_classes;
    std::vector<std::unique_ptr<PackedBatch>> packed;
};

namespace {
thread_local std::string g_err;
void find_active(orfxh_workload* w) {
    w->active.clear();
    for (int bi = 0; bi < (int)w->tr.bundles.size(); bi++)
        if (w->tr.bundles[bi].has_template && w->tr.bundles[bi].has_query) w->active.push_back(bi);
}
}  // namespace

extern "C" {

const char* orfxh_last_error(void) { return g_err.c_str(); }

int orfxh_load(orfxh_workload** out, const char* query_gtf, const char* const* template_gtfs, int n_templates,
               const char* reference_fasta, int use_id, int keep_all_cds) {
    return orfxh_load_overhang(out, query_gtf, template_gtfs, n_templates, reference_fasta, use_id, keep_all_cds, 0, 0);
}

int orfxh_load_overhang(orfxh_workload** out, const char* query_gtf, const char* const* template_gtfs, int n_templates,
                        const char* reference_fasta, int use_id, int keep_all_cds, int overhang, int spliced_overhang) {
    if (!out || !query_gtf || !template_gtfs || n_templates <= 0) { g_err = "null or empty argument"; return ORFX_E_INVALID; }
    try {
        std::unique_ptr<orfxh_workload> w(new orfxh_workload());
        w->opt.query = query_gtf;
        for (int i = 0; i < n_templates; i++) w->opt.templates.push_back(template_gtfs[i]);
        if (reference_fasta) w->opt.reference = reference_fasta;
        w->opt.use_id = use_id != 0;
        w->opt.keep_all_cds = keep_all_cds != 0;
        w->opt.overhang = overhang;
        w->opt.spliced_overhang = spliced_overhang != 0;
        prepare_transcriptome(w->opt, w->fasta, w->tr, false);
        find_active(w.get());
        build_tid_classes(w->tr, w->tid_classes);
        *out = w.release();
        return ORFX_OK;
    } catch (const std::exception& e) { g_err = e.what(); return ORFX_E_INVALID; }
}

int64_t orfxh_count(const orfxh_workload* w, int what) {
    if (!w) return ORFX_E_INVALID;
    const Transcriptome& tr = w->tr;
    switch (what) {
        case 0: return (int64_t)tr.bundles.size();
        case 1: return (int64_t)w->active.size();
        case 2: case 5: {
            int64_t n = 0;
            for (int bi : w->active) {
                const Bundle& b = tr.bundles[bi];
                for (int i = 0; i < b.count; i++) {
                    const Tx& t = tr.txs[b.first + i];
                    if (t.is_templ || (w->opt.keep_all_cds && t.is_coding)) continue;
                    n += what == 2 ? 1 : chain_len(t.exons);
                }
            }
            return n;
        }
        case 3: { int64_t n = 0; for (const Tx& t : tr.txs) n += t.is_templ; return n; }
        case 4: { int64_t n = 0; for (const Tx& t : tr.txs) n += !t.is_templ; return n; }
    }
    g_err = "unknown counter";
    return ORFX_E_INVALID;
}

int orfxh_pack(orfxh_workload* w, int shard, int n_shards, int layout, void* (*alloc)(size_t), void (*dealloc)(void*), orfx_batch* out) {
    if (!w || !out || n_shards <= 0 || shard < 0 || shard >= n_shards) { g_err = "bad shard arguments"; return ORFX_E_INVALID; }
    if ((alloc == nullptr) != (dealloc == nullptr)) { g_err = "alloc and dealloc must be given together"; return ORFX_E_INVALID; }
    try {
        std::vector<std::vector<int>> shards;
        shard_bundles(w->tr, w->active, n_shards, shards);
        if (w->packed.size() <= (size_t)shard) w->packed.resize((size_t)shard + 1);
        w->packed[(size_t)shard].reset(new PackedBatch());
        pack_bundles(w->tr, shards[(size_t)shard], w->opt.keep_all_cds, w->tid_classes, *w->packed[(size_t)shard], alloc, dealloc, layout != 0, PACK_EVAL, layout == 2,
                     w->opt.overhang, w->opt.spliced_overhang, layout == 3, layout == 4 || layout == 5, layout == 5);
        *out = w->packed[(size_t)shard]->view();
        return ORFX_OK;
    } catch (const std::exception& e) { g_err = e.what(); return ORFX_E_INVALID; }
}

int orfxh_truncate(orfxh_workload* w, int64_t n_active_bundles) {
    if (!w || n_active_bundles < 0) { g_err = "bad argument"; return ORFX_E_INVALID; }
    if ((size_t)n_active_bundles < w->active.size()) w->active.resize((size_t)n_active_bundles);
    return ORFX_OK;
}

const char* orfxh_tx_id(const orfxh_workload* w, int shard, int32_t tx) {
    if (!w || shard < 0 || (size_t)shard >= w->packed.size() || !w->packed[(size_t)shard]) return nullptr;
    const PackedBatch& pb = *w->packed[(size_t)shard];
    if (tx < 0 || (size_t)tx >= pb.tx_global.size()) return nullptr;
    return w->tr.txs[(size_t)pb.tx_global[(size_t)tx]].tid.c_str();
}

int64_t orfxh_tx_global(const orfxh_workload* w, int shard, int32_t* out, int64_t capacity) {
    if (!w || !out || shard < 0 || (size_t)shard >= w->packed.size() || !w->packed[(size_t)shard]) { g_err = "no packed batch for this shard"; return ORFX_E_INVALID; }
    const PackedBatch& pb = *w->packed[(size_t)shard];
    if ((int64_t)pb.tx_global.size() > capacity) { g_err = "output too small"; return ORFX_E_INVALID; }
    for (size_t i = 0; i < pb.tx_global.size(); i++) out[i] = pb.tx_global[i];
    return (int64_t)pb.tx_global.size();
}

void orfxh_free(orfxh_workload* w) { delete w; }

}  // extern "C"
