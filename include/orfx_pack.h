/*
 * orfx_pack.h — C-ABI of the host-side locus-bundle batcher.
 *
 * The reference has no such interface: loading, template preparation and bundling are the first
 * sixty lines of run() (orfanage.cpp:313-367; Transcriptome::add / build_cds_chains / sort /
 * set_cds_as_exons / remove_non_coding / deduplicate / bundleup, Transcriptome.cpp:866-1150).
 * This header exposes that prefix plus the batcher so that callers that are not
 * the `orfanage` binary — the benchmark, the parity tests, a binding in another language — can
 * obtain the flat orfx_batch that orfx_submit / orfx_upload (orfx.h) consume.
 *
 * The library behind it (orfanage_b200/liborfx_host.so) is plain C++ and has no CUDA dependency;
 * it never evaluates anything.
 */
#ifndef ORFX_PACK_H
#define ORFX_PACK_H

#include "orfx.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orfxh_workload orfxh_workload;

const char* orfxh_last_error(void);

/* Loads the template GTF/GFFs, the query GTF and (optionally, may be NULL) the genome FASTA,
 * prepares the templates and bundles everything exactly as orfanage.cpp:313-367 does.
 * use_id / keep_all_cds are the CLI flags of the same names. */
int orfxh_load(orfxh_workload** out, const char* query_gtf, const char* const* template_gtfs, int n_templates,
               const char* reference_fasta, int use_id, int keep_all_cds);

/* The same with --overhang N [--spliced_overhang]: bundle windows are padded by N (Transcriptome.cpp:978-991) and
 * the batches of orfxh_pack carry tx_ext_lo / tx_ext_hi (orfx.h) and the flank bases. */
int orfxh_load_overhang(orfxh_workload** out, const char* query_gtf, const char* const* template_gtfs, int n_templates,
                        const char* reference_fasta, int use_id, int keep_all_cds, int overhang, int spliced_overhang);

/* Counts over the whole workload.  what: 0 = bundles, 1 = bundles holding a template and a query
 * ("active"), 2 = query transcripts in active bundles, 3 = template transcripts, 4 = all query
 * transcripts, 5 = spliced query nucleotides in active bundles. */
int64_t orfxh_count(const orfxh_workload* w, int what);

/* Splits the active bundles over n_shards (longest-processing-time greedy) and packs
 * shard `shard`.  layout: 0 = whole locus windows, 1 = islands (only the bases under query exons; see
 * orfx_batch.seq in orfx.h), 2 = islands as 4-bit base codes (ORFX_SEQ_NT4), 3 = islands as raw bytes that also cover the template
 * CDS segments (what orfx_params::align_enabled needs), 4 = islands as 2-bit base codes plus an exception list (ORFX_SEQ_NT2), 5 = layout 4 with the segment dictionary (orfx_batch::seg_ref;
 * the per-segment arrays are still there for host readers, the device library takes the chains from the dictionary).  The arrays behind *out are owned by the workload and stay valid until the next
 * orfxh_pack with the same shard index or orfxh_free.  alloc/dealloc (both NULL = malloc/free) let
 * the caller place the arrays in pinned memory (orfx_host_alloc / orfx_host_free of orfx.h). */
int orfxh_pack(orfxh_workload* w, int shard, int n_shards, int layout, void* (*alloc)(size_t), void (*dealloc)(void*), orfx_batch* out);

/* Keeps only the first n active bundles of the workload (bounded samples for CPU baselines). */
int orfxh_truncate(orfxh_workload* w, int64_t n_active_bundles);

/* transcript_id of transcript `tx` of the batch last packed for `shard`; NULL when out of range. */
const char* orfxh_tx_id(const orfxh_workload* w, int shard, int32_t tx);

/* Workload-wide index (position in the loaded, sorted transcriptome: the same on every rank that loaded the same files)
 * of every transcript of the batch last packed for `shard`: out[n_tx].  Lets results of different shards be merged or
 * compared without going through transcript_id strings.  Returns n_tx or a negative error. */
int64_t orfxh_tx_global(const orfxh_workload* w, int shard, int32_t* out, int64_t capacity);

void orfxh_free(orfxh_workload* w);

#ifdef __cplusplus
}
#endif
#endif /* ORFX_PACK_H */
