// Batcher, multi-GPU locus sharding and GTF / stats emission around the C-ABI.
// Replaces run() of the reference (orfanage.cpp:307-822); the evaluation itself (the body of the
// bundle loop) happens behind Backend::submit / wait.
#include "orfx_host.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <functional>
#include <future>
#include <thread>
#include <unordered_map>

namespace orfx {

std::string float_to_string(float v) { char b[64]; snprintf(b, sizeof b, "%f", (double)v); return b; }
std::string float_ostream(float v) { char b[64]; snprintf(b, sizeof b, "%g", (double)v); return b; }

char nt_complement(char c) {   // ntComplement (gclib), IUPAC, case preserving
    static char table[256]; static bool init = false;
    if (!init) {
        const char* defs = "AaCcTtGgUuMmRrWwSsYyKkVvHhDdBbNnXx-*";
        const char* comp = "TtGgAaCcAaKkYyWwSsRrMmBbDdHhVvNnXx-*";
        for (int i = 0; i < 256; i++) table[i] = (char)i;
        for (int i = 0; defs[i]; i++) table[(unsigned char)defs[i]] = comp[i];
        init = true;
    }
    return table[(unsigned char)c];
}

void PackedBatch::reset() {
    if (arena) { if (dealloc) dealloc(arena); else free(arena); }
    if (dict_arena) { if (dealloc) dealloc(dict_arena); else free(dict_arena); }
    dict_arena = nullptr; seg_ref = nullptr; bundle_dict_off = nullptr; dict_segs = nullptr; dict_phase = nullptr; dict_seq_off = nullptr; n_dict = 0;
    arena = nullptr; arena_bytes = 0; dealloc = nullptr;
    bundle_ids.clear(); tx_global.clear();
    n_tx = 0; n_segs = 0; seq_bytes = 0; n_queries = 0; seq_format = ORFX_SEQ_RAW;
    tx_ext_lo = tx_ext_hi = nullptr; tx_ext_flags = nullptr;
    seq_exc_pos.clear(); seq_exc_code.clear();
}

orfx_batch PackedBatch::view() const {
    orfx_batch b;
    memset(&b, 0, sizeof b);
    b.n_bundles = (int32_t)bundle_ids.size();
    b.n_tx = n_tx;
    b.n_segs = n_segs;
    b.seq_bytes = seq_bytes;
    b.seq = seq_bytes ? seq : nullptr;
    b.seq_off = seq_off;
    b.bundle_start = bundle_start;
    b.bundle_tx_off = bundle_tx_off;
    b.tx_kind = tx_kind;
    b.tx_strand = tx_strand;
    b.tx_seg_off = tx_seg_off;
    b.tx_cds_start = tx_cds_start;
    b.tx_cds_end = tx_cds_end;
    b.tx_tid_class = tx_tid_class;
    b.segs = segs;
    b.seg_phase = seg_phase;
    b.seg_seq_off = seq_bytes ? seg_seq_off : nullptr;
    b.seq_format = seq_format;
    b.tx_ext_lo = tx_ext_lo;
    b.tx_ext_hi = tx_ext_hi;
    b.tx_ext_flags = tx_ext_flags;
    b.seq_exc_pos = seq_exc_pos.empty() ? nullptr : seq_exc_pos.data();
    b.seq_exc_code = seq_exc_code.empty() ? nullptr : seq_exc_code.data();
    b.n_seq_exc = (int64_t)seq_exc_pos.size();
    if (seg_ref && n_dict > 0) {
        b.seg_ref = seg_ref; b.bundle_dict_off = bundle_dict_off; b.dict_segs = dict_segs; b.dict_phase = dict_phase;
        b.dict_seq_off = b.seg_seq_off ? dict_seq_off : nullptr; b.n_dict = n_dict;
    }
    return b;
}

void build_tid_classes(const Transcriptome& tr, TidClasses& out) {
    std::unordered_map<std::string, int> ids;
    ids.reserve(tr.txs.size() * 2);
    out.resize(tr.txs.size());
    for (size_t i = 0; i < tr.txs.size(); i++) out[i] = ids.emplace(tr.txs[i].tid, (int)ids.size()).first->second;
}

namespace {
// ORFX_SEQ_NT4 codes of the 256 byte values (include/orfx.h)
struct Nt4Table {
    uint8_t code[256];
    Nt4Table() {
        memset(code, ORFX_NT4_OTHER, sizeof code);
        const char* letters = "ACGTURY";
        const uint8_t codes[] = {ORFX_NT4_A, ORFX_NT4_C, ORFX_NT4_G, ORFX_NT4_T, ORFX_NT4_U, ORFX_NT4_R, ORFX_NT4_Y};
        for (int i = 0; letters[i]; i++) { code[(unsigned char)letters[i]] = codes[i]; code[(unsigned char)(letters[i] | 0x20)] = codes[i]; }
    }
};
const Nt4Table NT4;
// ORFX_SEQ_NT2 codes: A C G T (either case) -> 0..3, every other byte 0xFF (listed as an exception with its NT4 code)
struct Nt2Table {
    uint8_t code[256];
    Nt2Table() {
        memset(code, 0xFF, sizeof code);
        const char* letters = "ACGT";
        for (int i = 0; letters[i]; i++) { code[(unsigned char)letters[i]] = (uint8_t)i; code[(unsigned char)(letters[i] | 0x20)] = (uint8_t)i; }
    }
};
const Nt2Table NT2;
// Islands of one bundle: the union of its query exons (what the path reads from the genome when
// --overhang is 0).  Returns the merged intervals, ascending.
inline uint8_t tx_kind_of(const Tx& t, bool keep_all_cds, bool clean_mode) {
    if (clean_mode) return (t.is_coding && !t.cds.empty()) ? ORFX_TX_QUERY : ORFX_TX_SKIP;
    return t.is_templ ? ORFX_TX_TEMPLATE : ((keep_all_cds && t.is_coding) ? ORFX_TX_SKIP : ORFX_TX_QUERY);
}
// --overhang flanks of one query transcript, as the batch carries them (orfx_batch::tx_ext_lo / tx_ext_hi / tx_ext_flags).
struct QFlank {
    int32_t lo = 0, hi = 0;      // bases stored right before the first / right after the last exon's bases
    uint8_t flags = 0;           // ORFX_EXT_*
    bool private_copy = false;   // strand '.' with a 5' flank: the transcript gets its own copy of its bases (see below)
    int fake_from = 0;           // ... and the `hi` bases after its last exon are genomic [fake_from, fake_from + hi)
};
// ext (per transcript of the bundle, may be null): the flanks widen the terminal exons; transcripts with a private copy
// are left out of the shared islands.
void bundle_islands(const Transcriptome& tr, const Bundle& b, bool keep_all_cds, bool clean_mode, std::vector<std::pair<int, int>>& isl,
                    const QFlank* ext = nullptr, bool template_seq = false) {
    isl.clear();
    for (int i = 0; i < b.count; i++) {
        const Tx& t = tr.txs[b.first + i];
        const uint8_t kind = tx_kind_of(t, keep_all_cds, clean_mode);
        if (template_seq && kind == ORFX_TX_TEMPLATE) { for (const Seg& s : t.cds) isl.emplace_back(s.s, s.e); continue; }   // the alignment stage translates the templates
        if (kind != ORFX_TX_QUERY) continue;
        if (ext && ext[i].private_copy) continue;
        const size_t first = isl.size();
        for (const Seg& s : t.exons) isl.emplace_back(s.s, s.e);
        if (ext && isl.size() > first) { isl[first].first -= ext[i].lo; isl.back().second += ext[i].hi; }
    }
    std::sort(isl.begin(), isl.end());
    size_t k = 0;
    for (size_t i = 0; i < isl.size(); i++) {
        if (k > 0 && isl[i].first <= isl[k - 1].second + 1) isl[k - 1].second = std::max(isl[k - 1].second, isl[i].second);
        else isl[k++] = isl[i];
    }
    isl.resize(k);
}
// Bases [pos0, pos0 + len) (0-based) of a record; positions outside it read the record's first / last base.  The only such
// reads are the one-base "flank" of a transcript that touches the end of its chromosome (see query_flanks).
void fetch_clamped(const Fasta& fa, const Fasta::Seq& chrom, int64_t pos0, int64_t len, char* dst) {
    int64_t a = pos0, n = len;
    while (n > 0 && a < 0) { *dst++ = fa.at(chrom, 0); a++; n--; }
    const int64_t inside = std::max<int64_t>(0, std::min<int64_t>(n, chrom.length - a));
    if (inside > 0) fa.fetch(chrom, a, inside, dst);
    for (int64_t i = inside; i < n; i++) dst[i] = fa.at(chrom, chrom.length - 1);
}
// The string TX::extend_seq (Transcriptome.cpp:342-399) adds on one end of `t`, described by where its bases lie.
//   side    high when (strand == '+' && towards_stop) || (strand == '-' && !towards_stop), low otherwise -- so a strand '.'
//           transcript takes BOTH flanks from the low side, exactly as written (neither '+' / '-' test fires for it);
//   clamp   [min(end + 1, bend), min(end + N, bend)] resp. [max(start - N, bstart), max(start - 1, bstart)]: a transcript that
//           touches the (padded) bundle window -- only possible at an end of the chromosome -- re-reads its own edge base as a
//           one-base flank, a position that does not exist (`virt`);
//   splice  with --spliced_overhang the strand-corrected, upper-cased string is cut before its first GT (towards the stop) or
//           after its last AG (towards the start).
// n bases survive; `nearest` says whether they are the n bases next to the exon (every case but the 5' flank of a strand '.'
// transcript, whose reverse-complemented low-side string is cut from the far end).
struct RefFlank { int es = 0, ee = -1, n = 0; bool high = false, nearest = true, virt = false; };
RefFlank reference_flank(const Fasta& fa, const Fasta::Seq& chrom, const Bundle& b, const Tx& t, bool towards_stop, int overhang, bool spliced) {
    RefFlank f;
    if (overhang <= 0 || t.exons.empty()) return f;
    f.high = (t.strand == '+' && towards_stop) || (t.strand == '-' && !towards_stop);
    if (f.high) { f.es = std::min(t.exons.back().e + 1, b.bend); f.ee = std::min(t.exons.back().e + overhang, b.bend); f.virt = t.exons.back().e >= b.bend; }
    else { f.es = std::max(t.exons.front().s - overhang, b.bstart); f.ee = std::max(t.exons.front().s - 1, b.bstart); f.virt = t.exons.front().s <= b.bstart; }
    const int len = f.ee - f.es + 1;
    f.n = len;
    f.nearest = towards_stop || t.strand == '+' || t.strand == '-';
    if (!spliced || len <= 0) return f;
    std::string g((size_t)len, 'N');
    fa.fetch(chrom, f.es - 1, len, &g[0]);
    std::string s5;   // 5'->3' on the transcript's strand
    if (t.strand == '+') s5 = g; else for (int i = len - 1; i >= 0; i--) s5 += nt_complement(g[(size_t)i]);
    for (char& c : s5) c = (char)toupper((unsigned char)c);
    if (towards_stop) { const size_t at = s5.find("GT"); if (at != std::string::npos) f.n = (int)at; }
    else { const size_t at = s5.rfind("AG"); if (at != std::string::npos) f.n = len - (int)(at + 2); }
    return f;
}
// What the batch carries for one query (include/orfx.h, orfx_batch::tx_ext_*).  The device reads a flank as the bases stored
// next to the terminal exon on the side the strand byte implies ('+': 5' below, 3' above; any other strand: 5' above, 3'
// below).  That is where the reference takes them from for '+' and '-'.  A strand '.' transcript takes its 5' flank from the
// LOW side: it gets a private copy of its bases in which those low-side bases are stored above its last exon, in ascending
// order -- read downwards and complemented they are the string the reference prepends.
QFlank query_flanks(const Fasta& fa, const Fasta::Seq& chrom, const Bundle& b, const Tx& t, int overhang, bool spliced) {
    QFlank q;
    const RefFlank f3 = reference_flank(fa, chrom, b, t, true, overhang, spliced), f5 = reference_flank(fa, chrom, b, t, false, overhang, spliced);
    (f3.high ? q.hi : q.lo) = f3.n;
    if (f3.virt && f3.n > 0) q.flags |= f3.high ? ORFX_EXT_HI_VIRTUAL : ORFX_EXT_LO_VIRTUAL;
    if (f5.high != f3.high) (f5.high ? q.hi : q.lo) = f5.n;   // '+' and '-': the other side, bases next to the exon
    else if (f5.n > 0) { q.hi = f5.n; q.private_copy = true; q.fake_from = f5.es; }   // strand '.': the far end of the low-side string
    return q;
}
}  // namespace

// Two passes: size everything, take one arena from the allocator, fill.  Sequence bytes (the bulk) are
// copied by several threads.  islands = true uploads only the bases under query exons.
void pack_bundles(const Transcriptome& tr, const std::vector<int>& bundle_ids, bool keep_all_cds,
                  const TidClasses& tid_classes, PackedBatch& out,
                  void* (*alloc)(size_t), void (*dealloc)(void*), bool islands, int mode, bool nt4, int overhang, bool spliced_overhang, bool template_seq,
                  bool nt2, bool dict) {
    const bool clean_mode = mode == PACK_CLEAN, compare_mode = mode == PACK_COMPARE;
    template_seq = template_seq && mode == PACK_EVAL;
    if (template_seq) nt4 = nt2 = false;   // the translation needs the letters themselves
    out.reset();
    out.bundle_ids = bundle_ids;
    const size_t nb = bundle_ids.size();
    const bool has_seq = tr.fasta != nullptr && !compare_mode;
    islands = islands && has_seq;
    nt2 = nt2 && islands;
    nt4 = nt4 && islands && !nt2;
    if (!has_seq) overhang = 0;
    size_t total_seq = 0, total_tx = 0, total_segs = 0;
    std::vector<int64_t> bundle_bytes(nb, 0);
    std::vector<std::pair<int, int>> isl;
    // --overhang: flanks of every query, per bundle in transcript order
    std::vector<std::vector<QFlank>> fl(overhang > 0 ? nb : 0);
    if (overhang > 0) {
        for (size_t j = 0; j < nb; j++) {
            const Bundle& b = tr.bundles[bundle_ids[j]];
            const Fasta::Seq* chrom = tr.fasta->find(tr.seqnames[(size_t)b.seqid]);
            fl[j].assign((size_t)b.count, QFlank());
            for (int i = 0; i < b.count; i++) {
                const Tx& t = tr.txs[b.first + i];
                if (tx_kind_of(t, keep_all_cds, clean_mode) != ORFX_TX_QUERY) continue;
                fl[j][(size_t)i] = query_flanks(*tr.fasta, *chrom, b, t, overhang, spliced_overhang);
                if (fl[j][(size_t)i].private_copy && !islands)
                    throw std::runtime_error("--overhang with strand '.' transcripts needs the islands layout (their 5' flank is a private copy of low-side bases)");
            }
        }
    }
    auto flk = [&](size_t j) { return overhang > 0 ? fl[j].data() : nullptr; };
    // bases of the private copy of transcript i of bundle j: [lo flank][exons][hi "flank"]
    auto private_len = [&](size_t j, int i, const Tx& t) { return (int64_t)fl[j][(size_t)i].lo + chain_len(t.exons) + fl[j][(size_t)i].hi; };
    // whole-window layout with --overhang: one more base on either side, so that the one-base flank of a transcript that
    // touches the end of its chromosome (a re-read of the edge base) has a place
    const int wpad = (!islands && overhang > 0) ? 1 : 0;
    // sizes per bundle (bases, transcripts, segments), by several threads: the island merge is a sort per bundle
    std::vector<int32_t> bundle_ntx(nb, 0), bundle_nseg(nb, 0);
    auto over_bundles = [&](const std::function<void(size_t, size_t)>& f) {
        const unsigned nthr = nb > 2000 ? std::min(8u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
        if (nthr <= 1) { f(0, nb); return; }
        std::vector<std::thread> th;
        for (unsigned k = 0; k < nthr; k++) th.emplace_back(f, nb * k / nthr, nb * (k + 1) / nthr);
        for (auto& t : th) t.join();
    };
    over_bundles([&](size_t j0, size_t j1) {
        std::vector<std::pair<int, int>> il;
        for (size_t j = j0; j < j1; j++) {
            const Bundle& b = tr.bundles[bundle_ids[j]];
            if (has_seq) {
                if (islands) {
                    bundle_islands(tr, b, keep_all_cds, clean_mode, il, flk(j), template_seq);
                    for (auto& iv : il) bundle_bytes[j] += iv.second - iv.first + 1;
                    if (overhang > 0) for (int i = 0; i < b.count; i++) if (fl[j][(size_t)i].private_copy) bundle_bytes[j] += private_len(j, i, tr.txs[b.first + i]);
                    if (nt4) bundle_bytes[j] += bundle_bytes[j] & 1;   // bundles start on a byte boundary
                    if (nt2) bundle_bytes[j] = (bundle_bytes[j] + 3) & ~(int64_t)3;
                } else bundle_bytes[j] = b.bend - b.bstart + 1 + 2 * wpad;
            }
            bundle_ntx[j] = b.count;
            int64_t ns = 0;
            for (int i = 0; i < b.count; i++) { const Tx& t = tr.txs[b.first + i]; ns += (int64_t)(((t.is_templ && !clean_mode) || compare_mode) ? t.cds.size() : t.exons.size()); }
            bundle_nseg[j] = (int32_t)std::min<int64_t>(ns, INT32_MAX);
        }
    });
    for (size_t j = 0; j < nb; j++) { total_seq += (size_t)bundle_bytes[j]; total_tx += (size_t)bundle_ntx[j]; total_segs += (size_t)bundle_nseg[j]; }
    if (total_tx > (size_t)INT32_MAX || total_segs > (size_t)INT32_MAX) throw std::runtime_error("batch too large for 32-bit transcript / segment indices: shard it");
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    size_t o_seq = 0, o_seq_off = o_seq + up((nt2 ? (total_seq + 3) / 4 : nt4 ? (total_seq + 1) / 2 : total_seq) + 64), o_bstart = o_seq_off + up((nb + 1) * 8), o_btx = o_bstart + up(nb * 4 + 4),
           o_kind = o_btx + up((nb + 1) * 4), o_strand = o_kind + up(total_tx + 4), o_segoff = o_strand + up(total_tx + 4),
           o_cs = o_segoff + up((total_tx + 1) * 4), o_ce = o_cs + up(total_tx * 4 + 4), o_tid = o_ce + up(total_tx * 4 + 4),
           o_segs = o_tid + up(total_tx * 4 + 4), o_phase = o_segs + up(total_segs * 8 + 8), o_sso = o_phase + up(total_segs + 4),
           o_xlo = o_sso + up(islands ? total_segs * 4 + 4 : 4), o_xhi = o_xlo + up(overhang > 0 ? total_tx * 4 + 4 : 4),
           o_xfl = o_xhi + up(overhang > 0 ? total_tx * 4 + 4 : 4), o_end = o_xfl + up(overhang > 0 ? total_tx + 4 : 4);
    out.arena = alloc ? alloc(o_end) : malloc(o_end);
    if (!out.arena) throw std::runtime_error("cannot allocate " + std::to_string(o_end) + " bytes for the packed batch");
    out.arena_bytes = o_end;
    out.dealloc = alloc ? dealloc : nullptr;
    char* base = (char*)out.arena;
    out.seq = (uint8_t*)(base + o_seq); out.seq_off = (int64_t*)(base + o_seq_off); out.bundle_start = (int32_t*)(base + o_bstart);
    out.bundle_tx_off = (int32_t*)(base + o_btx); out.tx_kind = (uint8_t*)(base + o_kind); out.tx_strand = (uint8_t*)(base + o_strand);
    out.tx_seg_off = (int32_t*)(base + o_segoff); out.tx_cds_start = (int32_t*)(base + o_cs); out.tx_cds_end = (int32_t*)(base + o_ce);
    out.tx_tid_class = (int32_t*)(base + o_tid); out.segs = (orfx_seg*)(base + o_segs); out.seg_phase = (int8_t*)(base + o_phase);
    out.seg_seq_off = islands ? (int32_t*)(base + o_sso) : nullptr;
    out.tx_ext_lo = overhang > 0 ? (int32_t*)(base + o_xlo) : nullptr;
    out.tx_ext_hi = overhang > 0 ? (int32_t*)(base + o_xhi) : nullptr;
    out.tx_ext_flags = overhang > 0 ? (uint8_t*)(base + o_xfl) : nullptr;
    out.n_tx = (int32_t)total_tx; out.n_segs = (int64_t)total_segs; out.seq_bytes = (int64_t)total_seq;
    out.seq_format = nt2 ? ORFX_SEQ_NT2 : nt4 ? ORFX_SEQ_NT4 : ORFX_SEQ_RAW;
    out.tx_global.resize(total_tx);

    // offsets per bundle, then the per-transcript / per-segment arrays by several threads
    std::vector<int32_t> x_at(nb + 1, 0), sg_at(nb + 1, 0);
    out.seq_off[0] = 0; out.bundle_tx_off[0] = 0; out.tx_seg_off[0] = 0;
    for (size_t j = 0; j < nb; j++) {
        out.seq_off[j + 1] = out.seq_off[j] + bundle_bytes[j];
        x_at[j + 1] = x_at[j] + bundle_ntx[j];
        sg_at[j + 1] = sg_at[j] + bundle_nseg[j];
        out.bundle_tx_off[j + 1] = x_at[j + 1];
    }
    std::vector<int64_t> nq_part(nb, 0);
    over_bundles([&](size_t j0, size_t j1) {
        for (size_t j = j0; j < j1; j++) {
            const Bundle& b = tr.bundles[bundle_ids[j]];
            out.bundle_start[j] = b.bstart - wpad;
            int32_t x = x_at[j], sg = sg_at[j];
            for (int i = 0; i < b.count; i++, x++) {
                const Tx& t = tr.txs[b.first + i];
                out.tx_global[(size_t)x] = b.first + i;
                const uint8_t kind = tx_kind_of(t, keep_all_cds, clean_mode);
                const bool as_templ = (t.is_templ && !clean_mode) || compare_mode;   // carries its CDS chain and bounds
                if (kind == ORFX_TX_QUERY) nq_part[j]++;
                out.tx_kind[x] = kind;
                out.tx_strand[x] = (uint8_t)t.strand;
                const Chain& c = as_templ ? t.cds : t.exons;
                for (const Seg& sgm : c) { out.segs[sg] = orfx_seg{sgm.s, sgm.e}; out.seg_phase[sg] = as_templ ? sgm.phase : 0; sg++; }
                out.tx_seg_off[x + 1] = sg;
                out.tx_cds_start[x] = (as_templ || (clean_mode && kind == ORFX_TX_QUERY)) ? t.cds_start : 0;
                out.tx_cds_end[x] = (as_templ || (clean_mode && kind == ORFX_TX_QUERY)) ? t.cds_end : 0;
                out.tx_tid_class[x] = tid_classes[(size_t)(b.first + i)];
                if (overhang > 0) { const QFlank& f = fl[j][(size_t)i]; out.tx_ext_lo[x] = f.lo; out.tx_ext_hi[x] = f.hi; out.tx_ext_flags[x] = f.flags; }
            }
        }
    });
    for (size_t j = 0; j < nb; j++) out.n_queries += nq_part[j];
    if (has_seq && total_seq) {   // Bundle::load_seq, cpp:821-830
        std::vector<const Fasta::Seq*> chrom(tr.seqnames.size(), nullptr);
        for (size_t k = 0; k < chrom.size(); k++) chrom[k] = tr.fasta->find(tr.seqnames[k]);
        std::vector<std::vector<int64_t>> exc_pos_part(16);
        std::vector<std::vector<uint8_t>> exc_code_part(16);
        auto copy_range = [&](size_t j0, size_t j1, unsigned part) {
            std::vector<std::pair<int, int>> il;
            std::vector<int32_t> il_off;
            std::vector<char> raw;
            for (size_t j = j0; j < j1; j++) {
                const Bundle& b = tr.bundles[bundle_ids[j]];
                const Fasta::Seq& src = *chrom[(size_t)b.seqid];
                uint8_t* dst = nt2 ? out.seq + out.seq_off[j] / 4 : nt4 ? out.seq + out.seq_off[j] / 2 : out.seq + out.seq_off[j];
                if (nt2) memset(dst, 0, (size_t)(bundle_bytes[j] / 4));
                if (!islands) {
                    if (wpad) fetch_clamped(*tr.fasta, src, (int64_t)b.bstart - 2, (int64_t)b.bend - b.bstart + 3, (char*)dst);
                    else tr.fasta->fetch(src, b.bstart - 1, b.bend - b.bstart + 1, (char*)dst);
                    continue;
                }
                int32_t o = 0;
                // appends genomic [g0, g0 + len) (1-based; clamped reads past a chromosome end) at base offset o of the bundle
                auto append = [&](int64_t g0, int len) {
                    if (len <= 0) return;
                    raw.resize((size_t)len);
                    fetch_clamped(*tr.fasta, src, g0 - 1, len, raw.data());
                    if (nt2) {
                        const uint8_t* sp = (const uint8_t*)raw.data();
                        auto one = [&](int i, int p) {
                            const uint8_t c2 = NT2.code[sp[i]];
                            if (c2 < 4) dst[p >> 2] = (uint8_t)(dst[p >> 2] | (c2 << ((p & 3) * 2)));
                            else { exc_pos_part[part].push_back(out.seq_off[j] + p); exc_code_part[part].push_back(NT4.code[sp[i]]); }
                        };
                        int i = 0, p = o;
                        for (; i < len && (p & 3); i++, p++) one(i, p);
                        for (; i + 4 <= len; i += 4, p += 4) {   // whole bytes
                            const uint8_t c0 = NT2.code[sp[i]], c1 = NT2.code[sp[i + 1]], c2 = NT2.code[sp[i + 2]], c3 = NT2.code[sp[i + 3]];
                            if (!((c0 | c1 | c2 | c3) & 0x80)) dst[p >> 2] = (uint8_t)(c0 | (c1 << 2) | (c2 << 4) | (c3 << 6));
                            else for (int k = 0; k < 4; k++) one(i + k, p + k);
                        }
                        for (; i < len; i++, p++) one(i, p);
                    } else if (!nt4) memcpy((char*)dst + o, raw.data(), (size_t)len);
                    else {
                        const uint8_t* sp = (const uint8_t*)raw.data();
                        int i = 0, p = o;
                        if (p & 1) { dst[p >> 1] = (uint8_t)(dst[p >> 1] | (NT4.code[sp[0]] << 4)); i = 1; p++; }
                        for (; i + 1 < len; i += 2, p += 2) dst[p >> 1] = (uint8_t)(NT4.code[sp[i]] | (NT4.code[sp[i + 1]] << 4));
                        if (i < len) dst[p >> 1] = NT4.code[sp[i]];   // low nibble; the high one is OR-ed in by whatever follows
                    }
                    o += len;
                };
                bundle_islands(tr, b, keep_all_cds, clean_mode, il, flk(j), template_seq);
                il_off.resize(il.size());
                for (size_t k = 0; k < il.size(); k++) { il_off[k] = o; append(il[k].first, il[k].second - il[k].first + 1); }
                // offset of every query segment: the island holding its start, or the transcript's private copy
                for (int xx = out.bundle_tx_off[j]; xx < out.bundle_tx_off[j + 1]; xx++) {
                    const bool isq = out.tx_kind[xx] == ORFX_TX_QUERY;
                    const QFlank* f = overhang > 0 ? &fl[j][(size_t)(xx - out.bundle_tx_off[j])] : nullptr;
                    if (isq && f && f->private_copy) {
                        const int k0 = out.tx_seg_off[xx], k1 = out.tx_seg_off[xx + 1];
                        append((int64_t)out.segs[k0].start - f->lo, f->lo);
                        for (int k = k0; k < k1; k++) { out.seg_seq_off[k] = o; append(out.segs[k].start, out.segs[k].end - out.segs[k].start + 1); }
                        append(f->fake_from, f->hi);
                        continue;
                    }
                    const bool ist = template_seq && out.tx_kind[xx] == ORFX_TX_TEMPLATE;
                    for (int k = out.tx_seg_off[xx]; k < out.tx_seg_off[xx + 1]; k++) {
                        if (!isq && !ist) { out.seg_seq_off[k] = 0; continue; }
                        const int st = out.segs[k].start;
                        size_t lo = 0, hi = il.size() - 1;
                        while (lo < hi) { size_t mid = (lo + hi + 1) >> 1; if (il[mid].first <= st) lo = mid; else hi = mid - 1; }
                        out.seg_seq_off[k] = il_off[lo] + (st - il[lo].first);
                    }
                }
            }
        };
        unsigned nthr = (total_seq > (32u << 20) || (islands && nb > 2000)) ? std::min(8u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
        nthr = std::min(nthr, 16u);
        if (nthr <= 1) copy_range(0, nb, 0);
        else {
            std::vector<std::thread> th;
            size_t j0 = 0;
            for (unsigned k = 0; k < nthr; k++) {   // equal byte shares
                const int64_t want = (int64_t)(total_seq * (k + 1) / nthr);
                size_t j1 = j0;
                while (j1 < nb && out.seq_off[j1 + 1] <= want) j1++;
                if (k == nthr - 1) j1 = nb;
                th.emplace_back(copy_range, j0, j1, k);
                j0 = j1;
            }
            for (auto& t : th) t.join();
        }
        memset(out.seq + (nt2 ? (total_seq + 3) / 4 : nt4 ? (total_seq + 1) / 2 : total_seq), 0, 64);
        for (size_t k = 0; k < exc_pos_part.size(); k++) {   // parts hold consecutive bundle ranges: positions stay ascending
            out.seq_exc_pos.insert(out.seq_exc_pos.end(), exc_pos_part[k].begin(), exc_pos_part[k].end());
            out.seq_exc_code.insert(out.seq_exc_code.end(), exc_code_part[k].begin(), exc_code_part[k].end());
        }
    }
    if (dict && out.n_segs > 0) build_segment_dictionary(out, alloc);
}

bool build_segment_dictionary(PackedBatch& out, void* (*alloc)(size_t)) {
    struct Rec {
        int32_t s, e, off; int8_t ph;
        bool operator<(const Rec& o) const { return s != o.s ? s < o.s : e != o.e ? e < o.e : ph != o.ph ? ph < o.ph : off < o.off; }
        bool operator==(const Rec& o) const { return s == o.s && e == o.e && ph == o.ph && off == o.off; }
    };
    const size_t nb = out.bundle_ids.size();
    const bool with_off = out.seq_bytes > 0 && out.seg_seq_off;
    std::vector<uint16_t> ref((size_t)out.n_segs);
    const unsigned nthr = nb > 2000 ? std::min(8u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
    std::vector<std::vector<Rec>> part_recs(nthr);
    std::vector<int32_t> count(nb + 1, 0);
    std::atomic<bool> overflow{false};
    auto work = [&](size_t j0, size_t j1, unsigned part) {
        std::vector<Rec> all, uniq;
        for (size_t j = j0; j < j1 && !overflow.load(std::memory_order_relaxed); j++) {
            const int k0 = out.tx_seg_off[out.bundle_tx_off[j]], k1 = out.tx_seg_off[out.bundle_tx_off[j + 1]];
            all.clear();
            for (int k = k0; k < k1; k++) all.push_back(Rec{out.segs[k].start, out.segs[k].end, with_off ? out.seg_seq_off[k] : 0, out.seg_phase[k]});
            uniq = all;
            std::sort(uniq.begin(), uniq.end());
            uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
            if (uniq.size() > 65536) { overflow = true; return; }
            for (int k = k0; k < k1; k++) ref[(size_t)k] = (uint16_t)(std::lower_bound(uniq.begin(), uniq.end(), all[(size_t)(k - k0)]) - uniq.begin());
            count[j] = (int32_t)uniq.size();
            part_recs[part].insert(part_recs[part].end(), uniq.begin(), uniq.end());
        }
    };
    if (nthr <= 1) work(0, nb, 0);
    else {
        std::vector<std::thread> th;
        for (unsigned k = 0; k < nthr; k++) th.emplace_back(work, nb * k / nthr, nb * (k + 1) / nthr, k);
        for (auto& t : th) t.join();
    }
    if (overflow) return false;
    int64_t nd = 0;
    for (size_t j = 0; j < nb; j++) nd += count[j];
    auto up = [](size_t x) { return (x + 63) & ~(size_t)63; };
    const size_t o_ref = 0, o_bdo = o_ref + up((size_t)out.n_segs * 2 + 8), o_ds = o_bdo + up((nb + 1) * 4), o_dp = o_ds + up((size_t)nd * 8 + 8),
                 o_do = o_dp + up((size_t)nd + 8), o_end = o_do + up((size_t)nd * 4 + 8);
    void* a = alloc ? alloc(o_end) : malloc(o_end);
    if (!a) throw std::runtime_error("cannot allocate " + std::to_string(o_end) + " bytes for the segment dictionary");
    char* base = (char*)a;
    out.dict_arena = a;
    out.seg_ref = (uint16_t*)(base + o_ref); out.bundle_dict_off = (int32_t*)(base + o_bdo); out.dict_segs = (orfx_seg*)(base + o_ds);
    out.dict_phase = (int8_t*)(base + o_dp); out.dict_seq_off = (int32_t*)(base + o_do); out.n_dict = nd;
    memcpy(out.seg_ref, ref.data(), (size_t)out.n_segs * 2);
    int32_t at = 0;
    for (size_t j = 0; j < nb; j++) { out.bundle_dict_off[j] = at; at += count[j]; }
    out.bundle_dict_off[nb] = at;
    size_t d = 0;
    for (const auto& recs : part_recs)   // parts hold consecutive bundle ranges
        for (const Rec& r : recs) { out.dict_segs[d] = orfx_seg{r.s, r.e}; out.dict_phase[d] = r.ph; out.dict_seq_off[d] = r.off; d++; }
    return true;
}

void shard_bundles(const Transcriptome& tr, const std::vector<int>& active, int n, std::vector<std::vector<int>>& shards) {
    shards.assign((size_t)n, std::vector<int>());
    if (n == 1) { shards[0] = active; return; }
    // cost ~ (#queries x #templates) pair tests + spliced query bases
    std::vector<std::pair<double, int>> cost;
    cost.reserve(active.size());
    for (int bi : active) {
        const Bundle& b = tr.bundles[bi];
        // Calibrated against the measured stage times of configs[1] and configs[3] on one B200 (profiles/r02_bench_1gpu.json,
        // r02_bench_c4_1gpu.json): the sequence stages cost ~1 unit per spliced query base; the row stages (candidate rows,
        // rescue + score, select) ~150 units per (query, template) pair and exon of the query; rows of heavy (titin-class)
        // bundles, which take a warp each, 2.5 times that.  (Round 1's model, pairs x 64 + bases, put a third of the time of
        // a configs[3] shard -- its titin-class loci -- at a sixth of the cost: LPT imbalance 1.11 on 8 GPUs.)
        double nq = 0, nt = 0, nts = 0, nex = 0;
        bool heavy = false;
        for (int i = 0; i < b.count; i++) {
            const Tx& t = tr.txs[b.first + i];
            const size_t segs = t.is_templ ? t.cds.size() : t.exons.size();
            heavy = heavy || segs > 32;
            if (t.is_templ) nt++; else { nq++; nts += chain_len(t.exons); nex += (double)t.exons.size(); }
        }
        cost.emplace_back(nts + 150.0 * nt * nex * (heavy ? 2.5 : 1.0), bi);
    }
    std::sort(cost.begin(), cost.end(), [](const std::pair<double, int>& a, const std::pair<double, int>& b) { return a.first != b.first ? a.first > b.first : a.second < b.second; });
    std::vector<double> load((size_t)n, 0.0);
    for (auto& c : cost) {
        int best = 0;
        for (int k = 1; k < n; k++) if (load[k] < load[best]) best = k;
        load[best] += c.first;
        shards[best].push_back(c.second);
    }
    for (auto& s : shards) std::sort(s.begin(), s.end());   // keep bundle order inside a shard
}

namespace {
const char* status_note(int st) {
    switch (st) {
        case ORFX_ST_OK: return "-";
        case ORFX_ST_NO_OVERLAP: return "no-overlap";
        case ORFX_ST_SEGMENT_LT3: return "segment_len<3";
        case ORFX_ST_CDS_LT3: return "cds_len<3";
        case ORFX_ST_NOT_RESCUED: return "not_rescued";
        case ORFX_ST_MISSING_STOP: return "missing_stop";
        case ORFX_ST_MISSING_START: return "missing_start";
        case ORFX_ST_MINLEN: return "minlen";
        case ORFX_ST_DUP: return "dup";
    }
    return "?";
}
const char* DASH19 = "-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-";
const char* KEEP19 = "-\t-\tkeep_cds\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-";

// The CDS chain of an emitted ORF: the query's exon chain clipped to [cs,ce] with GTF phases
// (what build_cds / trim_to_pos / extend leave behind; Transcriptome.cpp:92-122, 675-681).
void set_orf(Tx& q, int cs, int ce) {
    q.cds.clear();
    q.cds_start = cs;
    q.cds_end = ce;
    for (const Seg& e : q.exons) {
        if (e.e < cs || e.s > ce) continue;
        Seg c;
        c.s = std::max(e.s, cs);
        c.e = std::min(e.e, ce);
        q.cds.push_back(c);
    }
    assign_phase(q.cds, q.strand, 0);
}

struct Shard {
    PackedBatch batch;
    orfx_ctx* ctx = nullptr;
    orfx_results res;
    int rc = 0;
    std::string err;
    double ms_create = 0, ms_pack = 0, ms_eval = 0;
};
}  // namespace

// Everything upstream of the bundle loop (orfanage.cpp:313-367): load, prepare templates, bundle.
// Transcriptome::clean_cds (Transcriptome.cpp:1063-1100): bundle the templates, then per coding template
// load_seq (+ rescue_cds) and the start / stop / length checks — that per-transcript body runs behind
// Backend::clean on the device; the host applies the verdicts (TX::remove_cds or the rescued CDS bounds).
static size_t clean_cds(const Options& opt, Transcriptome& tr, const Backend& be) {
    if (!be.clean) throw std::runtime_error("--cleant needs a backend with template cleaning (orfx_clean_templates)");
    tr.bundleup(opt.use_id, (uint32_t)opt.overhang);
    std::vector<int> ids;
    for (int bi = 0; bi < (int)tr.bundles.size(); bi++) ids.push_back(bi);
    TidClasses tid_classes;
    build_tid_classes(tr, tid_classes);
    PackedBatch pb;
    pack_bundles(tr, ids, false, tid_classes, pb, be.host_alloc, be.host_free, true, PACK_CLEAN, true, opt.overhang, opt.spliced_overhang);
    orfx_params params;
    fill_params(opt, false, params);
    orfx_ctx* ctx = nullptr;
    if (be.create(&ctx, opt.devices.empty() ? 0 : opt.devices[0], &params)) throw std::runtime_error(std::string("template cleaning: ") + be.last_error());
    std::vector<uint8_t> keep((size_t)pb.n_queries + 1);
    std::vector<int32_t> cs((size_t)pb.n_queries + 1), ce((size_t)pb.n_queries + 1);
    orfx_clean_out out{keep.data(), cs.data(), ce.data()};
    int64_t n = 0;
    const orfx_batch view = pb.view();
    const int rc = be.clean(ctx, &view, opt.rescue ? 1 : 0, &out, &n);
    const std::string err = rc ? be.last_error() : "";
    be.destroy(ctx);
    if (rc) throw std::runtime_error("template cleaning: " + err);
    if (n != pb.n_queries) throw std::runtime_error("template cleaning: result count mismatch");
    size_t dirty = 0, k = 0;
    for (int32_t x = 0; x < pb.n_tx; x++) {
        if (pb.tx_kind[x] != ORFX_TX_QUERY) continue;
        Tx& t = tr.txs[(size_t)pb.tx_global[(size_t)x]];
        if (!keep[k]) { t.remove_cds(); dirty++; }
        else if (cs[k] != t.cds_start || ce[k] != t.cds_end) {
            // --overhang: a rescued CDS may reach into the flank; the terminal exons grow with it (cpp:418-429, 461-474;
            // the template's exon chain is replaced by its CDS chain right after, orfanage.cpp:351)
            if (!t.exons.empty()) { t.exons.front().s = std::min(t.exons.front().s, cs[k]); t.exons.back().e = std::max(t.exons.back().e, ce[k]); }
            t.cds_start = cs[k]; t.cds_end = ce[k]; t.cds_phase = 0; t.build_cds();
        }
        k++;
    }
    return dirty;
}

void prepare_transcriptome(const Options& opt, Fasta& fasta, Transcriptome& tr, bool verbose, const Backend* be) {
    if (opt.overhang < 0) throw std::runtime_error("--overhang must not be negative");
    if (opt.clean_templ && !be) throw std::runtime_error("--cleant needs a compute backend");
    typedef std::chrono::steady_clock clk;
    auto last = clk::now();
    std::string timing;
    auto say = [&](const char* m) {
        if (verbose) fprintf(stderr, "%s\n", m);
        if (opt.timing) {   // time since the previous message = the stage that message announced
            const auto now = clk::now();
            char b[64]; snprintf(b, sizeof b, " %.0f", std::chrono::duration<double, std::milli>(now - last).count());
            timing += b; last = now;
        }
    };
    // The query GTF is parsed and converted to transcripts on the side while the templates are loaded, sorted, cleaned and deduplicated (the two files meet
    // in bundleup); its records join the transcriptome where the reference loads them, and a parse error of the query surfaces
    // there too -- after any error of the template stages, as in the reference.
    struct QueryTxs { ParsedTxs parsed; std::exception_ptr err; };
    std::future<QueryTxs> query_parse = std::async(std::launch::async, [&opt]() {
        QueryTxs q;
        try {
            std::vector<GtfRecord> recs;
            read_gtf(opt.query, recs);
            convert_records(recs, false, false, q.parsed);
        } catch (...) { q.err = std::current_exception(); }
        return q;
    });
    say("loading reference genome");
    if (!opt.reference.empty()) {
        fasta.load(opt.reference);
        if (fasta.write_fai_if_missing(opt.reference)) say("FASTA index rebuilt.");
        tr.fasta = &fasta;
    }
    say("loading reference transcriptomes");
    for (const std::string& k : opt.templates) tr.add(k, true, true);
    tr.build_cds_chains();
    if (opt.clean_templ || !opt.reference.empty()) tr.correct_chain_len();
    if (opt.clean_templ) tr.clean_short_orfs(opt.minlen);
    say("sorting reference transcriptome");
    tr.sort(false);
    if (opt.clean_templ) clean_cds(opt, tr, *be);
    tr.set_cds_as_exons();
    tr.remove_non_coding();
    tr.sort(false);
    say("start removing duplicates");
    tr.deduplicate(opt.use_id);
    say("loading query transcriptome");
    {
        QueryTxs q = query_parse.get();
        if (q.err) std::rethrow_exception(q.err);
        tr.append(std::move(q.parsed));
    }
    say("bundling transcriptome");
    tr.bundleup(opt.use_id, (uint32_t)opt.overhang);
    if (opt.timing) {
        say("");
        fprintf(stderr, "[timing] prepare (ms): start / genome index / template GTFs + CDS chains / sort + cleaning / dedup / query GTF / bundling:%s\n", timing.c_str());
    }
}

// Formatting of one bundle's output records (orfanage.cpp:383-432, 662-815), appended to per-chunk buffers so that the
// bundles can be formatted by several threads and written in bundle order.
namespace {
struct EmitBuf { std::string gtf, stats; };
inline void put_line(std::string& o, const std::string& rec) { o += rec; o += '\n'; }
}  // namespace

int run_orfanage(const Options& opt, const Backend& be) {
    typedef std::chrono::steady_clock clk;
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    auto t0 = clk::now();
    const bool mode_all = !opt.modes.empty() && opt.modes[0] == ORFX_MODE_ALL;
    const bool want_rows = opt.stats_set || mode_all;
    orfx_params params;
    fill_params(opt, want_rows, params);
    const int n_shards = (int)opt.devices.size();
    // The device contexts (CUDA initialisation: several hundred milliseconds each) come up while the annotation loads.
    std::vector<Shard> shards((size_t)n_shards);   // Shard is not copyable (PackedBatch owns an arena)
    std::vector<std::thread> ctx_threads;
    for (int k = 0; k < n_shards; k++)
        ctx_threads.emplace_back([&, k]() {
            Shard& s = shards[(size_t)k];
            auto c0 = clk::now();
            s.rc = be.create(&s.ctx, opt.devices[(size_t)k], &params);
            if (s.rc) s.err = be.last_error();
            s.ms_create = ms(c0, clk::now());
        });
    struct JoinAll { std::vector<std::thread>& t; ~JoinAll() { for (auto& x : t) if (x.joinable()) x.join(); } } join_ctx{ctx_threads};
    // The loaded annotation (several hundred thousand transcripts with their strings, maps and chains) and the mapped genome
    // are left to the operating system when the process ends: tearing them down one object at a time costs a few hundred
    // milliseconds that a command-line run has no use for.
    Fasta& fasta = *new Fasta();
    Transcriptome& tr = *new Transcriptome();
    prepare_transcriptome(opt, fasta, tr, true, &be);
    auto t1 = clk::now();

    FILE* gtf = fopen(opt.output.c_str(), "w");
    if (!gtf) throw std::runtime_error("Cannot open output file: " + opt.output);
    FILE* stats = nullptr;
    if (opt.stats_set) {
        stats = fopen(opt.stats.c_str(), "w");
        if (!stats) throw std::runtime_error("Cannot open stats file: " + opt.stats);
        fputs("query_id\ttemplate_id\tsegment\tnotes\tnum_templates\tquery_len\ttemplate_len\tunion_len\tpass\tlen_match\tlen_inframe\t"
              "len_outframe\tlen_extra\tlen_missing\tlength_pi\tmatch_length_pi\tinframe_length_pi\talignment_match\tstart_match\tstop_match\tpi", stats);
        // --pi (this build): the four integers of the protein alignment behind every scored row
        fputs(opt.pi != 0 && !opt.reference.empty() ? "\taln_max\taln_score\taln_mqe\taln_mte\n" : "\n", stats);
    }

    fprintf(stderr, "starting main evaluation\n");
    std::vector<int> active;
    for (int bi = 0; bi < (int)tr.bundles.size(); bi++) if (tr.bundles[bi].has_template && tr.bundles[bi].has_query) active.push_back(bi);
    TidClasses tid_classes;
    build_tid_classes(tr, tid_classes);

    std::vector<std::vector<int>> shard_ids;
    shard_bundles(tr, active, n_shards, shard_ids);
    auto work = [&](int k) {
        Shard& s = shards[(size_t)k];
        try {
            memset(&s.res, 0, sizeof s.res);
            auto a = clk::now();
            // one batch per process: page-locking the arena would cost more than the one upload saves
            if (!shard_ids[(size_t)k].empty())
                pack_bundles(tr, shard_ids[(size_t)k], opt.keep_all_cds, tid_classes, s.batch, nullptr, nullptr, true, PACK_EVAL, false, opt.overhang, opt.spliced_overhang,
                             params.align_enabled != 0, true, true);
            auto b = clk::now();
            s.ms_pack = ms(a, b);
            ctx_threads[(size_t)k].join();   // this shard's context (the contexts of one process come up one after the other)
            if (s.rc || shard_ids[(size_t)k].empty()) return;
            b = clk::now();
            orfx_batch v = s.batch.view();
            s.rc = be.submit(s.ctx, &v);
            if (s.rc) { s.err = be.last_error(); return; }
            s.rc = be.wait(s.ctx, &s.res);
            if (s.rc) { s.err = be.last_error(); return; }
            s.ms_eval = ms(b, clk::now());
        } catch (const std::exception& e) { s.rc = -100; s.err = e.what(); }
    };
    if (n_shards == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int k = 0; k < n_shards; k++) th.emplace_back(work, k);
        for (auto& t : th) t.join();
    }
    for (Shard& s : shards) if (s.rc) {
        const std::string err = s.err;
        for (Shard& z : shards) if (z.ctx) { be.destroy(z.ctx); z.ctx = nullptr; }
        throw std::runtime_error("device evaluation failed: " + err);
    }
    auto t2 = clk::now();

    // bundle -> (shard, index of its first query among the shard's results)
    std::vector<int> b_shard(tr.bundles.size(), -1);
    std::vector<int> b_firstq(tr.bundles.size(), 0);
    for (int k = 0; k < n_shards; k++) {
        int qi = 0;
        const PackedBatch& pb = shards[(size_t)k].batch;
        for (size_t j = 0; j < pb.bundle_ids.size(); j++) {
            b_shard[(size_t)pb.bundle_ids[j]] = k;
            b_firstq[(size_t)pb.bundle_ids[j]] = qi;
            for (int x = pb.bundle_tx_off[j]; x < pb.bundle_tx_off[j + 1]; x++) if (pb.tx_kind[x] == ORFX_TX_QUERY) qi++;
        }
    }

    auto stats_score = [&](std::string& o, const orfx_results& r, int64_t row) {
        char buf[512];
        int n = snprintf(buf, sizeof buf, "%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%s\t%s\t%s\t0\t%d\t0\t0", r.row_qlen[row], r.row_tlen[row], r.row_ulen[row],
                         (int)r.row_pass[row], r.row_match[row], r.row_inframe[row], r.row_outframe[row], r.row_extra[row], r.row_missing[row],
                         float_ostream(r.row_lpi[row]).c_str(), float_ostream(r.row_mlpi[row]).c_str(), float_ostream(r.row_ilpi[row]).c_str(),
                         (int)r.row_start_match[row]);
        o.append(buf, (size_t)n);
        if (r.row_aln_max) { n = snprintf(buf, sizeof buf, "\t%d\t%d\t%d\t%d", r.row_aln_max[row], r.row_aln_score[row], r.row_aln_mqe[row], r.row_aln_mte[row]); o.append(buf, (size_t)n); }
        o += '\n';
    };
    auto add_found_attrs = [&](Tx& fq, const Tx& t, int duplicity, float ilpi, float lpi, float mlpi, int start_match) {
        fq.add_attribute("orfanage_status", "1");
        fq.add_attribute("orfanage_template_source", t.source);
        fq.add_attribute("orfanage_template", t.tid);
        fq.add_attribute("orfanage_duplicity", std::to_string(duplicity));
        fq.add_attribute("orfanage_ilpi", float_to_string(ilpi));
        fq.add_attribute("orfanage_lpi", float_to_string(lpi));
        fq.add_attribute("orfanage_mlpi", float_to_string(mlpi));
        fq.add_attribute("orfanage_start_match", start_match ? "1" : "0");
        bool stop_match = fq.strand == '+' ? fq.cds_end == t.cds_end : fq.cds_start == t.cds_start;
        fq.add_attribute("orfanage_stop_match", stop_match ? "1" : "0");
    };
    const bool with_stats = stats != nullptr;
    auto emit_bundle = [&](int bi, EmitBuf& out) {
        const Bundle& b = tr.bundles[(size_t)bi];
        const std::string& seqname = tr.seqnames[(size_t)b.seqid];
        if (!(b.has_template && b.has_query)) {   // orfanage.cpp:383-432
            for (int i = 0; i < b.count; i++) {
                Tx& q = tr.txs[(size_t)(b.first + i)];
                if (q.is_templ) continue;
                if ((opt.keep_all_cds || opt.keep_cds_if_not_found) && q.is_coding) q.build_cds();
                put_line(out.gtf, q.gtf(seqname));
                if (with_stats) { out.stats += q.tid; out.stats += '\t'; out.stats += DASH19; out.stats += '\n'; }
            }
            return;
        }
        const Shard& sh = shards[(size_t)b_shard[(size_t)bi]];
        const orfx_results& r = sh.res;
        const std::vector<int>& txg = sh.batch.tx_global;
        int qi = b_firstq[(size_t)bi];
        char buf[256];
        for (int i = 0; i < b.count; i++) {
            Tx& q = tr.txs[(size_t)(b.first + i)];
            if (q.is_templ) continue;
            Tx og_q;
            bool use_og = false;
            if (opt.keep_cds_if_not_found && q.is_coding) { og_q = q; og_q.build_cds(); use_og = true; }
            if (opt.keep_all_cds && q.is_coding) {   // orfanage.cpp:458-491
                q.build_cds();
                put_line(out.gtf, q.gtf(seqname));
                if (with_stats) { out.stats += q.tid; out.stats += '\t'; out.stats += KEEP19; out.stats += '\n'; }
                continue;
            }
            q.remove_cds();
            q.cds_source = "ORFanage";
            const orfx_best& best = r.best[qi++];
            if (best.t_tx < 0) {   // orfanage.cpp:662-691
                Tx* o = &q;
                if (use_og) { o = &og_q; o->build_cds(); }
                o->add_attribute("orfanage_status", "0");
                put_line(out.gtf, o->gtf(seqname));
                if (with_stats) { out.stats += o->tid; out.stats += '\t'; out.stats += KEEP19; out.stats += '\n'; }
                continue;
            }
            if (!mode_all) {   // orfanage.cpp:753-791
                Tx fq = q;
                if (opt.overhang > 0 && !fq.exons.empty()) { fq.exons.front().s -= best.ext_lo; fq.exons.back().e += best.ext_hi; }   // cpp:418-429, 461-474
                set_orf(fq, best.cds_start, best.cds_end);
                const Tx& t = tr.txs[(size_t)txg[(size_t)best.t_tx]];
                add_found_attrs(fq, t, best.duplicity, best.ilpi, best.lpi, best.mlpi, best.start_match);
                put_line(out.gtf, fq.gtf(seqname));
            }
            if (!want_rows) continue;
            // walk the key groups in print order (orfanage.cpp:693-752, 793-814)
            int64_t o0 = best.row_off, o1 = best.row_off + best.n_rows;
            for (int64_t o = o0; o < o1;) {
                int64_t front = r.print_order[o];
                int ks = r.row_cds_start[front], ke = r.row_cds_end[front];
                int64_t oe = o;
                while (oe < o1 && r.row_cds_start[r.print_order[oe]] == ks && r.row_cds_end[r.print_order[oe]] == ke) oe++;
                std::string tid = q.tid;
                if (mode_all) { tid += "."; tid += std::to_string(ks); tid += "-"; tid += std::to_string(ke); }
                const char* front_note = nullptr;
                if (mode_all) {
                    if (r.row_pass[front]) {
                        Tx fq = q;
                        if (r.row_ext_lo && !fq.exons.empty()) { fq.exons.front().s -= r.row_ext_lo[front]; fq.exons.back().e += r.row_ext_hi[front]; }
                        set_orf(fq, ks, ke);
                        fq.tid = tid;
                        const Tx& t = tr.txs[(size_t)txg[(size_t)r.row_t_tx[front]]];
                        add_found_attrs(fq, t, r.row_n_dup[front], r.row_ilpi[front], r.row_lpi[front], r.row_mlpi[front], r.row_start_match[front]);
                        front_note = front == best.best_row ? "best_gtf" : "gtf";
                        put_line(out.gtf, fq.gtf(seqname));
                    }
                } else if (front == best.best_row) front_note = "best_gtf";
                if (with_stats) {
                    for (int64_t k = o; k < oe; k++) {
                        int64_t row = r.print_order[k];
                        const Tx& t = tr.txs[(size_t)txg[(size_t)r.row_t_tx[row]]];
                        const char* note = (k == o && front_note) ? front_note : status_note(r.row_status[row]);
                        out.stats += tid; out.stats += '\t'; out.stats += t.tid;
                        int n;
                        if (r.row_status[row] == ORFX_ST_NO_OVERLAP) n = snprintf(buf, sizeof buf, "\t4294967295-4294967295\t%s\t%d\t", note, r.row_n_dup[row] + 1);
                        else n = snprintf(buf, sizeof buf, "\t%u-%u\t%s\t%d\t", (uint32_t)r.row_seg_start[row], (uint32_t)r.row_seg_end[row], note, r.row_n_dup[row] + 1);
                        out.stats.append(buf, (size_t)n);
                        stats_score(out.stats, r, row);
                    }
                }
                o = oe;
            }
        }
    };
    // Bundles are independent: chunks of consecutive bundles are formatted by a pool of threads into their own buffers
    // and written in bundle order as they complete (the reference prints under `omp critical`, orfanage.cpp:658).
    {
        const int nb = (int)tr.bundles.size();
        const int chunk = 128;
        const int n_chunks = (nb + chunk - 1) / chunk;
        unsigned nthr = std::max(1u, std::min(opt.threads > 1 ? (unsigned)opt.threads : std::thread::hardware_concurrency(), 32u));
        if (n_chunks < 4) nthr = 1;
        std::vector<EmitBuf> bufs((size_t)n_chunks);
        std::vector<std::atomic<int>> done((size_t)n_chunks);
        for (auto& d : done) d.store(0);
        std::atomic<int> next(0);
        std::atomic<bool> failed(false);
        std::string fail_msg;
        auto worker = [&]() {
            try {
                for (;;) {
                    const int c = next.fetch_add(1);
                    if (c >= n_chunks || failed.load()) break;
                    EmitBuf& out = bufs[(size_t)c];
                    for (int bi = c * chunk; bi < std::min(nb, (c + 1) * chunk); bi++) emit_bundle(bi, out);
                    done[(size_t)c].store(1, std::memory_order_release);
                }
            } catch (const std::exception& e) {
                if (!failed.exchange(true)) fail_msg = e.what();
            }
        };
        std::vector<std::thread> th;
        for (unsigned k = 1; k < nthr; k++) th.emplace_back(worker);
        if (nthr == 1) worker();
        for (int c = 0; c < n_chunks && !failed.load(); c++) {   // the writer: this thread
            while (!done[(size_t)c].load(std::memory_order_acquire)) {
                if (failed.load()) break;
                if (nthr == 1) break;
                std::this_thread::yield();
            }
            if (!done[(size_t)c].load(std::memory_order_acquire)) break;
            EmitBuf& out = bufs[(size_t)c];
            if (!out.gtf.empty()) fwrite(out.gtf.data(), 1, out.gtf.size(), gtf);
            if (stats && !out.stats.empty()) fwrite(out.stats.data(), 1, out.stats.size(), stats);
            std::string().swap(out.gtf); std::string().swap(out.stats);
        }
        for (auto& t : th) t.join();
        if (failed.load()) { fclose(gtf); if (stats) fclose(stats); throw std::runtime_error(fail_msg); }
    }
    fclose(gtf);
    if (stats) fclose(stats);
    for (Shard& s : shards) if (s.ctx) { be.release(s.ctx); be.destroy(s.ctx); }
    fprintf(stderr, "Done\n");
    if (opt.timing) {
        auto t3 = clk::now();
        fprintf(stderr, "[timing] load+prep %.1f ms, evaluate %.1f ms, emit %.1f ms, total %.1f ms\n", ms(t0, t1), ms(t1, t2), ms(t2, t3), ms(t0, t3));
        for (int k = 0; k < n_shards; k++)
            fprintf(stderr, "[timing] shard %d: %zu bundles, context %.1f ms (beside the load), pack %.1f ms, submit+wait %.1f ms (h2d %.2f, kernels %.2f, d2h %.2f)\n", k,
                    shards[(size_t)k].batch.bundle_ids.size(), shards[(size_t)k].ms_create, shards[(size_t)k].ms_pack, shards[(size_t)k].ms_eval, shards[(size_t)k].res.ms_h2d,
                    shards[(size_t)k].res.ms_kernels, shards[(size_t)k].res.ms_d2h);
    }
    return 0;
}

}  // namespace orfx
