// Interval chains, transcripts, GTF loading, template preparation and locus bundling.
// Host mirror of the reference's Transcriptome layer; each function cites what it replaces.
#include "orfx_host.h"
#include <thread>
#include <functional>
#include <unordered_map>
#include <algorithm>
#include <chrono>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace orfx {

int chain_len(const Chain& c) {
    int l = 0;
    for (const Seg& s : c) l += (int)(uint32_t)(s.e + 1 - s.s);
    return l;
}

int genome2nt(const Chain& c, int g, char strand) {
    int acc = 0;
    bool found = false;
    for (const Seg& s : c) {
        if (g >= s.s && g <= s.e) { acc += g - s.s; found = true; break; }
        acc += s.len();
    }
    if (!found) return -1;
    return strand == '+' ? acc : (chain_len(c) - acc) - 1;
}

int nt2genome(const Chain& c, int off, char strand) {
    const int n = (int)c.size();
    if (off >= 0) {
        if (strand == '+') {
            for (int i = 0; i < n; i++) { int l = c[i].len(); if (off < l) return c[i].s + off; off -= l; }
        } else {
            for (int i = n - 1; i >= 0; i--) { int l = c[i].len(); if (off < l) return c[i].e - off; off -= l; }
        }
    }
    // over-requested (or negative, which the reference compares as unsigned): last position
    return strand == '+' ? c.back().e : c.front().s;
}

void assign_phase(Chain& c, char strand, int start_phase) {
    int acc = start_phase;
    if (strand == '-') {
        for (int i = (int)c.size() - 1; i >= 0; i--) { c[i].phase = (int8_t)((3 - acc % 3) % 3); acc += c[i].e - c[i].s + 1; }
    } else {
        for (Seg& s : c) { s.phase = (int8_t)((3 - acc % 3) % 3); acc += s.e - s.s + 1; }
    }
}

void cut_chain(Chain& c, int lo, int hi) {
    if (c.empty()) return;
    if (lo <= c.front().s && hi >= c.back().e) return;
    Chain tmp;
    for (Seg s : c) {
        if (s.s <= lo && s.e >= lo) { s.s = lo; s.phase = -1; }
        if (s.e >= hi) { s.e = hi; s.phase = -1; tmp.push_back(s); break; }
        if (s.e < lo || s.s > hi) continue;
        tmp.push_back(s);
    }
    c.swap(tmp);
}

// ---- Tx -------------------------------------------------------------------------------------
void Tx::build_cds() {
    cds.clear();
    if (cds_start == 0 && cds_end == 0) return;
    if (cds_phase != 0) {
        if (strand == '+') cds_start = nt2genome(exons, genome2nt(exons, cds_start, strand) + cds_phase, strand);
        else cds_end = nt2genome(exons, genome2nt(exons, cds_end, strand) + cds_phase, strand);
        cds_phase = 0;
    }
    for (Seg e : exons) {
        if (cds_end < e.s || cds_start > e.e) continue;
        if (cds_start >= e.s && cds_start <= e.e) e.s = cds_start;
        if (cds_end >= e.s && cds_end <= e.e) e.e = cds_end;
        e.phase = 0;
        cds.push_back(e);
    }
    assign_phase(cds, strand, 0);
}

void Tx::correct_chain_len() {
    int clen = chain_len(cds);
    int num_trim = clen % 3;
    if (num_trim > 0) {
        int a = nt2genome(cds, 0, strand);
        int b = nt2genome(cds, clen - num_trim - 1, strand);
        cut_chain(cds, std::min(a, b), std::max(a, b));
        assign_phase(cds, strand, 0);
        cds_start = cds.front().s;
        cds_end = cds.back().e;
    }
}

void Tx::remove_cds() {
    cds.clear();
    cds_start = cds_end = cds_phase = 0;
    is_coding = false;
}

std::string Tx::gtf(const std::string& seqname) const {
    std::string o;
    o.reserve(256 + 96 * (exons.size() + cds.size()));
    char num[64];
    auto coords = [&](int s, int e) { snprintf(num, sizeof num, "%d\t%d", s, e); o += num; };
    o += seqname; o += '\t'; o += source; o += "\ttranscript\t";
    coords(start(), end());
    o += "\t.\t"; o += strand; o += "\t.\ttranscript_id \""; o += tid; o += "\"; gene_id \""; o += gid; o += "\"; ";
    {
        std::string a;
        for (const auto& kv : attrs) { a += kv.first; a += " \""; a += kv.second; a += "\"; "; }
        if (!a.empty()) a.pop_back();
        o += a;
    }
    o += '\n';
    for (const Seg& e : exons) {
        o += seqname; o += '\t'; o += source; o += "\texon\t";
        coords(e.s, e.e);
        o += "\t.\t"; o += strand; o += "\t.\ttranscript_id \""; o += tid; o += "\";\n";
    }
    for (const Seg& c : cds) {
        o += seqname; o += '\t'; o += cds_source; o += "\tCDS\t";
        coords(c.s, c.e);
        o += "\t.\t"; o += strand; o += '\t';
        snprintf(num, sizeof num, "%u", (uint32_t)(int32_t)c.phase);
        o += num;
        o += "\ttranscript_id \""; o += tid; o += "\";\n";
    }
    o.pop_back();   // the caller appends the record-terminating newline (std::endl in the reference)
    return o;
}

bool tx_less(const Tx& a, const Tx& b) {
    if (a.seqid != b.seqid) return a.seqid < b.seqid;
    if (a.strand != b.strand) return a.strand < b.strand;
    if (a.start() != b.start()) return a.start() < b.start();
    if (a.end() != b.end()) return a.end() < b.end();
    return false;
}

bool tx_same(const Tx& a, const Tx& b) {
    if (a.seqid != b.seqid || a.strand != b.strand || a.start() != b.start() || a.end() != b.end()) return false;
    if (a.exons.size() != b.exons.size()) return false;
    for (size_t i = 0; i < a.exons.size(); i++)
        if (a.exons[i].s != b.exons[i].s || a.exons[i].e != b.exons[i].e || a.exons[i].phase != b.exons[i].phase) return false;
    return true;
}

// ---- GTF reader --------------------------------------------------------------------------------
// Accepts `transcript`/`mRNA`, `exon` and `CDS` rows keyed by transcript_id; transcripts come out
// in order of first appearance, exons ascending.  CDS-only records use their CDS rows as exons.
namespace {
struct Span { const char* p; size_t n; };
inline bool eq(const Span& s, const char* lit) { size_t l = strlen(lit); return s.n == l && memcmp(s.p, lit, l) == 0; }

void parse_attr_column(const char* p, const char* end, std::vector<std::pair<Span, Span>>& out) {
    out.clear();
    while (p < end) {
        while (p < end && (*p == ' ' || *p == ';' || *p == '\t')) p++;
        if (p >= end) break;
        const char* k0 = p;
        while (p < end && *p != ' ' && *p != '\t' && *p != ';') p++;
        Span key{k0, (size_t)(p - k0)};
        while (p < end && (*p == ' ' || *p == '\t')) p++;
        Span val{p, 0};
        if (p < end && *p == '"') {
            p++;
            const char* v0 = p;
            while (p < end && *p != '"') p++;
            val = Span{v0, (size_t)(p - v0)};
            if (p < end) p++;
        } else {
            const char* v0 = p;
            while (p < end && *p != ';') p++;
            const char* v1 = p;
            while (v1 > v0 && v1[-1] == ' ') v1--;
            val = Span{v0, (size_t)(v1 - v0)};
        }
        out.push_back(std::make_pair(key, val));
    }
}
}  // namespace

namespace {
struct GtfWork { std::vector<std::pair<uint32_t, uint32_t>> cds; std::vector<char> frames; };
// what one parser thread makes of its slice of the file: records in order of first appearance in the slice
struct GtfSlice {
    std::vector<GtfRecord> out;
    std::vector<GtfWork> work;
    std::vector<uint8_t> has_tx_row;     // the record's `source` and attrs come from a transcript row of this slice
    std::unordered_map<std::string, size_t> by_id;
};

void parse_gtf_slice(const char* p, const char* fend, GtfSlice& S) {
    std::vector<GtfRecord>& out = S.out;
    std::vector<std::pair<Span, Span>> av;
    std::string tid;
    size_t last = (size_t)-1;   // rows of one transcript are usually consecutive
    while (p < fend) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(fend - p));
        if (!eol) eol = fend;
        const char* le = eol;
        while (le > p && (le[-1] == '\r')) le--;
        if (le > p && *p != '#') {
            Span col[9];
            int nc = 0;
            const char* q = p;
            while (nc < 9) {
                const char* t = nc < 8 ? (const char*)memchr(q, '\t', (size_t)(le - q)) : nullptr;
                if (!t) { col[nc++] = Span{q, (size_t)(le - q)}; break; }
                col[nc++] = Span{q, (size_t)(t - q)};
                q = t + 1;
            }
            if (nc == 9) {
                bool is_tx = eq(col[2], "transcript") || eq(col[2], "mRNA");
                bool is_exon = eq(col[2], "exon");
                bool is_cds = eq(col[2], "CDS");
                if (is_tx || is_exon || is_cds) {
                    parse_attr_column(col[8].p, col[8].p + col[8].n, av);
                    const Span* tspan = nullptr;
                    const Span* gspan = nullptr;
                    for (auto& kv : av) {
                        if (!tspan && eq(kv.first, "transcript_id")) tspan = &kv.second;
                        else if (!gspan && eq(kv.first, "gene_id")) gspan = &kv.second;
                    }
                    if (tspan && tspan->n) {
                        size_t ri;
                        if (last != (size_t)-1 && out[last].tid.size() == tspan->n && memcmp(out[last].tid.data(), tspan->p, tspan->n) == 0) ri = last;
                        else {
                            tid.assign(tspan->p, tspan->n);
                            auto it = S.by_id.find(tid);
                            if (it == S.by_id.end()) {
                                ri = out.size();
                                S.by_id[tid] = ri;
                                out.emplace_back();
                                S.work.emplace_back();
                                S.has_tx_row.push_back(0);
                                GtfRecord& r = out.back();
                                r.tid = tid;
                                r.seqname.assign(col[0].p, col[0].n);
                                r.source.assign(col[1].p, col[1].n);
                                r.strand = col[6].n ? col[6].p[0] : '.';
                            } else ri = it->second;
                            last = ri;
                        }
                        GtfRecord& r = out[ri];
                        GtfWork& w = S.work[ri];
                        if (gspan && !r.has_gid) { r.has_gid = true; r.gid.assign(gspan->p, gspan->n); }
                        uint32_t s = (uint32_t)strtoul(col[3].p, nullptr, 10);
                        uint32_t e = (uint32_t)strtoul(col[4].p, nullptr, 10);
                        if (s > e) std::swap(s, e);
                        if (is_tx) {
                            S.has_tx_row[ri] = 1;
                            r.source.assign(col[1].p, col[1].n);
                            for (auto& kv : av) {
                                if (eq(kv.first, "transcript_id")) continue;
                                r.attrs.emplace_back(std::string(kv.first.p, kv.first.n), std::string(kv.second.p, kv.second.n));
                            }
                        } else if (is_exon) {
                            r.exons.emplace_back(s, e);
                        } else {
                            r.has_cds = true;
                            w.cds.emplace_back(s, e);
                            w.frames.push_back(col[7].n ? col[7].p[0] : '.');
                        }
                    }
                }
            }
        }
        p = eol + 1;
    }
}
}  // namespace

// The file is cut at line starts into one slice per hardware thread; the slices are parsed concurrently and
// stitched in file order, so that "first appearance" (record order, seqname / strand / gene_id of a record)
// and the order of a record's rows are what a single pass over the file gives.
void read_gtf(const std::string& path, std::vector<GtfRecord>& out) {
    int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("Couldn't open the GTF file: " + path);
    struct stat st;
    fstat(fd, &st);
    size_t size = (size_t)st.st_size;
    const char* data = size ? (const char*)mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0) : nullptr;
    if (size && data == MAP_FAILED) { close(fd); throw std::runtime_error("Couldn't map the GTF file: " + path); }

    unsigned nthr = size > (8u << 20) ? std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
    if (const char* e = getenv("ORFX_GTF_SLICES")) nthr = (unsigned)std::max(1, std::min(64, atoi(e)));   // tests: force slicing of small files
    std::vector<const char*> cut(nthr + 1, data + size);
    cut[0] = data;
    for (unsigned k = 1; k < nthr; k++) {
        const char* p = data + size / nthr * k;
        const char* nl = (const char*)memchr(p, '\n', (size_t)(data + size - p));
        cut[k] = nl ? nl + 1 : data + size;
    }
    std::vector<GtfSlice> slices(nthr);
    if (nthr == 1) parse_gtf_slice(data, data + size, slices[0]);
    else {
        std::vector<std::thread> th;
        for (unsigned k = 0; k < nthr; k++) th.emplace_back(parse_gtf_slice, cut[k], cut[k + 1], std::ref(slices[k]));
        for (auto& t : th) t.join();
    }
    if (data) munmap((void*)data, size);
    close(fd);

    const size_t base = out.size();
    std::vector<GtfWork> work;
    auto in_parallel = [&](unsigned n, const std::function<void(unsigned)>& f) {
        if (n <= 1) { for (unsigned k = 0; k < n; k++) f(k); return; }
        std::vector<std::thread> th;
        for (unsigned k = 0; k < n; k++) th.emplace_back(f, k);
        for (auto& t : th) t.join();
    };
    if (nthr == 1) {
        for (auto& r : slices[0].out) out.push_back(std::move(r));
        work = std::move(slices[0].work);
    } else {
        // A transcript whose rows straddle a cut (or are scattered over the file) appears in several slices: its first
        // appearance keeps the record, later ones are folded into it in file order.  Every slice looks its ids up in the
        // slices before it (concurrently); the records are then moved to their final places concurrently as well.
        struct Dup { int slice; size_t idx; };
        std::vector<std::vector<Dup>> first(nthr);        // per record: where it appeared first, slice -1 = here
        std::vector<std::vector<size_t>> place(nthr);     // per record: its index among the kept records of the slice
        std::vector<size_t> kept(nthr, 0);
        in_parallel(nthr, [&](unsigned k) {
            GtfSlice& S = slices[k];
            first[k].assign(S.out.size(), Dup{-1, 0});
            place[k].assign(S.out.size(), 0);
            size_t n = 0;
            for (size_t i = 0; i < S.out.size(); i++) {
                for (unsigned j = 0; j < k; j++) {
                    auto it = slices[j].by_id.find(S.out[i].tid);
                    if (it != slices[j].by_id.end()) { first[k][i] = Dup{(int)j, it->second}; break; }
                }
                if (first[k][i].slice < 0) place[k][i] = n++;
            }
            kept[k] = n;
        });
        std::vector<size_t> at(nthr + 1, 0);
        for (unsigned k = 0; k < nthr; k++) at[k + 1] = at[k] + kept[k];
        out.resize(base + at[nthr]);
        work.resize(at[nthr]);
        in_parallel(nthr, [&](unsigned k) {
            GtfSlice& S = slices[k];
            for (size_t i = 0; i < S.out.size(); i++) {
                if (first[k][i].slice >= 0) continue;
                out[base + at[k] + place[k][i]] = std::move(S.out[i]);
                work[at[k] + place[k][i]] = std::move(S.work[i]);
            }
        });
        for (unsigned k = 1; k < nthr; k++) {   // the (rare) later appearances, in file order
            GtfSlice& S = slices[k];
            for (size_t i = 0; i < S.out.size(); i++) {
                const Dup d = first[k][i];
                if (d.slice < 0) continue;
                const size_t gi = at[(size_t)d.slice] + place[(size_t)d.slice][d.idx];
                GtfRecord& g = out[base + gi];
                GtfWork& gw = work[gi];
                GtfRecord& r = S.out[i];
                if (!g.has_gid && r.has_gid) { g.has_gid = true; g.gid = r.gid; }
                if (S.has_tx_row[i]) { g.source = r.source; for (auto& a : r.attrs) g.attrs.push_back(std::move(a)); }
                g.exons.insert(g.exons.end(), r.exons.begin(), r.exons.end());
                if (r.has_cds) g.has_cds = true;
                gw.cds.insert(gw.cds.end(), S.work[i].cds.begin(), S.work[i].cds.end());
                gw.frames.insert(gw.frames.end(), S.work[i].frames.begin(), S.work[i].frames.end());
            }
        }
        in_parallel(nthr, [&](unsigned k) { GtfSlice().by_id.swap(slices[k].by_id); std::vector<GtfRecord>().swap(slices[k].out); std::vector<GtfWork>().swap(slices[k].work); });
    }
    const size_t n_new = out.size() - base;
    in_parallel(n_new > 20000 ? nthr : 1u, [&](unsigned k) {
        const unsigned parts = n_new > 20000 ? nthr : 1u;
        for (size_t i = base + n_new * k / parts; i < base + n_new * (k + 1) / parts; i++) {
            GtfRecord& r = out[i];
            GtfWork& w = work[i - base];
            if (r.exons.empty()) r.exons = w.cds;
            std::sort(r.exons.begin(), r.exons.end());
            if (r.has_cds) {
                uint32_t mn = UINT32_MAX, mx = 0;
                size_t imn = 0, imx = 0;
                for (size_t c = 0; c < w.cds.size(); c++) {
                    if (w.cds[c].first < mn) { mn = w.cds[c].first; imn = c; }
                    if (w.cds[c].second > mx) { mx = w.cds[c].second; imx = c; }
                }
                r.cd_start = mn;
                r.cd_end = mx;
                r.cd_phase = r.strand == '-' ? w.frames[imx] : w.frames[imn];
            }
        }
    });
}

// ---- Transcriptome -----------------------------------------------------------------------------
void Transcriptome::add(const std::string& gtf_path, bool is_templ, bool coding_only) {
    std::vector<GtfRecord> recs;
    read_gtf(gtf_path, recs);
    add_records(recs, is_templ, coding_only);
}
void Transcriptome::add_records(const std::vector<GtfRecord>& recs, bool is_templ, bool coding_only) {
    ParsedTxs p;
    convert_records(recs, is_templ, coding_only, p);
    append(std::move(p));
}
// TX::TX(GffObj*), cpp:11-82, for every record that is kept; the conversions are independent of each other.  No Transcriptome
// state is touched (the query file is converted beside the template stages): Tx::seqid indexes out.seqnames, the names of ALL
// records in first-seen order (Transcriptome::load_seqids, cpp:855-865, registers them before any record is dropped).
void convert_records(const std::vector<GtfRecord>& recs, bool is_templ, bool coding_only, ParsedTxs& out) {
    std::unordered_map<std::string, int> local;
    std::vector<int> seqid_of(recs.size());
    for (size_t i = 0; i < recs.size(); i++) {
        auto ins = local.insert(std::make_pair(recs[i].seqname, (int)out.seqnames.size()));
        if (ins.second) out.seqnames.push_back(recs[i].seqname);
        seqid_of[i] = ins.first->second;
    }
    std::vector<size_t> keep;
    keep.reserve(recs.size());
    for (size_t i = 0; i < recs.size(); i++) if (!(coding_only && !recs[i].has_cds) && !recs[i].exons.empty()) keep.push_back(i);
    std::vector<Tx>& txs = out.txs;
    txs.resize(keep.size());
    auto convert = [&](size_t k0, size_t k1, std::string* err) {
        try {
            for (size_t k = k0; k < k1; k++) {
                const GtfRecord& r = recs[keep[k]];
                Tx& t = txs[k];
                if (r.has_cds) {
                    t.is_coding = true;
                    t.cds_start = (int)r.cd_start;
                    t.cds_end = (int)r.cd_end;
                    switch (r.cd_phase) {
                        case '0': t.cds_phase = 0; break;
                        case '1': t.cds_phase = 1; break;
                        case '2': t.cds_phase = 2; break;
                        case '3': t.cds_phase = 3; break;
                        case '.': t.cds_phase = 0; break;
                        default: throw std::runtime_error("Unknown CDS phase for the initial segment");
                    }
                }
                t.is_templ = is_templ;
                t.id = (int)k;                      // Transcriptome::append adds the number of transcripts before it
                t.tid = r.tid;
                t.gid = r.has_gid ? r.gid : r.tid;
                t.seqid = seqid_of[keep[k]];        // local; Transcriptome::append maps it
                t.strand = r.strand;
                t.source = r.source;
                t.cds_source = r.source;
                t.exons.reserve(r.exons.size());
                for (auto& e : r.exons) { Seg s; s.s = (int)e.first; s.e = (int)e.second; s.phase = 0; t.exons.push_back(s); }
                bool tr_id = false;
                for (const auto& kv : r.attrs) {
                    if (kv.second.empty()) continue;
                    std::string name = kv.first;
                    if (name == "transcriptID") { if (tr_id) continue; tr_id = true; }
                    if (name == "transcript_id" && !tr_id) { name = "transcriptID"; tr_id = true; }
                    if (name == "geneID" || name == "gene_id") continue;
                    if (!t.attrs.insert(std::make_pair(name, kv.second)).second)
                        throw std::runtime_error("Attribute: " + name + " already exists");
                }
            }
        } catch (const std::exception& e) { *err = e.what(); }
    };
    const unsigned nthr = keep.size() > 20000 ? std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
    std::vector<std::string> errs(nthr);
    if (nthr == 1) convert(0, keep.size(), &errs[0]);
    else {
        std::vector<std::thread> th;
        for (unsigned k = 0; k < nthr; k++) th.emplace_back(convert, keep.size() * k / nthr, keep.size() * (k + 1) / nthr, &errs[k]);
        for (auto& t : th) t.join();
    }
    for (const std::string& e : errs) if (!e.empty()) throw std::runtime_error(e);   // the first failing slice in file order
}
void Transcriptome::append(ParsedTxs&& p) {
    bundles.clear();
    std::vector<int> global(p.seqnames.size());
    for (size_t i = 0; i < p.seqnames.size(); i++) {   // Transcriptome::load_seqids, cpp:855-865
        auto ins = seqids_.insert(std::make_pair(p.seqnames[i], (int)seqnames.size()));
        if (ins.second) seqnames.push_back(p.seqnames[i]);
        global[i] = ins.first->second;
    }
    const size_t base = txs.size();
    if (base == 0) txs = std::move(p.txs);
    else { txs.reserve(base + p.txs.size()); for (Tx& t : p.txs) txs.emplace_back(std::move(t)); }
    for (size_t k = base; k < txs.size(); k++) { txs[k].id = (int)k; txs[k].seqid = global[(size_t)txs[k].seqid]; }
}

void Transcriptome::build_cds_chains() {
    for (Tx& t : txs) if (t.is_coding && t.is_templ) t.build_cds();
}
void Transcriptome::correct_chain_len() {
    for (Tx& t : txs) if (t.is_coding && t.is_templ) t.correct_chain_len();
}
void Transcriptome::sort(bool use_id) {
    // std::sort on purpose: the reference sorts with libstdc++'s introsort, whose order among equal keys decides the
    // bundle-internal transcript order (SURVEY.md §7 hard parts).  The permutation introsort produces depends only on the
    // initial order and the comparison outcomes, so it is computed on small key records (the fields TX::operator< reads,
    // Transcriptome.h:668-684) and the transcripts -- strings, maps, chains -- are moved once, into place.
    struct Key { int seqid, start, end; char strand; uint32_t idx; const std::string* gid; };
    std::vector<Key> keys(txs.size());
    for (size_t i = 0; i < txs.size(); i++) keys[i] = Key{txs[i].seqid, txs[i].start(), txs[i].end(), txs[i].strand, (uint32_t)i, &txs[i].gid};
    auto less = [](const Key& a, const Key& b) {
        if (a.seqid != b.seqid) return a.seqid < b.seqid;
        if (a.strand != b.strand) return a.strand < b.strand;
        if (a.start != b.start) return a.start < b.start;
        if (a.end != b.end) return a.end < b.end;
        return false;
    };
    if (use_id) {
        std::sort(keys.begin(), keys.end(), [&](const Key& a, const Key& b) {
            if (a.seqid != b.seqid) return a.seqid < b.seqid;
            if (a.strand != b.strand) return a.strand < b.strand;
            if (*a.gid != *b.gid) return *a.gid < *b.gid;
            return less(a, b);
        });
    } else {
        std::sort(keys.begin(), keys.end(), less);
    }
    // the objects move once, into place: destination i takes source keys[i].idx (a permutation, so every object is read by
    // one thread and written by one); the copy is bound by the scattered reads, which several threads overlap
    const size_t n = txs.size();
    const unsigned nthr = n > 50000 ? std::min(8u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
    if (nthr <= 1) {
        std::vector<Tx> sorted;
        sorted.reserve(n);
        for (size_t i = 0; i < n; i++) sorted.emplace_back(std::move(txs[keys[i].idx]));
        txs.swap(sorted);
    } else {
        std::vector<Tx> sorted(n);
        std::vector<std::thread> th;
        for (unsigned k = 0; k < nthr; k++)
            th.emplace_back([&, k]() { for (size_t i = n * k / nthr, e = n * (k + 1) / nthr; i < e; i++) sorted[i] = std::move(txs[keys[i].idx]); });
        for (auto& t : th) t.join();
        txs.swap(sorted);
    }
}
void Transcriptome::set_cds_as_exons() {
    for (Tx& t : txs) {
        if (t.is_coding && !t.cds.empty()) {
            t.exons = t.cds;
            for (Seg& s : t.exons) s.phase = 0;
        }
    }
}
void Transcriptome::remove_non_coding() {
    txs.erase(std::remove_if(txs.begin(), txs.end(), [](const Tx& t) { return !t.is_coding; }), txs.end());
}
size_t Transcriptome::deduplicate(bool use_id) {
    size_t before = txs.size();
    if (use_id) txs.erase(std::unique(txs.begin(), txs.end(), [](const Tx& a, const Tx& b) { return a.gid == b.gid && tx_same(a, b); }), txs.end());
    else txs.erase(std::unique(txs.begin(), txs.end(), tx_same), txs.end());
    return before - txs.size();
}

size_t Transcriptome::clean_short_orfs(int minlen) {
    size_t n = 0;
    for (Tx& t : txs) if (chain_len(t.cds) <= minlen) { t.remove_cds(); n++; }
    return n;
}

size_t Transcriptome::bundleup(bool use_id, uint32_t overhang) {
    sort(use_id);
    bool any_missing = false;
    if (fasta) for (const std::string& name : seqnames) if (!fasta->find(name)) { any_missing = true; break; }
    if (fasta && any_missing) {   // (every sequence present: nothing to filter, the transcripts stay where they are)
        std::vector<Tx> valid;
        valid.reserve(txs.size());
        int cur = -1;
        bool ok = false;
        for (Tx& t : txs) {
            if (t.seqid != cur) { cur = t.seqid; ok = fasta->find(seqnames[cur]) != nullptr; }
            if (!ok) {
                if (!t.is_templ) throw std::runtime_error("Reference FASTA is missing sequence for query: " + seqnames[cur]);
            } else valid.push_back(std::move(t));
        }
        txs.swap(valid);
    }
    bundles.clear();
    if (txs.empty()) return 0;
    auto close_bundle = [&](Bundle& b) {
        if (!fasta) return;
        const Fasta::Seq* seq = fasta->find(seqnames[b.seqid]);
        int seqlen = seq ? (int)seq->length : 0;
        int ns = std::max(b.bstart - (int)overhang, 1);
        uint32_t ne = std::min((uint32_t)b.bend + overhang, (uint32_t)seqlen);
        b.bstart = std::min(b.bstart, ns); b.bend = std::max(b.bend, ns);
        b.bstart = std::min(b.bstart, (int)ne); b.bend = std::max(b.bend, (int)ne);
        uint32_t blen = (uint32_t)(b.bend + 1) - (uint32_t)b.bstart;
        if ((int64_t)b.bstart + (int64_t)blen - 1 > (int64_t)seqlen || b.bstart < 0)
            throw std::runtime_error("Bundle + stop codon extends past the end of the reference sequence");
    };
    Bundle cur;
    cur.first = 0;
    for (int i = 0; i < (int)txs.size(); i++) {
        const Tx& t = txs[i];
        bool can = cur.count == 0;
        if (!can) {
            if (use_id) can = cur.gid == t.gid;
            else can = cur.seqid == t.seqid && cur.strand == t.strand && (t.start() <= cur.bend && t.end() >= cur.bstart);
        }
        if (!can) {
            close_bundle(cur);
            bundles.push_back(cur);
            cur = Bundle();
            cur.first = i;
        }
        cur.count++;
        cur.seqid = t.seqid;
        cur.strand = t.strand;
        cur.bstart = std::min(cur.bstart, t.start());
        cur.bend = std::max(cur.bend, t.end());
        cur.gid = use_id ? t.gid : std::string();
        if (t.is_templ) cur.has_template = true; else cur.has_query = true;
    }
    close_bundle(cur);
    bundles.push_back(cur);
    return bundles.size();
}

}  // namespace orfx
