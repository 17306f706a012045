// TEST INFRASTRUCTURE ONLY — CPU restatement ("oracle") of the reference's hot path.
//
// This file restates, in plain scalar C++, what the body of the reference's bundle loop computes
// (alevar/ORFanage v1.2.0, orfanage.cpp:376-817 and the SEGTP/CHAIN/TX/SEQ methods it calls) on the
// flat batch layout of include/orfx.h.  It deliberately keeps the reference's procedural shape —
// chains are mutated step by step, the spliced sequence is a string, codons go through a map — so
// that it is an independent check of the CUDA path, which works on bit masks instead.
//
// It is the checker, never the product: only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load it.  Parity pin: its GTF / stats output through the
// test driver (oracle/orfanage_oracle_cli) is compared byte for byte with oracle/_ref/orfanage_ref,
// the reference's own sources compiled against the gclib stand-in (tests/test_oracle_vs_ref.py), and
// with the committed goldens under tests/golden/.  The protein-alignment part (ksw2 extension DP)
// has no call site and no source in the reference tree: for that part parity is UNPINNED (see
// orfx_oracle_align below and DESIGN.md).
#include "../include/orfx.h"
#include "../include/orfx_codon_table.h"
#include "../include/orfx_mat50.h"

#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <cstring>
#include <map>
#include <memory>
#include <set>
#include <stdexcept>
#include <string>
#include <tuple>
#include <unordered_set>
#include <vector>

namespace {

thread_local std::string g_err;

// ---- SEGTP / CHAIN (Transcriptome.h:23-563) ------------------------------------------------------
struct S {
    uint32_t s = UINT32_MAX, e = UINT32_MAX;
    int ph = -1;
    bool empty() const { return s == UINT32_MAX && e == UINT32_MAX; }
    uint32_t slen() const { return (e + 1) - s; }
};
typedef std::vector<S> C;

S mk(uint32_t s, uint32_t e, int ph = -1) { S x; x.s = s; x.e = e; x.ph = ph; return x; }

int clen(const C& c) { int l = 0; for (const S& x : c) l += (int)x.slen(); return l; }

// SEGTP::intersect, h:93-102
int seg_intersect(const S& a, const S& b, S& res) {
    int is = a.s != UINT32_MAX ? (int)std::max(a.s, b.s) : (int)b.s;
    int ie = (int)std::min(a.e, b.e);
    if (is <= ie) { res = mk(is, ie); return ie - is + 1; }
    res = S();
    return 0;
}
// SEGTP::unite / get_union, h:127-139
void seg_unite(S& a, const S& b) {
    int is = (int)std::min(a.s, b.s);
    int ie = a.e != UINT32_MAX ? (int)std::max(a.e, b.e) : (int)b.e;
    if (is <= ie) a = mk(is, ie); else a = S();
}
// SEGTP::split, h:103-126
void seg_split(const S& self, const S& other, S& left, S& inter, S& right) {
    left = S(); inter = S(); right = S();
    int il = seg_intersect(self, other, inter);
    if (il > 0) {
        if (inter.s > self.s) left = mk(self.s, inter.s - 1, self.ph);
        if (inter.e < self.e) right = mk(inter.e + 1, self.e, self.ph);
    } else {
        if (self.s < other.s) left = mk(self.s, self.e, self.ph);
        else right = mk(self.s, self.e, self.ph);
    }
}
// SEGTP::get_phase(pos,strand), h:80-89
int seg_phase_at(const S& c, uint32_t pos, char strand) {
    uint32_t r = ((uint32_t)c.ph + (strand == '+' ? 3 - (pos - c.s) % 3 : 3 - (c.e - pos) % 3) % 3) % 3;
    return (int)r;
}
// CHAIN::genome2nt, h:447-462
int g2n(const C& c, int pos, char strand) {
    int acc = 0; bool found = false;
    for (const S& x : c) {
        if (pos >= (int)x.s && pos <= (int)x.e) { found = true; acc += pos - (int)x.s; break; }
        acc += (int)x.slen();
    }
    if (!found) return -1;
    return strand == '+' ? acc : (clen(c) - acc) - 1;
}
// CHAIN::nt2genome, h:463-481 (the int/size_t comparison makes negative offsets "over-requested")
int n2g(const C& c, int off, char strand) {
    int n = (int)c.size();
    int i0 = strand == '+' ? 0 : n - 1, i1 = strand == '+' ? n : -1, st = strand == '+' ? 1 : -1;
    int left = off;
    for (int i = i0; i != i1; i += st) {
        size_t l = c[i].slen();
        if ((size_t)(long long)left < l && left >= 0) return strand == '+' ? (int)c[i].s + left : (int)c[i].e - left;
        if (left < 0) continue;
        left -= (int)l;
    }
    return strand == '+' ? (int)c.back().e : (int)c.front().s;
}
// CHAIN::assign_phase, h:432-446
void assign_phase(C& c, char strand, int start_phase) {
    int acc = start_phase;
    if (strand == '-') for (int i = (int)c.size() - 1; i >= 0; i--) { c[i].ph = (3 - acc % 3) % 3; acc += (int)(c[i].e - c[i].s + 1); }
    else for (S& x : c) { x.ph = (3 - acc % 3) % 3; acc += (int)(x.e - x.s + 1); }
}
// CHAIN::cut, h:268-289
int cut(C& c, int start, int end) {
    if (start <= (int)c.front().s && end >= (int)c.back().e) return clen(c);
    C tmp;
    for (S x : c) {
        if ((int)x.s <= start && (int)x.e >= start) { x.s = start; x.ph = -1; }
        if ((int)x.e >= end) { x.e = end; x.ph = -1; tmp.push_back(x); break; }
        if ((int)x.e < start || (int)x.s > end) continue;
        tmp.push_back(x);
    }
    c = tmp;
    return clen(c);
}
// CHAIN::trim_to_pos, h:483-494
void trim_to_pos(C& c, int start_nt, int end_nt, char strand) {
    int a = n2g(c, start_nt, strand), b = n2g(c, end_nt, strand);
    cut(c, std::min(a, b), std::max(a, b));
    assign_phase(c, strand, start_nt % 3);
}
// CHAIN::extend, h:365-397
void extend(C& c, const C& along, int numbp, bool forward) {
    if (numbp <= 0) return;
    int rem = numbp, n = (int)along.size();
    int i0 = forward ? 0 : n - 1, i1 = forward ? n : -1, st = forward ? 1 : -1;
    for (int i = i0; i != i1; i += st) {
        const S& e = along[i];
        int lp = forward ? (int)c.back().e : (int)c.front().s;
        int es = (int)e.s, ee = (int)e.e;
        if (forward && lp >= ee) continue;
        if (!forward && lp <= es) continue;
        int cur = forward ? std::max(lp + 1, es) : std::min(lp - 1, ee);
        int avail = forward ? (ee - cur + 1) : (cur - es + 1);
        if (avail > 0) {
            int take = std::min(rem, avail);
            if (forward) {
                int end_pos = cur + take - 1;
                if (cur - lp > 1) c.push_back(mk(cur, end_pos)); else { c.back().e = end_pos; c.back().ph = -1; }
            } else {
                int start_pos = cur - take + 1;
                if (lp - cur > 1) c.insert(c.begin(), mk(start_pos, cur)); else { c.front().s = start_pos; c.front().ph = -1; }
            }
            rem -= take;
            if (rem == 0) break;
        }
    }
}
// CHAIN::intersection, h:321-361
int intersection(const C& c1, const C& c2, C& res, bool segments) {
    bool found_yet = false, continue_prev = false;
    int start = 0, full = 0;
    S inter;
    for (const S& s1 : c1) {
        found_yet = false;
        for (int j = start; j < (int)c2.size(); j++) {
            int l = seg_intersect(s1, c2[j], inter);
            if (l > 0) {
                full += l;
                found_yet = true;
                start = j;
                if (inter.e == c2[j].e) start++;
                if (segments) {
                    bool ends_match = inter.e == s1.e && s1.e == c2[j].e;
                    bool starts_match = inter.s == s1.s && s1.s == c2[j].s;
                    if (!continue_prev || !starts_match) res.push_back(S());
                    seg_unite(res.back(), inter);
                    continue_prev = ends_match;
                } else res.push_back(inter);
            } else {
                if (found_yet) break;
                continue_prev = false;
            }
        }
        if (!found_yet && continue_prev) continue_prev = false;
    }
    return full;
}
// CHAIN::get_phase, h:496-512
int chain_phase_at(const C& c, int pos, char strand) {
    int phase = strand == '+' ? c.front().ph : c.back().ph;
    for (const S& x : c) if (pos >= (int)x.s && pos <= (int)x.e) { phase = seg_phase_at(x, (uint32_t)pos, strand); break; }
    return phase;
}
// CHAIN::get_union, h:291-319 (lossy on purpose)
int union_len(const C& a, const C& b) {
    C res;
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size()) {
        const S& c1 = a[i]; const S& c2 = b[j];
        if (c1.e < c2.s) { res.push_back(c1); i++; }
        else if (c2.e < c1.s) { res.push_back(c2); j++; }
        else { S u = c1; seg_unite(u, c2); res.push_back(u); i++; j++; }
    }
    return clen(res);
}
// CHAIN::compare, h:514-558
void compare(const C& q, const C& t, C& res) {
    res.clear();
    C iv;
    for (const S& v : q) iv.push_back(mk(v.s, v.e, -1));
    for (const S& v : t) iv.push_back(mk(v.s, v.e, 1));
    if (iv.empty()) return;
    std::sort(iv.begin(), iv.end(), [](const S& a, const S& b) { return a.s == b.s ? a.e < b.e : a.s < b.s; });
    res.push_back(iv[0]);
    S left, inter, right, tmp;
    for (size_t i = 1; i < iv.size(); i++) {
        seg_split(res.back(), iv[i], left, inter, right);
        if (right.empty()) seg_split(iv[i], res.back(), tmp, inter, right);
        res.pop_back();
        if (!left.empty()) res.push_back(left);
        if (!inter.empty()) { inter.ph = 0; res.push_back(inter); }
        if (!right.empty()) res.push_back(right);
    }
}

// ---- SEQ (common.hpp:300-496) -------------------------------------------------------------------
struct CodonMap {
    std::map<std::string, char> m;
    CodonMap() { for (int i = 0; i < ORFX_CODON_TABLE_N; i++) m[std::string(ORFX_CODON_TABLE[i], 3)] = ORFX_CODON_TABLE[i][3]; }
    char aa(const std::string& nt, size_t i) const {   // codon_map[nts.substr(i,3)]
        auto it = m.find(nt.substr(i, 3));
        return it == m.end() ? '\0' : it->second;
    }
};
const CodonMap& codons() { static CodonMap c; return c; }

char nt_complement(char c) {   // gclib ntComplement as provided by oracle/shim (IUPAC, case preserving)
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

struct Seq {
    std::string nt, aa;
    uint32_t cds_start = 0, cds_end = 0;
    bool empty() const { return nt.empty(); }
    uint32_t cds_nt_len() const { return (cds_end + 1) - cds_start; }
    static std::string translate(const std::string& nts, uint32_t a, int b) {   // common.hpp:477-495
        b = b == 0 ? (int)nts.size() : b;
        std::string r;
        for (int i = (int)a; i + 2 <= b; i += 3) r += codons().aa(nts, (size_t)i);
        return r;
    }
    void set_cds(uint32_t a, uint32_t b) { cds_start = a; cds_end = b; aa = translate(nt, cds_start, (int)cds_end); }
    bool trim_to_stop() {   // common.hpp:389-401
        size_t k = aa.find('.');
        if (k == std::string::npos) return false;
        cds_end = cds_start + (uint32_t)(k * 3) - 1;
        aa.erase(aa.begin() + (long)k, aa.end());
        return true;
    }
    bool trim_to_start() {  // common.hpp:402-418
        size_t k = aa.find('M');
        if (k == 0 || k == std::string::npos) return false;
        cds_start += (uint32_t)(k * 3);
        aa.erase(0, k);
        return true;
    }
    // common.hpp:420-450 (uint wrap-around ends the backward walk)
    uint32_t find_inframe_codon(char c, char stop_c, std::vector<uint32_t>& pos, uint32_t start_idx, uint32_t stop_idx, bool forward, bool first) const {
        pos.clear();
        uint32_t i = start_idx;
        while (true) {
            if (!(i < nt.size())) break;
            char cur = codons().aa(nt, i);
            if (cur == stop_c) break;
            if (cur == c) { pos.push_back(i); if (first) break; }
            bool term = forward ? i > stop_idx : i < stop_idx;
            if (term) break;
            i = forward ? i + 3 : i - 3;
        }
        return (uint32_t)pos.size();
    }
    void extend_to_pos(uint32_t p) {  // common.hpp:451-465
        if (p >= cds_start && p <= cds_end) return;
        cds_start = std::min(p, cds_start);
        cds_end = std::max(p, cds_end);
        aa = translate(nt, cds_start, (int)cds_end);
    }
};

// ---- TX (Transcriptome.h:567-838, Transcriptome.cpp:92-762) ----------------------------------
struct Bn {   // what a TX needs from its Bundle (Transcriptome.cpp:770-782)
    const uint8_t* seq = nullptr;
    int bstart = 0;
    int64_t len = 0;   // bases behind `seq`
    // Positions outside the stored bases read as 'N'.  The reference has real genome there (its bundle string spans the whole
    // window) but can only reach them as 5' UTR of an exon that `extend_len - starts[0]` grew past the flank (cpp:461-474):
    // bases upstream of the start codon it has just chosen, which nothing reads again.
    std::string nts(int s, int e) const {
        std::string r((size_t)std::max(e + 1 - s, 0), 'N');
        const int64_t a = std::max<int64_t>((int64_t)s - bstart, 0), z = std::min<int64_t>((int64_t)e - bstart, len - 1);
        if (z >= a) memcpy(&r[(size_t)(a - ((int64_t)s - bstart))], seq + a, (size_t)(z - a + 1));
        return r;
    }
};

struct T {
    char strand = '.';
    int cds_start = 0, cds_end = 0, cds_phase = 0;
    C exons, cds;
    Seq seq;
    const Bn* bn = nullptr;
    // --overhang: genomic bases usable beyond the low / high end of the exon chain (orfx_batch::tx_ext_lo / tx_ext_hi:
    // the flank of TX::extend_seq after its clamp to the bundle window and its --spliced_overhang cut, cpp:342-399)
    int ext_lo = 0, ext_hi = 0;
    // ORFX_EXT_*_VIRTUAL: the one-base flank on that side re-reads the transcript's own edge base at the end of the
    // chromosome (clamp of cpp:352-364).  Positions beyond [real_lo, real_hi] do not exist: Bundle::get_nts (cpp:776-782)
    // returns a string that ends early for them (std::string::substr clamps the length; below the window start the
    // reference asserts / throws instead -- mirrored here as the same short read).
    bool virt_lo = false, virt_hi = false;
    int real_lo = 0, real_hi = 0;
    std::string exon_nts(const S& c) const {
        int s_ = (int)c.s, e_ = (int)c.e;
        if (virt_lo) s_ = std::max(s_, real_lo);
        if (virt_hi) e_ = std::min(e_, real_hi);
        return e_ < s_ ? std::string() : bn->nts(s_, e_);
    }

    int cds_len() const { return clen(cds); }
    // the string TX::extend_seq adds (cpp:342-399), 5'->3' on the transcript's strand
    std::string extension(bool towards_stop) const {
        const bool high = (strand == '+') == towards_stop;
        const int n = high ? ext_hi : ext_lo;
        if (n <= 0) return "";
        std::string g = high ? bn->nts((int)exons.back().e + 1, (int)exons.back().e + n) : bn->nts((int)exons.front().s - n, (int)exons.front().s - 1);
        if (strand != '+') {
            std::string r;
            for (int i = (int)g.size() - 1; i >= 0; i--) r += nt_complement(g[(size_t)i]);
            g = r;
        }
        for (char& ch : g) ch = (char)toupper((unsigned char)ch);
        return g;
    }
    void build_cds() {   // cpp:92-122
        cds.clear();
        if (cds_start == 0 && cds_end == 0) return;
        if (cds_phase != 0) {
            if (strand == '+') cds_start = n2g(exons, g2n(exons, cds_start, strand) + cds_phase, strand);
            else cds_end = n2g(exons, g2n(exons, cds_end, strand) + cds_phase, strand);
            cds_phase = 0;
        }
        for (S e : exons) {
            if (cds_end < (int)e.s || cds_start > (int)e.e) continue;
            if (cds_start >= (int)e.s && cds_start <= (int)e.e) e.s = (uint32_t)cds_start;
            if (cds_end >= (int)e.s && cds_end <= (int)e.e) e.e = (uint32_t)cds_end;
            cds.push_back(mk(e.s, e.e, 0));
        }
        assign_phase(cds, strand, 0);
    }
    void correct_chain_len() {   // cpp:123-130
        int nt = clen(cds) % 3;
        if (nt > 0) {
            trim_to_pos(cds, 0, clen(cds) - nt - 1, strand);
            cds_start = (int)cds.front().s;
            cds_end = (int)cds.back().e;
        }
    }
    void remove_cds() { cds.clear(); cds_start = cds_end = cds_phase = 0; seq = Seq(); }   // cpp:131-138
    void load_seq() {   // cpp:683-697, 155-171
        seq = Seq();
        std::string nt;
        if (strand == '+') for (const S& c : exons) nt += exon_nts(c);
        else for (int ci = (int)exons.size() - 1; ci >= 0; ci--) {
            std::string tmp = exon_nts(exons[ci]);
            for (int i = (int)tmp.size() - 1; i >= 0; i--) nt += nt_complement(tmp[i]);
        }
        for (char& ch : nt) ch = (char)toupper((unsigned char)ch);
        seq.nt = nt;
        int ss = strand == '+' ? cds_start : cds_end, se = strand == '+' ? cds_end : cds_start;
        seq.set_cds((uint32_t)g2n(exons, ss, strand), (uint32_t)g2n(exons, se, strand));
    }
    int adjust_stop() {   // cpp:182-201
        if (seq.empty()) return 0;
        bool trimmed = seq.trim_to_stop();
        if (seq.cds_nt_len() < 3) { remove_cds(); return -1; }
        if (trimmed) {
            trim_to_pos(cds, 0, (int)seq.cds_nt_len() - 1, strand);
            cds_start = (int)cds.front().s; cds_end = (int)cds.back().e;
        }
        return 0;
    }
    void extend_to_stop(int extend_len) {   // cpp:400-442
        if (seq.aa[seq.aa.size() - 1] == '.') return;
        if (extend_len > 0) seq.nt += extension(true);   // SEQ::append_seq, common.hpp:311-314
        std::vector<uint32_t> stops;
        uint32_t res = seq.find_inframe_codon('.', ';', stops, seq.cds_end + 1, (uint32_t)seq.nt.size(), true, true);
        if (res != 0) {
            if (extend_len > 0) {   // stop inside the extension: the terminal exon grows up to it (cpp:418-429)
                int extension_len = std::max(((int)(stops[0] + 2) - clen(exons)) + 1, 0);
                if (strand == '+') exons.back().e += (uint32_t)extension_len;
                else exons.front().s -= (uint32_t)extension_len;
            }
            int ext = (int)(stops[0] + 2) - (int)seq.cds_end;
            seq.extend_to_pos(stops[0] + 2);
            extend(cds, exons, ext, strand == '+');
            cds_start = (int)cds.front().s; cds_end = (int)cds.back().e;
        }
        load_seq();
    }
    void extend_to_start_at(int new_start) {   // cpp:446-492, new_start given
        int ext = (int)seq.cds_start - new_start;
        seq.extend_to_pos((uint32_t)new_start);
        extend(cds, exons, ext, strand == '-');
        cds_start = (int)cds.front().s; cds_end = (int)cds.back().e;
        load_seq();
    }
    void extend_to_start_search(int extend_len) {   // cpp:446-492, new_start == -1
        if (seq.aa[0] == 'M') return;
        if (extend_len > 0) {   // SEQ::push_front, common.hpp:317-322
            const std::string e5 = extension(false);
            seq.cds_start += (uint32_t)e5.size(); seq.cds_end += (uint32_t)e5.size();
            seq.nt = e5 + seq.nt;
        }
        std::vector<uint32_t> starts;
        uint32_t res = seq.find_inframe_codon('M', '.', starts, seq.cds_start, 0, false, true);
        if (res == 0) { load_seq(); return; }
        if (extend_len > 0) {   // as written: measured against --overhang, not against the bases actually added (cpp:461-474)
            int extension_len = std::max(extend_len - (int)starts[0], 0);
            if (strand == '+') exons.front().s -= (uint32_t)extension_len;
            else exons.back().e += (uint32_t)extension_len;
        }
        extend_to_start_at((int)starts[0]);
    }
    bool extend_to_start_guided(const T& ref, bool non_aug) {   // cpp:496-620
        char start_codon = seq.aa[0];
        bool match_start = strand == '+' ? cds_start == ref.cds_start : cds_end == ref.cds_end;
        if ((start_codon == 'M' || non_aug) && match_start) return true;
        if (non_aug) {
            int ref_5p = strand == '+' ? ref.cds_start : ref.cds_end;
            int this_5p = strand == '+' ? cds_start : cds_end;
            int ref_in_this = g2n(exons, ref_5p, strand);
            int q_in_this = g2n(exons, this_5p, strand);
            if (ref_in_this >= 0 && ref_in_this % 3 == q_in_this % 3) {
                std::vector<uint32_t> stops;
                uint32_t res = seq.find_inframe_codon('.', ';', stops, seq.cds_start, (uint32_t)ref_in_this, false, true);
                if (res == 0) { extend_to_start_at(ref_in_this); return true; }
            }
        }
        std::vector<uint32_t> tstarts;
        uint32_t res = seq.find_inframe_codon('M', '.', tstarts, seq.cds_start, 0, false, false);
        if (res == 0) return false;
        int max_ext = (int)seq.cds_start - (int)tstarts.back();
        std::vector<uint32_t> gstarts;
        for (uint32_t s : tstarts) gstarts.push_back((uint32_t)n2g(exons, (int)s, strand));
        int old_start = strand == '+' ? ref.cds_start : ref.cds_end;
        for (size_t si = 0; si < gstarts.size(); si++)
            if ((int)gstarts[si] == old_start) { extend_to_start_at((int)tstarts[si]); return true; }
        extend(cds, exons, max_ext, strand == '-');
        C ic;
        intersection(cds, ref.cds, ic, false);
        uint32_t best = (uint32_t)gstarts.size() - 1;
        auto visit = [&](const S& c) {
            uint32_t t_tpos = (uint32_t)g2n(ref.cds, (int)(strand == '+' ? c.s : c.e), strand);
            uint32_t q_tpos = (uint32_t)g2n(cds, (int)(strand == '+' ? c.s : c.e), strand);
            for (int n = (int)best; n >= 0; n--) { if (tstarts[(size_t)n] <= q_tpos) best = (uint32_t)n; else break; }
            return (t_tpos % 3) == (q_tpos % 3);
        };
        if (strand == '+') { for (size_t i = 0; i < ic.size(); i++) if (visit(ic[i])) break; }
        else { for (int i = (int)ic.size() - 1; i >= 0; i--) if (visit(ic[(size_t)i])) break; }
        int start_nt = (int)tstarts[best] - (int)tstarts.back();
        trim_to_pos(cds, start_nt, clen(cds), strand);
        seq.extend_to_pos(tstarts[best]);
        cds_start = (int)cds.front().s; cds_end = (int)cds.back().e;
        return true;
    }
    int adjust_start() {   // cpp:202-227
        if (seq.empty()) return 0;
        uint32_t old_start = seq.cds_start;
        bool trimmed = seq.trim_to_start();
        if (trimmed) {
            uint32_t end_nt = seq.cds_end - old_start;
            uint32_t start_nt = (end_nt + 1) - seq.cds_nt_len();
            trim_to_pos(cds, (int)start_nt, (int)end_nt, strand);
            cds_start = (int)cds.front().s; cds_end = (int)cds.back().e;
        }
        return 0;
    }
    int rescue_unguided(bool non_aug, int extend_len = 0) {   // cpp:641-677 with t == nullptr (Transcriptome::clean_cds, cpp:1075)
        if (seq.cds_nt_len() < 3) { remove_cds(); return 0; }
        adjust_stop();
        if (cds_len() == 0) { remove_cds(); return 0; }
        extend_to_stop(extend_len);
        if (!non_aug) extend_to_start_search(extend_len);
        adjust_start();
        cds_phase = 0;
        assign_phase(cds, strand, 0);
        return 1;
    }
    int rescue(bool non_aug, const T& t, int extend_len = 0) {   // cpp:641-677
        if (seq.cds_nt_len() < 3) { remove_cds(); return 0; }
        adjust_stop();
        if (cds_len() == 0) { remove_cds(); return 0; }
        extend_to_stop(extend_len);
        bool found = extend_to_start_guided(t, non_aug);
        if (!found) extend_to_start_search(extend_len);
        bool run_adjust = true;
        if (non_aug) run_adjust = !(strand == '+' ? cds_start == t.cds_start : cds_end == t.cds_end);
        if (run_adjust) adjust_start();
        cds_phase = 0;
        assign_phase(cds, strand, 0);
        return 1;
    }
};

struct Sc {   // Score, common.hpp:32-65
    int qlen = 0, tlen = 0, ulen = 0, query_start = 0;
    bool start_match = false, pass = true;
    int extra = 0, missing = 0, match = 0, outframe = 0, inframe = 0;
    float lpi = 0, ilpi = 0, mlpi = 0;
};

Sc score(const T& q, const T& t) {   // TX::score, cpp:699-762
    Sc s;
    s.qlen = clen(q.cds);
    s.tlen = clen(t.cds);
    s.ulen = union_len(q.cds, t.cds);
    C mods;
    compare(q.cds, t.cds, mods);
    int tf = 0, qf = 0;
    auto visit = [&](const S& m) {
        if (m.ph == -1) { s.extra += (int)m.slen(); qf += (int)m.slen(); }
        else if (m.ph == 1) { s.missing += (int)m.slen(); tf += (int)m.slen(); }
        else { s.match += (int)m.slen(); if (qf % 3 == tf % 3) s.inframe += (int)m.slen(); else s.outframe += (int)m.slen(); }
    };
    if (q.strand == '+') for (size_t i = 0; i < mods.size(); i++) visit(mods[i]);
    else for (int i = (int)mods.size() - 1; i >= 0; i--) visit(mods[(size_t)i]);
    s.lpi = 100.0 * ((float)s.qlen / (float)s.ulen);
    s.ilpi = 100.0 * ((float)s.inframe / (float)s.ulen);
    s.mlpi = 100.0 * ((float)s.match / (float)s.ulen);
    s.query_start = q.strand == '+' ? q.cds_start - (int)q.exons.front().s : (int)q.exons.back().e - q.cds_end;
    s.start_match = q.strand == t.strand && (q.strand == '+' ? q.cds_start : q.cds_end) == (q.strand == '+' ? t.cds_start : t.cds_end);
    return s;
}

struct Row {
    int q_tx, t_tx;
    S seg;
    int key_s, key_e;
    Sc sc;
    int status;
    int n_dup = 0;
    int ext_lo = 0, ext_hi = 0;   // growth of the row's query copy (--overhang)
    int32_t aln[4] = {0, 0, 0, 0};   // ez.max, ez.score, ez.mqe, ez.mte (params.align_enabled; 0 when the row was not scored)
    int64_t idx;
};

struct Ctx {
    orfx_params p;
    std::vector<orfx_best> best;
    std::vector<int32_t> ext_lo, ext_hi, aln[4];
    std::vector<int32_t> q_tx, t_tx, seg_s, seg_e, cds_s, cds_e, n_dup, qlen, tlen, ulen, qstart, extra, missing, match, outframe, inframe;
    std::vector<uint8_t> status, pass, start_match;
    std::vector<float> lpi, ilpi, mlpi;
    std::vector<int64_t> order;
    bool submitted = false;
};

bool score_gt(const orfx_params& p, const Sc& l, const Sc& r) {   // orfanage.cpp:204-245
    if (l.pass != r.pass) return l.pass > r.pass;
    for (int i = 0; i < p.n_modes; i++) {
        switch (p.mode_array[i]) {
            case ORFX_MODE_START_MATCH: if (l.start_match != r.start_match) return l.start_match > r.start_match; break;
            case ORFX_MODE_ALL: break;
            case ORFX_MODE_LONGEST: if (l.qlen != r.qlen) return l.qlen > r.qlen; break;
            case ORFX_MODE_LONGEST_MATCH: if (l.inframe != r.inframe) return l.inframe > r.inframe; break;
            case ORFX_MODE_BEST: if (l.ilpi != r.ilpi) return l.ilpi > r.ilpi; break;
            case ORFX_MODE_FIRST: if (l.query_start != r.query_start) return l.query_start < r.query_start; break;
            default: break;
        }
    }
    return l.tlen > r.tlen;
}

// Segment dictionary (orfx_batch::seg_ref, include/orfx.h): when present it is the only source of the chains.  A copy of the
// batch header whose segs / seg_phase / seg_seq_off point at the expanded records.
struct Expanded {
    orfx_batch b;
    std::vector<orfx_seg> segs;
    std::vector<int8_t> phase;
    std::vector<int32_t> sso;
    explicit Expanded(const orfx_batch* in) : b(*in) {
        if (!in->seg_ref || in->n_segs <= 0) return;
        if (!in->bundle_dict_off || !in->dict_segs || !in->dict_phase || in->n_dict <= 0) throw std::runtime_error("seg_ref without a dictionary");
        segs.resize((size_t)in->n_segs); phase.resize((size_t)in->n_segs);
        if (in->dict_seq_off) sso.resize((size_t)in->n_segs);
        for (int bi = 0; bi < in->n_bundles; bi++) {
            const int64_t d0 = in->bundle_dict_off[bi], d1 = in->bundle_dict_off[bi + 1];
            if (d0 < 0 || d1 < d0 || d1 > in->n_dict) throw std::runtime_error("bundle_dict_off not monotone / beyond n_dict");
            for (int x = in->bundle_tx_off[bi]; x < in->bundle_tx_off[bi + 1]; x++)
                for (int k = in->tx_seg_off[x]; k < in->tx_seg_off[x + 1]; k++) {
                    const int64_t d = d0 + in->seg_ref[k];
                    if (d >= d1) throw std::runtime_error("seg_ref outside its bundle's dictionary records");
                    segs[(size_t)k] = in->dict_segs[d]; phase[(size_t)k] = in->dict_phase[d];
                    if (in->dict_seq_off) sso[(size_t)k] = in->dict_seq_off[d];
                }
        }
        b.segs = segs.data(); b.seg_phase = phase.data(); b.seg_seq_off = in->dict_seq_off ? sso.data() : nullptr;
    }
};

T load_tx(const orfx_batch* b, int tx) {
    T t;
    t.strand = (char)b->tx_strand[tx];
    for (int k = b->tx_seg_off[tx]; k < b->tx_seg_off[tx + 1]; k++) t.exons.push_back(mk((uint32_t)b->segs[k].start, (uint32_t)b->segs[k].end, 0));
    if (b->tx_kind[tx] == ORFX_TX_TEMPLATE) {
        t.cds = t.exons;
        int k0 = b->tx_seg_off[tx];
        for (size_t i = 0; i < t.cds.size(); i++) t.cds[i].ph = b->seg_phase[k0 + (int)i];
        t.cds_start = b->tx_cds_start[tx];
        t.cds_end = b->tx_cds_end[tx];
    } else if (b->tx_kind[tx] == ORFX_TX_QUERY && b->tx_ext_lo && b->tx_ext_hi) {
        t.ext_lo = b->tx_ext_lo[tx];
        t.ext_hi = b->tx_ext_hi[tx];
        if (b->tx_ext_flags && !t.exons.empty()) {
            t.virt_lo = (b->tx_ext_flags[tx] & ORFX_EXT_LO_VIRTUAL) != 0; t.virt_hi = (b->tx_ext_flags[tx] & ORFX_EXT_HI_VIRTUAL) != 0;
            t.real_lo = (int)t.exons.front().s; t.real_hi = (int)t.exons.back().e;
        }
    }
    return t;
}

// The bundle's sequence as the reference holds it (Bundle::bundle_seq: the whole window).  With the
// islands layout of orfx_batch only the bases under query exons were packed: the window is rebuilt
// here with 'N' everywhere else, so that the rest of this file indexes it exactly as the reference
// does (an access outside the query exons would then show up as a mismatch in the parity tests).
// A strand '.' query with a 5' flank carries a private copy of its bases (include/orfx.h, orfx_batch::tx_ext_flags): the
// bases stored above its last exon are not the genome's.  It is left out of the bundle's window and read through a window
// of its own.
bool has_private_copy(const orfx_batch* b, int x) {
    return b->seg_seq_off && b->tx_ext_hi && b->tx_kind[x] == ORFX_TX_QUERY && b->tx_strand[x] != '+' && b->tx_strand[x] != '-' && b->tx_ext_hi[x] > 0;
}
struct BundleSeq {
    std::string rebuilt;
    Bn bn;
    // only_tx >= 0: the window of that one (private-copy) transcript, padded so that exon growth beyond the stored flank
    // (growth is measured against --overhang, cpp:461-474) stays inside it
    // with_templates: the batch also carries the bases under the template CDS segments (params.align_enabled)
    BundleSeq(const orfx_batch* b, int bi, int only_tx = -1, int pad = 0, bool with_templates = false) {
        bn.bstart = b->bundle_start[bi];
        if (!b->seq) { bn.seq = nullptr; return; }
        if (b->seq_format != ORFX_SEQ_RAW && !b->seg_seq_off) throw std::runtime_error("ORFX_SEQ_NT4 / ORFX_SEQ_NT2 are defined for the islands layout only");
        if (!b->seg_seq_off) { bn.seq = b->seq + b->seq_off[bi]; bn.len = b->seq_off[bi + 1] - b->seq_off[bi]; return; }
        int hi = bn.bstart;
        for (int x = b->bundle_tx_off[bi]; x < b->bundle_tx_off[bi + 1]; x++)
            if (b->tx_seg_off[x + 1] > b->tx_seg_off[x]) {
                hi = std::max(hi, (int)b->segs[b->tx_seg_off[x + 1] - 1].end + (b->tx_ext_hi ? b->tx_ext_hi[x] : 0));
                if (b->tx_ext_lo && b->tx_kind[x] == ORFX_TX_QUERY)   // a one-base flank below the chromosome's first base (ORFX_EXT_LO_VIRTUAL)
                    bn.bstart = std::min(bn.bstart, (int)b->segs[b->tx_seg_off[x]].start - b->tx_ext_lo[x]);
            }
        if (only_tx >= 0) {
            bn.bstart = (int)b->segs[b->tx_seg_off[only_tx]].start - (b->tx_ext_lo ? b->tx_ext_lo[only_tx] : 0) - pad - 1;
            hi = (int)b->segs[b->tx_seg_off[only_tx + 1] - 1].end + b->tx_ext_hi[only_tx] + pad + 1;
        }
        rebuilt.assign((size_t)(hi - bn.bstart + 1), 'N');
        for (int x = b->bundle_tx_off[bi]; x < b->bundle_tx_off[bi + 1]; x++) {
            if (b->tx_kind[x] != ORFX_TX_QUERY && !(with_templates && b->tx_kind[x] == ORFX_TX_TEMPLATE)) continue;
            if (only_tx >= 0 ? x != only_tx : has_private_copy(b, x)) continue;
            for (int k = b->tx_seg_off[x]; k < b->tx_seg_off[x + 1]; k++) {
                orfx_seg sg = b->segs[k];
                int64_t at = b->seg_seq_off[k];
                const bool isq = b->tx_kind[x] == ORFX_TX_QUERY;
                if (isq && b->tx_ext_lo && k == b->tx_seg_off[x]) { sg.start -= b->tx_ext_lo[x]; at -= b->tx_ext_lo[x]; }   // the --overhang flanks
                if (isq && b->tx_ext_hi && k == b->tx_seg_off[x + 1] - 1) sg.end += b->tx_ext_hi[x];
                if (b->seq_format == ORFX_SEQ_NT4) {   // decode the 4-bit codes back to letters ("other" -> N)
                    static const char LET[16] = {'N', 'A', 'C', 'U', 'G', 'R', 'N', 'N', 'T', 'N', 'Y', 'N', 'N', 'N', 'N', 'N'};
                    const int64_t base0 = b->seq_off[bi] + at;
                    for (int i = 0; i <= sg.end - sg.start; i++) {
                        const int64_t p = base0 + i;
                        rebuilt[(size_t)(sg.start - bn.bstart + i)] = LET[(b->seq[p >> 1] >> ((p & 1) * 4)) & 15];
                    }
                    continue;
                }
                if (b->seq_format == ORFX_SEQ_NT2) {   // 2-bit codes, then the listed exceptions of this segment on top
                    static const char LET[16] = {'N', 'A', 'C', 'U', 'G', 'R', 'N', 'N', 'T', 'N', 'Y', 'N', 'N', 'N', 'N', 'N'};
                    const int64_t base0 = b->seq_off[bi] + at, n = sg.end - sg.start + 1;
                    char* out = &rebuilt[(size_t)(sg.start - bn.bstart)];
                    for (int64_t i = 0; i < n; i++) {
                        const int64_t p = base0 + i;
                        out[i] = "ACGT"[(b->seq[p >> 2] >> ((p & 3) * 2)) & 3];
                    }
                    const int64_t* e0 = b->seq_exc_pos, * e1 = b->seq_exc_pos + b->n_seq_exc;
                    for (const int64_t* e = std::lower_bound(e0, e1, base0); e != e1 && *e < base0 + n; ++e)
                        out[*e - base0] = LET[b->seq_exc_code[e - e0] & 15];
                    continue;
                }
                memcpy(&rebuilt[(size_t)(sg.start - bn.bstart)], b->seq + b->seq_off[bi] + at, (size_t)(sg.end - sg.start + 1));
            }
        }
        bn.seq = (const uint8_t*)rebuilt.data();
        bn.len = (int64_t)rebuilt.size();
    }
};

// ---- alignment stage (params.align_enabled): Finder::align (aln.hpp:27-41) with the build definition ----------------
const int8_t MAT50[24 * 24] = ORFX_MAT50_LITERAL;   // common.hpp:239-265
// Finder::align's encoding loop (aln.hpp:35-36) through aa_table (common.hpp:268-277: the 24 letters in both cases, every
// other byte 23).  The strings come from SEQ::translate: '.' for a stop, '\0' for a triplet codon_map does not know.
std::vector<uint8_t> encode_protein(const std::string& aa) {
    static int8_t table[256]; static bool init = false;
    if (!init) {
        for (int i = 0; i < 256; i++) table[i] = 23;
        const char* order = ORFX_AA_ORDER;
        for (int i = 0; order[i]; i++) { table[(unsigned char)order[i]] = (int8_t)i; table[(unsigned char)tolower(order[i])] = (int8_t)i; }
        init = true;
    }
    std::vector<uint8_t> r(aa.size());
    for (size_t i = 0; i < aa.size(); i++) r[i] = (uint8_t)table[(unsigned char)aa[i]];
    return r;
}
// protein of a template: SEQ::translate over its spliced CDS chain from the first base (load_seq + set_cds of a transcript
// whose exon chain is its CDS chain, orfanage.cpp:351)
std::vector<uint8_t> template_protein(const T& t, const Bn& bn) {
    T c = t;
    c.bn = &bn;
    c.seq = Seq();
    std::string nt;
    if (c.strand == '+') for (const S& e : c.exons) nt += c.exon_nts(e);
    else for (int ci = (int)c.exons.size() - 1; ci >= 0; ci--) { std::string tmp = c.exon_nts(c.exons[(size_t)ci]); for (int i = (int)tmp.size() - 1; i >= 0; i--) nt += nt_complement(tmp[(size_t)i]); }
    for (char& ch : nt) ch = (char)toupper((unsigned char)ch);
    return encode_protein(nt.size() >= 3 ? Seq::translate(nt, 0, (int)nt.size() - 1) : std::string());
}
}  // namespace
extern "C" int orfx_oracle_align(const uint8_t* query, int qlen, const uint8_t* target, int tlen, const int8_t* mat, int m, int gapo, int gape,
                                 int32_t* out_max, int32_t* out_score, int32_t* out_mqe, int32_t* out_mte);
namespace {

void evaluate(Ctx& cx, const orfx_batch* b) {
    const orfx_params& p = cx.p;
    const bool align = p.align_enabled && p.has_reference && b->seq;
    const int gapo = p.align_gapo ? p.align_gapo : 12, gape = p.align_gape ? p.align_gape : 2;   // aln.hpp:146
    std::vector<Row> rows_all;
    cx.best.clear();
    cx.order.clear();
    for (int bi = 0; bi < b->n_bundles; bi++) {
        BundleSeq bseq(b, bi, -1, 0, align);
        const Bn& bn = bseq.bn;
        int x0 = b->bundle_tx_off[bi], x1 = b->bundle_tx_off[bi + 1];
        std::vector<T> txs;
        for (int x = x0; x < x1; x++) txs.push_back(load_tx(b, x));
        std::map<int, std::vector<uint8_t>> t_prot;   // proteins of the bundle's templates, on first use
        for (int qx = x0; qx < x1; qx++) {
            if (b->tx_kind[qx] != ORFX_TX_QUERY) continue;
            T q = txs[(size_t)(qx - x0)];
            std::unique_ptr<BundleSeq> own;
            if (has_private_copy(b, qx)) own.reset(new BundleSeq(b, bi, qx, p.overhang));
            q.bn = own ? &own->bn : &bn;
            q.remove_cds();
            std::vector<std::vector<Row>> stats;
            for (int tx = x0; tx < x1; tx++) {   // orfanage.cpp:495-637
                if (b->tx_kind[tx] != ORFX_TX_TEMPLATE) continue;
                const T& t = txs[(size_t)(tx - x0)];
                if ((int)q.exons.front().s > t.cds_end || (int)q.exons.back().e < t.cds_start) continue;
                C segments;
                int intlen = intersection(q.exons, t.cds, segments, true);
                stats.push_back(std::vector<Row>());
                if (intlen == 0) {
                    Row r; r.q_tx = qx; r.t_tx = tx; r.seg = S(); r.key_s = q.cds_start; r.key_e = q.cds_end; r.status = ORFX_ST_NO_OVERLAP; r.sc.pass = false;
                    stats.back().push_back(r);
                    continue;
                }
                std::unordered_set<int> orf_starts;
                for (S& s : segments) {
                    Row r; r.q_tx = qx; r.t_tx = tx; r.seg = s; r.key_s = q.cds_start; r.key_e = q.cds_end; r.status = ORFX_ST_OK;
                    stats.back().push_back(r);
                    Row& cur = stats.back().back();
                    int seg_phase = chain_phase_at(t.cds, t.strand == '+' ? (int)s.s : (int)s.e, t.strand);
                    if (s.slen() < 3) { cur.status = ORFX_ST_SEGMENT_LT3; cur.sc.pass = false; continue; }
                    s.ph = seg_phase;
                    T qseg = q;
                    qseg.cds_start = (int)s.s; qseg.cds_end = (int)s.e; qseg.cds_phase = s.ph;
                    qseg.build_cds();
                    if (qseg.cds_len() < 3) { cur.key_s = qseg.cds_start; cur.key_e = qseg.cds_end; cur.status = ORFX_ST_CDS_LT3; cur.sc.pass = false; continue; }
                    qseg.correct_chain_len();
                    if (qseg.cds_len() < 3) { cur.key_s = qseg.cds_start; cur.key_e = qseg.cds_end; cur.status = ORFX_ST_CDS_LT3; cur.sc.pass = false; continue; }
                    if (p.has_reference) {
                        qseg.load_seq();
                        int ret = qseg.rescue(p.non_aug != 0, t, p.overhang);
                        cur.key_s = qseg.cds_start; cur.key_e = qseg.cds_end;
                        if (!qseg.exons.empty()) { cur.ext_lo = (int)q.exons.front().s - (int)qseg.exons.front().s; cur.ext_hi = (int)qseg.exons.back().e - (int)q.exons.back().e; }
                        if (!ret) { cur.status = ORFX_ST_NOT_RESCUED; cur.sc.pass = false; continue; }
                        const std::string& aa = qseg.seq.aa;
                        bool at_t5 = qseg.strand == '+' ? qseg.cds_start == t.cds_start : qseg.cds_end == t.cds_end;
                        if (aa.back() != '.' || (!p.non_aug && aa.front() != 'M') || (p.non_aug && aa.front() != 'M' && !at_t5)) {
                            cur.sc.pass = false;
                            cur.status = aa.back() != '.' ? ORFX_ST_MISSING_STOP : ORFX_ST_MISSING_START;
                            continue;
                        }
                    }
                    if (clen(qseg.cds) < p.cds_minlen) { cur.status = ORFX_ST_MINLEN; cur.sc.pass = false; continue; }
                    Sc sc = score(qseg, t);
                    cur.sc = sc;
                    if (align) {   // the row's ORF as SEQ holds it, against its template's protein
                        if (!t_prot.count(tx)) t_prot[tx] = template_protein(t, bn);
                        const std::vector<uint8_t> qa = encode_protein(qseg.seq.aa);
                        const std::vector<uint8_t>& ta = t_prot[tx];
                        orfx_oracle_align(qa.data(), (int)qa.size(), ta.data(), (int)ta.size(), MAT50, 24, gapo, gape, &cur.aln[0], &cur.aln[1], &cur.aln[2], &cur.aln[3]);
                    }
                    if (sc.lpi < p.lpi || sc.ilpi < p.ilpi || sc.mlpi < p.mlpi) { cur.sc.pass = false; continue; }
                    if (!orf_starts.insert(qseg.cds_start).second) { cur.status = ORFX_ST_DUP; cur.sc.pass = false; continue; }
                    cur.key_s = qseg.cds_start; cur.key_e = qseg.cds_end;
                }
            }
            // flatten (orfanage.cpp:250-283) and find_best_stat (286-305)
            int64_t row_off = (int64_t)rows_all.size();
            int n_rows = 0;
            for (auto& g : stats) for (Row& r : g) { r.idx = row_off + n_rows; n_rows++; }
            std::map<std::pair<int, int>, std::vector<Row>> flat;
            for (auto& g : stats) for (Row& r : g) flat[std::make_pair(r.key_s, r.key_e)].push_back(r);
            for (auto& kv : flat) std::sort(kv.second.begin(), kv.second.end(), [&](const Row& a, const Row& c) { return score_gt(p, a.sc, c.sc); });
            orfx_best be;
            memset(&be, 0, sizeof be);
            be.q_tx = qx; be.t_tx = -1; be.best_row = -1; be.n_rows = n_rows; be.row_off = row_off;
            std::pair<int, int> best_se(-1, -1);
            if (!flat.empty()) {
                best_se = flat.begin()->first;
                Sc bs = flat.begin()->second.front().sc;
                for (auto& kv : flat) if (score_gt(p, kv.second.front().sc, bs)) { bs = kv.second.front().sc; best_se = kv.first; }
                if (!bs.pass) best_se = std::make_pair(-1, -1);
            }
            bool found = !flat.empty() && best_se.first >= 0;
            bool mode_all = p.n_modes > 0 && p.mode_array[0] == ORFX_MODE_ALL;
            if (found) {
                if (mode_all) {   // orfanage.cpp:700-709
                    for (auto& kv : flat) {
                        if (!kv.second.front().sc.pass) continue;
                        if (kv.second.size() > 1) { std::set<int> d; for (Row& v : kv.second) d.insert(b->tx_tid_class[v.t_tx]); kv.second.front().n_dup = (int)d.size(); }
                    }
                } else {          // orfanage.cpp:755-766
                    auto& v = flat[best_se];
                    std::set<int> d;
                    for (size_t k = 1; k < v.size(); k++) d.insert(b->tx_tid_class[v[k].t_tx]);
                    v.front().n_dup = (int)d.size();
                }
                const Row& w = flat[best_se].front();
                be.t_tx = w.t_tx; be.cds_start = w.key_s; be.cds_end = w.key_e; be.duplicity = w.n_dup; be.best_row = w.idx;
                be.lpi = w.sc.lpi; be.ilpi = w.sc.ilpi; be.mlpi = w.sc.mlpi; be.start_match = w.sc.start_match;
                be.ext_lo = w.ext_lo; be.ext_hi = w.ext_hi;
            }
            cx.best.push_back(be);
            // rows in candidate order, with n_dup copied back from the flattened copies
            size_t base = rows_all.size();
            for (auto& g : stats) for (Row& r : g) rows_all.push_back(r);
            for (auto& kv : flat) for (Row& r : kv.second) { rows_all[(size_t)r.idx].n_dup = r.n_dup; cx.order.push_back(r.idx); }
            (void)base;
        }
    }
    size_t n = rows_all.size();
    auto rs = [&](std::vector<int32_t>& v) { v.resize(n); };
    rs(cx.q_tx); rs(cx.t_tx); rs(cx.seg_s); rs(cx.seg_e); rs(cx.cds_s); rs(cx.cds_e); rs(cx.n_dup); rs(cx.qlen); rs(cx.tlen); rs(cx.ulen);
    rs(cx.ext_lo); rs(cx.ext_hi);
    for (int k = 0; k < 4; k++) rs(cx.aln[k]);
    rs(cx.qstart); rs(cx.extra); rs(cx.missing); rs(cx.match); rs(cx.outframe); rs(cx.inframe);
    cx.status.resize(n); cx.pass.resize(n); cx.start_match.resize(n); cx.lpi.resize(n); cx.ilpi.resize(n); cx.mlpi.resize(n);
    for (size_t i = 0; i < n; i++) {
        const Row& r = rows_all[i];
        cx.q_tx[i] = r.q_tx; cx.t_tx[i] = r.t_tx; cx.seg_s[i] = (int32_t)r.seg.s; cx.seg_e[i] = (int32_t)r.seg.e;
        cx.cds_s[i] = r.key_s; cx.cds_e[i] = r.key_e; cx.n_dup[i] = r.n_dup; cx.status[i] = (uint8_t)r.status; cx.pass[i] = r.sc.pass;
        cx.start_match[i] = r.sc.start_match; cx.qlen[i] = r.sc.qlen; cx.tlen[i] = r.sc.tlen; cx.ulen[i] = r.sc.ulen; cx.qstart[i] = r.sc.query_start;
        cx.extra[i] = r.sc.extra; cx.missing[i] = r.sc.missing; cx.match[i] = r.sc.match; cx.outframe[i] = r.sc.outframe; cx.inframe[i] = r.sc.inframe;
        cx.lpi[i] = r.sc.lpi; cx.ilpi[i] = r.sc.ilpi; cx.mlpi[i] = r.sc.mlpi;
        cx.ext_lo[i] = r.ext_lo; cx.ext_hi[i] = r.ext_hi;
        for (int k = 0; k < 4; k++) cx.aln[k][i] = r.aln[k];
    }
}

}  // namespace

extern "C" {

const char* orfx_oracle_last_error(void) { return g_err.c_str(); }

int orfx_oracle_create(orfx_ctx** out, int device, const orfx_params* params) {
    (void)device;
    if (!out || !params) { g_err = "null argument"; return ORFX_E_INVALID; }
    if (params->overhang < 0) { g_err = "negative overhang"; return ORFX_E_INVALID; }
    Ctx* c = new Ctx();
    c->p = *params;
    *out = reinterpret_cast<orfx_ctx*>(c);
    return ORFX_OK;
}

int orfx_oracle_submit(orfx_ctx* ctx, const orfx_batch* batch) {
    Ctx* c = reinterpret_cast<Ctx*>(ctx);
    if (!c || !batch) { g_err = "null argument"; return ORFX_E_INVALID; }
    if (c->p.overhang > 0 && c->p.has_reference && batch->n_tx > 0 && (!batch->tx_ext_lo || !batch->tx_ext_hi)) { g_err = "--overhang needs tx_ext_lo / tx_ext_hi"; return ORFX_E_INVALID; }
    try { const Expanded ex(batch); evaluate(*c, &ex.b); } catch (const std::exception& e) { g_err = e.what(); return ORFX_E_INVALID; }
    c->submitted = true;
    return ORFX_OK;
}

int orfx_oracle_wait(orfx_ctx* ctx, orfx_results* r) {
    Ctx* c = reinterpret_cast<Ctx*>(ctx);
    if (!c || !r) { g_err = "null argument"; return ORFX_E_INVALID; }
    if (!c->submitted) { g_err = "wait without submit"; return ORFX_E_STATE; }
    memset(r, 0, sizeof *r);
    r->n_queries = (int32_t)c->best.size();
    r->n_rows = (int64_t)c->q_tx.size();
    r->best = c->best.data();
    r->row_q_tx = c->q_tx.data(); r->row_t_tx = c->t_tx.data(); r->row_seg_start = c->seg_s.data(); r->row_seg_end = c->seg_e.data();
    r->row_cds_start = c->cds_s.data(); r->row_cds_end = c->cds_e.data(); r->row_status = c->status.data(); r->row_pass = c->pass.data();
    r->row_start_match = c->start_match.data(); r->row_n_dup = c->n_dup.data(); r->row_qlen = c->qlen.data(); r->row_tlen = c->tlen.data();
    r->row_ulen = c->ulen.data(); r->row_query_start = c->qstart.data(); r->row_extra = c->extra.data(); r->row_missing = c->missing.data();
    r->row_match = c->match.data(); r->row_outframe = c->outframe.data(); r->row_inframe = c->inframe.data();
    r->row_lpi = c->lpi.data(); r->row_ilpi = c->ilpi.data(); r->row_mlpi = c->mlpi.data(); r->print_order = c->order.data();
    if (c->p.overhang > 0) { r->row_ext_lo = c->ext_lo.data(); r->row_ext_hi = c->ext_hi.data(); }
    if (c->p.align_enabled && c->p.has_reference && r->n_rows > 0) {
        r->row_aln_max = c->aln[0].data(); r->row_aln_score = c->aln[1].data(); r->row_aln_mqe = c->aln[2].data(); r->row_aln_mte = c->aln[3].data();
    }
    return ORFX_OK;
}

int orfx_oracle_release(orfx_ctx* ctx) { (void)ctx; return ORFX_OK; }
void orfx_oracle_destroy(orfx_ctx* ctx) { delete reinterpret_cast<Ctx*>(ctx); }

// Template cleaning oracle: the per-transcript body of Transcriptome::clean_cds (Transcriptome.cpp:1063-1100)
// on a cleaning batch (see orfx_clean_templates in include/orfx.h).
int orfx_oracle_clean_templates(orfx_ctx* ctx, const orfx_batch* b_in, int rescue, orfx_clean_out* out, int64_t* n_out) {
    Ctx* c = reinterpret_cast<Ctx*>(ctx);
    if (!c || !b_in || !out) { g_err = "null argument"; return ORFX_E_INVALID; }
    std::unique_ptr<Expanded> ex;
    try { ex.reset(new Expanded(b_in)); } catch (const std::exception& e) { g_err = e.what(); return ORFX_E_INVALID; }
    const orfx_batch* b = &ex->b;
    if (!b->seq && b->seq_bytes) { g_err = "cleaning needs the sequence"; return ORFX_E_INVALID; }
    const bool non_aug = c->p.non_aug != 0;
    int64_t k = 0;
    for (int bi = 0; bi < b->n_bundles; bi++) {
        BundleSeq bs(b, bi);
        for (int x = b->bundle_tx_off[bi]; x < b->bundle_tx_off[bi + 1]; x++) {
            if (b->tx_kind[x] != ORFX_TX_QUERY) continue;
            T t = load_tx(b, x);
            std::unique_ptr<BundleSeq> own;
            if (has_private_copy(b, x)) own.reset(new BundleSeq(b, bi, x, c->p.overhang));
            t.bn = own ? &own->bn : &bs.bn;
            t.cds_start = b->tx_cds_start[x]; t.cds_end = b->tx_cds_end[x]; t.cds_phase = 0;
            t.build_cds();   // the chain the template carries at this point: exons clipped to the CDS bounds
            bool keep = !t.cds.empty();
            if (keep) {
                t.load_seq();
                bool alive = true;
                if (rescue) alive = t.rescue_unguided(non_aug, c->p.overhang) != 0;
                if (!alive || t.seq.aa.empty()) keep = false;   // rescue removed the CDS (the reference then drops it, cpp:1083-1089)
                else {
                    char start_codon = t.seq.aa[0], stop_codon = t.seq.aa[t.seq.aa.size() - 1];
                    if ((start_codon != 'M' && !non_aug) || stop_codon != '.') keep = false;
                    else if (t.cds_len() % 3 != 0) keep = false;
                }
            }
            out->keep[k] = keep ? 1 : 0;
            out->cds_start[k] = keep ? t.cds_start : 0;
            out->cds_end[k] = keep ? t.cds_end : 0;
            k++;
        }
    }
    if (n_out) *n_out = k;
    return ORFX_OK;
}

// orfcompare oracle: TX::score (Transcriptome.cpp:699-762) of every (query, template) pair of every bundle of a
// compare batch (see orfx_score_pairs in include/orfx.h); results live in the context until the next call.
int orfx_oracle_score_pairs(orfx_ctx* ctx, const orfx_batch* b_in, orfx_results* r) {
    Ctx* c = reinterpret_cast<Ctx*>(ctx);
    if (!c || !b_in || !r) { g_err = "null argument"; return ORFX_E_INVALID; }
    std::unique_ptr<Expanded> ex;
    try { ex.reset(new Expanded(b_in)); } catch (const std::exception& e) { g_err = e.what(); return ORFX_E_INVALID; }
    const orfx_batch* b = &ex->b;
    c->q_tx.clear(); c->t_tx.clear(); c->qlen.clear(); c->tlen.clear(); c->ulen.clear(); c->extra.clear(); c->missing.clear();
    c->match.clear(); c->outframe.clear(); c->inframe.clear(); c->lpi.clear(); c->ilpi.clear(); c->mlpi.clear();
    int nq = 0;
    auto load = [&](int x) {
        T t;
        t.strand = (char)b->tx_strand[x];
        for (int k = b->tx_seg_off[x]; k < b->tx_seg_off[x + 1]; k++) t.cds.push_back(mk((uint32_t)b->segs[k].start, (uint32_t)b->segs[k].end, 0));
        t.exons = t.cds;
        t.cds_start = b->tx_cds_start[x]; t.cds_end = b->tx_cds_end[x];
        return t;
    };
    for (int bi = 0; bi < b->n_bundles; bi++) {
        for (int x = b->bundle_tx_off[bi]; x < b->bundle_tx_off[bi + 1]; x++) {
            if (b->tx_kind[x] != ORFX_TX_QUERY) continue;
            nq++;
            T q = load(x);
            for (int y = b->bundle_tx_off[bi]; y < b->bundle_tx_off[bi + 1]; y++) {
                if (b->tx_kind[y] != ORFX_TX_TEMPLATE) continue;
                T t = load(y);
                Sc sc;
                if (!q.cds.empty()) sc = score(q, t);
                else { sc.tlen = clen(t.cds); sc.missing = sc.tlen; sc.lpi = sc.ilpi = sc.mlpi = 100.0 * (0.0f / 0.0f); }
                c->q_tx.push_back(x); c->t_tx.push_back(y); c->qlen.push_back(sc.qlen); c->tlen.push_back(sc.tlen); c->ulen.push_back(sc.ulen);
                c->extra.push_back(sc.extra); c->missing.push_back(sc.missing); c->match.push_back(sc.match); c->outframe.push_back(sc.outframe);
                c->inframe.push_back(sc.inframe); c->lpi.push_back(sc.lpi); c->ilpi.push_back(sc.ilpi); c->mlpi.push_back(sc.mlpi);
            }
        }
    }
    memset(r, 0, sizeof *r);
    r->n_queries = nq;
    r->n_rows = (int64_t)c->q_tx.size();
    r->row_q_tx = c->q_tx.data(); r->row_t_tx = c->t_tx.data(); r->row_qlen = c->qlen.data(); r->row_tlen = c->tlen.data(); r->row_ulen = c->ulen.data();
    r->row_extra = c->extra.data(); r->row_missing = c->missing.data(); r->row_match = c->match.data(); r->row_outframe = c->outframe.data();
    r->row_inframe = c->inframe.data(); r->row_lpi = c->lpi.data(); r->row_ilpi = c->ilpi.data(); r->row_mlpi = c->mlpi.data();
    return ORFX_OK;
}

// Codon-scan oracle: stop / ATG masks of one spliced transcript, bit p = codon starting at spliced
// position p (5'->3').  Follows SEQ::translate / find_inframe_codon (common.hpp:420-450, 477-495):
// upper-cased bases, reverse-complemented for non-'+' strands (Transcriptome.cpp:155-171).
int64_t orfx_oracle_masks(const orfx_batch* b_in, int32_t tx, uint32_t* stop_bits, uint32_t* start_bits, int64_t max_words) {
    std::unique_ptr<Expanded> ex;
    try { ex.reset(new Expanded(b_in)); } catch (const std::exception&) { return ORFX_E_INVALID; }
    const orfx_batch* b = &ex->b;
    int bi = 0;
    while (bi < b->n_bundles && !(tx >= b->bundle_tx_off[bi] && tx < b->bundle_tx_off[bi + 1])) bi++;
    if (bi >= b->n_bundles) return ORFX_E_INVALID;
    BundleSeq bseq(b, bi, has_private_copy(b, tx) ? tx : -1);
    const Bn& bn = bseq.bn;
    T t = load_tx(b, tx);
    t.bn = &bn;
    if (!t.exons.empty()) { t.exons.front().s -= (uint32_t)t.ext_lo; t.exons.back().e += (uint32_t)t.ext_hi; }   // the scanned chain covers the --overhang flanks
    std::string nt;
    if (t.strand == '+') for (const S& c : t.exons) nt += bn.nts((int)c.s, (int)c.e);
    else for (int ci = (int)t.exons.size() - 1; ci >= 0; ci--) { std::string tmp = bn.nts((int)t.exons[ci].s, (int)t.exons[ci].e); for (int i = (int)tmp.size() - 1; i >= 0; i--) nt += nt_complement(tmp[i]); }
    for (char& ch : nt) ch = (char)toupper((unsigned char)ch);
    int64_t L = (int64_t)nt.size(), words = (L + 31) / 32;
    if (words > max_words) return ORFX_E_INVALID;
    for (int64_t w = 0; w < words; w++) { stop_bits[w] = 0; start_bits[w] = 0; }
    for (int64_t p = 0; p < L; p++) {
        char a = codons().aa(nt, (size_t)p);
        if (a == '.') stop_bits[p >> 5] |= 1u << (p & 31);
        if (a == 'M') start_bits[p >> 5] |= 1u << (p & 31);
    }
    return L;
}

// Protein alignment oracle — PARITY UNPINNED.  The reference's aln.hpp:27-41 calls ksw_extz2_sse of
// lh3/ksw2 (un-vendored, un-pinned submodule, .gitmodules:7-9; no call site, no test).  This is the
// published scalar recurrence of ksw2's ksw_extz (score-only branch) with the build definition of
// m = 24, mat = mat50, unbanded (w = -1), no z-drop, end_bonus 0.  `mat` is the
// 24x24 matrix, row-major.  Returns max / score / mqe / mte exactly as ksw_extz_t would hold them.
int orfx_oracle_align(const uint8_t* query, int qlen, const uint8_t* target, int tlen, const int8_t* mat, int m,
                      int gapo, int gape, int32_t* out_max, int32_t* out_score, int32_t* out_mqe, int32_t* out_mte) {
    const int32_t NEG_INF = -0x40000000;
    if (qlen <= 0 || tlen <= 0) { *out_max = 0; *out_score = NEG_INF; *out_mqe = NEG_INF; *out_mte = NEG_INF; return 0; }
    std::vector<int32_t> H((size_t)qlen + 1), E((size_t)qlen + 1);
    int32_t gapoe = gapo + gape;
    uint32_t ez_max = 0;
    int32_t mqe = NEG_INF, mte = NEG_INF, sc = NEG_INF;
    H[0] = 0; E[0] = -gapoe - gapoe;
    for (int j = 1; j <= qlen; j++) { H[(size_t)j] = -(gapoe + gape * (j - 1)); E[(size_t)j] = -(gapoe + gapoe + gape * j); }
    for (int i = 0; i < tlen; i++) {
        int32_t f, h1, mx = NEG_INF;
        const int8_t* row = mat + (size_t)target[i] * (size_t)m;
        h1 = -(gapoe + gape * i);
        f = -(gapoe + gapoe + gape * i);
        for (int j = 0; j < qlen; j++) {
            int32_t h = H[(size_t)j], e = E[(size_t)j];
            H[(size_t)j] = h1;
            h += row[query[j]];
            h = h >= e ? h : e;
            h = h >= f ? h : f;
            h1 = h;
            mx = mx > h ? mx : h;
            h -= gapoe;
            e -= gape;
            e = e > h ? e : h;
            E[(size_t)j] = e;
            f -= gape;
            f = f > h ? f : h;
        }
        H[(size_t)qlen] = h1; E[(size_t)qlen] = NEG_INF;
        if (H[(size_t)qlen] > mqe) mqe = H[(size_t)qlen];
        if (i == tlen - 1) for (int j = 0; j < qlen; j++) if (H[(size_t)j + 1] > mte) mte = H[(size_t)j + 1];
        if (mx > (int32_t)ez_max) ez_max = (uint32_t)mx;
        if (i == tlen - 1) sc = H[(size_t)qlen];
    }
    *out_max = (int32_t)ez_max; *out_score = sc; *out_mqe = mqe; *out_mte = mte;
    return 0;
}

// The same recurrence over a batch of pairs, OpenMP over pairs (CPU baseline of the alignment benchmark;
// sequences concatenated as in orfx_align_batch).  threads <= 0: all cores.  Returns the threads used.
int orfx_oracle_align_batch(const uint8_t* aa, const int64_t* off, const int32_t* pair_q, const int32_t* pair_t, int64_t n_pairs,
                            const int8_t* mat, int m, int gapo, int gape, int threads, int32_t* out_max, int32_t* out_score,
                            int32_t* out_mqe, int32_t* out_mte) {
    int used = 1;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
    used = threads;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
#endif
    for (int64_t p = 0; p < n_pairs; p++) {
        const int64_t q0 = off[pair_q[p]], q1 = off[pair_q[p] + 1], t0 = off[pair_t[p]], t1 = off[pair_t[p] + 1];
        orfx_oracle_align(aa + q0, (int)(q1 - q0), aa + t0, (int)(t1 - t0), mat, m, gapo, gape, out_max + p, out_score + p, out_mqe + p,
                          out_mte + p);
    }
    return used;
}

}  // extern "C"
