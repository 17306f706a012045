// Indexed FASTA access (replaces the gclib GFastaDb / GFaSeqGet / GFastaIndex use at
// Transcriptome.cpp:821-830, 849-853, 907-918 and orfanage.cpp:981-1012).
#include "orfx_host.h"
#include <algorithm>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace orfx {

Fasta::~Fasta() { if (data_) munmap((void*)data_, size_); }

void Fasta::load(const std::string& path) {
    int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("Reference FASTA file does not exist: " + path);
    struct stat st;
    fstat(fd, &st);
    size_ = (size_t)st.st_size;
    data_ = size_ ? (const char*)mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd, 0) : nullptr;
    close(fd);
    if (size_ && data_ == MAP_FAILED) { data_ = nullptr; throw std::runtime_error("Cannot map FASTA: " + path); }
    fai_was_read_ = read_fai(path + ".fai");
    if (!fai_was_read_) scan();
    for (size_t i = 0; i < recs_.size(); i++) idx_[recs_[i].name] = i;
}

// name, length, offset, line bases, line bytes; rejected (-> scan) when it does not fit the file
bool Fasta::read_fai(const std::string& fai) {
    FILE* f = fopen(fai.c_str(), "r");
    if (!f) return false;
    char name[4096];
    long long len, off;
    int lb, lw;
    std::vector<Seq> recs;
    bool ok = true;
    while (fscanf(f, "%4095s %lld %lld %d %d", name, &len, &off, &lb, &lw) == 5) {
        Seq s;
        s.name = name; s.length = len; s.file_off = off; s.line_bases = lb; s.line_bytes = lw;
        if (len < 0 || off < 0 || lb <= 0 || lw < lb) { ok = false; break; }
        const int64_t last = len ? off + (len - 1) / lb * lw + (len - 1) % lb : off;
        if ((size_t)last >= size_ + (len ? 0 : 1)) { ok = false; break; }
        recs.push_back(std::move(s));
    }
    fclose(f);
    if (!ok || recs.empty()) return false;
    recs_.swap(recs);
    return true;
}

void Fasta::scan() {
    if (size_) madvise((void*)data_, size_, MADV_SEQUENTIAL);
    const char* p = data_;
    const char* end = data_ + size_;
    Seq* cur = nullptr;
    bool ragged = false, short_seen = false;
    const char* rec_begin = nullptr;
    auto finish = [&](const char* rec_end) {
        if (!cur || !ragged) return;
        cur->in_mem = true;   // lines of different lengths: the index arithmetic does not apply
        cur->mem.reserve((size_t)cur->length);
        for (const char* q = rec_begin; q < rec_end;) {
            const char* eol = (const char*)memchr(q, '\n', (size_t)(rec_end - q));
            if (!eol) eol = rec_end;
            const char* le = eol;
            while (le > q && le[-1] == '\r') le--;
            cur->mem.append(q, (size_t)(le - q));
            q = eol + 1;
        }
    };
    while (p < end) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        const char* le = eol;
        while (le > p && le[-1] == '\r') le--;
        const char* next = eol < end ? eol + 1 : end;
        if (le > p && *p == '>') {
            finish(p);
            const char* q = p + 1;
            while (q < le && *q != ' ' && *q != '\t') q++;
            recs_.emplace_back();
            cur = &recs_.back();
            cur->name.assign(p + 1, (size_t)(q - p - 1));
            cur->file_off = (int64_t)(next - data_);
            rec_begin = next;
            ragged = short_seen = false;
        } else if (cur && le > p) {
            const int nb = (int)(le - p), nw = (int)(next - p);
            if (cur->line_bases == 0) { cur->line_bases = nb; cur->line_bytes = nw; }
            else if (short_seen || nb > cur->line_bases || (nb == cur->line_bases && nw != cur->line_bytes)) ragged = true;
            if (nb < cur->line_bases) short_seen = true;   // only the last line of a record may be short
            cur->length += nb;
        } else if (cur && cur->length > 0) short_seen = true;   // an empty line inside a record
        p = next;
    }
    finish(end);
    if (size_) madvise((void*)data_, size_, MADV_RANDOM);
}

const Fasta::Seq* Fasta::find(const std::string& name) const {
    auto it = idx_.find(name);
    return it == idx_.end() ? nullptr : &recs_[it->second];
}

void Fasta::fetch(const Seq& s, int64_t pos0, int64_t len, char* dst) const {
    if (pos0 < 0 || len < 0 || pos0 + len > s.length) throw std::runtime_error("FASTA range outside of sequence " + s.name);
    if (s.in_mem) { memcpy(dst, s.mem.data() + pos0, (size_t)len); return; }
    int64_t line = pos0 / s.line_bases, col = pos0 % s.line_bases;
    const char* src = data_ + s.file_off + line * s.line_bytes + col;
    while (len > 0) {
        const int64_t take = std::min<int64_t>(len, s.line_bases - col);
        memcpy(dst, src, (size_t)take);
        dst += take; len -= take;
        src += take + (s.line_bytes - s.line_bases);
        col = 0;
    }
}

// samtools-style index next to the FASTA, as orfanage.cpp:989-1011 does through GFastaIndex
bool Fasta::write_fai_if_missing(const std::string& path) const {
    std::string fai = path + ".fai";
    if (fai_was_read_) return false;
    if (FILE* f = fopen(fai.c_str(), "r")) { fclose(f); return false; }   // present but unusable: left alone
    for (const Seq& r : recs_) if (r.in_mem) return false;                 // not expressible as an index
    FILE* f = fopen(fai.c_str(), "w");
    if (!f) throw std::runtime_error("Cannot create fasta index: " + fai);
    for (const Seq& r : recs_)
        fprintf(f, "%s\t%lld\t%lld\t%d\t%d\n", r.name.c_str(), (long long)r.length, (long long)r.file_off, r.line_bases, r.line_bytes);
    fclose(f);
    return true;
}

}  // namespace orfx
