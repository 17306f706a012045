// In-memory FASTA access (replaces the gclib GFastaDb / GFaSeqGet / GFastaIndex use at
// Transcriptome.cpp:821-830, 849-853, 907-918 and orfanage.cpp:981-1012).
#include "orfx_host.h"
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace orfx {

void Fasta::load(const std::string& path) {
    int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("Reference FASTA file does not exist: " + path);
    struct stat st;
    fstat(fd, &st);
    size_t size = (size_t)st.st_size;
    const char* data = size ? (const char*)mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0) : nullptr;
    if (size && data == MAP_FAILED) { close(fd); throw std::runtime_error("Cannot map FASTA: " + path); }
    if (size) madvise((void*)data, size, MADV_SEQUENTIAL);
    const char* p = data;
    const char* end = data + size;
    Rec* cur = nullptr;
    while (p < end) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        const char* le = eol;
        while (le > p && le[-1] == '\r') le--;
        if (le > p) {
            if (*p == '>') {
                const char* q = p + 1;
                while (q < le && *q != ' ' && *q != '\t') q++;
                recs_.emplace_back();
                cur = &recs_.back();
                cur->name.assign(p + 1, (size_t)(q - p - 1));
                cur->file_off = (int64_t)((eol < end ? eol + 1 : end) - data);
                // reserve the rest of the record in one go: distance to the next header
                const char* nxt = (const char*)memchr(eol, '>', (size_t)(end - eol));
                cur->seq.reserve((size_t)((nxt ? nxt : end) - eol));
            } else if (cur) {
                if (cur->line_bases == 0) { cur->line_bases = (int)(le - p); cur->line_bytes = (int)((eol < end ? eol + 1 : end) - p); }
                cur->seq.append(p, (size_t)(le - p));
            }
        }
        p = eol + 1;
    }
    if (data) munmap((void*)data, size);
    close(fd);
    for (size_t i = 0; i < recs_.size(); i++) { recs_[i].seq.shrink_to_fit(); idx_[recs_[i].name] = i; }
}

const std::string* Fasta::get(const std::string& name) const {
    auto it = idx_.find(name);
    return it == idx_.end() ? nullptr : &recs_[it->second].seq;
}

// samtools-style index next to the FASTA, as orfanage.cpp:989-1011 does through GFastaIndex
bool Fasta::write_fai_if_missing(const std::string& path) const {
    std::string fai = path + ".fai";
    if (FILE* f = fopen(fai.c_str(), "r")) { fclose(f); return false; }
    FILE* f = fopen(fai.c_str(), "w");
    if (!f) throw std::runtime_error("Cannot create fasta index: " + fai);
    for (const Rec& r : recs_)
        fprintf(f, "%s\t%zu\t%lld\t%d\t%d\n", r.name.c_str(), r.seq.size(), (long long)r.file_off, r.line_bases, r.line_bytes);
    fclose(f);
    return true;
}

}  // namespace orfx
