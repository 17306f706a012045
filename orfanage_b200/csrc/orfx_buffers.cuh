// Grow-only device / pinned-host buffers owned by a context (no cudaMalloc on the steady-state path).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdlib.h>

namespace orfx {

struct DBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void free_() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};
struct HBuf {
    void* p = nullptr;
    size_t cap = 0;
    bool pinned = true;
    // pageable = true: ordinary memory (a caller that evaluates one batch per process: page-locking a result buffer
    // costs more than the one download gains)
    cudaError_t ensure(size_t bytes, bool pageable = false) {
        if (bytes <= cap && pinned == !pageable) return cudaSuccess;
        free_();
        size_t want = bytes + bytes / 8 + 256;
        if (pageable) { p = malloc(want); if (!p) return cudaErrorMemoryAllocation; pinned = false; cap = want; return cudaSuccess; }
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) { cap = want; pinned = true; }
        return e;
    }
    void free_() { if (p) { if (pinned) cudaFreeHost(p); else free(p); } p = nullptr; cap = 0; pinned = true; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// One protein pair of the alignment stage: residues aa[q_off + i * q_stride], i < qlen (query) and
// aa[t_off + j * t_stride], j < tlen (target).  qlen < 0: the slot is not aligned and its outputs are left alone.
struct AlnPair { long long q_off, t_off; int qlen, tlen; };

// Buffers of the alignment stage (align.cu), owned by the context.
struct AlnWork {
    DBuf pairs, order, cls, hist, scratch, cta, counters, meta, out, aa, off, pq, pt;
    HBuf h_meta;
    void free_() {
        DBuf* d[] = {&pairs, &order, &cls, &hist, &scratch, &cta, &counters, &meta, &out, &aa, &off, &pq, &pt};
        for (DBuf* b : d) b->free_();
        h_meta.free_();
    }
};

}  // namespace orfx
