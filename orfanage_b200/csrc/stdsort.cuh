// Replays libstdc++'s std::sort (GCC 13: introsort with median-of-3 pivot, depth limit 2*floor(log2 n),
// heapsort fallback, threshold 16, final insertion sort) on an array of small items.
//
// Why: the reference orders the candidate rows of one ORF key with
//     std::sort(v.begin(), v.end(), score_gt)            (orfanage.cpp:276-280)
// and the front of each sorted vector is the ORF's representative template.  score_gt has ties, and
// std::sort is not stable beyond 16 elements, so "which of the equal rows ends up first" is defined
// only by the exact sequence of comparisons and moves the library performs.  This header performs
// that same sequence (the algorithm is a pure function of n and of the comparator's answers), on the
// device and on the host, so the chosen template and the stats order match the reference for any n.
// tests/test_stdsort.py checks it against the real std::sort on adversarial tie-heavy inputs.
#pragma once
#ifdef __CUDACC__
#define ORFX_HD __host__ __device__
#else
#define ORFX_HD
#endif

namespace orfx {
namespace stdsort {

template <class T, class Cmp>
ORFX_HD inline void unguarded_linear_insert(T* last, Cmp& comp) {
    T val = *last;
    T* next = last - 1;
    while (comp(val, *next)) { *last = *next; last = next; --next; }
    *last = val;
}
template <class T, class Cmp>
ORFX_HD inline void insertion_sort(T* first, T* last, Cmp& comp) {
    if (first == last) return;
    for (T* i = first + 1; i != last; ++i) {
        if (comp(*i, *first)) {
            T val = *i;
            for (T* p = i; p != first; --p) *p = *(p - 1);
            *first = val;
        } else unguarded_linear_insert(i, comp);
    }
}
template <class T, class Cmp>
ORFX_HD inline void push_heap_(T* first, long hole, long top, T value, Cmp& comp) {
    long parent = (hole - 1) / 2;
    while (hole > top && comp(first[parent], value)) { first[hole] = first[parent]; hole = parent; parent = (hole - 1) / 2; }
    first[hole] = value;
}
template <class T, class Cmp>
ORFX_HD inline void adjust_heap(T* first, long hole, long len, T value, Cmp& comp) {
    const long top = hole;
    long child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (comp(first[child], first[child - 1])) child--;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    push_heap_(first, hole, top, value, comp);
}
template <class T, class Cmp>
ORFX_HD inline void heap_sort(T* first, T* last, Cmp& comp) {   // __partial_sort(first,last,last)
    const long len = last - first;
    if (len >= 2) {
        long parent = (len - 2) / 2;
        while (true) {
            T v = first[parent];
            adjust_heap(first, parent, len, v, comp);
            if (parent == 0) break;
            parent--;
        }
    }
    while (last - first > 1) {
        --last;
        T v = *last;
        *last = *first;
        adjust_heap(first, 0L, (long)(last - first), v, comp);
    }
}
template <class T, class Cmp>
ORFX_HD inline void move_median_to_first(T* result, T* a, T* b, T* c, Cmp& comp) {
    T* pick;
    if (comp(*a, *b)) {
        if (comp(*b, *c)) pick = b; else if (comp(*a, *c)) pick = c; else pick = a;
    } else if (comp(*a, *c)) pick = a; else if (comp(*b, *c)) pick = c; else pick = b;
    T t = *result; *result = *pick; *pick = t;
}
template <class T, class Cmp>
ORFX_HD inline T* unguarded_partition(T* first, T* last, T* pivot, Cmp& comp) {
    while (true) {
        while (comp(*first, *pivot)) ++first;
        --last;
        while (comp(*pivot, *last)) --last;
        if (!(first < last)) return first;
        T t = *first; *first = *last; *last = t;
        ++first;
    }
}
// depth_override < 0: the library's own limit; >= 0 is a test hook to reach the heapsort branch
template <class T, class Cmp>
ORFX_HD inline void sort(T* first, T* last, Cmp comp, int depth_override = -1) {
    const long n = last - first;
    if (n <= 0) return;
    if (n <= 16) { insertion_sort(first, last, comp); return; }   // the introsort loop does nothing below the threshold
    int lg = 0;
    for (long m = n; m > 1; m >>= 1) lg++;
    int depth0 = depth_override >= 0 ? depth_override : 2 * lg;
    // __introsort_loop with its recursion (right part) turned into an explicit stack
    struct Frame { T* f; T* l; int d; };
    Frame stack[72];
    int sp = 0;
    stack[sp++] = Frame{first, last, depth0};
    while (sp > 0) {
        Frame fr = stack[--sp];
        T* f = fr.f; T* l = fr.l; int d = fr.d;
        while (l - f > 16) {
            if (d == 0) { heap_sort(f, l, comp); break; }
            --d;
            T* mid = f + (l - f) / 2;
            move_median_to_first(f, f + 1, mid, l - 1, comp);
            T* cut = unguarded_partition(f + 1, l, f, comp);
            // the library recurses on [cut,l) first, then continues with [f,cut); both halves are
            // independent, so deferring the right half on a stack performs the same moves per half
            if (sp < 72) stack[sp++] = Frame{cut, l, d};
            l = cut;
        }
    }
    if (n > 16) {
        insertion_sort(first, first + 16, comp);
        for (T* i = first + 16; i != last; ++i) unguarded_linear_insert(i, comp);
    } else insertion_sort(first, last, comp);
}

}  // namespace stdsort
}  // namespace orfx
