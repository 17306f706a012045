// Differential check of orfanage_b200/csrc/stdsort.cuh (host instantiation) against the real
// libstdc++ std::sort on tie-heavy inputs: the final permutation must be identical for every n,
// because the reference's choice among equal candidate rows is whatever std::sort leaves in front
// (orfanage.cpp:276-280).  Prints "ok <cases>" or the first mismatch.
#include "../orfanage_b200/csrc/stdsort.cuh"
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>

static uint64_t rng_state = 88172645463325252ULL;
static uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }

struct ByKey {
    const std::vector<int>* key;
    bool operator()(int a, int b) const { return (*key)[a] > (*key)[b]; }   // "greater", like score_gt
};

int main() {
    long cases = 0;
    for (int n = 0; n <= 700; n += (n < 70 ? 1 : 37)) {
        for (int distinct : {1, 2, 3, 5, 17, 1000}) {
            for (int rep = 0; rep < 12; rep++) {
                std::vector<int> key(n);
                int pattern = rep % 4;
                for (int i = 0; i < n; i++) {
                    if (pattern == 0) key[i] = (int)(rnd() % distinct);
                    else if (pattern == 1) key[i] = i % distinct;                 // sorted-ish runs
                    else if (pattern == 2) key[i] = (n - i) % distinct;
                    else key[i] = (int)((rnd() % 4 == 0) ? rnd() % distinct : 0);   // mostly equal
                }
                std::vector<int> a(n), b(n);
                for (int i = 0; i < n; i++) a[i] = b[i] = i;
                ByKey cmp{&key};
                std::sort(a.begin(), a.end(), cmp);
                orfx::stdsort::sort(b.data(), b.data() + n, cmp);
                if (a != b) { printf("MISMATCH n=%d distinct=%d rep=%d\n", n, distinct, rep); return 1; }
                cases++;
            }
        }
    }
    // the heapsort fallback: organ-pipe / median-of-3 killer style input makes introsort hit depth 0
    for (int n : {200, 1000, 5000}) {
        std::vector<int> key(n);
        for (int i = 0; i < n; i++) key[i] = i < n / 2 ? i : n - i;
        std::vector<int> a(n), b(n);
        for (int i = 0; i < n; i++) a[i] = b[i] = i;
        ByKey cmp{&key};
        std::sort(a.begin(), a.end(), cmp);
        orfx::stdsort::sort(b.data(), b.data() + n, cmp);
        if (a != b) { printf("MISMATCH organ-pipe n=%d\n", n); return 1; }
        cases++;
    }
    printf("ok %ld\n", cases);
    return 0;
}
