// `orfcompare` — the reference's companion CLI (orfcompare.cpp:173-279) over the CUDA pair-scoring kernel.
#include "orfx_host.h"
#include <cstdio>

int main(int argc, char** argv) {
    try {
        orfx::Options opt;
        if (!orfx::parse_compare_cli(argc, argv, opt)) return 0;
        orfx::Backend be{orfx_create, orfx_submit, orfx_wait, orfx_release, orfx_destroy, orfx_last_error, orfx_host_alloc, orfx_host_free, orfx_clean_templates,
                         orfx_score_pairs};
        return orfx::run_orfcompare(opt, be);
    } catch (const std::exception& e) {
        fprintf(stderr, "Error: %s\n", e.what());
        return 1;
    }
}
