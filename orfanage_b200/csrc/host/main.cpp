// `orfanage` — the reference's CLI surface (orfanage.cpp:831-1125) over the CUDA hot path.
#include "orfx_host.h"
#include <cstdio>

int main(int argc, char** argv) {
    try {
        orfx::Options opt;
        if (!orfx::parse_cli(argc, argv, opt)) return 0;
        orfx::Backend be{orfx_create, orfx_submit, orfx_wait, orfx_release, orfx_destroy, orfx_last_error, orfx_host_alloc, orfx_host_free, orfx_clean_templates, orfx_score_pairs};
        return orfx::run_orfanage(opt, be);
    } catch (const std::exception& e) {
        fprintf(stderr, "Error: %s\n", e.what());
        return 1;
    }
}
