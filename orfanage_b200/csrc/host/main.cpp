// `orfanage` — the reference's CLI surface (orfanage.cpp:831-1125) over the CUDA hot path.
#include "orfx_host.h"
#include <cstdio>
#include <unistd.h>

int main(int argc, char** argv) {
    try {
        orfx::Options opt;
        if (!orfx::parse_cli(argc, argv, opt)) return 0;
        orfx::Backend be{orfx_create, orfx_submit, orfx_wait, orfx_release, orfx_destroy, orfx_last_error, orfx_host_alloc, orfx_host_free, orfx_clean_templates, orfx_score_pairs};
        const int rc = orfx::run_orfanage(opt, be);
        fflush(nullptr);
        _exit(rc);   // outputs are closed and the device contexts destroyed: skip the static destructors and the CUDA runtime's teardown
    } catch (const std::exception& e) {
        fprintf(stderr, "Error: %s\n", e.what());
        return 1;
    }
}
