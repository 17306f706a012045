// TEST INFRASTRUCTURE ONLY.  Drives the product's host front-end (loading, bundling, emission)
// with the CPU oracle behind the five backend entry points, so that the oracle's results can be
// compared byte for byte with the reference binary (oracle/_ref/orfanage_ref) on the same inputs,
// and so that the host front-end itself is checked on machines without a GPU.
// The shipped `orfanage` binary (orfanage_b200/csrc/host/main.cpp) links only the CUDA library.
#include "../orfanage_b200/csrc/host/orfx_host.h"
#include <cstdio>

extern "C" {
const char* orfx_oracle_last_error(void);
int orfx_oracle_create(orfx_ctx**, int, const orfx_params*);
int orfx_oracle_submit(orfx_ctx*, const orfx_batch*);
int orfx_oracle_wait(orfx_ctx*, orfx_results*);
int orfx_oracle_release(orfx_ctx*);
void orfx_oracle_destroy(orfx_ctx*);
int orfx_oracle_clean_templates(orfx_ctx*, const orfx_batch*, int, orfx_clean_out*, int64_t*);
int orfx_oracle_score_pairs(orfx_ctx*, const orfx_batch*, orfx_results*);
}

int main(int argc, char** argv) {
    try {
        orfx::Options opt;
        orfx::Backend be{orfx_oracle_create, orfx_oracle_submit, orfx_oracle_wait, orfx_oracle_release, orfx_oracle_destroy, orfx_oracle_last_error, nullptr, nullptr,
                         orfx_oracle_clean_templates, orfx_oracle_score_pairs};
        if (argc > 1 && std::string(argv[1]) == "orfcompare") {   // oracle-backed orfcompare front-end: `orfanage_oracle_cli orfcompare --query ...`
            if (!orfx::parse_compare_cli(argc - 1, argv + 1, opt)) return 0;
            opt.devices.assign(1, 0);
            return orfx::run_orfcompare(opt, be);
        }
        if (!orfx::parse_cli(argc, argv, opt)) return 0;
        opt.devices.assign(1, 0);
        return orfx::run_orfanage(opt, be);
    } catch (const std::exception& e) {
        fprintf(stderr, "Error: %s\n", e.what());
        return 1;
    }
}
