// pic_main.cpp -- `pic_run`, the drop-in for the reference's out/main (src/main.cu:8-46):
// same positional arguments
//   <mode 30|31|32|33|test|bench> <verbose> <init n> <max t> <block size> <max n> <sleep time> <poisson_timestep>
// same stdout lines, same "./src/cross_section.txt" input relative to the working directory.
// Differences: arguments are parsed with strtol (no Boost), a missing argument prints usage
// and returns 64 instead of crashing, and PICB200_GPUS / PICB200_TABLE / PICB200_SEED are read
// from the environment so the eight positionals stay byte-compatible.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/picb200.h"
#include "pic_harness.h"

static bool parse_int(const char* s, int* out) {
    char* end = nullptr;
    long v = strtol(s, &end, 10);
    if (end == s || *end != '\0') return false;
    *out = (int)v;
    return true;
}

static int usage(const char* argv0) {
    fprintf(stderr,
            "usage: %s <mode> <verbose> <init n> <max t> <block size> <max n> <sleep time> <poisson_timestep>\n"
            "  mode: 30=Dynamic 31=CPU Sync 32=Naive 33=Dynamic Old, test, bench\n"
            "  env:  PICB200_GPUS=1|2|4|8  PICB200_TABLE=<path>  PICB200_SEED=<n>\n",
            argv0);
    return 64;
}

int main(int argc, char** argv) {
    if (argc < 2) return usage(argv[0]);
    if (!strcmp(argv[1], "bench")) return picb200_harness_bench(argc > 2 ? argv[2] : "out/data/mobility_timesteps_nodet.csv");
    if (argc < 9) return usage(argv[0]);
    int v[7];
    for (int k = 0; k < 7; ++k)
        if (!parse_int(argv[2 + k], &v[k])) { fprintf(stderr, "bad integer argument: %s\n", argv[2 + k]); return usage(argv[0]); }
    const int verbose = v[0], init_n = v[1], max_t = v[2], block_size = v[3];
    auto start = std::chrono::high_resolution_clock::now();              // src/main.cu:19
    const int max_n = v[4], sleep_time = v[5], poisson_timestep = v[6];

    picb200_config cfg;
    picb200_config_default(&cfg);
    cfg.verbose = verbose;
    cfg.init_n = init_n;
    cfg.poisson_steps = max_t;
    cfg.block_size = block_size;
    cfg.capacity = max_n;
    cfg.sleep_time_ns = sleep_time;
    cfg.poisson_timestep = poisson_timestep;
    cfg.print_stdout = 1;
    cfg.table_path = getenv("PICB200_TABLE");
    if (getenv("PICB200_SEED")) cfg.seed = strtoull(getenv("PICB200_SEED"), nullptr, 10);
    const int gpus = getenv("PICB200_GPUS") ? atoi(getenv("PICB200_GPUS")) : 1;

    int rc = 0;
    int mode = -1;
    if (!strcmp(argv[1], "30")) mode = PICB200_MODE_DYNAMIC;
    else if (!strcmp(argv[1], "31")) mode = PICB200_MODE_CPU_SYNC;
    else if (!strcmp(argv[1], "32")) mode = PICB200_MODE_NAIVE;
    else if (!strcmp(argv[1], "33")) mode = PICB200_MODE_DYNAMIC_OLD;

    if (mode >= 0) {
        cfg.mode = mode;
        rc = picb200_harness_run(&cfg, gpus);
    } else if (!strcmp(argv[1], "test")) {
        rc = picb200_harness_unit_test(&cfg);
    }   // any other mode string: run nothing, still print the CPU time (src/main.cu:26-46)
    if (rc != 0) fprintf(stdout, "%s", picb200_last_error());
    auto end = std::chrono::high_resolution_clock::now();
    printf("CPU time of program: %f ms\n", std::chrono::duration<double, std::milli>(end - start).count());
    return rc;
}
