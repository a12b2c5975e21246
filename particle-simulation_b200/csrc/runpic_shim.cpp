// runpic_shim.cpp -- the reference's operator `runPIC` (src/pic.h:9, defined at src/pic.cu:359) on top of
// libpicb200.so, so that the reference's own harnesses -- main (src/main.cu:8-46), runBenchmark
// (src/test.cu:4-41) and runUnitTest (src/test.cu:43-100) -- link and run UNCHANGED against the new engine.
//
// This file is the only translation unit a maintainer of the reference adds (it replaces src/pic.cu and the
// four kernel files below it); it includes the reference's headers and is therefore compiled only where the
// reference checkout is present:  `make -C oracle shim`  ->  oracle/_ref_shim/main  =
//   /root/reference/src/{main,test,utility,cell}.cu  +  this file  +  -lpicb200.
// tests/test_ref_shim.py runs that binary on the GPU box (the reference's `run test ...` sort-and-compare and
// its `run 30 ...` CLI, stdout compared with pic_run's).
#include "pic.h"            // RunData, TimingData (src/utility.h:15-31), Electron (src/electron.h:12-18)

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

#include "picb200.h"

static_assert(sizeof(Electron) == sizeof(picb200_electron), "same 64-byte record");
static_assert(offsetof(Electron, velocity) == offsetof(picb200_electron, velocity) &&
              offsetof(Electron, acceleration) == offsetof(picb200_electron, acceleration) &&
              offsetof(Electron, timestamp) == offsetof(picb200_electron, timestamp), "same field offsets");

RunData runPIC(int init_n, int capacity, int poisson_steps, int poisson_timestep, int mode, int verbose, int block_size,
               int sleep_time_ns) {
    picb200_config cfg;
    picb200_config_default(&cfg);
    cfg.mode = mode;                       // 0..3 exactly as src/main.cu:26-40 maps "30".."33"
    cfg.verbose = verbose;
    cfg.init_n = init_n;
    cfg.capacity = capacity;
    cfg.poisson_steps = poisson_steps;
    cfg.poisson_timestep = poisson_timestep;
    cfg.block_size = block_size;
    cfg.sleep_time_ns = sleep_time_ns;
    cfg.print_stdout = 1;                  // header, scheduler name, state dumps, summary lines of src/pic.cu:360-576
    cfg.table_path = getenv("PICB200_TABLE");      // NULL: "./src/cross_section.txt" relative to CWD (src/pic.cu:381)
    // src/pic.cu:583 allocates with new[] and every caller releases with free() (src/test.cu:31,69,87): malloc here.
    // The population is returned in two steps so that only final_n records are ever allocated.
    picb200_summary s;
    picb200_sim* sim = nullptr;
    Electron* electrons = nullptr;
    int rc = picb200_run_keep(&cfg, &s, &sim);
    if (!rc) {
        electrons = static_cast<Electron*>(malloc(sizeof(Electron) * (size_t)(s.final_n > 0 ? s.final_n : 1)));
        int64_t got = 0;
        rc = picb200_download_aos(sim, reinterpret_cast<picb200_electron*>(electrons), s.final_n, &got);
    }
    if (sim) picb200_destroy(sim);
    if (rc) {
        printf("%s", picb200_last_error());               // "CUDA error: ...\nAt: ...\n"   (src/utility.cu:111-112)
        free(electrons);
        throw std::runtime_error(picb200_last_error());   // src/utility.cu:113
    }
    RunData run_data;
    run_data.timing_data.init_n = init_n;                 // src/pic.cu:363-368
    run_data.timing_data.iterations = poisson_steps;
    run_data.timing_data.block_size = block_size;
    run_data.timing_data.sleep_time = sleep_time_ns;
    run_data.timing_data.mobility_steps = poisson_timestep;
    run_data.timing_data.split_chance = 0.f;              // never assigned by the reference (src/utility.h:23): garbage there, 0 here
    run_data.timing_data.final_n = (int)s.final_n;
    run_data.timing_data.function = (mode == 0 ? "Dynamic" : mode == 1 ? "CPU Sync" : mode == 2 ? "Naive" : "Dynamic Old");   // src/pic.cu:457-474
    run_data.timing_data.time = s.gpu_ms;                 // src/pic.cu:572
    run_data.final_n = (int)s.final_n;                    // src/pic.cu:582
    run_data.electrons = electrons;                       // src/pic.cu:583-584
    return run_data;
}
