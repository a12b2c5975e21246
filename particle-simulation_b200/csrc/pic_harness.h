// pic_harness.h -- the callers around runPIC that the reference keeps in src/main.cu and
// src/test.cu, rebuilt as clients of the C ABI (include/picb200.h).
#pragma once
#include "../../include/picb200.h"

// `run 30|31|32|33 ...`: one simulation (src/main.cu:26-40).  gpus > 1 shards the electrons
// over that many devices of this box, one host thread and one picb200 rank per GPU.
int picb200_harness_run(const picb200_config* cfg, int gpus);
// `run test ...`: cross-scheduler equivalence test (src/test.cu:43-100).
int picb200_harness_unit_test(const picb200_config* cfg);
// `run bench`: the timing sweep and its CSV (src/test.cu:4-41, src/utility.cu:87-106).
int picb200_harness_bench(const char* csv_path);
