// pic_harness.cpp -- see pic_harness.h.  Pure host C++ on top of the C ABI.
#include "pic_harness.h"

#include <algorithm>
#include <cstdio>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

static const char* kSchedulerName[] = {"Dynamic", "CPU Sync", "Naive", "Dynamic Old"};

int picb200_harness_run(const picb200_config* cfg_in, int gpus) {
    if (gpus <= 1) {
        picb200_summary sum;
        const int rc = picb200_run(cfg_in, &sum, nullptr, 0);
        // PICB200_PRINT_PUSHES=1: the push count, on stderr (stdout keeps the reference's lines only)
        if (!rc && getenv("PICB200_PRINT_PUSHES")) fprintf(stderr, "pushes %lld\n", (long long)sum.pushes);
        return rc;
    }
    // Particle slabs: rank r owns initial ids [r*n/k, (r+1)*n/k) and capacity max_n/k.
    picb200_config base = *cfg_in;
    printf("PIC with\ninit n: %d\ncapacity: %d\npoisson steps: %d\npoisson_timestep: %d\nblock size: %d\nsleep time: %d\n",
           base.init_n, base.capacity, base.poisson_steps, base.poisson_timestep, base.block_size, base.sleep_time_ns);
    if (picb200_device_count() < gpus) {
        printf("CUDA error: %d GPUs requested, %d visible\nAt: Setup\n", gpus, picb200_device_count());
        return 1;
    }
    // One host thread and one picb200 rank per GPU.  The ranks exchange peer handles (threads of one
    // process: a shared array and a barrier) so that the charge grid is summed inside the persistent
    // kernel over NVLink; PICB200_NCCL=1 selects the NCCL all-reduce path instead.
    const bool use_nccl = getenv("PICB200_NCCL") && atoi(getenv("PICB200_NCCL")) != 0;
    char uid[128];
    if (use_nccl && picb200_nccl_unique_id(uid)) return 1;
    std::vector<picb200_summary> sums((size_t)gpus);
    std::vector<int> rcs((size_t)gpus, 0);
    std::vector<std::string> errs((size_t)gpus);
    std::vector<unsigned char> handles((size_t)gpus * PICB200_PEER_HANDLE_BYTES, 0);
    std::mutex mu;
    std::condition_variable cv;
    int arrived = 0, generation = 0;
    auto barrier = [&]() {
        std::unique_lock<std::mutex> lk(mu);
        const int gen = generation;
        if (++arrived == gpus) { arrived = 0; ++generation; cv.notify_all(); }
        else cv.wait(lk, [&] { return generation != gen; });
    };
    std::vector<std::thread> th;
    for (int r = 0; r < gpus; ++r) {
        th.emplace_back([&, r]() {
            picb200_config c = base;
            c.rank = r;
            c.nranks = gpus;
            c.device = r;
            c.capacity = base.capacity / gpus;
            c.nccl_unique_id = use_nccl ? uid : nullptr;
            c.print_stdout = 0;
            int& rc = rcs[(size_t)r];
            if (use_nccl) {
                rc = picb200_run(&c, &sums[(size_t)r], nullptr, 0);
                if (rc) errs[(size_t)r] = picb200_last_error();
                return;
            }
            picb200_sim* sim = nullptr;
            rc = picb200_create(&c, &sim);
            if (!rc) rc = picb200_peer_export(sim, handles.data() + (size_t)r * PICB200_PEER_HANDLE_BYTES);
            if (rc) errs[(size_t)r] = picb200_last_error();
            barrier();                                  // every handle is in place (or a rank failed)
            bool all_ok = true;
            for (int q = 0; q < gpus; ++q) all_ok = all_ok && rcs[(size_t)q] == 0;
            if (all_ok) {
                rc = picb200_peer_connect(sim, handles.data());
                // particle slab of this rank: ids [r*n/k, (r+1)*n/k)
                const long long lo = (long long)base.init_n * r / gpus, hi = (long long)base.init_n * (r + 1) / gpus;
                if (!rc) rc = picb200_init_uniform(sim, (int32_t)(hi - lo), (uint64_t)lo);
                if (rc) errs[(size_t)r] = picb200_last_error();
            }
            barrier();                                  // nobody launches unless everybody connected
            all_ok = true;
            for (int q = 0; q < gpus; ++q) all_ok = all_ok && rcs[(size_t)q] == 0;
            if (all_ok) {
                rc = picb200_run_steps(sim, base.poisson_steps);
                if (!rc) rc = picb200_get_summary(sim, &sums[(size_t)r]);
                if (rc) errs[(size_t)r] = picb200_last_error();
            }
            barrier();                                  // no rank frees a grid a peer may still map
            if (sim) picb200_destroy(sim);
        });
    }
    for (auto& t : th) t.join();
    for (int r = 0; r < gpus; ++r)
        if (rcs[(size_t)r]) { printf("%s", errs[(size_t)r].c_str()); return 1; }
    picb200_summary tot = sums[0];
    for (int r = 1; r < gpus; ++r) {
        tot.final_n += sums[(size_t)r].final_n;
        tot.total_added += sums[(size_t)r].total_added;
        tot.total_removed += sums[(size_t)r].total_removed;
        tot.pushes += sums[(size_t)r].pushes;
        tot.gpu_ms = std::max(tot.gpu_ms, sums[(size_t)r].gpu_ms);
        tot.overflowed += sums[(size_t)r].overflowed;
    }
    printf("Multiprocessor count: %d\nBlocks per multiprocessor: %d\n", tot.sm_count, tot.blocks_per_sm);
    printf("%s\n", kSchedulerName[base.mode]);
    if (tot.overflowed) printf("\n\nOVERFLOW FROM ADDING PARTICLES\n\n\n");
    if (tot.final_n == 0) printf("Hit 0\n");
    printf("Final amount of particles: %lld\n", (long long)tot.final_n);
    printf("Particles added: %lld\n", (long long)tot.total_added);
    printf("Particles removed: %lld\n", (long long)tot.total_removed);
    printf("GPU time of program: %f ms\n", tot.gpu_ms);
    if (getenv("PICB200_PRINT_PUSHES")) fprintf(stderr, "pushes %lld\n", (long long)tot.pushes);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// `run test ...` and `run bench`: the two harness modes of src/test.cu, as clients of the C ABI.
// What is shared with the reference is its CONTRACT -- the printf strings a user (or a script) sees, the order of
// the comparison key (src/electron.h:28-41), the sweep grid and the CSV columns that analyse/plot_pic_*.py read
// (src/utility.cu:87-93) -- and those must match character for character; the code around them is this
// repository's own (the reference's own harness, unchanged, runs on this engine too: oracle/_ref_shim/main,
// csrc/runpic_shim.cpp).
// ---------------------------------------------------------------------------------------------
namespace {

// src/electron.h:28-30: (timestamp, y, x, z, vy, vx, vz)
using SortKey = std::tuple<int, double, double, double, double, double, double>;
SortKey sort_key(const picb200_electron& e) {
    return SortKey(e.timestamp, e.position[1], e.position[0], e.position[2], e.velocity[1], e.velocity[0], e.velocity[2]);
}
bool key_less(const picb200_electron& a, const picb200_electron& b) { return sort_key(a) < sort_key(b); }

void print_electron(int i, const picb200_electron& e) {       // src/electron.h:19-26
    printf("%d: (%.15f, %.15f, %.15f) (%.15f, %.15f, %.15f) ((%.7f, %.7f, %.7f)) [%d]\n", i, e.position[0], e.position[1],
           e.position[2], e.velocity[0], e.velocity[1], e.velocity[2], e.acceleration[0], e.acceleration[1], e.acceleration[2], e.timestamp);
}

// One scheduler's run with its final population as 64-byte records (src/electron.h:12-18).
struct SortedRun {
    int final_n = 0;
    std::vector<picb200_electron> electrons;
};
int run_one(picb200_config cfg, int scheduler, SortedRun& out) {
    cfg.mode = scheduler;
    picb200_summary sum;
    picb200_sim* sim = nullptr;
    if (picb200_run_keep(&cfg, &sum, &sim)) return 1;
    out.final_n = (int)sum.final_n;
    out.electrons.resize((size_t)std::max(out.final_n, 1));
    int64_t got = 0;
    const int rc = picb200_download_aos(sim, out.electrons.data(), out.final_n, &got);
    picb200_destroy(sim);
    if (rc) return 1;
    out.electrons.resize((size_t)out.final_n);
    return 0;
}

}  // namespace

int picb200_harness_unit_test(const picb200_config* cfg_in) {
    // src/test.cu:43-100 compares schedulers 0, 1, 3 with scheduler 1 (CPU Sync) and leaves Naive out, because the
    // reference's per-slot RNG makes Naive diverge.  Here the RNG is keyed on the particle id, all four schedulers
    // must agree, and Naive (2) is compared as well (reported last, so the first three lines read as the reference's).
    constexpr int kBaseline = 1;
    const int candidates[] = {0, 1, 3, 2};
    struct Verdict { int scheduler; bool ok; int final_n; };
    std::vector<Verdict> verdicts;

    picb200_config cfg = *cfg_in;
    cfg.print_stdout = 1;
    SortedRun base;
    if (run_one(cfg, kBaseline, base)) return 1;
    printf("Sorting base...\n");
    std::sort(base.electrons.begin(), base.electrons.end(), key_less);
    printf("Done sorting base\n");

    for (int scheduler : candidates) {
        SortedRun cur;
        if (run_one(cfg, scheduler, cur)) return 1;
        Verdict v{scheduler, true, cur.final_n};
        if (cur.final_n != base.final_n) {
            printf("\n\nFinal n does not match in %d. Base: %d, test: %d\n\n\n", scheduler, base.final_n, cur.final_n);
            v.ok = false;
        } else {
            printf("Sorting %d...\n", scheduler);
            std::sort(cur.electrons.begin(), cur.electrons.end(), key_less);
            printf("Done sorting\n");
            for (int i = 0; i < cur.final_n; ++i) {
                if (sort_key(base.electrons[(size_t)i]) != sort_key(cur.electrons[(size_t)i])) {
                    printf("Mismatch in %d!\n", scheduler);
                    print_electron(i, base.electrons[(size_t)i]);
                    print_electron(i, cur.electrons[(size_t)i]);
                    v.ok = false;
                    break;
                }
                if (cfg.verbose != 0 && i % cfg.verbose == 0) print_electron(i, base.electrons[(size_t)i]);
            }
        }
        verdicts.push_back(v);
    }
    printf("\nTests done with following results as compared to %d (%d):\n", kBaseline, base.final_n);
    int failures = 0;
    for (const Verdict& v : verdicts) {
        printf("%d: %s (%d)\n", v.scheduler, v.ok ? "success" : "failure", v.final_n);
        failures += v.ok ? 0 : 1;
    }
    return failures ? 3 : 0;
}

int picb200_harness_bench(const char* csv_path) {
    // The sweep of src/test.cu:4-41: init n 1e6, max n 5e7, 10 Poisson steps, sleep 100 ns; mobility steps 10..100 by
    // 10 then 200..1000 by 100; block sizes 128..1024; all four schedulers, Naive only up to block 256.
    // PICB200_BENCH_QUICK=1 (tests): mobility steps {10, 100} x block sizes {256, 1024} -- the cells
    // analyse/plot_pic_mobility.py:8-12 selects.  PICB200_BENCH_INIT_N overrides the population.
    const bool quick = getenv("PICB200_BENCH_QUICK") && atoi(getenv("PICB200_BENCH_QUICK")) != 0;
    const int init_n = getenv("PICB200_BENCH_INIT_N") ? atoi(getenv("PICB200_BENCH_INIT_N")) : 1000000;
    std::vector<int> mobility_grid, block_grid;
    if (quick) {
        mobility_grid = {10, 100};
        block_grid = {256, 1024};
    } else {
        for (int m = 10; m <= 100; m += 10) mobility_grid.push_back(m);
        for (int m = 200; m <= 1000; m += 100) mobility_grid.push_back(m);
        block_grid = {128, 256, 512, 1024};
    }
    struct CsvRow { const char* func; int init_n, iterations, mobility_steps, block_size, sleep_time, final_n; double time_ms; };
    std::vector<CsvRow> table;
    for (int mobility_steps : mobility_grid)
        for (int block_size : block_grid)
            for (int scheduler = 0; scheduler < 4; ++scheduler) {
                if (scheduler == PICB200_MODE_NAIVE && block_size > 256) continue;      // src/test.cu:22
                picb200_config cfg;
                picb200_config_default(&cfg);
                cfg.mode = scheduler;
                cfg.init_n = init_n;
                cfg.capacity = 50000000;
                cfg.poisson_steps = 10;
                cfg.poisson_timestep = mobility_steps;
                cfg.block_size = block_size;
                cfg.sleep_time_ns = 100;
                cfg.print_stdout = 1;
                cfg.table_path = getenv("PICB200_TABLE");
                picb200_summary sum;
                if (picb200_run(&cfg, &sum, nullptr, 0)) { printf("%s", picb200_last_error()); return 1; }
                if (sum.final_n >= cfg.capacity) { printf("\n\n\nIllegal!!!\n\n\n"); continue; }      // src/test.cu:25-29
                table.push_back(CsvRow{kSchedulerName[scheduler], cfg.init_n, cfg.poisson_steps, mobility_steps, block_size, cfg.sleep_time_ns, (int)sum.final_n, sum.gpu_ms});
            }
    FILE* os = fopen(csv_path, "w");
    if (!os) { fprintf(stderr, "cannot write %s\n", csv_path); return 2; }
    // header and row format of src/utility.cu:87-93; `split chance` is never assigned by the reference
    // (src/utility.h:23: uninitialised) -- 0 here
    fprintf(os, "func,init n,iterations,mobility steps,block size,sleep time,split chance,final n,time\n");
    for (const CsvRow& r : table)
        fprintf(os, "%s,%d,%d,%d,%d,%d,%f,%d,%f\n", r.func, r.init_n, r.iterations, r.mobility_steps, r.block_size, r.sleep_time, 0.0, r.final_n, r.time_ms);
    fclose(os);
    return 0;
}
