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

// Sort key and comparison of src/electron.h:28-41.
static std::tuple<int, double, double, double, double, double, double> key_of(const picb200_electron& e) {
    return std::make_tuple(e.timestamp, e.position[1], e.position[0], e.position[2], e.velocity[1], e.velocity[0], e.velocity[2]);
}
static void print_electron(int i, const picb200_electron& e) {
    printf("%d: (%.15f, %.15f, %.15f) (%.15f, %.15f, %.15f) ((%.7f, %.7f, %.7f)) [%d]\n", i, e.position[0], e.position[1],
           e.position[2], e.velocity[0], e.velocity[1], e.velocity[2], e.acceleration[0], e.acceleration[1], e.acceleration[2], e.timestamp);
}

int picb200_harness_unit_test(const picb200_config* cfg_in) {
    // src/test.cu:43-100: base scheduler 1 (CPU Sync), then every scheduler against it.  The
    // reference leaves Naive out because its per-slot RNG makes Naive diverge; with the RNG
    // keyed on the particle id all four schedulers must agree, so Naive (2) is tested as well.
    const int base_function = 1;
    const int test_functions[] = {0, 1, 3, 2};
    const int amnt = (int)(sizeof(test_functions) / sizeof(int));
    std::vector<bool> broken((size_t)amnt, false);
    std::vector<int> final_ns((size_t)amnt, 0);

    picb200_config cfg = *cfg_in;
    cfg.print_stdout = 1;
    std::vector<picb200_electron> base((size_t)cfg.capacity), cur((size_t)cfg.capacity);
    picb200_summary sum;
    cfg.mode = base_function;
    if (picb200_run(&cfg, &sum, base.data(), cfg.capacity)) return 1;
    const int base_final_n = (int)sum.final_n;
    auto less = [](const picb200_electron& a, const picb200_electron& b) { return key_of(a) < key_of(b); };
    printf("Sorting base...\n");
    std::sort(base.begin(), base.begin() + base_final_n, less);
    printf("Done sorting base\n");

    for (int fi = 0; fi < amnt; ++fi) {
        const int function = test_functions[fi];
        cfg.mode = function;
        if (picb200_run(&cfg, &sum, cur.data(), cfg.capacity)) return 1;
        const int final_n = (int)sum.final_n;
        final_ns[(size_t)fi] = final_n;
        if (final_n != base_final_n) {
            printf("\n\nFinal n does not match in %d. Base: %d, test: %d\n\n\n", function, base_final_n, final_n);
            broken[(size_t)fi] = true;
            continue;
        }
        printf("Sorting %d...\n", function);
        std::sort(cur.begin(), cur.begin() + final_n, less);
        printf("Done sorting\n");
        for (int i = 0; i < final_n; ++i) {
            if (key_of(base[(size_t)i]) != key_of(cur[(size_t)i])) {
                printf("Mismatch in %d!\n", function);
                print_electron(i, base[(size_t)i]);
                print_electron(i, cur[(size_t)i]);
                broken[(size_t)fi] = true;
                break;
            } else if (cfg.verbose != 0 && i % cfg.verbose == 0) {
                print_electron(i, base[(size_t)i]);
            }
        }
    }
    printf("\nTests done with following results as compared to %d (%d):\n", base_function, base_final_n);
    int bad = 0;
    for (int fi = 0; fi < amnt; ++fi) {
        printf("%d: %s (%d)\n", test_functions[fi], broken[(size_t)fi] ? "failure" : "success", final_ns[(size_t)fi]);
        bad += broken[(size_t)fi] ? 1 : 0;
    }
    return bad ? 3 : 0;
}

int picb200_harness_bench(const char* csv_path) {
    // Sweep of src/test.cu:4-41: init n 1e6, max n 5e7, 10 Poisson steps, mobility steps
    // 10..100 by 10 then 200..1000 by 100, block 128..1024, all four schedulers (Naive only
    // up to block 256).  CSV columns of src/utility.cu:87-93, read by analyse/plot_pic_*.py.
    struct Row { std::string func; int init_n, iterations, mobility_steps, block_size, sleep_time, final_n; double time; };
    std::vector<Row> rows;
    const int block_sizes[] = {128, 256, 512, 1024};
    int incr = 10;
    for (int mobility_steps = incr; mobility_steps <= 1000; mobility_steps += incr) {
        if (mobility_steps == incr * 10) incr *= 10;
        for (int block_size : block_sizes) {
            for (int function = 0; function < 4; ++function) {
                if (function == 2 && block_size > 256) continue;
                picb200_config cfg;
                picb200_config_default(&cfg);
                cfg.mode = function;
                cfg.init_n = 1000000;
                cfg.capacity = 50000000;
                cfg.poisson_steps = 10;
                cfg.poisson_timestep = mobility_steps;
                cfg.block_size = block_size;
                cfg.sleep_time_ns = 100;
                cfg.print_stdout = 1;
                picb200_summary sum;
                if (picb200_run(&cfg, &sum, nullptr, 0)) { printf("%s", picb200_last_error()); return 1; }
                if (sum.final_n >= cfg.capacity) { printf("\n\n\nIllegal!!!\n\n\n"); continue; }
                rows.push_back(Row{kSchedulerName[function], cfg.init_n, cfg.poisson_steps, mobility_steps, block_size, 100, (int)sum.final_n, sum.gpu_ms});
            }
        }
    }
    FILE* os = fopen(csv_path, "w");
    if (!os) { fprintf(stderr, "cannot write %s\n", csv_path); return 2; }
    fprintf(os, "func,init n,iterations,mobility steps,block size,sleep time,split chance,final n,time\n");
    for (const Row& d : rows)
        fprintf(os, "%s,%d,%d,%d,%d,%d,%f,%d,%f\n", d.func.c_str(), d.init_n, d.iterations, d.mobility_steps, d.block_size, d.sleep_time, 0.0, d.final_n, d.time);
    fclose(os);
    return 0;
}
