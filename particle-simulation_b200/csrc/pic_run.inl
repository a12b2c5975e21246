// pic_run.inl -- picb200_run: the whole of runPIC (src/pic.cu:359-599) behind one C call,
// including the reference's stdout contract (header :360, device lines :404, scheduler name
// :457-474, state dumps src/utility.cu:124-137 + src/electron.h:19-26, overflow :544,
// "Hit 0" :557, summary :573-576).  Included at the end of pic_engine.cu.

namespace picb200 {

// log() of src/utility.cu:124-137: "Time t, amount n" + one line per electron + the PNG snapshot with its
// "Saved image" line (written when ./out/visualization exists: the reference crashes without it) + blank line.
// PICB200_DUMP_DIR=<dir> adds a binary snapshot <dir>/state_%04d.bin per logged step (picb200_save_state).
static int print_state(picb200_sim* sim, int t) {
    Sim& s = sim->s;
    if (s.sync_state()) return 1;
    const int amount = std::max(s.h_st->n, s.h_st->n_alloc);
    const size_t n = (size_t)std::min(amount, s.cfg.capacity);
    std::vector<float> f(n * 9 + 1);
    std::vector<int32_t> ts(n + 1);
    int got = 0;
    float* b = f.data();
    if (picb200_download(sim, &got, b, b + n, b + 2 * n, b + 3 * n, b + 4 * n, b + 5 * n, b + 6 * n, b + 7 * n, b + 8 * n, ts.data(), nullptr)) return 1;
    printf("Time %d, amount %d\n", t, amount);
    for (size_t i = 0; i < n; ++i) {
        printf("%d: (%.15f, %.15f, %.15f) (%.15f, %.15f, %.15f) ((%.7f, %.7f, %.7f)) [%d]\n", (int)i,
               (double)b[i], (double)b[n + i], (double)b[2 * n + i], (double)b[3 * n + i], (double)b[4 * n + i],
               (double)b[5 * n + i], b[6 * n + i], b[7 * n + i], b[8 * n + i], ts[i]);
    }
    char png[64];
    snprintf(png, sizeof png, "./out/visualization/test_%04d.png", t);       // src/utility.cu:69
    if (FILE* probe = fopen(png, "wb")) {
        fclose(probe);
        if (picb200_snapshot_png(b, b + n, (int64_t)n, png)) return 1;
        printf("Saved image to %s\n", png);                                  // src/utility.cu:34
    }
    printf("\n");
    if (const char* dir = getenv("PICB200_DUMP_DIR")) {
        char path[4096];
        snprintf(path, sizeof path, "%s/state_%04d.bin", dir, t);
        if (picb200_save_state(sim, path, 1)) return 1;
    }
    return 0;
}

static void naive_step_logger(Sim* s, void* user) {
    picb200_sim* sim = static_cast<picb200_sim*>(user);
    // src/pic.cu:528: the naive loop logs after every launch with the *Poisson* index
    const int t = s->h_st->poisson_index;
    if (s->cfg.verbose != 0 && t % s->cfg.verbose == 0) print_state(sim, t);
}

}  // namespace picb200

// The whole of runPIC; on success *keep (if given) receives the still-live simulation, so that a caller that does not
// know final_n in advance can size its buffer first and fetch the population with picb200_download_aos afterwards.
static int run_impl(const picb200_config* cfg_in, picb200_summary* out, picb200_electron* electrons_out, int64_t electrons_cap,
                    picb200_sim** keep) {
    if (keep) *keep = nullptr;
    if (!cfg_in) return fail("null argument", "picb200_run");
    picb200_config cfg = *cfg_in;
    const bool say = cfg.print_stdout != 0;
    if (say)
        printf("PIC with\ninit n: %d\ncapacity: %d\npoisson steps: %d\npoisson_timestep: %d\nblock size: %d\nsleep time: %d\n",
               cfg.init_n, cfg.capacity, cfg.poisson_steps, cfg.poisson_timestep, cfg.block_size, cfg.sleep_time_ns);
    picb200_sim* sim = nullptr;
    if (picb200_create(&cfg, &sim)) return 1;
    Sim& s = sim->s;
    int rc = 1;
    do {
        if (say) {
            printf("Multiprocessor count: %d\nBlocks per multiprocessor: %d\n", s.sm_count, s.blocks_per_sm);
            static const char* names[] = {"Dynamic", "CPU Sync", "Naive", "Dynamic Old"};
            printf("%s\n", names[cfg.mode]);
        }
        // particle slab of this rank: ids [rank*init_n/nranks, (rank+1)*init_n/nranks)
        const long long lo = (long long)cfg.init_n * cfg.rank / cfg.nranks;
        const long long hi = (long long)cfg.init_n * (cfg.rank + 1) / cfg.nranks;
        // PICB200_RESTART=<file>: continue the run a snapshot was taken from (single GPU) instead of seeding a new one
        const char* restart = cfg.nranks == 1 ? getenv("PICB200_RESTART") : nullptr;
        if (restart ? picb200_load_state(sim, restart) : s.init_uniform((int)(hi - lo), (unsigned long long)lo)) break;
        if (cudaEventRecord(s.ev_start, s.stream) != cudaSuccess) { fail("event record", "Setup"); break; }

        const int t_begin = restart ? s.h_st->poisson_index : 0;
        const int P = std::max(0, cfg.poisson_steps - t_begin);     // Poisson steps still to do
        const bool verbose = say && cfg.verbose != 0;
        const bool host_in_loop = verbose || (cfg.mode == PICB200_MODE_CPU_SYNC && !s.device_windows()) || cfg.nranks > 1;     // (mode 32 keeps its step windows on the device)
        bool failed = false;
        if (verbose && cfg.mode == PICB200_MODE_NAIVE) { s.naive_hook = naive_step_logger; s.naive_hook_user = sim; }
        if (!host_in_loop) {
            // modes 30/33: the whole run is enqueued; the host reads nothing until the end
            if (s.fused) failed = s.fused_steps(P) != 0;
            else for (int t = 0; t < P && !failed; ++t) failed = s.poisson_step(false) != 0;
            if (failed) break;
            const size_t n_log = (size_t)std::max(1, std::min(cfg.poisson_steps, s.log_len));
            std::vector<StepLog> lg(n_log);
            // (a kernel that died -- e.g. a grid barrier that timed out -- makes either call fail: sync_state has the message)
            if (cudaMemcpyAsync(lg.data(), s.log, sizeof(StepLog) * n_log, cudaMemcpyDeviceToHost, s.stream) != cudaSuccess) {
                if (!s.sync_state()) fail("log readback", "Mobility steps");
                break;
            }
            if (s.sync_state()) break;
            if (say) {
                for (int t = t_begin; t < s.h_st->poisson_index && (size_t)t < n_log; ++t) {
                    if (lg[(size_t)t].n_after_mobility >= cfg.capacity) printf("\n\nOVERFLOW FROM ADDING PARTICLES\n\n\n");
                    if (lg[(size_t)t].n_after_compaction == 0) { printf("Hit 0\n"); break; }
                }
            }
        } else {
            for (int t = t_begin; t < cfg.poisson_steps; ++t) {
                if (verbose && t % cfg.verbose == 0 && print_state(sim, t)) { failed = true; break; }
                if (s.poisson_step(true)) { failed = true; break; }
                if (say && s.h_st->last_n_after_mobility >= cfg.capacity) printf("\n\nOVERFLOW FROM ADDING PARTICLES\n\n\n");
                if (s.h_st->hit_zero) { if (say) printf("Hit 0\n"); break; }
            }
            if (failed) break;
        }
        if (verbose && cfg.poisson_steps % cfg.verbose == 0 && print_state(sim, cfg.poisson_steps)) break;     // src/pic.cu:561
        if (cudaEventRecord(s.ev_stop, s.stream) != cudaSuccess || cudaEventSynchronize(s.ev_stop) != cudaSuccess) {
            fail(cudaGetErrorString(cudaGetLastError()), "After sim");
            break;
        }
        float ms = 0;
        cudaEventElapsedTime(&ms, s.ev_start, s.ev_stop);
        s.gpu_ms = ms;
        picb200_summary sum;
        if (picb200_get_summary(sim, &sum)) break;
        // PICB200_DUMP_FINAL=<file>: binary snapshot of the final state (no text dump needed; restart with PICB200_RESTART)
        if (const char* fin = getenv("PICB200_DUMP_FINAL")) { if (picb200_save_state(sim, fin, 1)) break; }
        if (electrons_out && electrons_cap > 0) {
            int64_t got = 0;
            if (picb200_download_aos(sim, electrons_out, electrons_cap, &got)) break;
        }
        if (say) {
            printf("Final amount of particles: %lld\n", (long long)sum.final_n);
            printf("Particles added: %lld\n", (long long)sum.total_added);
            printf("Particles removed: %lld\n", (long long)sum.total_removed);
            printf("GPU time of program: %f ms\n", sum.gpu_ms);
        }
        if (out) *out = sum;
        rc = 0;
    } while (0);
    if (rc == 0 && keep) *keep = sim;
    else picb200_destroy(sim);
    return rc;
}

extern "C" int picb200_run(const picb200_config* cfg, picb200_summary* out, picb200_electron* electrons_out, int64_t electrons_cap) {
    return run_impl(cfg, out, electrons_out, electrons_cap, nullptr);
}
extern "C" int picb200_run_keep(const picb200_config* cfg, picb200_summary* out, picb200_sim** sim_out) {
    if (!sim_out) return fail("null argument", "picb200_run_keep");
    return run_impl(cfg, out, nullptr, 0, sim_out);
}
