/*
 * picb200.h -- C ABI of libpicb200.so, the B200-native PIC-MCC electron engine.
 *
 * This is the drop-in boundary for the hot path of MagnusMouritzen/particle-simulation:
 * plain pointers and sizes, no C++ or torch types.  Each entry point names the reference
 * interface it replaces (paths relative to the reference checkout).  INTEGRATION.md shows
 * the few lines a maintainer of the reference would add to bind them.
 *
 * Threading: one caller thread per picb200_sim.  All functions return 0 on success and a
 * non-zero code on failure; picb200_last_error() then returns a thread-local message of
 * the form the reference prints ("CUDA error: ...\nAt: ...", src/utility.cu:108-115).
 * There is no CPU fallback: without a CUDA device every compute call fails.
 */
#ifndef PICB200_H
#define PICB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PICB200_ABI_VERSION 1

/* Scheduler selector = the reference's `mode` argument of runPIC (src/pic.cu:359;
 * CLI strings "30".."33" in src/main.cu:26-40 map to 0..3). */
enum {
    PICB200_MODE_DYNAMIC     = 0, /* "30": persistent device-side work queue, one launch per Poisson step */
    PICB200_MODE_CPU_SYNC    = 1, /* "31": one launch per generation of children, host reads n in between  */
    PICB200_MODE_NAIVE       = 2, /* "32": one launch per mobility step, host reads n after every step     */
    PICB200_MODE_DYNAMIC_OLD = 3  /* "33": persistent kernel, block-granular claiming                      */
};

#define PICB200_DEAD (-2)          /* src/electron.h:8  */
#define PICB200_TABLE_ROWS 10000   /* src/cross_section.h:9 (N_STEPS) */
#define PICB200_GRID 512           /* src/cell.cu:3 (Grid_Size) */

/* Arguments of runPIC (src/pic.cu:359, src/pic.h:9) plus what the reference hard-codes. */
typedef struct picb200_config {
    int32_t  abi_version;        /* PICB200_ABI_VERSION */
    int32_t  mode;               /* PICB200_MODE_* */
    int32_t  verbose;            /* 0 = quiet; k = print state every k-th Poisson step (src/utility.cu:124-125) */
    int32_t  init_n;             /* initial electrons (global, over all ranks) */
    int32_t  poisson_steps;      /* "max t" */
    int32_t  block_size;         /* threads per logical block */
    int32_t  capacity;           /* "max n" (per rank) */
    int32_t  sleep_time_ns;      /* only mode 33 polls; kept for CLI compatibility */
    int32_t  poisson_timestep;   /* mobility steps per Poisson step */
    int32_t  device;             /* CUDA device ordinal; -1 = current device (reference: device 0, src/pic.cu:394) */
    uint64_t seed;               /* 0 = 39587, the reference's seed (src/random.cu:4) */
    const char* table_path;      /* NULL = "./src/cross_section.txt" relative to CWD (src/pic.cu:381) */
    const float* table_rows;     /* optional: PICB200_TABLE_ROWS x {split_chance, remove_chance}; overrides table_path */
    /* multi-GPU (the reference has none, src/pic.cu:394): one picb200_sim per rank, particle slabs,
     * one integer all-reduce of the charge box per Poisson step. */
    int32_t  rank;               /* 0-based */
    int32_t  nranks;             /* 1 = single GPU */
    const void* nccl_unique_id;  /* 128-byte ncclUniqueId shared by all ranks; NULL = no NCCL (use picb200_peer_connect) */
    int32_t  print_stdout;       /* 1 = print the reference's stdout lines (header, scheduler name, states, summary) */
    int32_t  reserved;
} picb200_config;

/* Per-run summary: what runPIC prints at src/pic.cu:573-576 and returns in RunData (src/utility.h:27-31). */
typedef struct picb200_summary {
    int64_t final_n;            /* min(n, capacity) */
    int64_t total_added;        /* "Particles added"   */
    int64_t total_removed;      /* "Particles removed" */
    int64_t pushes;             /* number of updateParticle evaluations (src/particle_move.cu:83) */
    int64_t poisson_steps_done;
    double  gpu_ms;             /* cudaEvent span around the Poisson loop, as src/pic.cu:476 .. :563 */
    int32_t overflowed;         /* any step hit capacity ("OVERFLOW FROM ADDING PARTICLES", src/pic.cu:544) */
    int32_t hit_zero;           /* population died out ("Hit 0", src/pic.cu:556-559) */
    int32_t sm_count;           /* "Multiprocessor count" */
    int32_t blocks_per_sm;      /* "Blocks per multiprocessor" */
} picb200_summary;

/* The reference's 64-byte AoS record (src/electron.h:12-18), for callers that consume RunData.electrons. */
typedef struct picb200_electron {
    double  position[3];
    double  velocity[3];
    float   acceleration[3];
    int32_t timestamp;
} picb200_electron;

/* Device time per phase of the Poisson loop (CUDA events on the launching stream) and the number
 * of kernels the engine launched, accumulated since picb200_set_profiling(sim, 1). */
typedef struct picb200_profile {
    double  deposit_ms;      /* particlesToGrid             (src/grid_operations.cu:15-26)  */
    double  gather_ms;       /* updateGrid + gridToParticles (src/grid_operations.cu:29-72) */
    double  clear_ms;        /* resetGrid                   (src/grid_operations.cu:6-12)   */
    double  mover_ms;        /* the scheduler kernels       (src/pic.cu:509-537)            */
    double  compact_ms;      /* remove_dead_particles       (src/pic.cu:547-551)            */
    double  collective_ms;   /* multi-GPU charge all-reduce (no reference counterpart)      */
    int64_t launches;        /* kernels launched by the engine */
    int64_t mover_launches;  /* of which scheduler kernels */
} picb200_profile;

typedef struct picb200_sim picb200_sim;

/* ---- library ------------------------------------------------------------------------------- */
int         picb200_abi_version(void);
const char* picb200_last_error(void);
int         picb200_device_count(void);          /* 0 when no CUDA device is usable */
void        picb200_config_default(picb200_config* cfg);

/* ---- whole run: replaces runPIC (src/pic.cu:359-599) ------------------------------------------
 * electrons_out may be NULL; otherwise it receives min(final_n, electrons_cap) records in slot order. */
int picb200_run(const picb200_config* cfg, picb200_summary* out,
                picb200_electron* electrons_out, int64_t electrons_cap);
/* The same run, but the simulation is handed back alive (*sim_out; release with picb200_destroy): a caller that
 * returns RunData.electrons = new Electron[final_n] (src/pic.cu:582-584) learns final_n from *out, allocates, and
 * fetches the records with picb200_download_aos.  Used by the runPIC shim (csrc/runpic_shim.cpp). */
int picb200_run_keep(const picb200_config* cfg, picb200_summary* out, picb200_sim** sim_out);

/* ---- session API: the phases of runPIC's loop, callable one by one -------------------------- */
/* allocations + table upload + persistent-grid sizing (src/pic.cu:380-454) */
int picb200_create(const picb200_config* cfg, picb200_sim** out);
void picb200_destroy(picb200_sim* sim);
/* setup_particles (src/particle_move.cu:7-19, launch src/pic.cu:425); ids = id_base + i */
int picb200_init_uniform(picb200_sim* sim, int32_t init_n, uint64_t id_base);
/* replace the population from host SoA arrays (n <= capacity); any pointer except x,y,z may be NULL (-> 0, ts -1, id i) */
int picb200_upload(picb200_sim* sim, int32_t n,
                   const float* x, const float* y, const float* z,
                   const float* vx, const float* vy, const float* vz,
                   const float* ax, const float* ay, const float* az,
                   const int32_t* ts, const uint64_t* id);
/* copy min(n, capacity) electrons to host SoA arrays (any pointer may be NULL); returns count in *n_out */
int picb200_download(picb200_sim* sim, int32_t* n_out,
                     float* x, float* y, float* z, float* vx, float* vy, float* vz,
                     float* ax, float* ay, float* az, int32_t* ts, uint64_t* id);
int picb200_download_aos(picb200_sim* sim, picb200_electron* out, int64_t cap, int64_t* n_out);
/* one iteration of the Poisson loop (src/pic.cu:487-560) without printing */
int picb200_poisson_step(picb200_sim* sim);
/* `steps` iterations enqueued back to back; stops early at "Hit 0" */
int picb200_run_steps(picb200_sim* sim, int32_t steps);
int picb200_get_summary(picb200_sim* sim, picb200_summary* out);
int picb200_set_poisson_index(picb200_sim* sim, int32_t index);
/* per-phase device timing (replaces the single cudaEvent pair of src/pic.cu:374-376, :476, :563) */
/* on: 0 = off, 1 = time the scheduler kernels only, 2 = time every phase */
int picb200_set_profiling(picb200_sim* sim, int32_t on);
int picb200_get_profile(picb200_sim* sim, picb200_profile* out);
/* page-locked host buffers for picb200_upload / picb200_download (the reference uses pageable calloc, src/pic.cu:420) */
/* benchmarking aid: rewrite the (all-zero) 512 MiB charge grid so the 126 MB L2 holds none of the electrons */
int picb200_flush_l2(picb200_sim* sim);
void* picb200_alloc_pinned(size_t bytes);
void  picb200_free_pinned(void* p);

/* ---- single phases, for kernel-level parity tests ------------------------------------------- */
/* resetGrid + particlesToGrid (src/grid_operations.cu:6-26); multi-GPU: followed by the all-reduce */
int picb200_op_deposit(picb200_sim* sim);
/* updateGrid + gridToParticles fused (src/grid_operations.cu:29-72); leaves the charge grid intact */
int picb200_op_gather(picb200_sim* sim);
/* zero the charge grid again (the engine keeps it all-zero between Poisson steps) */
int picb200_op_clear_grid(picb200_sim* sim);
/* the mobility phase of the configured mode (src/pic.cu:509-537) */
int picb200_op_mobility(picb200_sim* sim);
/* remove_dead_particles (src/pic.cu:320-357, launch :547-551) */
int picb200_op_compact(picb200_sim* sim);
/* dense copy of charge counts in the inclusive cell box [lo,hi], x fastest */
int picb200_charge_box(picb200_sim* sim, const int32_t lo[3], const int32_t hi[3], int32_t* out);
/* picb200_charge_box for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same arguments and result as picb200_charge_box; the packing kernel's scratch is a fixed device array of
 * PICB200_NOALLOC_MAX_CELLS cells (4 MiB, held by every GPU that loads the library until exit) instead of a cudaMalloc
 * per call: the call allocates and frees nothing.  A box with more cells is rejected with a clear error (use
 * picb200_charge_box).  All calls on one GPU share the scratch, which keeps the last call's contents (every cell of the
 * requested box is overwritten before it is read, nothing needs to start at zero): calls that overlap on the GPU race. */
#define PICB200_NOALLOC_MAX_CELLS (1 << 20)
int picb200_charge_box_noalloc(picb200_sim* sim, const int32_t lo[3], const int32_t hi[3], int32_t* out);
/* cell bounding box of the last deposit (lo > hi when empty) */
int picb200_charge_bbox(picb200_sim* sim, int32_t lo[3], int32_t hi[3]);

/* ---- pieces of the arithmetic, evaluated ON THE DEVICE, for known-answer tests --------------- */
/* Philox4x32-10: n blocks; ctr = n x 4 words, key = 2 words, out = n x 4 words */
int picb200_dev_philox(int32_t device, int64_t n, const uint32_t* ctr, const uint32_t* key, uint32_t* out);
/* energyToIndex (src/cross_section.cu:32-35) for n fp32 energies */
int picb200_dev_energy_index(int32_t device, int64_t n, const float* energy, int32_t* out);
/* energyToIndex on the device for every stride-th non-negative fp32 bit pattern (stride 1: all 2^31, incl. +inf and NaNs),
 * compared with a bisection of the exact bin-edge table; returns the number of mismatches and one offending bit pattern */
int picb200_dev_energy_index_sweep(int32_t device, uint32_t stride, int64_t* mismatches, uint32_t* first_bad_bits);
/* host-side loader (src/cross_section.cu:3-28): returns rows read, -1 if the file cannot be opened */
int picb200_load_table(const char* path, float* rows2);

/* ---- snapshots ------------------------------------------------------------ */
/* image() + save_png() of src/utility.cu:4-74 without libpng: 500 x 500 RGB, an 11 x 11 black square per
 * electron at pixel (x + i, 500 - (y + j)); host only, no GPU needed.  The CLI writes
 * ./out/visualization/test_%04d.png and the "Saved image to ..." line at every logged step, as log() does
 * (src/utility.cu:133), when that directory exists. */
int picb200_snapshot_png(const float* x, const float* y, int64_t n, const char* path);
/* Binary snapshot of the population (SoA, optionally sorted by particle id so that equal runs give equal
 * files) and of the run counters; the reference can only print its state as text (src/utility.cu:124-137)
 * and cannot restart.  picb200_load_state replaces the population of `sim` (same seed and poisson_timestep
 * required) and restores Poisson index, totals and push count: the run continues bit for bit. */
int picb200_save_state(picb200_sim* sim, const char* path, int32_t sorted_by_id);
int picb200_load_state(picb200_sim* sim, const char* path);

/* ---- multi-GPU over NVLink peer memory (no reference counterpart: device 0 is hard-coded, src/pic.cu:394)
 * With nranks > 1 the modes 30/33 kernel exchanges the charge grid itself, by adding the rows of its local
 * grid into every rank's summed grid over NVLink, so that a whole run stays one launch per GPU.  Every rank
 * exports a handle, the caller gathers the nranks handles in rank order (threads: shared array; processes: any
 * all-gather) and every rank connects.  Works across threads of one process (peer access) and across processes
 * of one node (CUDA IPC).  Where the GPUs sit behind an NVSwitch the adds go through a multicast object (one
 * multimem.red reaches all ranks; PICB200_NVLS=0 keeps one bulk reduce-add per peer): rank 0 creates it in
 * picb200_peer_export, the others take it over in picb200_peer_connect, which is therefore COLLECTIVE -- all
 * ranks must be inside the call at the same time (it waits up to 60 s for the others).
 * Without this call the engine falls back to one NCCL all-reduce per Poisson step (needs nccl_unique_id). */
#define PICB200_PEER_HANDLE_BYTES 256
int picb200_peer_export(picb200_sim* sim, void* handle_out /* PICB200_PEER_HANDLE_BYTES */);
int picb200_peer_connect(picb200_sim* sim, const void* handles /* nranks x PICB200_PEER_HANDLE_BYTES */);

/* NCCL bootstrap helper: writes a 128-byte ncclUniqueId (call on rank 0, broadcast it yourself) */
int picb200_nccl_unique_id(void* out128);

#ifdef __cplusplus
}
#endif
#endif /* PICB200_H */
