// pic_engine.cu -- host side of libpicb200.so: the Poisson-loop driver that replaces runPIC
// (src/pic.cu:359-599 of MagnusMouritzen/particle-simulation) and the C ABI of include/picb200.h.
//
// Host C++ only orchestrates: it allocates the SoA electron store and the charge grid once,
// stages the collision tables, and enqueues kernels that size themselves from a device-resident
// control block (DevState).  In modes 30/33 a whole run is enqueued without a single host read;
// modes 31/32 keep the host round trips that define them in the reference.
#include "pic_device.cuh"
#include "pic_nccl.h"
#include "pic_nvls.h"
#include "../../include/picb200.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <string>
#include <vector>
#include <unistd.h>

namespace picb200 {

static thread_local std::string g_last_error;

static int fail(const std::string& what, const std::string& where) {
    g_last_error = "CUDA error: " + what + "\nAt: " + where + "\n";   // src/utility.cu:111
    return 1;
}
#define CK(call, where)                                                                  \
    do {                                                                                 \
        cudaError_t _e = (call);                                                         \
        if (_e != cudaSuccess) return fail(cudaGetErrorString(_e), where);               \
    } while (0)
#define CK_LAST(where) CK(cudaGetLastError(), where)

// ---- cross-section loader: src/cross_section.cu:3-28 (two floats per row, stop at a short read)
static int load_table_file(const char* path, std::vector<float>& rows2) {
    rows2.assign((size_t)kRows * 2, 0.f);
    std::ifstream in(path);
    if (!in) return -1;
    int i = 0;
    for (; i < kRows; ++i) {
        float a, b;
        if (!(in >> a >> b)) break;
        rows2[2 * (size_t)i] = a;
        rows2[2 * (size_t)i + 1] = b;
    }
    return i;
}

// The reference's energyToIndex before clamping (src/cross_section.cu:33), on an fp32 energy.
static inline double ref_index_value(float e) { return std::trunc((std::log10((double)e) + 6) * kRows / 22); }

// edge[j] (1 <= j < kRows) = smallest fp32 energy whose reference index is >= j;
// edge[0] = 0, edge[kRows] = +inf.  energy_index() on the device probes exactly one entry.
static void build_energy_edges(float* edge) {
    edge[0] = 0.f;
    for (int j = 1; j < kRows; ++j) {
        float e = (float)std::pow(10.0, -6.0 + 22.0 * (double)j / (double)kRows);
        while (ref_index_value(e) >= (double)j) e = std::nextafterf(e, 0.f);
        while (ref_index_value(e) < (double)j) e = std::nextafterf(e, std::numeric_limits<float>::infinity());
        edge[j] = e;
    }
    edge[kRows] = std::numeric_limits<float>::infinity();
}

// bnd[k] (1 <= k < kGrid) = smallest fp32 coordinate whose cell index int(double(x)/1e-2)
// (src/grid_operations.cu:21-23) is >= k; bnd[0] = -inf, bnd[kGrid] = +inf.
// The collider draws r = fmul(fma(float(w), 2^-32, 2^-33), 100) from a Philox word w and tests `r < chance`
// (src/particle_move.cu:61-79, src/random.cu:7-12).  r never decreases as float(w) grows, so the test is
// `float(w) < T(chance)` with T the smallest fp32 in [0, 2^32] whose r is not below the chance (0: never, also for a
// NaN chance; +inf: always).  The tables hold T, which saves the kernels an FMA and a multiply per push; equality
// with the literal form is checked over every value float(w) can take (tests/test_table.py).
static float draw_threshold(float chance) {
    auto r_of = [](float f) { return std::fmaf(f, 0x1p-32f, 0x1p-33f) * 100.0f; };
    if (!(r_of(0.0f) < chance)) return 0.0f;
    if (r_of(4294967296.0f) < chance) return INFINITY;
    uint32_t lo = 0u, hi = 0x4f800000u;          // bit patterns of 0.0f (r < chance) and 2^32 (not); non-negative floats order like their bits
    while (hi - lo > 1u) {
        const uint32_t mid = lo + (hi - lo) / 2u;
        float f;
        memcpy(&f, &mid, 4);
        if (r_of(f) < chance) lo = mid; else hi = mid;
    }
    float t;
    memcpy(&t, &hi, 4);
    return t;
}

static void build_cell_bounds(float* bnd) {
    bnd[0] = -std::numeric_limits<float>::infinity();
    for (int k = 1; k < kGrid; ++k) {
        float x = (float)((double)k * 1e-2);
        while ((double)x / 1e-2 >= (double)k) x = std::nextafterf(x, 0.f);
        while ((double)x / 1e-2 < (double)k) x = std::nextafterf(x, std::numeric_limits<float>::infinity());
        bnd[k] = x;
    }
    bnd[kGrid] = std::numeric_limits<float>::infinity();
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Page-locked words the kernels of a device write when a grid barrier times out (barrier_timeout_abort): one set per
// device for the life of the process -- they must stay readable after the CUDA context has failed.
static int* g_host_abort[64] = {nullptr};
static int* abort_words(int device) { return (device >= 0 && device < 64) ? g_host_abort[device] : nullptr; }
static std::string abort_message(int device) {
    const int* w = abort_words(device);
    if (!w || w[0] == 0) return std::string();
    char buf[256];
    snprintf(buf, sizeof buf, "grid barrier %d of the persistent kernel timed out: block %d of %d waited, not every block arrived (%s)",
             w[2], w[1], w[3], w[0] == 1 ? "k_mover" : w[0] == 2 ? "k_poisson_fused" : "peer exchange");
    return std::string(buf);
}

struct Sim {
    picb200_config cfg{};
    int device = 0;
    int sm_count = 0;
    int mover_threads = 1024;
    int fused_smem_pad = 0;          // PICB200_SMEM_PAD: extra dynamic shared memory for the fused kernel (carve-out experiments)
    int blocks_per_sm = 1;
    int aux_blocks = 0;
    int gather_blocks[2] = {0, 0};   // grid of k_gather<false/true>, sized by the configured population
    int deposit_blocks = 0;          // grid of k_deposit, likewise
    uint2 seed{};
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    Soa soa{};
    Soa park{};             // parking area of the fused kernel's lazy compaction (step_end_fused)
    void* park_base = nullptr;
    int park_cap = 0;
    int compact_policy = 0;
    void* soa_base = nullptr;
    float4* field = nullptr;             // fused kernel, one rank: dense per-cell field for dense plasmas (field_rows), 2 GiB
    int field_ratio = 2;                 // electrons per cell of the occupied box from which it is used (PICB200_FIELD_RATIO; < 0: never)
    int field_min_n = 0;                 // ... and electrons on this rank (PICB200_FIELD_MIN_N)
    bool field_launch = false;           // modes 31/32: k_field_box is part of the grid phase (decided at create, see there)
    int* grid[kGrids] = {nullptr, nullptr, nullptr};   // three charge grids in rotation: gather from one, deposit the next step into the second, empty the third off the critical path
    int cur = 0;                         // host view of DevState::cur (valid after every sync)
    bool charge_valid = false;           // grid[cur] holds the deposit of the current population
    bool generic_ts = false;             // the population came with the caller's own timestamps (upload, state file): the next fused launch must handle any value
    int* holes = nullptr;
    int* movers = nullptr;
    size_t list_cap = 0;
    float* tables = nullptr;
    DevState* st = nullptr;
    DevState* h_st = nullptr;        // pinned mirror
    StepLog* log = nullptr;
    int log_len = 0;
    double efc = 0;
    double gpu_ms = 0;
    // multi-GPU
    NcclApi* nccl = nullptr;
    void* comm = nullptr;
    int* box_buf = nullptr;
    size_t box_buf_cells = 0;
    int* d_bbox_red = nullptr;
    int* h_bbox_red = nullptr;
    // multi-GPU, peer mode: the exchange runs inside the fused kernel over NVLink peer mappings
    PeerSync* psync = nullptr;           // arrival counters and union-box words, written by the peers' remote atomics
    PeerLocal* pl = nullptr;
    PeerLocal* h_pl = nullptr;           // pinned mirror
    PeerArgs pa{};
    bool peer_connected = false;
    std::vector<void*> ipc_opened;
    // ... with the reduction in the NVSwitch (pic_nvls.h): grids 0/1 are ONE virtual-memory allocation bound to a multicast object
    bool nvls = false;
    size_t nvls_bytes = 0;
    CUmemGenericAllocationHandle nvls_mem = 0, nvls_mc = 0;
    CUdeviceptr nvls_uc = 0, nvls_mv = 0;      // the rank's own mapping of its grids / the multicast mapping
    bool nvls_mc_owner = false, nvls_bound = false;
    int nvls_fd = -1;                          // rank 0: the exported descriptor of the multicast object
    char nvls_shm[40] = {0};                   // rank 0: name of the set-up barrier's shared-memory segment
    int nvls_alloc_grids();
    int peer_export(void* out);
    int peer_connect(const void* handles);
    // profiling: CUDA-event spans per phase on the launching stream + a count of kernel launches
    enum Phase { kPhDeposit = 0, kPhGather, kPhClear, kPhMover, kPhCompact, kPhCollective, kPhCount };
    struct Span { cudaEvent_t a, b; int phase; };
    int profiling = 0;             // 0 off, 1 scheduler kernels only, 2 every phase
    std::vector<Span> spans;
    size_t spans_used = 0;
    bool span_open = false;
    double phase_ms[kPhCount] = {0, 0, 0, 0, 0, 0};
    long long launches = 0, mover_launches = 0;
    // mode 32 is launch-bound (2 * poisson_timestep + 7 kernels per Poisson step): its Poisson step is captured once
    // into a CUDA graph and replayed (nothing in it depends on a value the host would have to read)
    cudaGraphExec_t step_graph = nullptr;
    long long graph_launches = 0, graph_mover_launches = 0;     // kernels in one replay
    void span_begin(int phase);
    void span_end();
    int collect_spans();
    int build_step_graph();
    // mode 31 with its generation windows kept on the device (else: the reference's host loop, src/pic.cu:514-523)
    bool device_windows() const { return cfg.nranks == 1 && cfg.poisson_timestep <= 128 && !getenv("PICB200_HOST_SYNC"); }
    // optional callback after every launch of the naive scheduler (verbose state dumps, src/pic.cu:528)
    void (*naive_hook)(Sim*, void*) = nullptr;
    void* naive_hook_user = nullptr;

    ~Sim() { release(); }
    void release();
    int create(const picb200_config* c);
    int sync_state();
    int reset_state();
    int init_uniform(int init_n, unsigned long long id_base);
    int op_deposit();
    int op_gather();
    int op_clear(bool reset_box = true);
    int op_mobility_begin(int reset_box = -1);
    int op_mobility();
    int op_compact();
    int op_step_end();
    int poisson_step(bool sync);
    int fused_steps(int steps);
    FusedArgs fused_args(int per_launch, bool peer);
    int launch_fused(FusedArgs& a, bool peer, int variant);
    int warm_kernels();
    bool fused = false;                  // modes 30/33: whole Poisson steps inside one cooperative kernel
    int launch_mover_window(int lo, int hi, int t_from, int t_to, bool stepwise);
    int allreduce_charge();
};

void Sim::span_begin(int phase) {
    span_open = profiling >= 2 || (profiling == 1 && phase == kPhMover);
    if (!span_open) return;
    if (spans_used == spans.size()) {
        Span sp{nullptr, nullptr, 0};
        cudaEventCreate(&sp.a);
        cudaEventCreate(&sp.b);
        spans.push_back(sp);
    }
    spans[spans_used].phase = phase;
    cudaEventRecord(spans[spans_used].a, stream);
}
void Sim::span_end() {
    if (!span_open) return;
    span_open = false;
    cudaEventRecord(spans[spans_used].b, stream);
    ++spans_used;
}
// call after the stream has been synchronised
int Sim::collect_spans() {
    for (size_t k = 0; k < spans_used; ++k) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, spans[k].a, spans[k].b), "Profile");
        phase_ms[spans[k].phase] += ms;
    }
    spans_used = 0;
    return 0;
}

void Sim::release() {
    for (auto& sp : spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    spans.clear();
    if (device >= 0 && (soa_base || st)) cudaSetDevice(device);
    if (comm && nccl) { nccl->commDestroy(comm); comm = nullptr; }
    if (stream) { cudaStreamSynchronize(stream); }
    for (void* q : ipc_opened) cudaIpcCloseMemHandle(q);
    ipc_opened.clear();
    cudaFree(psync); psync = nullptr;
    cudaFree(pl); pl = nullptr;
    if (h_pl) { cudaFreeHost(h_pl); h_pl = nullptr; }
    cudaFree(soa_base); soa_base = nullptr;
    cudaFree(park_base); park_base = nullptr;
    if (nvls) {
        DrvApi* d = drv_api();
        if (d) {
            if (nvls_mv) { d->cuMemUnmap_(nvls_mv, nvls_bytes); d->cuMemAddressFree_(nvls_mv, nvls_bytes); nvls_mv = 0; }
            if (nvls_uc) { d->cuMemUnmap_(nvls_uc, nvls_bytes); d->cuMemAddressFree_(nvls_uc, nvls_bytes); nvls_uc = 0; }
            if (nvls_mc) { d->cuMemRelease_(nvls_mc); nvls_mc = 0; }
            if (nvls_mem) { d->cuMemRelease_(nvls_mem); nvls_mem = 0; }
        }
        if (nvls_fd >= 0) { close(nvls_fd); nvls_fd = -1; }
        if (nvls_shm[0]) { shm_unlink(nvls_shm); nvls_shm[0] = 0; }
        grid[0] = grid[1] = nullptr;
        nvls = false;
    }
    for (int g = 0; g < kGrids; ++g) { cudaFree(grid[g]); grid[g] = nullptr; }
    cudaFree(field); field = nullptr;
    cudaFree(holes); holes = nullptr;
    cudaFree(movers); movers = nullptr;
    cudaFree(tables); tables = nullptr;
    cudaFree(st); st = nullptr;
    cudaFree(log); log = nullptr;
    cudaFree(box_buf); box_buf = nullptr;
    cudaFree(d_bbox_red); d_bbox_red = nullptr;
    if (h_st) { cudaFreeHost(h_st); h_st = nullptr; }
    if (h_bbox_red) { cudaFreeHost(h_bbox_red); h_bbox_red = nullptr; }
    if (ev_start) { cudaEventDestroy(ev_start); ev_start = nullptr; }
    if (ev_stop) { cudaEventDestroy(ev_stop); ev_stop = nullptr; }
    if (step_graph) { cudaGraphExecDestroy(step_graph); step_graph = nullptr; }
    if (stream) { cudaStreamDestroy(stream); stream = nullptr; }
}

int Sim::create(const picb200_config* c) {
    cfg = *c;
    if (cfg.abi_version != PICB200_ABI_VERSION) return fail("ABI version mismatch", "picb200_create");
    if (cfg.capacity <= 0 || cfg.poisson_timestep < 0 || cfg.mode < 0 || cfg.mode > 3 || cfg.nranks < 1 ||
        cfg.rank < 0 || cfg.rank >= cfg.nranks)
        return fail("invalid configuration", "picb200_create");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail("no CUDA device: libpicb200 has no CPU fallback", "picb200_create");
    if (cfg.device >= 0) { device = cfg.device; } else { CK(cudaGetDevice(&device), "Get device"); }
    CK(cudaSetDevice(device), "Set device");
    CK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device), "Device attributes");
    const unsigned long long sd = cfg.seed ? cfg.seed : 39587ull;      // src/random.cu:4
    seed = make_uint2((unsigned)sd, (unsigned)(sd >> 32));
    // logical blocks of `block size` threads share one CTA (and one copy of the shared-memory tables)
    int bs = std::max(32, std::min(1024, (cfg.block_size / 32) * 32));
    blocks_per_sm = std::max(1, 1024 / bs);
    mover_threads = bs * blocks_per_sm;
    if (const char* mt = getenv("PICB200_MOVER_THREADS")) mover_threads = std::max(32, std::min(1024, (atoi(mt) / 32) * 32));   // (experiments: how the fused kernel's rate depends on the warps per SM)
    aux_blocks = sm_count * 8;
    // k_gather is latency-bound and likes few trips per thread (1e7 electrons: 8 blocks per SM 3.62 ms, 32: 3.26, 64: 3.19 for
    // cfg5 mode 32 at poisson_timestep = 1; 1e6 electrons: 8 is best by 1 %): one trip of four electrons per thread for the
    // configured population, within [8, 64] blocks per SM (the step's graph is captured once)
    {
        const long long want = ((long long)cfg.init_n + 1024LL * sm_count - 1) / (1024LL * sm_count);
        const int per_sm = getenv("PICB200_GATHER_BLOCKS") ? std::max(1, atoi(getenv("PICB200_GATHER_BLOCKS"))) : (int)std::max(8LL, std::min(64LL, want));
        gather_blocks[0] = gather_blocks[1] = sm_count * per_sm;
        // k_deposit likewise, up to 32 per SM (every block ends with six atomics on the box words): 3.18 -> 3.11 ms there
        deposit_blocks = sm_count * (getenv("PICB200_DEPOSIT_BLOCKS") ? std::max(1, atoi(getenv("PICB200_DEPOSIT_BLOCKS"))) : (int)std::max(8LL, std::min(32LL, want)));
    }
    efc = (-1.602176487e-19 * -1.602176487e-19) / (4 * 3.1415926536 * 8.8541878176E-12 * 1e-2 * 1e-2 * 9.1093837015e-31);  // src/cell.cu:5

    if (device < 64) {
        if (!g_host_abort[device]) { CK(cudaMallocHost(&g_host_abort[device], 4 * sizeof(int)), "Minor alloc"); memset(g_host_abort[device], 0, 4 * sizeof(int)); }
        int* dev_view = nullptr;
        CK(cudaHostGetDevicePointer(&dev_view, g_host_abort[device], 0), "Minor alloc");
        CK(cudaMemcpyToSymbol(g_abort_flag, &dev_view, sizeof(dev_view)), "Minor alloc");
        const unsigned long long ns = (unsigned long long)(getenv("PICB200_BARRIER_TIMEOUT_MS") ? std::max(1, atoi(getenv("PICB200_BARRIER_TIMEOUT_MS"))) : 30000) * 1000000ull;
        CK(cudaMemcpyToSymbol(g_barrier_timeout_ns, &ns, sizeof(ns)), "Minor alloc");
    }
    CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking), "Stream");
    CK(cudaEventCreate(&ev_start), "Event");
    CK(cudaEventCreate(&ev_stop), "Event");

    // ---- collision tables (src/pic.cu:380-386)
    std::vector<float> rows2;
    if (cfg.table_rows) {
        rows2.assign(cfg.table_rows, cfg.table_rows + (size_t)kRows * 2);
    } else {
        const char* path = cfg.table_path ? cfg.table_path : "./src/cross_section.txt";
        int r = load_table_file(path, rows2);
        if (r < 0) return fail(std::string("Failed to open file: ") + path, "CS Alloc");   // hard error (the reference only warns)
        if (r < kRows) fprintf(stderr, "Failed to read data for step %d\n", r);           // src/cross_section.cu:17
    }
    std::vector<unsigned char> img((size_t)kAllTableBytes, 0);
    float* thr = reinterpret_cast<float*>(img.data());
    build_cell_bounds(reinterpret_cast<float*>(img.data() + kTableSmemBytes));
    for (int i = 0; i < kRows; ++i) {
        thr[2 * i] = draw_threshold(rows2[2 * (size_t)i]);                                   // r < split_chance
        thr[2 * i + 1] = draw_threshold(rows2[2 * (size_t)i + 1] + rows2[2 * (size_t)i]);    // r < remove_chance + split_chance, summed in fp32 (src/particle_move.cu:76)
    }
    build_energy_edges(reinterpret_cast<float*>(img.data() + kThrBytes));
    CK(cudaMalloc(&tables, (size_t)kAllTableBytes), "CS Alloc");
    CK(cudaMemcpyAsync(tables, img.data(), (size_t)kAllTableBytes, cudaMemcpyHostToDevice, stream), "CS Alloc");
    CK(cudaStreamSynchronize(stream), "CS Alloc");

    // ---- electron store: one allocation, 11 arrays (src/pic.cu:420-424 allocates 2 x cap x 64 B AoS)
    const size_t cap = (size_t)cfg.capacity;
    const size_t f_bytes = align_up(cap * 4, 256), id_bytes = align_up(cap * 8, 256);
    CK(cudaMalloc(&soa_base, f_bytes * 10 + id_bytes), "Electron alloc");
    unsigned char* p = static_cast<unsigned char*>(soa_base);
    float** fields[9] = {&soa.x, &soa.y, &soa.z, &soa.vx, &soa.vy, &soa.vz, &soa.ax, &soa.ay, &soa.az};
    for (auto f : fields) { *f = reinterpret_cast<float*>(p); p += f_bytes; }
    soa.ts = reinterpret_cast<int*>(p); p += f_bytes;
    soa.id = reinterpret_cast<unsigned long long*>(p);
    CK(cudaMemsetAsync(soa.ts, 0, f_bytes, stream), "Electron alloc");
    {
        // the lazy compaction parks children when the slots above the holes run out before the reference's counter does; a
        // forced compaction comes when one slot in eight is a hole, so an eighth of the capacity bounds what can wait
        park_cap = (int)std::min<size_t>(cap, std::max<size_t>((size_t)1 << 20, cap / 8));
        const size_t pf = align_up((size_t)park_cap * 4, 256), pi = align_up((size_t)park_cap * 8, 256);
        CK(cudaMalloc(&park_base, pf * 10 + pi), "Electron alloc");
        unsigned char* q = static_cast<unsigned char*>(park_base);
        float** pfields[9] = {&park.x, &park.y, &park.z, &park.vx, &park.vy, &park.vz, &park.ax, &park.ay, &park.az};
        for (auto f : pfields) { *f = reinterpret_cast<float*>(q); q += pf; }
        park.ts = reinterpret_cast<int*>(q); q += pf;
        park.id = reinterpret_cast<unsigned long long*>(q);
        // PICB200_COMPACT=eager: remove_dead_particles after every Poisson step, as the reference does;
        // =noguard: lazy without the capacity guard (exercises the parking path in the tests)
        const char* pol = getenv("PICB200_COMPACT");
        compact_policy = !pol ? 0 : (!strcmp(pol, "eager") ? 1 : (!strcmp(pol, "noguard") ? 2 : 0));
    }

    // ---- compaction work lists: at most min(D, S-D) <= cap/2 entries each
    list_cap = cap / 2 + 1;
    CK(cudaMalloc(&holes, list_cap * 4), "Minor alloc");
    CK(cudaMalloc(&movers, list_cap * 4), "Minor alloc");

    // ---- charge grid 512^3 int32, zeroed once (src/pic.cu:448-450 allocates 2 GiB of 16-B cells)
    // several ranks, peer transport: the two summed grids come from the virtual-memory API when the box can reduce in the
    // NVSwitch (pic_nvls.h; PICB200_NVLS=0 keeps cudaMalloc and the per-peer bulk reduce-adds)
    if (cfg.nranks > 1 && !cfg.nccl_unique_id && !(getenv("PICB200_NVLS") && atoi(getenv("PICB200_NVLS")) == 0)) {
        if (nvls_alloc_grids()) return 1;
    }
    for (int g = 0; g < kGrids; ++g) {
        if (!grid[g]) CK(cudaMalloc(&grid[g], (size_t)kGrid * kGrid * kGrid * 4), "Grid alloc");
        CK(cudaMemsetAsync(grid[g], 0, (size_t)kGrid * kGrid * kGrid * 4, stream), "Grid alloc");
    }

    CK(cudaMalloc(&st, sizeof(DevState)), "Minor alloc");
    CK(cudaMemsetAsync(st, 0, sizeof(DevState), stream), "Minor alloc");
    k_reset_bbox<<<1, 32, 0, stream>>>(st, kGrids);
    CK(cudaMallocHost(&h_st, sizeof(DevState)), "Minor alloc");
    memset(h_st, 0, sizeof(DevState));
    log_len = std::max(1, cfg.poisson_steps);
    CK(cudaMalloc(&log, sizeof(StepLog) * (size_t)log_len), "Minor alloc");
    CK(cudaMemsetAsync(log, 0, sizeof(StepLog) * (size_t)log_len, stream), "Minor alloc");

    if (const char* pad = getenv("PICB200_SMEM_PAD")) fused_smem_pad = std::max(0, std::min(atoi(pad), 227 * 1024 - 512 - kFusedSmemBytes(1024)));
    CK(cudaFuncSetAttribute(k_mover<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMoverSmemBytes(1024)), "Mover attr");
    CK(cudaFuncSetAttribute(k_poisson_fused<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmemBytes(1024) + fused_smem_pad), "Mover attr");
    CK(cudaFuncSetAttribute(k_poisson_fused<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmemBytes(1024) + fused_smem_pad), "Mover attr");
    CK(cudaFuncSetAttribute(k_poisson_fused<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmemBytes(1024) + fused_smem_pad), "Mover attr");
    CK(cudaFuncSetAttribute(k_poisson_fused<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmemBytes(1024) + fused_smem_pad), "Mover attr");
    CK(cudaFuncSetAttribute(k_poisson_fused<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmemBytes(1024, true)), "Mover attr");
    CK(cudaFuncSetAttribute(k_poisson_fused<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmemBytes(1024, true)), "Mover attr");
    fused = (cfg.mode == PICB200_MODE_DYNAMIC || cfg.mode == PICB200_MODE_DYNAMIC_OLD) && !getenv("PICB200_NO_FUSE");
    if (const char* fr = getenv("PICB200_FIELD_RATIO")) field_ratio = atoi(fr);
    // several ranks: the pass over the union box and its grid barrier are paid per rank whatever the slab holds: large slabs only
    field_min_n = getenv("PICB200_FIELD_MIN_N") ? atoi(getenv("PICB200_FIELD_MIN_N")) : (cfg.nranks > 1 ? 3000000 : 0);
    // modes 31/32: the per-cell field is one more launch per Poisson step (2 % of a step when it finds a sparse plasma and
    // returns); the six-point gather is bound by the L1's tag stage, one sector per cycle and SM (1e6 electrons: 29 us,
    // against 11 us for field + one sector per electron).  The step's graph is captured once, so the configured population
    // decides: from 5e5 electrons, which the initial 62^3-cell cube holds at 2 per cell (the device's own density rule)
    field_launch = !fused && field_ratio >= 0 && cfg.nranks == 1 &&
                   (getenv("PICB200_FIELD_LAUNCH") ? atoi(getenv("PICB200_FIELD_LAUNCH")) != 0 : cfg.init_n >= 500000);
    if (field_ratio >= 0 && ((cfg.nranks == 1 && (fused || field_launch)) || (cfg.nranks > 1 && fused && !cfg.nccl_unique_id)))      // 16-byte field records over the dense grid; only cells of the occupied box are ever written or read
        CK(cudaMalloc(&field, (size_t)kGrid * kGrid * kGrid * sizeof(float4)), "Grid alloc");
    CK(cudaFuncSetAttribute(k_mover<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMoverSmemBytes(1024)), "Mover attr");
    CK(cudaFuncSetAttribute(k_mover<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMoverSmemBytes(1024)), "Mover attr");

    if (cfg.nranks > 1) {
        if (cfg.nranks > kMaxRanks) return fail("at most 8 ranks", "picb200_create");
        {
            PeerSync init;
            memset(&init, 0, sizeof(init));
            for (int p = 0; p < 2; ++p)
                for (int q = 0; q < kMaxRanks; ++q) init.box[p][q] = 1023ull << 54;      // exchange number 1023: not before 1023 exchanges, by when it has been overwritten
            PeerLocal pl0;
            memset(&pl0, 0, sizeof(pl0));
            for (int p = 0; p < 2; ++p)
                for (int k = 0; k < 3; ++k) { pl0.sbox_lo[p][k] = kGrid; pl0.sbox_hi[p][k] = -1; }
            CK(cudaMalloc(&psync, sizeof(PeerSync)), "Minor alloc");
            CK(cudaMemcpy(psync, &init, sizeof(init), cudaMemcpyHostToDevice), "Minor alloc");
            CK(cudaMalloc(&pl, sizeof(PeerLocal)), "Minor alloc");
            CK(cudaMemcpy(pl, &pl0, sizeof(pl0), cudaMemcpyHostToDevice), "Minor alloc");
        }
        CK(cudaMallocHost(&h_pl, sizeof(PeerLocal)), "Minor alloc");
        memset(h_pl, 0, sizeof(PeerLocal));
        if (cfg.nccl_unique_id) {            // NCCL path (baseline / no peer access): one communicator per rank
            nccl = load_nccl(g_last_error);
            if (!nccl) return 1;
            NcclUniqueId id;
            memcpy(&id, cfg.nccl_unique_id, sizeof(id));
            int r = nccl->commInitRank(&comm, cfg.nranks, id, cfg.rank);
            if (r != 0) return fail(std::string("ncclCommInitRank: ") + nccl->getErrorString(r), "NCCL init");
            CK(cudaMalloc(&d_bbox_red, 8 * sizeof(int)), "Minor alloc");
            CK(cudaMallocHost(&h_bbox_red, 8 * sizeof(int)), "Minor alloc");
        }
    }
    CK(cudaStreamSynchronize(stream), "Setup");
    CK_LAST("Setup");
    if (build_step_graph()) return 1;
    return warm_kernels();
}

// CUDA loads a kernel's code on its first launch, and the first launch that wants 190 KB of shared memory also
// reconfigures the SMs: ~0.3 ms for the fused kernel, which would land inside the timed span of the FIRST run of every
// process (a whole cfg1 run takes 0.2 ms).  Like the allocations, it belongs to picb200_create: the kernels of the
// configured path are launched once with nothing to do (zero Poisson steps / an empty population).
int Sim::warm_kernels() {
    if (getenv("PICB200_NO_WARM")) return 0;
    k_clear_box<<<1, 32, 0, stream>>>(st, grid[2], 2);                      // (all boxes are empty)
    k_deposit<<<1, 256, 0, stream>>>(soa, st, grid[0], 0, tables + kTableSmemBytes / 4);    // n == 0
    k_mobility_begin<<<1, 1, 0, stream>>>(st, cfg.capacity, -1);
    if (fused && cfg.nranks == 1) {
        const bool adopt = cfg.poisson_timestep >= (getenv("PICB200_ADOPT_STEPS") ? atoi(getenv("PICB200_ADOPT_STEPS")) : kAdoptMinSteps);
        const bool streaming = !adopt && mover_threads == 1024 && cfg.poisson_timestep <= (getenv("PICB200_STREAM_STEPS") ? atoi(getenv("PICB200_STREAM_STEPS")) : kStreamMaxSteps);
        const int variants[2] = {adopt ? 1 : (streaming ? 2 : 0), 0};        // (the plain one serves uploads that bring their own timestamps)
        for (int k = 0; k < (variants[0] == 0 ? 1 : 2); ++k) {
            FusedArgs a = fused_args(0, false);
            if (launch_fused(a, false, variants[k])) return 1;
        }
    }
    if (step_graph) CK(cudaGraphLaunch(step_graph, stream), "Warm-up");     // modes 31/32: one Poisson step over the empty population
    CK(cudaStreamSynchronize(stream), "Warm-up");
    CK_LAST("Warm-up");
    return reset_state();                                                   // (the counters were touched)
}

// What one rank tells the others (picb200_peer_export -> picb200_peer_connect): where its two local
// summed grids and its sync block (arrival counters, union-box words) live.  Same process (one host thread per GPU): raw pointers + peer access.
// Other process (one process per GPU, e.g. torchrun): CUDA IPC handles.
struct PeerHandle {
    unsigned magic;
    int device;
    long long pid;
    void* raw[3];                    // summed grid 0, summed grid 1, sync block
    int sm_count;                    // blocks of the persistent kernel: every rank counts nranks x blocks arrivals per exchange
    int nvls;                        // this rank's summed grids can be bound to a multicast object (pic_nvls.h)
    cudaIpcMemHandle_t ipc_sync;
    union {
        cudaIpcMemHandle_t ipc_grid[2];          // !nvls: the summed grids as CUDA IPC handles
        struct {                                 // nvls, rank 0: the multicast object and the set-up barrier
            unsigned long long mc;               //   the handle itself (ranks inside rank 0's process)
            int fd;                              //   its POSIX file descriptor in process `pid` (other processes: pidfd_getfd)
            char shm[40];
        } nv;
    };
};
static_assert(sizeof(PeerHandle) <= PICB200_PEER_HANDLE_BYTES, "handle fits the ABI blob");

#define CKD(call, where)                                                                 \
    do {                                                                                 \
        CUresult _r = (call);                                                            \
        if (_r != CUDA_SUCCESS) return fail(drv_error(d, _r) + " (" #call ")", where);   \
    } while (0)

static constexpr size_t kGridBytes = (size_t)kGrid * kGrid * kGrid * 4;

// grids 0 and 1 as one 1 GiB virtual-memory allocation, when the device can join a multicast object; otherwise nothing
// happens and the caller falls back to cudaMalloc
int Sim::nvls_alloc_grids() {
    DrvApi* d = drv_api();
    if (!d) return 0;
    CUdevice dev;
    if (d->cuDeviceGet_(&dev, device) != CUDA_SUCCESS) return 0;
    int ok = 0;
    if (d->cuDeviceGetAttribute_(&ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS || !ok) return 0;
    nvls_bytes = 2 * kGridBytes;
    CUmemAllocationProp ap;
    memset(&ap, 0, sizeof ap);
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = device;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CKD(d->cuMemCreate_(&nvls_mem, nvls_bytes, &ap, 0), "Grid alloc (NVLS)");
    nvls = true;
    CKD(d->cuMemAddressReserve_(&nvls_uc, nvls_bytes, kGridBytes, 0, 0), "Grid alloc (NVLS)");
    CKD(d->cuMemMap_(nvls_uc, nvls_bytes, 0, nvls_mem, 0), "Grid alloc (NVLS)");
    CUmemAccessDesc ad;
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = device;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CKD(d->cuMemSetAccess_(nvls_uc, nvls_bytes, &ad, 1), "Grid alloc (NVLS)");
    grid[0] = reinterpret_cast<int*>(nvls_uc);
    grid[1] = reinterpret_cast<int*>(nvls_uc + kGridBytes);
    return 0;
}

int Sim::peer_export(void* out) {
    if (cfg.nranks < 2 || !psync) return fail("peer export needs nranks > 1", "Peer setup");
    CK(cudaSetDevice(device), "Set device");
    PeerHandle h;
    memset(&h, 0, sizeof(h));
    h.magic = 0x50494342u;
    h.device = device;
    h.pid = (long long)getpid();
    h.raw[0] = grid[0]; h.raw[1] = grid[1]; h.raw[2] = psync;
    h.sm_count = sm_count;
    h.nvls = nvls ? 1 : 0;
    CK(cudaIpcGetMemHandle(&h.ipc_sync, psync), "Peer setup (cudaIpcGetMemHandle)");
    if (!nvls) {
        for (int k = 0; k < 2; ++k) CK(cudaIpcGetMemHandle(&h.ipc_grid[k], h.raw[k]), "Peer setup (cudaIpcGetMemHandle)");
    } else if (cfg.rank == 0) {
        DrvApi* d = drv_api();
        if (!nvls_mc) {
            CUmulticastObjectProp mp;
            memset(&mp, 0, sizeof mp);
            mp.numDevices = (unsigned)cfg.nranks;
            mp.size = nvls_bytes;
            mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
            CKD(d->cuMulticastCreate_(&nvls_mc, &mp), "Peer setup (multicast object)");
            nvls_mc_owner = true;
            CKD(d->cuMemExportToShareableHandle_(&nvls_fd, nvls_mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "Peer setup (multicast object)");
            static std::atomic<int> serial{0};
            snprintf(nvls_shm, sizeof nvls_shm, "/picb200-%lld-%d", h.pid, serial.fetch_add(1));
            shm_unlink(nvls_shm);
            ShmBarrier* b = shm_barrier_open(nvls_shm, true);
            if (!b) return fail("cannot create the shared-memory segment of the set-up barrier", "Peer setup");
            munmap(b, 4096);
        }
        h.nv.mc = (unsigned long long)nvls_mc;
        h.nv.fd = nvls_fd;
        memcpy(h.nv.shm, nvls_shm, sizeof h.nv.shm);
    }
    memset(out, 0, PICB200_PEER_HANDLE_BYTES);
    memcpy(out, &h, sizeof(h));
    return 0;
}

// Collective: every rank of the job calls it at the same time (one host thread or process per rank).
int Sim::peer_connect(const void* handles) {
    if (cfg.nranks < 2 || !psync) return fail("peer connect needs nranks > 1", "Peer setup");
    CK(cudaSetDevice(device), "Set device");
    memset(&pa, 0, sizeof(pa));
    pa.rank = cfg.rank;
    pa.nranks = cfg.nranks;
    pa.pl = pl;
    PeerHandle h0;
    memset(&h0, 0, sizeof h0);
    for (int r = 0; r < cfg.nranks; ++r) {
        PeerHandle h;
        memcpy(&h, static_cast<const unsigned char*>(handles) + (size_t)r * PICB200_PEER_HANDLE_BYTES, sizeof(h));
        if (h.magic != 0x50494342u) return fail("bad peer handle", "Peer setup");
        if (h.sm_count != sm_count) return fail("the ranks' GPUs have different SM counts (the exchange counts one arrival per block of every rank)", "Peer setup");
        if ((h.nvls != 0) != nvls) return fail("the ranks disagree about the NVSwitch reduction (set PICB200_NVLS=0 on every rank)", "Peer setup");
        if (r == 0) h0 = h;
        void* p[3] = {nullptr, nullptr, nullptr};
        if (r == cfg.rank) {
            p[0] = grid[0]; p[1] = grid[1]; p[2] = psync;
        } else if (h.pid == (long long)getpid()) {
            int can = 0;
            CK(cudaDeviceCanAccessPeer(&can, device, h.device), "Peer setup");
            if (!can) return fail("GPUs cannot access each other's memory", "Peer setup");
            cudaError_t e = cudaDeviceEnablePeerAccess(h.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(cudaGetErrorString(e), "Peer setup (enable peer access)");
            cudaGetLastError();
            p[2] = h.raw[2];
            if (!nvls) { p[0] = h.raw[0]; p[1] = h.raw[1]; }
        } else {
            CK(cudaIpcOpenMemHandle(&p[2], h.ipc_sync, cudaIpcMemLazyEnablePeerAccess), "Peer setup (cudaIpcOpenMemHandle)");
            ipc_opened.push_back(p[2]);
            for (int k = 0; k < 2 && !nvls; ++k) {
                CK(cudaIpcOpenMemHandle(&p[k], h.ipc_grid[k], cudaIpcMemLazyEnablePeerAccess), "Peer setup (cudaIpcOpenMemHandle)");
                ipc_opened.push_back(p[k]);
            }
        }
        pa.sum[r][0] = static_cast<int*>(p[0]);          // nvls: the peers' grids are reached through the multicast mapping only
        pa.sum[r][1] = static_cast<int*>(p[1]);
        pa.sync[r] = static_cast<PeerSync*>(p[2]);
    }
    // connect before the first population is created (the grids are still all-zero then)
    if (charge_valid) return fail("picb200_peer_connect must precede the first Poisson step", "Peer setup");
    if (nvls && !nvls_bound) {
        DrvApi* d = drv_api();
        // a rank that cannot go on says so in the shared segment, so that the others give up at once instead of waiting
        ShmBarrier* bar = shm_barrier_open(h0.nv.shm, false);
        if (!bar) return fail("cannot open the shared-memory segment of the set-up barrier", "Peer setup");
        auto give_up = [&](const std::string& what, const char* where) { shm_barrier_fail(bar); munmap(bar, 4096); return fail(what, where); };
        CUdevice dev;
        CUresult r = d->cuDeviceGet_(&dev, device);
        if (r != CUDA_SUCCESS) return give_up(drv_error(d, r) + " (cuDeviceGet)", "Peer setup (multicast object)");
        if (const char* tf = getenv("PICB200_NVLS_TEST_FAIL"))              // tests: this rank's set-up fails (the callers' fall-back path)
            if (tf[0] && atoi(tf) == cfg.rank) return give_up("PICB200_NVLS_TEST_FAIL", "Peer setup (multicast object)");
        if (cfg.rank != 0) {
            if (h0.pid == (long long)getpid()) {
                nvls_mc = (CUmemGenericAllocationHandle)h0.nv.mc;          // (released by rank 0's Sim)
            } else {
                const int fd = steal_fd(h0.pid, h0.nv.fd);
                if (fd < 0) return give_up("cannot duplicate the multicast object's file descriptor out of rank 0's process (pidfd_getfd)", "Peer setup");
                r = d->cuMemImportFromShareableHandle_(&nvls_mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
                close(fd);
                if (r != CUDA_SUCCESS) return give_up(drv_error(d, r) + " (cuMemImportFromShareableHandle)", "Peer setup (multicast object)");
                nvls_mc_owner = true;
            }
        }
        r = d->cuMulticastAddDevice_(nvls_mc, dev);
        if (r != CUDA_SUCCESS) return give_up(drv_error(d, r) + " (cuMulticastAddDevice)", "Peer setup (multicast object)");
        if (!shm_barrier_wait(bar, 0, cfg.nranks)) return give_up("not every rank reached picb200_peer_connect with a multicast object (the call is collective)", "Peer setup");
        r = d->cuMulticastBindMem_(nvls_mc, 0, nvls_mem, 0, nvls_bytes, 0);
        if (r == CUDA_SUCCESS) r = d->cuMemAddressReserve_(&nvls_mv, nvls_bytes, kGridBytes, 0, 0);
        if (r == CUDA_SUCCESS) r = d->cuMemMap_(nvls_mv, nvls_bytes, 0, nvls_mc, 0);
        CUmemAccessDesc ad;
        ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ad.location.id = device;
        ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        if (r == CUDA_SUCCESS) r = d->cuMemSetAccess_(nvls_mv, nvls_bytes, &ad, 1);
        if (r != CUDA_SUCCESS) return give_up(drv_error(d, r) + " (binding / mapping the multicast object)", "Peer setup (multicast bind)");
        if (!shm_barrier_wait(bar, 1, cfg.nranks)) return give_up("not every rank bound its grids to the multicast object", "Peer setup");
        munmap(bar, 4096);
        if (cfg.rank == 0) {
            shm_unlink(nvls_shm); nvls_shm[0] = 0;
            close(nvls_fd); nvls_fd = -1;
        } else if (!nvls_mc_owner) {
            nvls_mc = 0;                                                    // rank 0's handle: mapped here, released there
        }
        nvls_bound = true;
    }
    if (nvls) {
        pa.mc[0] = reinterpret_cast<int*>(nvls_mv);
        pa.mc[1] = reinterpret_cast<int*>(nvls_mv + kGridBytes);
    }
    pa.debug = getenv("PICB200_PEER_DEBUG") ? atoi(getenv("PICB200_PEER_DEBUG")) : 0;
    pa.strict = (getenv("PICB200_PEER_FENCE") && !strcmp(getenv("PICB200_PEER_FENCE"), "sys")) ? 1 : 0;
    peer_connected = true;
    return 0;
}

int Sim::sync_state() {
    // a kernel that died makes whichever call comes next fail: say which barrier gave up, if one did (barrier_timeout_abort)
    cudaError_t e = cudaMemcpyAsync(h_st, st, sizeof(DevState), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess && pl) e = cudaMemcpyAsync(h_pl, pl, sizeof(PeerLocal), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) {
        const std::string why = abort_message(device);
        return fail(why.empty() ? std::string(cudaGetErrorString(e)) : why + " [" + cudaGetErrorString(e) + "]", why.empty() ? "State readback" : "Mobility steps");
    }
    cur = h_st->cur;
    if (h_st->error == 2) return fail("more children waited for a compaction than the parking area holds", "Mobility phase");
    if (h_st->error == 3) return fail("a spilled child found no slot below the capacity", "Mobility phase");
    if (h_st->error == 4) return fail("the streaming step loop met a timestamp other than -1 / DEAD", "Mobility phase");
    return collect_spans();
}

// A fresh population starts a fresh run: counters, Poisson index and totals go back to zero.
int Sim::reset_state() {
    // a deposit of the previous population may still sit in either grid: both boxes bound it
    // peer mode: grids 0/1 are the summed grids the peers add into.  They are never touched from the host:
    // the one that will receive the next exchange is already empty, the other one is emptied by the first
    // Poisson step after its gathers (see pic_peer.cuh); only the local grid 2 is this rank's alone.
    for (int g = peer_connected ? 2 : 0; g < kGrids; ++g) k_clear_box<<<aux_blocks, 256, 0, stream>>>(st, grid[g], g);
    CK(cudaMemsetAsync(st, 0, sizeof(DevState), stream), "Reset state");
    k_reset_bbox<<<1, 32, 0, stream>>>(st, kGrids);
    CK_LAST("Reset state");
    gpu_ms = 0;
    cur = 0;
    charge_valid = false;
    return 0;
}

int Sim::init_uniform(int init_n, unsigned long long id_base) {
    if (init_n < 0 || init_n > cfg.capacity) return fail("init n exceeds capacity", "Setup particles");
    CK(cudaSetDevice(device), "Set device");
    if (reset_state()) return 1;
    k_init_uniform<<<std::max(1, std::min(aux_blocks, (init_n + 255) / 256)), 256, 0, stream>>>(soa, st, init_n, id_base, seed);
    CK_LAST("Setup particles");
    return 0;
}

// ---- grid phase (src/pic.cu:501-504) ---------------------------------------------------------
int Sim::op_deposit() {
    span_begin(kPhDeposit);
    k_deposit<<<deposit_blocks, 256, 0, stream>>>(soa, st, grid[cur], cur, tables + kTableSmemBytes / 4);
    charge_valid = true;
    ++launches;
    span_end();
    CK_LAST("Deposit");
    if (cfg.nranks > 1) {
        span_begin(kPhCollective);
        const int rc = allreduce_charge();
        span_end();
        return rc;
    }
    return 0;
}
int Sim::op_gather() {
    span_begin(kPhGather);
    // dense plasma (decided on the device): the field once per cell of the occupied box, then one record per electron
    float4* const F = (cfg.nranks == 1 && field_launch) ? field : nullptr;
    if (F) { k_field_box<<<aux_blocks, 256, 0, stream>>>(st, grid[cur], F, efc, cur, field_ratio); ++launches; }
    if (F) k_gather<true><<<gather_blocks[1], 256, 0, stream>>>(soa, st, grid[cur], efc, tables + kTableSmemBytes / 4, F, cur, field_ratio);
    else k_gather<false><<<gather_blocks[0], 256, 0, stream>>>(soa, st, grid[cur], efc, tables + kTableSmemBytes / 4, nullptr, cur, field_ratio);
    ++launches;
    span_end();
    CK_LAST("Gather");
    return 0;
}
int Sim::op_clear(bool reset_box) {       // reset_box false: the caller's op_mobility_begin(cur) resets the box
    span_begin(kPhClear);
    k_clear_box<<<aux_blocks, 256, 0, stream>>>(st, grid[cur], cur);
    if (reset_box) k_reset_bbox<<<1, 32, 0, stream>>>(st, cur);
    charge_valid = false;
    launches += reset_box ? 2 : 1;
    span_end();
    CK_LAST("Clear grid");
    return 0;
}

// Sum the charge boxes of all ranks: agree on the union bounding box (max-reduce of
// {-lo, hi}), pack it contiguously, one integer ncclAllReduce over NVLink, unpack.
// Integer sums are order-independent, so every rank holds bit-identical counts.
int Sim::allreduce_charge() {
    if (!comm) return fail("this multi-GPU phase needs the NCCL path: pass nccl_unique_id at creation", "Charge all-reduce");
    k_bbox_encode<<<1, 32, 0, stream>>>(st, d_bbox_red, cur);
    launches += 4;   // encode, decode, pack, unpack (NCCL's own kernels are not counted)
    int r = nccl->allReduce(d_bbox_red, d_bbox_red, 8, kNcclInt32, kNcclMax, comm, stream);
    if (r != 0) return fail(std::string("ncclAllReduce: ") + nccl->getErrorString(r), "Charge all-reduce");
    k_bbox_decode<<<1, 32, 0, stream>>>(st, d_bbox_red, cur);
    CK(cudaMemcpyAsync(h_bbox_red, d_bbox_red, 8 * sizeof(int), cudaMemcpyDeviceToHost, stream), "Charge all-reduce");
    CK(cudaStreamSynchronize(stream), "Charge all-reduce");
    const int lx = -h_bbox_red[0], ly = -h_bbox_red[1], lz = -h_bbox_red[2];
    const int wx = h_bbox_red[3] - lx + 1, wy = h_bbox_red[4] - ly + 1, wz = h_bbox_red[5] - lz + 1;
    if (wx <= 0 || wy <= 0 || wz <= 0) return 0;
    const size_t cells = (size_t)wx * wy * wz;
    if (cells > box_buf_cells) {
        cudaFree(box_buf);
        box_buf = nullptr;
        box_buf_cells = cells + cells / 4;
        CK(cudaMalloc(&box_buf, box_buf_cells * 4), "Charge all-reduce");
    }
    k_pack_box<<<aux_blocks, 256, 0, stream>>>(grid[cur], box_buf, lx, ly, lz, wx, wy, wz, 0);
    r = nccl->allReduce(box_buf, box_buf, cells, kNcclInt32, kNcclSum, comm, stream);
    if (r != 0) return fail(std::string("ncclAllReduce: ") + nccl->getErrorString(r), "Charge all-reduce");
    k_pack_box<<<aux_blocks, 256, 0, stream>>>(grid[cur], box_buf, lx, ly, lz, wx, wy, wz, 1);
    CK_LAST("Charge all-reduce");
    return 0;
}

// ---- mobility phase (src/pic.cu:509-537) -----------------------------------------------------
int Sim::op_mobility_begin(int reset_box) {
    k_mobility_begin<<<1, 1, 0, stream>>>(st, cfg.capacity, reset_box);
    ++launches;
    CK_LAST("Mobility steps");
    return 0;
}

int Sim::launch_mover_window(int lo, int hi, int t_from, int t_to, bool stepwise) {
    if (hi >= 0 && hi <= lo) return 0;
    MoverArgs a{soa, st, tables, cfg.capacity, cfg.poisson_timestep, lo, hi, t_from, t_to, stepwise ? 1 : 0, seed};
    // hi < 0: the window ends at DevState::win_hi, which the host does not know: a full grid, grid-stride loop
    const int blocks = hi < 0 ? sm_count : std::max(1, std::min(sm_count, (hi - lo + mover_threads - 1) / mover_threads));
    span_begin(kPhMover);
    if (stepwise) k_mover<false, true><<<blocks, mover_threads, kMoverSmemBytes(mover_threads), stream>>>(a);
    else k_mover<false><<<blocks, mover_threads, kMoverSmemBytes(mover_threads), stream>>>(a);
    ++launches; ++mover_launches;
    span_end();
    CK_LAST("Mobility steps");
    return 0;
}

int Sim::op_mobility() {
    const int pt = cfg.poisson_timestep;
    switch (cfg.mode) {
        case PICB200_MODE_DYNAMIC:
        case PICB200_MODE_DYNAMIC_OLD: {   // one cooperative launch; generations advance on the device
            MoverArgs a{soa, st, tables, cfg.capacity, pt, 0, 0, 1, pt, 0, seed};
            void* args[] = {&a};
            span_begin(kPhMover);
            CK(cudaLaunchCooperativeKernel((const void*)k_mover<true>, dim3((unsigned)sm_count), dim3((unsigned)mover_threads),
                                           args, (size_t)kMoverSmemBytes(mover_threads), stream),
               "Mobility steps");
            ++launches; ++mover_launches;
            span_end();
            break;
        }
        case PICB200_MODE_CPU_SYNC: {      // src/pic.cu:514-523: relaunch over the new index range until n stops growing
            if (device_windows()) {
                // a child of generation g is born at step g or later: poisson_timestep + 1 passes always suffice, and
                // the passes that find no children return at once.  No host read, the step can be replayed as a graph.
                for (int g = 0; g <= pt; ++g) {
                    if (g) { k_freeze_window<<<1, 1, 0, stream>>>(st, cfg.capacity, 1); ++launches; }     // (the first window: k_mobility_begin)
                    if (launch_mover_window(-1, -1, 1, pt, false)) return 1;
                }
                break;
            }
            if (sync_state()) return 1;
            int last_n = 0, n_host = std::min(h_st->n_alloc, cfg.capacity);
            while (n_host != last_n) {
                if (launch_mover_window(last_n, n_host, 1, pt, false)) return 1;
                last_n = n_host;
                if (sync_state()) return 1;
                n_host = std::min(h_st->n_alloc, cfg.capacity);
            }
            break;
        }
        case PICB200_MODE_NAIVE: {         // src/pic.cu:524-532: one launch per mobility step
            // The reference copies n to the host before every launch; here the window of a step is frozen on the
            // device (k_freeze_window) and the host only looks when a logger wants the state.
            for (int t = 1; t <= pt; ++t) {
                if (t > 1) { k_freeze_window<<<1, 1, 0, stream>>>(st, cfg.capacity, 0); ++launches; }      // (the first window: k_mobility_begin)
                if (launch_mover_window(0, -1, t, t, true)) return 1;
                if (naive_hook) {
                    if (sync_state()) return 1;
                    naive_hook(this, naive_hook_user);
                }
            }
            break;
        }
    }
    return 0;
}

int Sim::op_compact() {
    span_begin(kPhCompact);
    k_compact_scan<<<aux_blocks, 256, 0, stream>>>(soa, st, cfg.capacity, holes, movers);
    k_compact_move<<<aux_blocks, 256, 0, stream>>>(soa, st, holes, movers);
    launches += 2;
    span_end();
    CK_LAST("Remove dead");
    return 0;
}
int Sim::op_step_end() {
    k_step_end<<<1, 1, 0, stream>>>(st, cfg.capacity, log, log_len, cfg.nranks == 1 ? 1 : 0);
    ++launches;
    CK_LAST("Remove dead");
    return 0;
}

// Modes 30/33: `steps` iterations of runPIC's loop (src/pic.cu:487-560) in one cooperative launch.
// Several ranks: one step per launch, with the charge all-reduce in between.
FusedArgs Sim::fused_args(int per_launch, bool peer) {
    return FusedArgs{soa, st, tables, {grid[0], grid[1], grid[2]}, holes, movers, log, log_len, cfg.capacity, cfg.poisson_timestep,
                     per_launch, cfg.nranks == 1 ? 1 : 0, efc, seed, nullptr, pa, (peer && !charge_valid) ? 1 : 0, park, park_cap, getenv("PICB200_ADOPT_MIN") ? atoi(getenv("PICB200_ADOPT_MIN")) : kAdoptMin, compact_policy,
                     std::max(1, std::min(kSS, getenv("PICB200_DRAIN_AT") ? atoi(getenv("PICB200_DRAIN_AT")) : kFusedDrainAt)),
                     (peer || cfg.nranks == 1) ? field : nullptr, field_ratio, field_min_n, getenv("PICB200_TEST_DESERT") ? 1 : 0, philox_round_keys(seed), (getenv("PICB200_WARP_FENCE") && atoi(getenv("PICB200_WARP_FENCE")) == 0) ? 0 : 1,
                     (getenv("PICB200_BALANCE") && atoi(getenv("PICB200_BALANCE")) == 0) ? 0 : 1};
}
// variant: 0 plain, 1 adopting, 2 streaming step loop (k_poisson_fused)
int Sim::launch_fused(FusedArgs& a, bool peer, int variant) {
    static const void* const kernels[2][3] = {{(const void*)k_poisson_fused<false, 0>, (const void*)k_poisson_fused<false, 1>, (const void*)k_poisson_fused<false, 2>},
                                              {(const void*)k_poisson_fused<true, 0>, (const void*)k_poisson_fused<true, 1>, (const void*)k_poisson_fused<true, 2>}};
    void* args[] = {&a};
    CK(cudaLaunchCooperativeKernel(kernels[peer ? 1 : 0][variant], dim3((unsigned)sm_count), dim3((unsigned)mover_threads), args,
                                   variant == 2 ? (size_t)kFusedSmemBytes(mover_threads, true) : (size_t)kFusedSmemBytes(mover_threads) + fused_smem_pad, stream),
       "Mobility steps");
    return 0;
}

int Sim::fused_steps(int steps) {
    const bool peer = cfg.nranks > 1 && peer_connected;      // exchange inside the kernel, whole run in one launch
    if (cfg.nranks > 1 && !peer && !comm)
        return fail("several ranks need picb200_peer_connect() or an nccl_unique_id", "Mobility steps");
    if (!charge_valid && !peer) {           // first step after init/upload: grid[cur] needs this population's deposit
        if (op_clear() || op_deposit()) return 1;
    }
    const int per_launch = (cfg.nranks > 1 && !peer) ? 1 : steps;
    for (int done = 0; done < steps; done += per_launch) {
        if (op_mobility_begin()) return 1;
        FusedArgs a = fused_args(per_launch, peer);
        charge_valid = true;
        // PICB200_TRACE=<file>: dump 16 globaltimer stamps per block per Poisson step of this launch (debugging aid)
        const char* trace_path = getenv("PICB200_TRACE");
        const size_t trace_words = (size_t)per_launch * (size_t)sm_count * 16;
        if (trace_path) {
            CK(cudaMalloc(&a.trace, trace_words * 8), "Trace alloc");
            CK(cudaMemsetAsync(a.trace, 0, trace_words * 8, stream), "Trace alloc");
        }
        span_begin(kPhMover);
        // which step loop: adopting from kAdoptMinSteps mobility steps per Poisson step, streaming up to kStreamMaxSteps (HBM-side)
        const bool adopt = cfg.poisson_timestep >= (getenv("PICB200_ADOPT_STEPS") ? atoi(getenv("PICB200_ADOPT_STEPS")) : kAdoptMinSteps);
        const bool streaming = !adopt && !generic_ts && mover_threads == 1024 && cfg.poisson_timestep <= (getenv("PICB200_STREAM_STEPS") ? atoi(getenv("PICB200_STREAM_STEPS")) : kStreamMaxSteps);
        if (launch_fused(a, peer, adopt ? 1 : (streaming ? 2 : 0))) return 1;
        ++launches; ++mover_launches;
        generic_ts = false;              // a launch leaves every live electron at timestamp -1
        span_end();
        if (trace_path) {
            std::vector<unsigned long long> h(trace_words);
            CK(cudaMemcpyAsync(h.data(), a.trace, trace_words * 8, cudaMemcpyDeviceToHost, stream), "Trace readback");
            CK(cudaStreamSynchronize(stream), "Trace readback");
            cudaFree(a.trace);
            const std::string tp = cfg.nranks > 1 ? std::string(trace_path) + "." + std::to_string(cfg.rank) : std::string(trace_path);
            if (FILE* f = fopen(tp.c_str(), "wb")) {
                const int hdr[2] = {per_launch, sm_count};
                fwrite(hdr, sizeof(int), 2, f);
                fwrite(h.data(), 8, trace_words, f);
                fclose(f);
            }
        }
        if (cfg.nranks > 1 && !peer) {      // NCCL path: the kernel deposited this rank's survivors into the next grid
            cur = (cur + 1) % kRot;
            span_begin(kPhCollective);
            const int rc = allreduce_charge();
            span_end();
            if (rc) return rc;
        }
    }
    return 0;
}

// One iteration of runPIC's loop (src/pic.cu:487-560).
int Sim::poisson_step(bool sync) {
    CK(cudaSetDevice(device), "Set device");
    if (fused) {
        if (fused_steps(1)) return 1;
        return sync ? sync_state() : 0;
    }
    if (step_graph && !profiling && !naive_hook) {         // mode 32: replay the captured Poisson step
        CK(cudaGraphLaunch(step_graph, stream), "Mobility steps");
        launches += graph_launches; mover_launches += graph_mover_launches;
        charge_valid = true;
        return sync ? sync_state() : 0;
    }
    if (op_deposit() || op_gather() || op_clear(false) || op_mobility_begin(cur) || op_mobility() || op_compact() || op_step_end()) return 1;
    if (sync) return sync_state();
    return 0;
}

// Capture one Poisson step of mode 32 into a CUDA graph (called once from create(): instantiation costs tens of
// milliseconds the first time).  Nothing is executed and no counter is touched.
int Sim::build_step_graph() {
    if (!(cfg.mode == PICB200_MODE_NAIVE || (cfg.mode == PICB200_MODE_CPU_SYNC && device_windows())) || cfg.nranks != 1 || fused || getenv("PICB200_NO_GRAPH")) return 0;
    const long long l0 = launches, m0 = mover_launches;
    const bool cv = charge_valid;
    cudaGraph_t captured = nullptr;
    CK(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal), "Graph capture");
    const int rc = op_deposit() || op_gather() || op_clear(false) || op_mobility_begin(cur) || op_mobility() || op_compact() || op_step_end();
    const cudaError_t e = cudaStreamEndCapture(stream, &captured);
    graph_launches = launches - l0; graph_mover_launches = mover_launches - m0;
    launches = l0; mover_launches = m0; charge_valid = cv;
    if (rc || e != cudaSuccess) { if (captured) cudaGraphDestroy(captured); return rc ? rc : fail(cudaGetErrorString(e), "Graph capture"); }
    const cudaError_t e2 = cudaGraphInstantiate(&step_graph, captured, 0);
    cudaGraphDestroy(captured);
    if (e2 != cudaSuccess) { step_graph = nullptr; return fail(cudaGetErrorString(e2), "Graph capture"); }
    CK(cudaGraphUpload(step_graph, stream), "Graph capture");
    CK(cudaStreamSynchronize(stream), "Graph capture");
    return 0;
}

}  // namespace picb200

// =============================================================================================
// C ABI
// =============================================================================================
using namespace picb200;
struct picb200_sim { Sim s; };

extern "C" {

int picb200_abi_version(void) { return PICB200_ABI_VERSION; }
const char* picb200_last_error(void) { return g_last_error.c_str(); }
int picb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
void picb200_config_default(picb200_config* c) {
    memset(c, 0, sizeof(*c));
    c->abi_version = PICB200_ABI_VERSION;
    c->mode = PICB200_MODE_DYNAMIC;
    c->block_size = 256;
    c->capacity = 1000000;
    c->poisson_timestep = 10;
    c->poisson_steps = 1;
    c->sleep_time_ns = 100;
    c->device = -1;
    c->nranks = 1;
}

int picb200_create(const picb200_config* cfg, picb200_sim** out) {
    if (!cfg || !out) return fail("null argument", "picb200_create");
    picb200_sim* h = new picb200_sim();
    if (h->s.create(cfg)) { delete h; *out = nullptr; return 1; }
    *out = h;
    return 0;
}
void picb200_destroy(picb200_sim* sim) { delete sim; }

int picb200_init_uniform(picb200_sim* sim, int32_t init_n, uint64_t id_base) {
    if (sim->s.init_uniform(init_n, id_base)) return 1;
    return sim->s.sync_state();
}

int picb200_upload(picb200_sim* sim, int32_t n, const float* x, const float* y, const float* z, const float* vx,
                   const float* vy, const float* vz, const float* ax, const float* ay, const float* az,
                   const int32_t* ts, const uint64_t* id) {
    Sim& s = sim->s;
    if (n < 0 || n > s.cfg.capacity || !x || !y || !z) return fail("bad upload arguments", "Upload");
    CK(cudaSetDevice(s.device), "Set device");
    const float* src[9] = {x, y, z, vx, vy, vz, ax, ay, az};
    float* dst[9] = {s.soa.x, s.soa.y, s.soa.z, s.soa.vx, s.soa.vy, s.soa.vz, s.soa.ax, s.soa.ay, s.soa.az};
    for (int k = 0; k < 9; ++k) {
        if (src[k]) CK(cudaMemcpyAsync(dst[k], src[k], (size_t)n * 4, cudaMemcpyHostToDevice, s.stream), "Upload");
        else CK(cudaMemsetAsync(dst[k], 0, (size_t)n * 4, s.stream), "Upload");
    }
    std::vector<int32_t> ts_def;
    std::vector<uint64_t> id_def;
    s.generic_ts = ts != nullptr;
    if (!ts) { ts_def.assign((size_t)n, -1); ts = ts_def.data(); }
    if (!id) { id_def.resize((size_t)n); for (int i = 0; i < n; ++i) id_def[(size_t)i] = (uint64_t)i; id = id_def.data(); }
    CK(cudaMemcpyAsync(s.soa.ts, ts, (size_t)n * 4, cudaMemcpyHostToDevice, s.stream), "Upload");
    CK(cudaMemcpyAsync(s.soa.id, id, (size_t)n * 8, cudaMemcpyHostToDevice, s.stream), "Upload");
    if (s.reset_state()) return 1;
    k_set_n<<<1, 1, 0, s.stream>>>(s.st, n);
    CK_LAST("Upload");
    return s.sync_state();
}

int picb200_download(picb200_sim* sim, int32_t* n_out, float* x, float* y, float* z, float* vx, float* vy, float* vz,
                     float* ax, float* ay, float* az, int32_t* ts, uint64_t* id) {
    Sim& s = sim->s;
    CK(cudaSetDevice(s.device), "Set device");
    if (s.sync_state()) return 1;
    const int n = std::min(std::max(s.h_st->n, s.h_st->n_alloc), s.cfg.capacity);
    float* dst[9] = {x, y, z, vx, vy, vz, ax, ay, az};
    float* src[9] = {s.soa.x, s.soa.y, s.soa.z, s.soa.vx, s.soa.vy, s.soa.vz, s.soa.ax, s.soa.ay, s.soa.az};
    for (int k = 0; k < 9; ++k)
        if (dst[k]) CK(cudaMemcpyAsync(dst[k], src[k], (size_t)n * 4, cudaMemcpyDeviceToHost, s.stream), "Download");
    if (ts) CK(cudaMemcpyAsync(ts, s.soa.ts, (size_t)n * 4, cudaMemcpyDeviceToHost, s.stream), "Download");
    if (id) CK(cudaMemcpyAsync(id, s.soa.id, (size_t)n * 8, cudaMemcpyDeviceToHost, s.stream), "Download");
    CK(cudaStreamSynchronize(s.stream), "Download");
    if (n_out) *n_out = n;
    return 0;
}

// RunData.electrons (src/pic.cu:580-584) / log() (src/utility.cu:124-137): the population as the reference's 64-byte
// records.  The transpose runs on the device (k_export_aos) in chunks of 4M electrons through two staging buffers, so the
// copy of one chunk to the host overlaps the transpose of the next; `out` may be pageable or pinned (picb200_alloc_pinned).
int picb200_download_aos(picb200_sim* sim, picb200_electron* out, int64_t cap, int64_t* n_out) {
    Sim& s = sim->s;
    CK(cudaSetDevice(s.device), "Set device");
    if (s.sync_state()) return 1;
    const int n_all = std::min(std::max(s.h_st->n, s.h_st->n_alloc), s.cfg.capacity);
    const int n = (int)std::min<int64_t>(n_all, std::max<int64_t>(cap, 0));
    if (n_out) *n_out = n;
    if (n == 0) return 0;
    if (!out) return fail("null output buffer", "Download");
    static_assert(sizeof(AosRecord) == sizeof(picb200_electron), "one record layout");
    const int chunk = std::min(n, 1 << 22);
    AosRecord* stage[2] = {nullptr, nullptr};
    cudaEvent_t done[2] = {nullptr, nullptr};
    cudaStream_t copy = nullptr;
    cudaError_t e = cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking);
    for (int k = 0; k < 2 && e == cudaSuccess; ++k) {
        if (k == 0 || n > chunk) e = cudaMalloc(&stage[k], (size_t)chunk * sizeof(AosRecord));
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming);
    }
    cudaEvent_t filled = nullptr;
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&filled, cudaEventDisableTiming);
    int b = 0;
    for (int first = 0; first < n && e == cudaSuccess; first += chunk, b ^= (n > chunk ? 1 : 0)) {
        const int cnt = std::min(chunk, n - first);
        e = cudaStreamWaitEvent(s.stream, done[b], 0);                 // the copy that last read this buffer (no-op the first time)
        if (e != cudaSuccess) break;
        k_export_aos<<<std::max(1, std::min(s.aux_blocks, (cnt + 255) / 256)), 256, 0, s.stream>>>(s.soa, first, cnt, stage[b]);
        ++s.launches;
        e = cudaEventRecord(filled, s.stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(copy, filled, 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(out + first, stage[b], (size_t)cnt * sizeof(AosRecord), cudaMemcpyDeviceToHost, copy);
        if (e == cudaSuccess) e = cudaEventRecord(done[b], copy);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(copy);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s.stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    for (int k = 0; k < 2; ++k) { cudaFree(stage[k]); if (done[k]) cudaEventDestroy(done[k]); }
    if (filled) cudaEventDestroy(filled);
    if (copy) cudaStreamDestroy(copy);
    CK(e, "Download");
    return 0;
}

int picb200_poisson_step(picb200_sim* sim) { return sim->s.poisson_step(true); }

int picb200_run_steps(picb200_sim* sim, int32_t steps) {
    Sim& s = sim->s;
    CK(cudaSetDevice(s.device), "Set device");
    CK(cudaEventRecord(s.ev_start, s.stream), "Setup");
    const bool host_in_loop = (s.cfg.mode == PICB200_MODE_CPU_SYNC && !s.device_windows()) || s.cfg.nranks > 1;
    if (s.fused) {
        if (s.fused_steps(steps)) return 1;
    } else {
        for (int t = 0; t < steps; ++t) {
            if (s.poisson_step(false)) return 1;
            if (host_in_loop && s.h_st->hit_zero) break;
        }
    }
    CK(cudaEventRecord(s.ev_stop, s.stream), "After sim");
    if (s.sync_state()) return 1;
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, s.ev_start, s.ev_stop), "After sim");
    s.gpu_ms += ms;
    return 0;
}

int picb200_get_summary(picb200_sim* sim, picb200_summary* out) {
    Sim& s = sim->s;
    if (s.sync_state()) return 1;
    memset(out, 0, sizeof(*out));
    out->final_n = std::min(s.h_st->n, s.cfg.capacity);
    out->total_added = s.h_st->total_added;
    out->total_removed = s.h_st->total_removed;
    out->pushes = (int64_t)s.h_st->pushes;
    out->poisson_steps_done = s.h_st->poisson_index;
    out->gpu_ms = s.gpu_ms;
    out->overflowed = s.h_st->overflow_steps;
    out->hit_zero = s.h_st->hit_zero;
    out->sm_count = s.sm_count;
    out->blocks_per_sm = s.blocks_per_sm;
    return 0;
}

int picb200_set_profiling(picb200_sim* sim, int32_t on) {
    Sim& s = sim->s;
    if (s.sync_state()) return 1;
    s.profiling = on;
    for (double& v : s.phase_ms) v = 0;
    s.launches = 0; s.mover_launches = 0; s.gpu_ms = 0;
    return 0;
}
int picb200_get_profile(picb200_sim* sim, picb200_profile* out) {
    Sim& s = sim->s;
    if (s.sync_state()) return 1;
    out->deposit_ms = s.phase_ms[Sim::kPhDeposit];
    out->gather_ms = s.phase_ms[Sim::kPhGather];
    out->clear_ms = s.phase_ms[Sim::kPhClear];
    out->mover_ms = s.phase_ms[Sim::kPhMover];
    out->compact_ms = s.phase_ms[Sim::kPhCompact];
    out->collective_ms = s.phase_ms[Sim::kPhCollective];
    out->launches = s.launches;
    out->mover_launches = s.mover_launches;
    return 0;
}
// Rewrites the whole 512 MiB charge grid with zeros (its value between Poisson steps anyway):
// 4x the 126 MB L2, so nothing of the electron store stays cached for the next timed run.
int picb200_flush_l2(picb200_sim* sim) {
    Sim& s = sim->s;
    CK(cudaSetDevice(s.device), "Set device");
    // peer mode: grids 0/1 are the summed grids the PEERS add into (never written from the host); grid 2 is this rank's alone
    CK(cudaMemsetAsync(s.grid[s.peer_connected ? 2 : (s.cur + 1) % kRot], 0, (size_t)kGrid * kGrid * kGrid * 4, s.stream), "Flush L2");
    CK(cudaStreamSynchronize(s.stream), "Flush L2");
    return 0;
}
void* picb200_alloc_pinned(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); fail("pinned allocation failed", "picb200_alloc_pinned"); return nullptr; }
    return p;
}
void picb200_free_pinned(void* p) { if (p) cudaFreeHost(p); }

int picb200_set_poisson_index(picb200_sim* sim, int32_t index) {
    Sim& s = sim->s;
    if (s.sync_state()) return 1;
    s.h_st->poisson_index = index;
    CK(cudaMemcpyAsync(s.st, s.h_st, sizeof(DevState), cudaMemcpyHostToDevice, s.stream), "Set index");
    CK(cudaStreamSynchronize(s.stream), "Set index");
    return 0;
}

int picb200_op_deposit(picb200_sim* sim) {
    Sim& s = sim->s;
    cudaSetDevice(s.device);
    if (s.sync_state()) return 1;
    if (s.charge_valid && s.op_clear()) return 1;     // e.g. after a fused step, which leaves the next deposit in place
    if (s.op_deposit()) return 1;
    return s.sync_state();
}
int picb200_op_gather(picb200_sim* sim) { Sim& s = sim->s; cudaSetDevice(s.device); if (s.op_gather()) return 1; return s.sync_state(); }
int picb200_op_clear_grid(picb200_sim* sim) { Sim& s = sim->s; cudaSetDevice(s.device); if (s.op_clear()) return 1; return s.sync_state(); }
int picb200_op_mobility(picb200_sim* sim) {
    Sim& s = sim->s;
    cudaSetDevice(s.device);
    if (s.op_mobility_begin() || s.op_mobility()) return 1;
    return s.sync_state();
}
int picb200_op_compact(picb200_sim* sim) {
    Sim& s = sim->s;
    cudaSetDevice(s.device);
    if (s.op_compact() || s.op_step_end()) return 1;
    return s.sync_state();
}

int picb200_charge_box(picb200_sim* sim, const int32_t lo[3], const int32_t hi[3], int32_t* out) {
    Sim& s = sim->s;
    cudaSetDevice(s.device);
    const int wx = hi[0] - lo[0] + 1, wy = hi[1] - lo[1] + 1, wz = hi[2] - lo[2] + 1;
    if (wx <= 0 || wy <= 0 || wz <= 0 || lo[0] < 0 || lo[1] < 0 || lo[2] < 0 || hi[0] >= kGrid || hi[1] >= kGrid || hi[2] >= kGrid)
        return fail("bad box", "Charge box");
    int* buf = nullptr;
    const size_t cells = (size_t)wx * wy * wz;
    CK(cudaMalloc(&buf, cells * 4), "Charge box");
    k_pack_box<<<s.aux_blocks, 256, 0, s.stream>>>(s.grid[s.cur], buf, lo[0], lo[1], lo[2], wx, wy, wz, 0);
    cudaError_t e = cudaMemcpyAsync(out, buf, cells * 4, cudaMemcpyDeviceToHost, s.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s.stream);
    cudaFree(buf);
    CK(e, "Charge box");
    return 0;
}
// ---- picb200_charge_box without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// k_pack_box needs a contiguous scratch buffer for the box; here that scratch is a file-scope device array of a fixed,
// modest size -- PICB200_NOALLOC_MAX_CELLS cells = 4 MiB, held by every GPU that loads the library until the process
// exits.  In host code the array's NAME is only a host-side stand-in, not a device address: its device address is asked
// for once per device, with that device current (cudaGetSymbolAddress), and that pointer is what the kernel gets.
// Every call on a device shares the array, and it keeps the last call's contents: k_pack_box overwrites every cell of the
// requested box before the copy reads it, so nothing has to start at zero -- but two calls that overlap on one GPU
// (two host threads, two streams) race on it.
}  // extern "C"
static __device__ __align__(256) int g_noalloc_box_scratch[PICB200_NOALLOC_MAX_CELLS];
static int* g_noalloc_box_ptr[64] = {nullptr};       // per device: the array's DEVICE address
extern "C" {
int picb200_charge_box_noalloc(picb200_sim* sim, const int32_t lo[3], const int32_t hi[3], int32_t* out) {
    if (!sim || !lo || !hi || !out) return fail("null argument", "Charge box (no allocation)");
    Sim& s = sim->s;
    CK(cudaSetDevice(s.device), "Set device");
    const int wx = hi[0] - lo[0] + 1, wy = hi[1] - lo[1] + 1, wz = hi[2] - lo[2] + 1;
    if (wx <= 0 || wy <= 0 || wz <= 0 || lo[0] < 0 || lo[1] < 0 || lo[2] < 0 || hi[0] >= kGrid || hi[1] >= kGrid || hi[2] >= kGrid)
        return fail("bad box", "Charge box (no allocation)");
    const size_t cells = (size_t)wx * wy * wz;
    if (cells > (size_t)PICB200_NOALLOC_MAX_CELLS)
        return fail("box of " + std::to_string(cells) + " cells exceeds the fixed scratch of " + std::to_string(PICB200_NOALLOC_MAX_CELLS) +
                    " cells (PICB200_NOALLOC_MAX_CELLS); use picb200_charge_box", "Charge box (no allocation)");
    if (s.device < 0 || s.device >= 64) return fail("device ordinal out of range", "Charge box (no allocation)");
    if (!g_noalloc_box_ptr[s.device]) {              // set-up, once per device, that device current
        int* ptr = nullptr;
        CK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_box_scratch), "Charge box (no allocation): cudaGetSymbolAddress");
        if (!ptr) return fail("no device address for the scratch array", "Charge box (no allocation)");
        g_noalloc_box_ptr[s.device] = ptr;
    }
    int* scratch = g_noalloc_box_ptr[s.device];
    k_pack_box<<<s.aux_blocks, 256, 0, s.stream>>>(s.grid[s.cur], scratch, lo[0], lo[1], lo[2], wx, wy, wz, 0);
    CK_LAST("Charge box (no allocation)");
    CK(cudaMemcpyAsync(out, scratch, cells * 4, cudaMemcpyDeviceToHost, s.stream), "Charge box (no allocation)");
    CK(cudaStreamSynchronize(s.stream), "Charge box (no allocation)");
    return 0;
}
int picb200_charge_bbox(picb200_sim* sim, int32_t lo[3], int32_t hi[3]) {
    Sim& s = sim->s;
    if (s.sync_state()) return 1;
    for (int k = 0; k < 3; ++k) { lo[k] = s.h_st->bbox_lo[s.cur][k]; hi[k] = s.h_st->bbox_hi[s.cur][k]; }
    return 0;
}

int picb200_dev_philox(int32_t device, int64_t n, const uint32_t* ctr, const uint32_t* key, uint32_t* out) {
    if (picb200_device_count() == 0) return fail("no CUDA device", "Philox test");
    if (device >= 0) CK(cudaSetDevice(device), "Set device");
    uint32_t *d_c = nullptr, *d_o = nullptr;
    CK(cudaMalloc(&d_c, (size_t)n * 16), "Philox test");
    CK(cudaMalloc(&d_o, (size_t)n * 16), "Philox test");
    CK(cudaMemcpy(d_c, ctr, (size_t)n * 16, cudaMemcpyHostToDevice), "Philox test");
    k_test_philox<<<64, 256>>>(n, d_c, make_uint2(key[0], key[1]), d_o);
    cudaError_t e = cudaMemcpy(out, d_o, (size_t)n * 16, cudaMemcpyDeviceToHost);
    cudaFree(d_c); cudaFree(d_o);
    CK(e, "Philox test");
    return 0;
}
int picb200_dev_energy_index(int32_t device, int64_t n, const float* energy, int32_t* out) {
    if (picb200_device_count() == 0) return fail("no CUDA device", "Energy index test");
    if (device >= 0) CK(cudaSetDevice(device), "Set device");
    std::vector<unsigned char> img((size_t)kTableSmemBytes, 0);
    build_energy_edges(reinterpret_cast<float*>(img.data() + kThrBytes));
    float *d_t = nullptr, *d_e = nullptr;
    int* d_o = nullptr;
    CK(cudaMalloc(&d_t, (size_t)kTableSmemBytes), "Energy index test");
    CK(cudaMalloc(&d_e, (size_t)n * 4), "Energy index test");
    CK(cudaMalloc(&d_o, (size_t)n * 4), "Energy index test");
    CK(cudaMemcpy(d_t, img.data(), (size_t)kTableSmemBytes, cudaMemcpyHostToDevice), "Energy index test");
    CK(cudaMemcpy(d_e, energy, (size_t)n * 4, cudaMemcpyHostToDevice), "Energy index test");
    k_test_energy_index<<<256, 256>>>(n, d_e, d_t, d_o);
    cudaError_t e = cudaMemcpy(out, d_o, (size_t)n * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_t); cudaFree(d_e); cudaFree(d_o);
    CK(e, "Energy index test");
    return 0;
}
int picb200_dev_energy_index_sweep(int32_t device, uint32_t stride, int64_t* mismatches, uint32_t* first_bad_bits) {
    if (picb200_device_count() == 0) return fail("no CUDA device", "Energy index sweep");
    if (device >= 0) CK(cudaSetDevice(device), "Set device");
    std::vector<unsigned char> img((size_t)kTableSmemBytes, 0);
    build_energy_edges(reinterpret_cast<float*>(img.data() + kThrBytes));
    float* d_t = nullptr;
    unsigned long long* d_o = nullptr;
    unsigned long long h[2] = {0, 0};
    CK(cudaMalloc(&d_t, (size_t)kTableSmemBytes), "Energy index sweep");
    CK(cudaMalloc(&d_o, sizeof(h)), "Energy index sweep");
    CK(cudaMemcpy(d_t, img.data(), (size_t)kTableSmemBytes, cudaMemcpyHostToDevice), "Energy index sweep");
    CK(cudaMemset(d_o, 0, sizeof(h)), "Energy index sweep");
    k_sweep_energy_index<<<148 * 8, 256>>>(stride ? stride : 1u, d_t, d_o);
    cudaError_t e = cudaMemcpy(h, d_o, sizeof(h), cudaMemcpyDeviceToHost);
    cudaFree(d_t); cudaFree(d_o);
    CK(e, "Energy index sweep");
    if (mismatches) *mismatches = (int64_t)h[0];
    if (first_bad_bits) *first_bad_bits = (uint32_t)h[1];
    return 0;
}
int picb200_load_table(const char* path, float* rows2) {
    std::vector<float> r;
    int n = load_table_file(path, r);
    memcpy(rows2, r.data(), sizeof(float) * 2 * (size_t)kRows);
    return n;
}
// host-side copy of the edge table, for the exhaustive CPU check of the device formula
void picb200_host_energy_edges(float* out_edges) { build_energy_edges(out_edges); }
// the collider's threshold on float(w) for one chance (draw_threshold)
float picb200_host_draw_threshold(float chance) { return draw_threshold(chance); }
// likewise the 513 cell-boundary coordinates cell_of_fast() probes
void picb200_host_cell_bounds(float* out_bounds) { build_cell_bounds(out_bounds); }

int picb200_peer_export(picb200_sim* sim, void* handle_out) { return sim->s.peer_export(handle_out); }
int picb200_peer_connect(picb200_sim* sim, const void* handles) { return sim->s.peer_connect(handles); }

int picb200_nccl_unique_id(void* out128) {
    NcclApi* api = load_nccl(g_last_error);
    if (!api) return 1;
    NcclUniqueId id;
    int r = api->getUniqueId(&id);
    if (r != 0) return fail(std::string("ncclGetUniqueId: ") + api->getErrorString(r), "NCCL init");
    memcpy(out128, &id, sizeof(id));
    return 0;
}

}  // extern "C"

#include "pic_state.inl"
#include "pic_run.inl"
