// pic_device.cuh -- device code of the B200-native PIC-MCC electron engine (sm_100a).
//
// Everything here is written from scratch for Blackwell; the reference lines cited are the
// *semantics* each kernel reproduces (paths relative to MagnusMouritzen/particle-simulation).
//
// Data layout in HBM: structure-of-arrays, fp32 state, one array per field, each `capacity`
// long and 256-byte aligned:  x y z | vx vy vz | ax ay az | ts (int32) | id (uint64).
// 48 B per electron, of which a mobility step must move 68 B (read 40, write 28; the id is
// bookkeeping the reference does not have).  The charge grid is 512^3 int32 (512 MiB), kept
// all-zero outside the bounding box of the last deposit, so no pass ever sweeps it whole.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace picb200 {

constexpr int kGrid = 512;              // src/cell.cu:3
constexpr int kRows = 10000;            // src/cross_section.h:9
constexpr int kDead = -2;               // src/electron.h:8
constexpr int kGrids = 3;               // charge grids in rotation (gather / next deposit / being emptied)
constexpr int kRot = 3;
constexpr int kBoxes = kGrids + 1;      // bounding-box slots in DevState
constexpr int kMaxBlocks = 160;         // blocks of the persistent kernel the quota balancer handles (5 per lane of one warp)
constexpr unsigned kFull = 0xffffffffu;
constexpr uint32_t kTagCollide = 0u, kTagInit = 1u;

// Shared-memory image of the collision tables: kRows x {T(split), T(split+remove)} -- the chances as thresholds on
// float(Philox word), draw_threshold() in pic_engine.cu -- then kRows+1 bin-edge energies (build_energy_edges()).
constexpr int kThrBytes = kRows * 8;
constexpr int kEdgeBytes = (kRows + 1) * 4;
constexpr int kTableSmemBytes = ((kThrBytes + kEdgeBytes + 15) / 16) * 16;

// The streaming instantiation of the fused kernel (small poisson_timestep, HBM-side) keeps only the first kStreamRows rows
// of the tables in shared memory -- |v|^2 < 10^(-6 + 22 * 2048 / 10000) = 0.032 m^2/s^2, sixty times what
// the BASELINE configurations reach -- and reads higher rows from the global image; the freed shared memory holds its
// chunk pipeline.  Layout of a window of R rows: thr[R] (8 B each), edge[R or kRows + 1] (4 B each), then the cell bounds.
constexpr int kStreamRows = 2048;
__host__ __device__ constexpr int smem_edge_off(int R) { return R == kRows ? kThrBytes : R * 8; }
__host__ __device__ constexpr int smem_bnd_off(int R) { return R == kRows ? kTableSmemBytes : R * 12; }

struct Soa {
    float *x, *y, *z, *vx, *vy, *vz, *ax, *ay, *az;
    int* ts;
    unsigned long long* id;
};

// Device-resident control block.  Every kernel sizes itself from here, so a whole Poisson
// step (or a whole run) is enqueued without the host ever reading a counter back
// (contrast src/pic.cu:495-497, :539, :551: three blocking copies per Poisson step).
struct DevState {
    // -- line 0: step bookkeeping (cold inside a mobility phase)
    int n;                  // live electrons = slots [0,n); all ts == -1 between Poisson steps
    int n_holes, n_movers;  // compaction work lists
    int gen_hi;             // end of the fall-back window agreed at the barrier (-1: none)
    int n_unfinished;       // children spilled to the global store before they were pushed to the end
    int bbox_lo[kBoxes][3], bbox_hi[kBoxes][3];  // cell bounding box of the counts deposited in each charge grid (peer mode: slots 2 and 3
                                                 // alternate as the box of the local deposit, slots 0/1 mirror the union boxes of S[0]/S[1])
    int cur;                // which of the charge grids holds the current deposit
    int poisson_index;      // Poisson steps completed
    int hit_zero;           // population reached 0 (src/pic.cu:556-559)
    int overflow_steps;     // steps whose n reached capacity (src/pic.cu:544)
    int last_n_after_mobility;
    int ticket;             // (spare work counter)
    long long total_added, total_removed;
    // -- lazy compaction inside the fused kernel (always back to hw == n when a launch ends)
    int hw;                 // slots in use: [0,hw) holds the n live electrons and hw-n dead ones
    int do_compact;         // this Poisson step ends with remove_dead_particles (else the holes stay)
    int scan_S, scan_D;     // its arguments: slots to look at, dead ones among them
    int parked;             // children of this step waiting for the compaction to make room (see FusedArgs::park)
    int error;              // 2: more children parked than the parking area holds, 3: a spilled child found no slot
    int win_lo, win_hi;     // modes 31/32: the index window of the next launch, kept on the device (k_freeze_window)
    // -- hot atomics, one 128-byte line each so they do not serialise behind one another in L2
    alignas(128) int n_alloc;                   // the reference's counter n during the mobility phase (may overshoot capacity)
    int top_alloc;                              // fused kernel: slots taken above the high-water mark; the two form ONE 64-bit
                                                // word so that a batch of children draws from both with a single atomic
    alignas(128) int n_dead;                    // electrons marked DEAD in this Poisson step
    alignas(128) unsigned long long pushes;     // updateParticle evaluations
    alignas(128) unsigned bar_count;            // grid barrier: arrivals
    alignas(128) unsigned bar_release;          // grid barrier: releases
    alignas(128) int n_parked;                  // parking cursor
    alignas(128) int pad_;
    // fused kernel, load balance: chunks per microsecond each block managed in its main phase (running mean), one set per
    // Poisson-step parity -- a block publishes into the set the NEXT step reads (k_poisson_fused, "quotas")
    float rate[2][kMaxBlocks];
};

struct StepLog { int n_after_mobility, n_after_compaction; };

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int ld_volatile_i32(const int* p) { return *(const volatile int*)p; }
// A word of the device state that nobody writes during THIS launch (an earlier kernel of the stream did), asked for by
// every thread: through the L1.  The volatile form sends 9472 warps to one L2 slice for the same sector -- k_clear_box
// took 16.7 us with an EMPTY box, the windowed mover launches 2 us each.
__device__ __forceinline__ int ld_state_i32(const int* p) { return __ldg(p); }

// A grid barrier of a persistent kernel that is never completed (a block left the kernel early, or the launch was
// not co-resident) must not spin for ever: after g_barrier_timeout_ns the waiting block records {code, block, barrier
// number} in page-locked HOST memory (readable after the context has failed) and traps; the host call then returns
// "CUDA error: ...\nAt: Mobility steps" (src/utility.cu:108-115) naming the barrier.  Set per device at picb200_create.
__device__ int* g_abort_flag = nullptr;
__device__ unsigned long long g_barrier_timeout_ns = 30ull * 1000ull * 1000ull * 1000ull;
__device__ __forceinline__ unsigned long long timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)::"memory");
    return t;
}
__device__ __noinline__ void barrier_timeout_abort(int code, unsigned seq) {
    int* flag = g_abort_flag;
    if (flag) {
        flag[1] = (int)blockIdx.x; flag[2] = (int)seq; flag[3] = (int)gridDim.x;
        __threadfence_system();
        *(volatile int*)flag = code;
        __threadfence_system();
    }
    __trap();
}
// spin until *word >= target (acquire), with the time-out above
__device__ __forceinline__ void barrier_spin(const unsigned* word, unsigned target, int code, unsigned seq) {
    if (ld_acquire_u32(word) >= target) return;
    const unsigned long long t0 = timer_ns(), limit = g_barrier_timeout_ns;
    unsigned spins = 0;
    while (ld_acquire_u32(word) < target) {
        __nanosleep(32);
        if ((++spins & 255u) == 0u && timer_ns() - t0 > limit) barrier_timeout_abort(code, seq);
    }
}

// Philox4x32-10 (Salmon et al. SC'11).  Replaces the per-slot XORWOW state of
// src/random.cu:3-12: no state array, keyed on (seed, particle id, step block).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        // one 32 x 32 -> 64 multiply per word pair (IMAD.WIDE.U32), not a high and a low multiply each
        const unsigned long long p0 = (unsigned long long)M0 * c.x, p1 = (unsigned long long)M1 * c.z;
        c = make_uint4((uint32_t)(p1 >> 32) ^ c.y ^ k.x, (uint32_t)p1, (uint32_t)(p0 >> 32) ^ c.w ^ k.y, (uint32_t)p0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}
// The same generator with its key schedule (round r uses key + r * Weyl constant) worked out once on the host and read
// from the kernel's parameter bank: two multiplies and two three-way XORs per round, nothing else.
struct PhiloxKeys { uint2 k[10]; };
inline PhiloxKeys philox_round_keys(uint2 key) {
    PhiloxKeys ks;
    for (int r = 0; r < 10; ++r) { ks.k[r] = key; key.x += 0x9E3779B9u; key.y += 0xBB67AE85u; }
    return ks;
}
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, const PhiloxKeys& ks) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned long long p0 = (unsigned long long)M0 * c.x, p1 = (unsigned long long)M1 * c.z;
        c = make_uint4((uint32_t)(p1 >> 32) ^ c.y ^ ks.k[r].x, (uint32_t)p1, (uint32_t)(p0 >> 32) ^ c.w ^ ks.k[r].y, (uint32_t)p0);
    }
    return c;
}
// curand_uniform's (0,1] mapping used by randFloat (src/random.cu:7-12).
__device__ __forceinline__ float uniform01(uint32_t w) {
    return __fmaf_rn(__uint2float_rn(w), 0x1p-32f, 0x1p-33f);
}
// Identity of a child spawned by `parent` at global mobility step g (splitmix64 finaliser).
__device__ __forceinline__ unsigned long long child_id(unsigned long long parent, unsigned long long g) {
    unsigned long long z = parent + g * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= (z >> 31);
    return z | 0x8000000000000000ull;
}

// energyToIndex (src/cross_section.cu:32-35) without the fp64 log10: an estimate of
// s = (log10(E)+6)*10000/22 from the SFU's log2 (lg2.approx: absolute error of s below 0.01, where 0.5 would do),
// rounded to the nearest bin boundary j, then one shared-memory probe of the exact boundary energy: edge[j] is the
// smallest fp32 E whose reference index (double formula) is >= j.  j is clamped to [0, kRows-1], so an energy beyond
// the last boundary needs no second clamp; E = 0 and subnormals (flushed) give lg2 = -inf -> j = 0; a NaN energy
// gives j = cvt(NaN) = 0 and fails the comparison -> 0, the reference's saturating conversion.
// The result equals the reference formula for EVERY fp32 E: the edge table is checked against the double formula on
// the CPU (tests/tools/check_energy_index.cpp), the estimate-and-probe against a bisection of that table on the
// device (k_sweep_energy_index, all 2^31 non-negative bit patterns).
__device__ __forceinline__ int energy_row_estimate(float e) {       // the row, or the one above it
    float lg;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(e));
    const float s = __fmaf_rn(lg, 136.83181621090054f, 2727.2727272727275f);   // *log10(2)*1e4/22, +6e4/22
    return max(0, min(__float2int_rn(s), kRows - 1));
}
__device__ __forceinline__ int energy_index(float e, const float* __restrict__ s_edge) {
    const int j = energy_row_estimate(e);
    return j - ((e < s_edge[j]) ? 1 : 0);
}
// the same index by bisection of the edge table (no estimate involved): the yardstick of the device-side sweep
__device__ __forceinline__ int energy_index_bisect(float e, const float* __restrict__ edge) {
    if (e != e) return 0;
    int lo = 0, hi = kRows;                // invariant: edge[lo] <= e (edge[0] = 0 <= e for e >= 0), edge[hi] > e or hi == kRows
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (e >= edge[mid]) lo = mid; else hi = mid;
    }
    return lo;
}

// cell index of a coordinate: int(pos / cell_size), double divide, C truncation
// (src/grid_operations.cu:21-23, :65-67), clamped for memory safety.
__device__ __forceinline__ int cell_of(float p) {
    int c = __double2int_rz((double)p / 1e-2);
    return max(0, min(c, kGrid - 1));
}
__device__ __forceinline__ size_t cell_index(int ix, int iy, int iz) {
    return ((size_t)iz * kGrid + (size_t)iy) * kGrid + (size_t)ix;
}

// ------------------------------------------------------------------------------------------
// K_init: setup_particles (src/particle_move.cu:7-19).  Three Philox words -> x,y,z in
// (2.26, 2.88]; velocity and acceleration start at 0 (uninitialised in the reference).
// ------------------------------------------------------------------------------------------
__global__ void k_init_uniform(Soa s, DevState* st, int init_n, unsigned long long id_base, uint2 seed) {
    const float lo = (float)((kGrid / 2 - 30) * 1e-2);
    const float hi = (float)((kGrid / 2 + 32) * 1e-2);
    const float range = __fsub_rn(hi, lo);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < init_n; i += gridDim.x * blockDim.x) {
        unsigned long long id = id_base + (unsigned long long)i;
        uint4 r = philox4x32_10(make_uint4((uint32_t)id, (uint32_t)(id >> 32), 0u, kTagInit), seed);
        s.x[i] = __fmaf_rn(uniform01(r.x), range, lo);
        s.y[i] = __fmaf_rn(uniform01(r.y), range, lo);
        s.z[i] = __fmaf_rn(uniform01(r.z), range, lo);
        s.vx[i] = 0.f; s.vy[i] = 0.f; s.vz[i] = 0.f;
        s.ax[i] = 0.f; s.ay[i] = 0.f; s.az[i] = 0.f;
        s.ts[i] = -1;
        s.id[i] = id;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { st->n = init_n; st->n_alloc = init_n; }
}

// (k_deposit, k_field_box and k_gather of modes 31/32 and the phase-level ABI: pic_mover.cuh, next to the fused kernel's forms)

// K_clear: resetGrid (src/grid_operations.cu:6-12) restricted to the deposited bounding box.
// Warp-granular form for the fused kernel: global warp `gw` of `nw` empties its x-rows of the box.
__device__ __forceinline__ void clear_box_rows(int* __restrict__ charge, const int* box /* lo xyz, hi xyz */, int gw, int nw, int lane) {
    const int lx = box[0], ly = box[1], lz = box[2];
    const int wx = box[3] - lx + 1, wy = box[4] - ly + 1, wz = box[5] - lz + 1;
    if (wx <= 0 || wy <= 0 || wz <= 0) return;
    const int rows = wy * wz;
    for (int r = gw; r < rows; r += nw) {
        const int yy = r % wy, zz = r / wy;
        int* row = charge + cell_index(lx, ly + yy, lz + zz);
        for (int xx = lane; xx < wx; xx += 32) __stcg(row + xx, 0);
    }
}
__device__ __forceinline__ void clear_box_body(const DevState* st, int* __restrict__ charge, int which) {
    const int lx = ld_state_i32(&st->bbox_lo[which][0]), ly = ld_state_i32(&st->bbox_lo[which][1]), lz = ld_state_i32(&st->bbox_lo[which][2]);
    const int wx = ld_state_i32(&st->bbox_hi[which][0]) - lx + 1, wy = ld_state_i32(&st->bbox_hi[which][1]) - ly + 1,
              wz = ld_state_i32(&st->bbox_hi[which][2]) - lz + 1;
    if (wx <= 0 || wy <= 0 || wz <= 0) return;
    // one warp per x-row of the box: contiguous stores, no 64-bit division
    const int lane = threadIdx.x & 31;
    const int rows = wy * wz, nw = (gridDim.x * blockDim.x) >> 5;
    for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows; r += nw) {
        const int yy = r % wy, zz = r / wy;
        int* row = charge + cell_index(lx, ly + yy, lz + zz);
        for (int xx = lane; xx < wx; xx += 32) __stcg(row + xx, 0);
    }
}
__global__ void k_clear_box(const DevState* st, int* __restrict__ charge, int which) { clear_box_body(st, charge, which); }
__global__ void k_set_n(DevState* st, int n) { st->n = n; st->n_alloc = n; st->top_alloc = n; }
// which: 0 .. kGrids-1 = that grid's box, kGrids = all slots
__global__ void k_reset_bbox(DevState* st, int which) {
    if (threadIdx.x < 3)
        for (int g = 0; g < kBoxes; ++g)
            if (which == g || which == kGrids) { st->bbox_lo[g][threadIdx.x] = kGrid; st->bbox_hi[g][threadIdx.x] = -1; }
}

// Multi-GPU: the union bounding box is one max-reduction of {-lo[3], hi[3], 0, 0}.
__global__ void k_bbox_encode(const DevState* st, int* red, int which) {
    if (threadIdx.x < 3) { red[threadIdx.x] = -st->bbox_lo[which][threadIdx.x]; red[3 + threadIdx.x] = st->bbox_hi[which][threadIdx.x]; }
    if (threadIdx.x >= 6 && threadIdx.x < 8) red[threadIdx.x] = 0;
}
__global__ void k_bbox_decode(DevState* st, const int* red, int which) {
    if (threadIdx.x < 3) { st->bbox_lo[which][threadIdx.x] = -red[threadIdx.x]; st->bbox_hi[which][threadIdx.x] = red[3 + threadIdx.x]; }
}

// pack / unpack the charge box into a contiguous buffer for the multi-GPU all-reduce
__global__ void k_pack_box(const int* __restrict__ charge, int* __restrict__ buf, int lx, int ly, int lz, int wx, int wy, int wz, int unpack) {
    const long long total = (long long)wx * wy * wz;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(k % wx);
        const long long r = k / wx;
        const int y = (int)(r % wy), z = (int)(r / wy);
        int* g = const_cast<int*>(charge) + cell_index(lx + x, ly + y, lz + z);
        if (unpack) *g = buf[k]; else buf[k] = *g;
    }
}

// ------------------------------------------------------------------------------------------
// Compaction: remove_dead_particles (src/pic.cu:320-357) done IN PLACE.  With S slots and D
// dead, the survivors must end up in [0, L), L = S - D.  Dead slots below L are "holes",
// live slots at or above L are "movers"; there are equally many of each.  Pass 1 lists both
// (warp-aggregated appends), pass 2 copies mover k into hole k.  Traffic is proportional to
// the number of dead electrons, not to n, and no second capacity-sized buffer exists.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void compact_scan_body(const Soa& s, DevState* st, int S, int D, int* __restrict__ holes, int* __restrict__ movers,
                                                  unsigned long long* trace = nullptr) {
    // Every block owns a contiguous range of 128-slot "quads" (one int4 of timestamps per lane).  A warp
    // loads up to kScanDepth quads back to back (all loads in flight at once), keeps the hole / mover
    // flags of its slots in two bit masks, and the block reserves list space with ONE pair of global
    // atomics per round -- per-warp atomics on the two list cursors serialised in L2 and cost 37 us per
    // Poisson step at 1.2e6 slots; re-reading the timestamps in a second pass cost another 8 us.
    constexpr int kScanDepth = 8;
    __shared__ int s_cnt[2][32];
    __shared__ int s_base[2];
    // S = slots in use, D = dead among them: read ONCE per block by the caller (every thread asking L2 for
    // the same word made that one L2 slice the bottleneck of the phase)
    if (D == 0) return;                                       // uniform over the grid
    const int L = S - D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int quads = (S + 127) >> 7;
    const int per_block = (quads + gridDim.x - 1) / gridDim.x;
    const int q_lo = min(quads, (int)blockIdx.x * per_block), q_hi = min(quads, q_lo + per_block);
    const int per_round = nwarps * kScanDepth;
    for (int q0 = q_lo; q0 < q_hi; q0 += per_round) {         // uniform over the block
        unsigned hm = 0, mm = 0;                              // bit 4k+j: slot j of this lane's int4 in quad k
#pragma unroll
        for (int k = 0; k < kScanDepth; ++k) {
            const int q = q0 + warp * kScanDepth + k;
            const int i = (q << 7) + (lane << 2);
            if (q < q_hi && i < S) {                          // i + 3 stays inside the 256-byte padded array
                const int4 t4 = __ldcg(reinterpret_cast<const int4*>(s.ts + i));
                const int tv[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const bool in = i + j < S, dead = tv[j] == kDead;
                    if (in && dead && i + j < L) hm |= 1u << (4 * k + j);
                    if (in && !dead && i + j >= L) mm |= 1u << (4 * k + j);
                }
            }
        }
        const int ch = __popc(hm), cm = __popc(mm);
        if (trace && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); trace[9] = t; }
        int ph = ch, pm = cm;                                 // inclusive scan over the lanes of the warp
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int th = __shfl_up_sync(kFull, ph, o), tm = __shfl_up_sync(kFull, pm, o);
            if (lane >= o) { ph += th; pm += tm; }
        }
        if (lane == 31) { s_cnt[0][warp] = ph; s_cnt[1][warp] = pm; }
        __syncthreads();
        if (warp == 0) {
            const int vh = lane < nwarps ? s_cnt[0][lane] : 0, vm = lane < nwarps ? s_cnt[1][lane] : 0;
            int wh = vh, wm = vm;                             // inclusive scan over the warps of the block
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int th = __shfl_up_sync(kFull, wh, o), tm = __shfl_up_sync(kFull, wm, o);
                if (lane >= o) { wh += th; wm += tm; }
            }
            const int tot_h = __shfl_sync(kFull, wh, 31), tot_m = __shfl_sync(kFull, wm, 31);
            if (lane == 0) {
                s_base[0] = tot_h ? atomicAdd(&st->n_holes, tot_h) : 0;
                s_base[1] = tot_m ? atomicAdd(&st->n_movers, tot_m) : 0;
            }
            if (lane < nwarps) { s_cnt[0][lane] = wh - vh; s_cnt[1][lane] = wm - vm; }
        }
        __syncthreads();
        if (trace && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); trace[10] = t; }
        int oh = s_base[0] + s_cnt[0][warp] + ph - ch, om = s_base[1] + s_cnt[1][warp] + pm - cm;
        const int i_base = ((q0 + warp * kScanDepth) << 7) + (lane << 2);
        while (hm) { const int b = __ffs(hm) - 1; hm &= hm - 1; __stcg(holes + oh++, i_base + ((b >> 2) << 7) + (b & 3)); }
        while (mm) { const int b = __ffs(mm) - 1; mm &= mm - 1; __stcg(movers + om++, i_base + ((b >> 2) << 7) + (b & 3)); }
        __syncthreads();                                      // s_cnt / s_base are reused by the next round
    }
}
__global__ void k_compact_scan(Soa s, DevState* st, int capacity, int* __restrict__ holes, int* __restrict__ movers) {
    compact_scan_body(s, st, min(ld_state_i32(&st->n_alloc), capacity), ld_state_i32(&st->n_dead), holes, movers);
}
__device__ __forceinline__ void compact_move_body(const Soa& s, int H, const int* holes, const int* movers) {
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < H; k += gridDim.x * blockDim.x) {
        const int src = __ldcg(movers + k), dst = __ldcg(holes + k);
        __stcg(s.x + dst, __ldcg(s.x + src)); __stcg(s.y + dst, __ldcg(s.y + src)); __stcg(s.z + dst, __ldcg(s.z + src));
        __stcg(s.vx + dst, __ldcg(s.vx + src)); __stcg(s.vy + dst, __ldcg(s.vy + src)); __stcg(s.vz + dst, __ldcg(s.vz + src));
        __stcg(s.ax + dst, __ldcg(s.ax + src)); __stcg(s.ay + dst, __ldcg(s.ay + src)); __stcg(s.az + dst, __ldcg(s.az + src));
        __stcg(s.ts + dst, __ldcg(s.ts + src));
        __stcg(s.id + dst, __ldcg(s.id + src));
    }
}
__global__ void k_compact_move(Soa s, const DevState* st, const int* __restrict__ holes, const int* __restrict__ movers) {
    compact_move_body(s, min(ld_state_i32(&st->n_holes), ld_state_i32(&st->n_movers)), holes, movers);
}

// Bookkeeping the reference does on the host after each phase (src/pic.cu:539-559), on the device.
// Mode 32 (src/pic.cu:524-532) launches one kernel per mobility step over the electrons that exist when the step
// starts; the reference reads that count back to the host before every launch, here it stays on the device.
// Mode 31 (src/pic.cu:514-523) relaunches over the children of the previous launch until none are born: `advance`
// moves the window on to [old end, n); an empty window makes the launch return at once.
__global__ void k_freeze_window(DevState* st, int capacity, int advance) {
    if (advance) st->win_lo = st->win_hi;
    st->win_hi = min(st->n_alloc, capacity);
}
// The window of the first launch is [0, min(n, capacity)) in modes 31 and 32 alike, so no k_freeze_window precedes it;
// reset_box >= 0: also k_reset_bbox for that grid (one single-thread launch instead of three per Poisson step).
__global__ void k_mobility_begin(DevState* st, int capacity, int reset_box) {
    if (reset_box >= 0)
        for (int k = 0; k < 3; ++k) { st->bbox_lo[reset_box][k] = kGrid; st->bbox_hi[reset_box][k] = -1; }
    st->n_alloc = st->n;
    st->win_lo = 0; st->win_hi = min(st->n, capacity);
    st->top_alloc = st->n;
    st->n_dead = 0;
    st->n_holes = 0; st->n_movers = 0;
    st->bar_count = 0; st->bar_release = 0;
    st->ticket = 0; st->n_unfinished = 0;
}
__device__ __forceinline__ void step_end_body(DevState* st_, int capacity, StepLog* log, int log_len, int stop_at_zero) {
    volatile DevState* st = st_;                   // also runs inside the fused kernel: never through a stale L1 line
    if (st->hit_zero) return;                      // the reference leaves its loop at "Hit 0"
    const int n0 = st->n;
    const int after = st->n_alloc;                 // may overshoot capacity, like the reference's n
    const int S = min(after, capacity);
    const int L = S - st->n_dead;
    st->total_added = st->total_added + (long long)(after - n0);    // src/pic.cu:540
    st->total_removed = st->total_removed + (long long)(after - L); // src/pic.cu:552
    st->last_n_after_mobility = after;
    if (after >= capacity) st->overflow_steps = st->overflow_steps + 1;   // src/pic.cu:544
    const int p = st->poisson_index;
    if (log && p < log_len) { log[p].n_after_mobility = after; log[p].n_after_compaction = L; }
    st->n = L;
    st->n_alloc = L;
    st->n_dead = 0; st->n_holes = 0; st->n_movers = 0;
    // src/pic.cu:556-559; with several ranks an empty slab keeps stepping (the collective needs it)
    if (L == 0 && stop_at_zero) st->hit_zero = 1;
    st->poisson_index = p + 1;
}
// The same bookkeeping inside the fused kernel, which does NOT compact after every Poisson step: a child
// moves into a slot whose electron has died (every warp keeps the free slots it comes across), and what
// holes remain stay where they are -- the chunks skip them -- until they are worth a compaction.  `n`, the
// totals and the overflow rule are the reference's: a child exists iff the reference's counter n (n_alloc,
// which starts every step at the number of LIVE electrons) is below capacity when it is born; only the slot
// it lands in differs (a free one, else the next above the high-water mark, top_alloc).  Should the slots
// above the mark run out while the reference's counter has not reached the capacity, the finished child is
// parked and placed by the compaction this step then ends with.  Publishes do_compact, scan_S, scan_D,
// parked, hw.
__device__ __forceinline__ void step_end_fused(DevState* st_, int capacity, StepLog* log, int log_len, int stop_at_zero, int hw_start,
                                               bool last_step, int policy /* 0 lazy, 1 every step, 2 lazy without the capacity guard (tests) */) {
    volatile DevState* st = st_;
    const int n0 = st->n;
    const int after = st->n_alloc;                 // may overshoot capacity, like the reference's n
    const int S = min(after, capacity);
    const int L = S - st->n_dead;
    const int parked = min(st->n_parked, capacity);
    st->total_added = st->total_added + (long long)(after - n0);    // src/pic.cu:540
    st->total_removed = st->total_removed + (long long)(after - L); // src/pic.cu:552
    st->last_n_after_mobility = after;
    if (after >= capacity) st->overflow_steps = st->overflow_steps + 1;   // src/pic.cu:544
    const int p = st->poisson_index;
    if (log && p < log_len) { log[p].n_after_mobility = after; log[p].n_after_compaction = L; }
    const int top = st->top_alloc;                 // may overshoot, too
    st->n = L;
    st->n_alloc = L;
    st->n_dead = 0; st->n_holes = 0; st->n_movers = 0; st->n_parked = 0;
    int hw = max(min(top, capacity), hw_start);
    const int holes = hw - (L - parked);
    const long long births = (long long)S - n0;
    // compact when the launch ends (the host sees slots [0,n) alive), when children wait for room, when
    // more than one slot in eight is dead, or when the next step could run into the capacity
    bool compact = last_step || policy == 1 || parked > 0 || holes > (L >> 3) || (policy != 2 && (long long)hw + 2 * births > (long long)capacity);
    if (holes == 0) compact = false;
    if (L == 0) { compact = false; hw = 0; }       // nobody left: nothing to move
    st->do_compact = compact ? 1 : 0;
    st->scan_S = hw; st->scan_D = holes; st->parked = compact ? parked : 0;
    st->hw = compact ? L : hw;
    st->top_alloc = compact ? L : hw;
    // src/pic.cu:556-559; with several ranks an empty slab keeps stepping (the collective needs it)
    if (L == 0 && stop_at_zero) st->hit_zero = 1;
    st->poisson_index = p + 1;
}
__global__ void k_step_end(DevState* st, int capacity, StepLog* log, int log_len, int stop_at_zero) {
    step_end_body(st, capacity, log, log_len, stop_at_zero);
}

// The reference's 64-byte AoS record (src/electron.h:12-18: double3 position, double3 velocity, float3 acceleration,
// int timestamp) built on the device from the SoA store: fp32 state widened to double, four 16-byte stores per electron
// (replaces a single-threaded transpose on the host; RunData.electrons of src/pic.cu:583 and log() of src/utility.cu:124-137
// consume this layout).
struct alignas(16) AosRecord { double px, py, pz, vx, vy, vz; float ax, ay, az; int ts; };
static_assert(sizeof(AosRecord) == 64, "src/electron.h:12-18");
__global__ void k_export_aos(Soa s, int first, int count, AosRecord* __restrict__ out) {
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const int i = first + k;
        double2* o = reinterpret_cast<double2*>(out + k);
        const float x = __ldcs(s.x + i), y = __ldcs(s.y + i), z = __ldcs(s.z + i);
        const float vx = __ldcs(s.vx + i), vy = __ldcs(s.vy + i), vz = __ldcs(s.vz + i);
        __stcs(o + 0, make_double2((double)x, (double)y));
        __stcs(o + 1, make_double2((double)z, (double)vx));
        __stcs(o + 2, make_double2((double)vy, (double)vz));
        __stcs(reinterpret_cast<float4*>(o + 3), make_float4(__ldcs(s.ax + i), __ldcs(s.ay + i), __ldcs(s.az + i), __int_as_float(__ldcs(s.ts + i))));
    }
}

#include "pic_mover.cuh"

// device-side known-answer hooks
__global__ void k_test_philox(long long n, const uint32_t* ctr, uint2 key, uint32_t* out) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        uint4 r = philox4x32_10(make_uint4(ctr[4 * i], ctr[4 * i + 1], ctr[4 * i + 2], ctr[4 * i + 3]), key);
        out[4 * i] = r.x; out[4 * i + 1] = r.y; out[4 * i + 2] = r.z; out[4 * i + 3] = r.w;
    }
}
// every stride-th non-negative fp32 bit pattern (0 .. 0x7fffffff: zero, subnormals, normals, +inf, NaNs): energy_index
// against the bisection of the same edge table; out[0] = mismatches, out[1] = bits of the first one seen
__global__ void k_sweep_energy_index(unsigned stride, const float* tables, unsigned long long* out) {
    const float* edge = reinterpret_cast<const float*>(reinterpret_cast<const unsigned char*>(tables) + kThrBytes);
    unsigned long long bad = 0;
    for (unsigned long long b = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * stride; b <= 0x7fffffffull;
         b += (unsigned long long)gridDim.x * blockDim.x * stride) {
        const float e = __uint_as_float((unsigned)b);
        if (energy_index(e, edge) != energy_index_bisect(e, edge)) { ++bad; out[1] = b; }
    }
    if (bad) atomicAdd(out, bad);
}
__global__ void k_test_energy_index(long long n, const float* e, const float* tables, int* out) {
    const float* edge = reinterpret_cast<const float*>(reinterpret_cast<const unsigned char*>(tables) + kThrBytes);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = energy_index(e[i], edge);
}

}  // namespace picb200
