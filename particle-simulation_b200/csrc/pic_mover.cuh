// pic_mover.cuh -- the scheduler kernels (included inside namespace picb200 by pic_device.cuh).
//
// K_mover: leapfrog + bounds + Monte-Carlo collision (updateParticle, src/particle_move.cu:22-87)
// under the scheduling disciplines of src/pic.cu:
//   kPersistent = true  (modes 30/33): ONE cooperative launch per Poisson step.  32-electron
//       chunks of the population are dealt round-robin to the grid's warps; an electron stays in
//       registers for all its remaining mobility steps.  Children are ranked into a warp-private
//       stage in shared memory and the SAME warp later pushes them to the end of the Poisson step
//       straight out of that stage, so a child normally reaches the global store exactly once, in
//       its final state, and no grid-wide synchronisation separates "generations".  Only if a
//       stage overflows are unfinished children spilled to the global store; a grid barrier at the
//       end detects that and a fall-back pass finishes them.  The host never sees any of it
//       (contrast src/pic.cu:70-211: block-wide spin lock, polling of timestamps, host reset of
//       four counters per Poisson step).
//   kPersistent = false: mode 31 (the host passes the index window of one generation of children,
//       src/pic.cu:514-523) and mode 32 (t_from == t_to: one mobility step per launch,
//       src/pic.cu:524-532).  Children are always flushed unfinished, as those modes require.

struct MoverArgs {
    Soa s;
    DevState* st;
    const float* tables;   // global image of the shared-memory tables (kTableSmemBytes)
    int capacity, pt;
    int lo, hi, t_from, t_to;   // kPersistent == false only
    int stepwise;               // mode 32: nobody revisits a child born at the last step, so it is born with ts = -1
    uint2 seed;
};

// Children are staged per warp in shared memory: 12 words x kSS entries, field-major, so both the
// ranked writes and the lane-indexed reads are bank-conflict free.  The unfused schedulers (k_mover) use the
// first 32 entries; the fused kernel uses all kSS = 44, so that it can wait for enough children to make up a batch
// of 32 with similar birth steps (fused_staged) -- a batch costs poisson_timestep - min(birth step) lockstep
// steps whatever its lanes hold.
constexpr int kStageWords = 12;
#ifndef PIC_KSS
#define PIC_KSS 44
#endif
constexpr int kSS = PIC_KSS;          // entries per field (stride in words): tables + 32 stages + pools must stay below the
                                      // 196 KB shared-memory carve-out -- the next one (228 KB) leaves the L1 28 KB, and the
                                      // memory-bound configurations lose far more to that than batching wins (measured)
constexpr int kStageBytesPerWarp = kStageWords * kSS * 4;
constexpr int kMoverSmemBytes(int threads) { return kTableSmemBytes + (threads / 32) * kStageBytesPerWarp; }
constexpr int kDrainThreshold = 16;   // k_mover: push staged children once this many wait
constexpr int kFusedDrainAt = kSS - 20 > 16 ? kSS - 20 : 16;   // fused kernel: ... once this many wait (PICB200_DRAIN_AT); 20 entries stay free for the next chunk

struct El {
    float x, y, z, vx, vy, vz, ax, ay, az;
    unsigned long long id;
};

struct MoverCtx {
    Soa s;
    DevState* st;
    const float2* thr;
    const float* edge;
    const float2* g_thr;   // the whole tables in global memory (rows beyond a shared-memory window of kStreamRows)
    const float* g_edge;
    uint32_t* stage;
    int stage_cap;       // entries of the stage in use: 32 (k_mover) or kSS (fused kernel)
    int cap, pt, lane, stepwise;
    int adopt_min;       // fused kernel, adopting loop: dead lanes in the warp from which children are handed over
    int* pool;           // fused kernel: free slots this warp has come across in the current Poisson step (shared memory)
    unsigned g0;
    uint2 seed;
    const PhiloxKeys* keys;   // fused kernel: the round keys of `seed`, in the kernel's parameter bank
};

// Reserve `count` consecutive slots of the global store: one atomic per call, honouring the
// reference's `if (*n < capacity)` pre-check (src/pic.cu:222).  Returns the first slot (>= cap: none).
__device__ __forceinline__ int reserve_slots(const MoverCtx& c, int count) {
    int base = 0;
    if (c.lane == 0) base = (ld_volatile_i32(&c.st->n_alloc) < c.cap) ? atomicAdd(&c.st->n_alloc, count) : c.cap;
    return __shfl_sync(kFull, base, 0);
}

// Fused kernel: `count` children are born, `from_top` of them need a slot above the high-water mark (the
// others move into free slots the warp holds).  ONE 64-bit atomic advances the reference's counter n (low
// word: decides which children exist, src/pic.cu:222) and the mark (high word).
constexpr int kPoolSlots = 64;
__device__ __forceinline__ void reserve_fused(const MoverCtx& c, int count, int from_top, int& vbase, int& tbase) {
    unsigned long long old = (unsigned long long)(unsigned)c.cap | ((unsigned long long)(unsigned)c.cap << 32);   // exhausted: nobody is born, no slot is taken
    if (c.lane == 0 && ld_volatile_i32(&c.st->n_alloc) < c.cap)
        old = atomicAdd(reinterpret_cast<unsigned long long*>(&c.st->n_alloc), (unsigned long long)(unsigned)count | ((unsigned long long)(unsigned)from_top << 32));
    old = __shfl_sync(kFull, old, 0);
    vbase = (int)(unsigned)(old & 0xffffffffull);
    tbase = (int)(unsigned)(old >> 32);
}

// A stage entry carries the parent's id and the birth step; the child's own id is derived when the
// entry is consumed (all lanes busy) instead of when it is written (one or two lanes busy).
__device__ __forceinline__ unsigned long long staged_child_id(const MoverCtx& c, const uint32_t* g) {
    const unsigned long long parent = ((unsigned long long)g[11 * kSS] << 32) | g[10 * kSS];
    return child_id(parent, (unsigned long long)(c.g0 + g[9 * kSS]));
}

// Write the staged children to the global store as they are (timestamp = birth step): they still
// need their remaining pushes.  kCount: tell the persistent kernel's fall-back pass they exist.
template <bool kCount>
__device__ __forceinline__ void stage_spill(const MoverCtx& c, int count) {
    __syncwarp();
    int vbase, sbase;
    if (c.pool) {            // fused kernel: the reference's index decides whether the child exists, it lives above the mark
        reserve_fused(c, count, count, vbase, sbase);
    } else {
        vbase = sbase = reserve_slots(c, count);
    }
    int written = 0;
#pragma unroll 1
    for (int h = 0; h < count; h += 32) {       // the fused kernel's stage holds up to kSS = 64 children
        const int e = h + c.lane;
        const int vslot = vbase + e, slot = sbase + e;
        if (c.pool) {
            if (e < count && vslot < c.cap && slot >= c.cap) *(volatile int*)&c.st->error = 3;   // unfinished children cannot be parked
            if (e < count && vslot >= c.cap && slot < c.cap) __stcg(c.s.ts + slot, kDead);       // a slot above the mark nobody was born into: a hole
        }
        const bool w = e < count && vslot < c.cap && slot < c.cap;
        if (w) {
            const uint32_t* g = c.stage + e;
            __stcg(c.s.x + slot, __uint_as_float(g[0 * kSS])); __stcg(c.s.y + slot, __uint_as_float(g[1 * kSS])); __stcg(c.s.z + slot, __uint_as_float(g[2 * kSS]));
            __stcg(c.s.vx + slot, __uint_as_float(g[3 * kSS])); __stcg(c.s.vy + slot, __uint_as_float(g[4 * kSS])); __stcg(c.s.vz + slot, __uint_as_float(g[5 * kSS]));
            __stcg(c.s.ax + slot, __uint_as_float(g[6 * kSS])); __stcg(c.s.ay + slot, __uint_as_float(g[7 * kSS])); __stcg(c.s.az + slot, __uint_as_float(g[8 * kSS]));
            const int tb = (int)g[9 * kSS];
            __stcg(c.s.ts + slot, (c.stepwise && tb == c.pt) ? -1 : tb);
            __stcg(c.s.id + slot, staged_child_id(c, g));
        }
        if (kCount) written += __popc(__ballot_sync(kFull, w));
    }
    if (kCount) {
        if (c.lane == 0 && written) atomicAdd(&c.st->n_unfinished, written);
    }
    __syncwarp();
}

// {T(split), T(split + remove)} of the table row of energy `en` (energyToIndex, src/cross_section.cu:32-35): estimate, one
// probe of the exact bin edge, one 8-byte row.  kSmemRows < kRows: shared memory holds the first kSmemRows rows only.
template <int kSmemRows>
__device__ __forceinline__ float2 collision_row(const MoverCtx& c, float en) {
    const int j = energy_row_estimate(en);
    if (kSmemRows < kRows && j >= kSmemRows) {
        const int idx = j - ((en < __ldg(c.g_edge + j)) ? 1 : 0);
        return __ldg(c.g_thr + idx);
    }
    const int idx = j - ((en < c.edge[j]) ? 1 : 0);
    return c.thr[idx];
}

// Push one lane's electron through mobility steps max(t_begin.., t0) .. t_to; all 32 lanes run the
// same step together.  Spawned children are ranked into the warp's stage.
template <bool kCountSpill, int kSmemRows = kRows>
__device__ __forceinline__ void advance(const MoverCtx& c, El& e, bool& alive, int t0, int t_to, int& staged,
                                        unsigned& pushes, int& dead, bool& touched) {
    const float dt = (float)1e-12, hdt = (float)(1e-12 / 2);      // src/pic.cu:371
    const float sim = (float)(kGrid * 1e-2);                      // Sim_Size, src/cell.cu:4
    const int t_begin = __reduce_min_sync(kFull, alive ? t0 : 0x7fffffff);
    uint4 rnd = make_uint4(0, 0, 0, 0);
    unsigned cur_blk = 0xffffffffu;
    for (int t = t_begin; t <= t_to; ++t) {
        bool spawn = false;
        float cvx = 0, cvy = 0, cvz = 0;
        const unsigned g = c.g0 + (unsigned)t;
        if (alive && t >= t0) {
            touched = true;
            ++pushes;
            // leapfrog, src/particle_move.cu:22-39
            e.vx = __fmaf_rn(-e.ax, hdt, e.vx); e.vy = __fmaf_rn(-e.ay, hdt, e.vy); e.vz = __fmaf_rn(-e.az, hdt, e.vz);
            e.x = __fmaf_rn(e.vx, dt, e.x); e.y = __fmaf_rn(e.vy, dt, e.y); e.z = __fmaf_rn(e.vz, dt, e.z);
            e.vx = __fmaf_rn(-e.ax, hdt, e.vx); e.vy = __fmaf_rn(-e.ay, hdt, e.vy); e.vz = __fmaf_rn(-e.az, hdt, e.vz);
            // checkOutOfBounds, src/particle_move.cu:41-52 (no draw consumed)
            // (written as the reference writes it: a NaN coordinate is NOT out of bounds)
            const bool oob = (e.x < 0.f) | (e.x >= sim) | (e.y < 0.f) | (e.y >= sim) | (e.z < 0.f) | (e.z >= sim);
            if (oob) {
                alive = false; ++dead;
            } else {
                // collider, src/particle_move.cu:55-80
                const unsigned blk = g >> 2;
                if (blk != cur_blk) {
                    rnd = philox4x32_10(make_uint4((uint32_t)e.id, (uint32_t)(e.id >> 32), blk, kTagCollide), c.seed);
                    cur_blk = blk;
                }
                const uint32_t w01 = (g & 1u) ? rnd.y : rnd.x, w23 = (g & 1u) ? rnd.w : rnd.z;
                const uint32_t w = (g & 2u) ? w23 : w01;
                const float r = __uint2float_rn(w);          // the rows hold thresholds on float(w): draw_threshold() on the host
                const float en = __fmaf_rn(e.vz, e.vz, __fmaf_rn(e.vy, e.vy, __fmul_rn(e.vx, e.vx)));
                const float2 th = collision_row<kSmemRows>(c, en);
                if (r < th.x) {              // ionisation: the child copies the pre-flip state
                    spawn = true;
                    cvx = e.vx; cvy = e.vy; cvz = e.vz;
                    e.vx = -e.vx; e.vy = -e.vy; e.vz = -e.vz;
                } else if (r < th.y) {       // attachment: electron removed
                    alive = false; ++dead;
                }
            }
        }
        const unsigned m = __ballot_sync(kFull, spawn);
        if (m) {
            const int cnt = __popc(m);
            if (staged + cnt > c.stage_cap) { stage_spill<kCountSpill>(c, staged); staged = 0; }
            if (spawn) {
                uint32_t* q = c.stage + staged + __popc(m & lanemask_lt());
                q[0 * kSS] = __float_as_uint(e.x); q[1 * kSS] = __float_as_uint(e.y); q[2 * kSS] = __float_as_uint(e.z);
                q[3 * kSS] = __float_as_uint(cvx); q[4 * kSS] = __float_as_uint(cvy); q[5 * kSS] = __float_as_uint(cvz);
                q[6 * kSS] = __float_as_uint(e.ax); q[7 * kSS] = __float_as_uint(e.ay); q[8 * kSS] = __float_as_uint(e.az);
                q[9 * kSS] = (uint32_t)t;
                q[10 * kSS] = (uint32_t)e.id; q[11 * kSS] = (uint32_t)(e.id >> 32);
            }
            staged += cnt;
        }
        if (!__any_sync(kFull, alive)) break;
    }
}

// One 32-electron chunk straight from the global store: slots [i0, i0+32) clipped to `end`.
// kLaterGen: fall-back pass of the persistent kernel -- only spilled children (ts >= 0) need work.
// In two halves, so that the windowed launches of modes 31/32 can have the NEXT chunk's loads in flight while they push.
__device__ __forceinline__ void chunk_load(const MoverCtx& c, int i, bool in, El& e, int& ts) {
    e = El{0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    ts = kDead;
    if (in) {
        e.x = __ldcg(c.s.x + i); e.y = __ldcg(c.s.y + i); e.z = __ldcg(c.s.z + i);
        e.vx = __ldcg(c.s.vx + i); e.vy = __ldcg(c.s.vy + i); e.vz = __ldcg(c.s.vz + i);
        e.ax = __ldcg(c.s.ax + i); e.ay = __ldcg(c.s.ay + i); e.az = __ldcg(c.s.az + i);
        ts = __ldcg(c.s.ts + i);
        e.id = __ldcg(c.s.id + i);
    }
}
template <bool kCountSpill, bool kLaterGen, int kSmemRows = kRows>
__device__ __forceinline__ void chunk_push_store(const MoverCtx& c, int i, bool in, El& e, int ts, int t_from, int t_to, int& staged,
                                                 unsigned& pushes, int& dead) {
    const bool was_alive = in && (kLaterGen ? ts >= 0 : ts != kDead);
    bool alive = was_alive;
    // first step of this electron: max(1, timestamp+1)  (src/pic.cu:218)
    const int t0 = max(t_from, max(1, ts + 1));
    bool touched = false;
    advance<kCountSpill, kSmemRows>(c, e, alive, t0, t_to, staged, pushes, dead, touched);
    if (was_alive) {
        if (touched) {
            __stcg(c.s.x + i, e.x); __stcg(c.s.y + i, e.y); __stcg(c.s.z + i, e.z);
            __stcg(c.s.vx + i, e.vx); __stcg(c.s.vy + i, e.vy); __stcg(c.s.vz + i, e.vz);
        }
        // survivors of the last mobility step are "alive since the start" again
        // (remove_dead_particles writes -1, src/pic.cu:356)
        const int ts1 = !alive ? kDead : (t_to == c.pt ? -1 : ts);
        if (ts1 != ts) __stcg(c.s.ts + i, ts1);
    }
}
template <bool kCountSpill, bool kLaterGen>
__device__ __forceinline__ void process_chunk(const MoverCtx& c, int i0, int end, int t_from, int t_to, int& staged,
                                              unsigned& pushes, int& dead) {
    const int i = i0 + c.lane;
    const bool in = i < end;
    El e;
    int ts;
    chunk_load(c, i, in, e, ts);
    chunk_push_store<kCountSpill, kLaterGen>(c, i, in, e, ts, t_from, t_to, staged, pushes, dead);
}

// Persistent kernel only: the warp takes the children waiting in its own stage into registers,
// reserves their final slots, pushes them to the end of the Poisson step (grandchildren go to the
// now empty stage) and writes each of them to the global store once, finished.
__device__ __forceinline__ void process_staged(const MoverCtx& c, int& staged, unsigned& pushes, int& dead) {
    const int k = staged;
    __syncwarp();
    El e{0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int tb = 0;
    if (c.lane < k) {
        const uint32_t* g = c.stage + c.lane;
        e.x = __uint_as_float(g[0 * kSS]); e.y = __uint_as_float(g[1 * kSS]); e.z = __uint_as_float(g[2 * kSS]);
        e.vx = __uint_as_float(g[3 * kSS]); e.vy = __uint_as_float(g[4 * kSS]); e.vz = __uint_as_float(g[5 * kSS]);
        e.ax = __uint_as_float(g[6 * kSS]); e.ay = __uint_as_float(g[7 * kSS]); e.az = __uint_as_float(g[8 * kSS]);
        tb = (int)g[9 * kSS];
        e.id = staged_child_id(c, g);
    }
    staged = 0;
    __syncwarp();
    const int slot = reserve_slots(c, k) + c.lane;
    const bool in = c.lane < k && slot < c.cap;      // children past the capacity are dropped (src/pic.cu:224)
    bool alive = in;
    bool touched = false;
    advance<true>(c, e, alive, tb + 1, c.pt, staged, pushes, dead, touched);
    if (in) {
        __stcg(c.s.x + slot, e.x); __stcg(c.s.y + slot, e.y); __stcg(c.s.z + slot, e.z);
        __stcg(c.s.vx + slot, e.vx); __stcg(c.s.vy + slot, e.vy); __stcg(c.s.vz + slot, e.vz);
        __stcg(c.s.ax + slot, e.ax); __stcg(c.s.ay + slot, e.ay); __stcg(c.s.az + slot, e.az);
        __stcg(c.s.ts + slot, alive ? -1 : kDead);
        __stcg(c.s.id + slot, e.id);
    }
}

// Grid barrier of the persistent kernel.  The last block to arrive decides whether a fall-back
// pass is needed: it publishes the end of the window that may hold spilled children, or -1.
__device__ __forceinline__ int mover_barrier_next_window(DevState* st, unsigned gen, int capacity) {
    __shared__ int s_hi;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned target = (gen + 1u) * gridDim.x;
        const unsigned prev = atomicAdd(&st->bar_count, 1u);
        if (prev + 1u == target) {
            __threadfence();
            const int unfinished = ld_volatile_i32(&st->n_unfinished);
            st->n_unfinished = 0;
            st->gen_hi = unfinished > 0 ? min(ld_volatile_i32(&st->n_alloc), capacity) : -1;
            __threadfence();
            atomicExch(&st->bar_release, gen + 1u);
        } else {
            barrier_spin(&st->bar_release, gen + 1u, 1, gen);
        }
        __threadfence();
        s_hi = ld_volatile_i32(&st->gen_hi);
    }
    __syncthreads();
    return s_hi;
}

template <bool kPersistent, bool kStepwise = false>
__global__ void __launch_bounds__(1024, 1) k_mover(MoverArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // hi < 0 / lo < 0: the window was frozen on the device by k_freeze_window (modes 31/32 without a host read per
    // launch); an empty one ends the launch before anything is staged
    const int w_hi = (!kPersistent && a.hi < 0) ? ld_state_i32(&a.st->win_hi) : a.hi;
    const int w_lo = (!kPersistent && a.lo < 0) ? ld_state_i32(&a.st->win_lo) : a.lo;
    if (!kPersistent && w_hi <= w_lo) return;

    MoverCtx c;
    c.s = a.s;
    c.st = a.st;
    c.thr = reinterpret_cast<const float2*>(smem_raw);
    c.edge = reinterpret_cast<const float*>(smem_raw + kThrBytes);
    c.g_thr = reinterpret_cast<const float2*>(a.tables);
    c.g_edge = reinterpret_cast<const float*>(reinterpret_cast<const unsigned char*>(a.tables) + kThrBytes);
    c.stage = reinterpret_cast<uint32_t*>(smem_raw + kTableSmemBytes) + (threadIdx.x >> 5) * (kStageWords * kSS);
    c.cap = a.capacity;
    c.pt = a.pt;
    c.lane = threadIdx.x & 31;
    c.stage_cap = 32;
    c.stepwise = a.stepwise;
    c.pool = nullptr;
    c.seed = a.seed;

    // kStepwise -- one launch per mobility step (mode 32; the tables are staged anew by every launch): the first chunk's loads go out
    // before the tables are staged, and only the first kStreamRows rows are -- the rest is read from the global image, as
    // in the streaming instantiation of the fused kernel
    constexpr bool stepwise = !kPersistent && kStepwise;      // (its own instantiation: the other two keep their code)
    const int span = kPersistent ? 0 : w_hi - w_lo;
    const int rounded = (span + 31) & ~31;
    const int stride = (int)(gridDim.x * blockDim.x);
    int k = (int)((blockIdx.x * blockDim.x + threadIdx.x) & ~31u);
    El e0{0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int ts0 = kDead;
    bool in0 = false;
    if (stepwise && k < rounded) {
        in0 = w_lo + k + c.lane < w_hi;
        chunk_load(c, w_lo + k + c.lane, in0, e0, ts0);
    }
    if (!stepwise) {   // stage the collision tables (120 KB) once per CTA, 128-bit copies
        const float4* src = reinterpret_cast<const float4*>(a.tables);
        float4* dst = reinterpret_cast<float4*>(smem_raw);
        for (int q = threadIdx.x; q < kTableSmemBytes / 16; q += blockDim.x) dst[q] = __ldg(src + q);
    } else {
        const unsigned char* g = reinterpret_cast<const unsigned char*>(a.tables);
        for (int q = threadIdx.x; q < kStreamRows * 8 / 16; q += blockDim.x) reinterpret_cast<float4*>(smem_raw)[q] = __ldg(reinterpret_cast<const float4*>(g) + q);
        for (int q = threadIdx.x; q < kStreamRows * 4 / 16; q += blockDim.x)
            reinterpret_cast<float4*>(smem_raw + kThrBytes)[q] = __ldg(reinterpret_cast<const float4*>(g + kThrBytes) + q);
    }
    const unsigned pidx = (unsigned)a.st->poisson_index;
    __syncthreads();
    c.g0 = pidx * (unsigned)a.pt;

    int staged = 0;            // warp-uniform: children waiting in this warp's stage
    unsigned pushes = 0;
    int dead = 0;

    if (kPersistent) {
        const int n0 = min(a.st->n, c.cap);
        const int chunks = (n0 + 31) >> 5;
        // generation 0: 32-electron chunks dealt round-robin over the grid's warps, SM-interleaved
        // (warp w of block b is global warp w*gridDim+b), so every SM gets the same number of
        // chunks to within one and no work counter is needed
        const int gw = (threadIdx.x >> 5) * gridDim.x + blockIdx.x, nw = gridDim.x * (blockDim.x >> 5);
        for (int chunk = gw; chunk < chunks; chunk += nw) {
            process_chunk<true, false>(c, chunk << 5, n0, 1, c.pt, staged, pushes, dead);
            if (staged >= kDrainThreshold) process_staged(c, staged, pushes, dead);
        }
        while (staged) process_staged(c, staged, pushes, dead);
        // fall-back passes: only if some stage overflowed and spilled unfinished children
        int lo = n0;
        for (unsigned gen = 0;; ++gen) {
            const int hi = mover_barrier_next_window(a.st, gen, c.cap);
            if (hi < 0) break;
            const int rounded = (hi - lo + 31) & ~31;
            for (int k = (blockIdx.x * blockDim.x + threadIdx.x) & ~31; k < rounded; k += gridDim.x * blockDim.x) {
                process_chunk<true, true>(c, lo + k, hi, 1, c.pt, staged, pushes, dead);
                if (staged >= kDrainThreshold) process_staged(c, staged, pushes, dead);
            }
            while (staged) process_staged(c, staged, pushes, dead);
            lo = hi;
        }
    } else {
        if (stepwise) {          // the next chunk's loads fly while this one is pushed
            while (k < rounded) {
                const int kn = k + stride;
                El e1{0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
                int ts1 = kDead;
                bool in1 = false;
                if (kn < rounded) {
                    in1 = w_lo + kn + c.lane < w_hi;
                    chunk_load(c, w_lo + kn + c.lane, in1, e1, ts1);
                }
                chunk_push_store<false, false, kStreamRows>(c, w_lo + k + c.lane, in0, e0, ts0, a.t_from, a.t_to, staged, pushes, dead);
                e0 = e1; ts0 = ts1; in0 = in1; k = kn;
            }
        } else {                 // a window through all its mobility steps (mode 31): slower with either (twelve more live
                                 // registers through ten steps, a window test in every step)
            for (; k < rounded; k += stride) process_chunk<false, false>(c, w_lo + k, w_hi, a.t_from, a.t_to, staged, pushes, dead);
        }
        if (staged) { stage_spill<false>(c, staged); staged = 0; }
    }

    // counters: one atomic per warp
    for (int o = 16; o > 0; o >>= 1) {
        pushes += __shfl_down_sync(kFull, pushes, o);
        dead += __shfl_down_sync(kFull, dead, o);
    }
    if (!kPersistent) {          // 4736 warps and two counters: one atomic per BLOCK (the launches of a small population are short)
        __shared__ unsigned s_pushes;
        __shared__ int s_dead;
        if (threadIdx.x == 0) { s_pushes = 0; s_dead = 0; }
        __syncthreads();
        if (c.lane == 0) {
            if (pushes) atomicAdd(&s_pushes, pushes);
            if (dead) atomicAdd(&s_dead, dead);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            if (s_pushes) atomicAdd(&a.st->pushes, (unsigned long long)s_pushes);
            if (s_dead) atomicAdd(&a.st->n_dead, s_dead);
        }
        return;
    }
    if (c.lane == 0) {
        if (pushes) atomicAdd(&a.st->pushes, (unsigned long long)pushes);
        if (dead) atomicAdd(&a.st->n_dead, dead);
    }
}

// =============================================================================================
// K_poisson_fused (modes 30/33, the headline path): whole Poisson steps in ONE cooperative launch.
// Per step, inside the kernel:
//   generation 0   each claimed chunk gathers its field from charge grid A (updateGrid +
//                  gridToParticles fused, src/grid_operations.cu:29-72), runs all mobility steps
//                  (src/particle_move.cu:22-87), and deposits its survivors into charge grid B
//                  (particlesToGrid of the NEXT step, src/grid_operations.cu:15-26);
//   children       pushed out of the warp's stage, then deposited and stored once;
//   barrier 1      (+ fall-back pass for spilled children, normally none)
//   clear A        (resetGrid restricted to A's bounding box) + list holes/movers
//   barrier 2      compaction moves (remove_dead_particles, src/pic.cu:320-357)
//   barrier 3      the last block to arrive does the host's bookkeeping (src/pic.cu:539-559) and
//                  swaps the roles of the two grids.
// The reference needs 6 kernels, 3 memsets, 1 H2D and 2 blocking D2H copies for the same work
// (src/pic.cu:495-553); here the host enqueues one kernel for any number of Poisson steps.
// =============================================================================================
#include "pic_peer.cuh"

struct FusedArgs {
    Soa s;
    DevState* st;
    const float* tables;      // collision tables + cell boundary table
    int* grid[kGrids];
    int* holes;
    int* movers;
    StepLog* log;
    int log_len;
    int capacity, pt, steps, stop_at_zero;
    double efc;
    uint2 seed;
    unsigned long long* trace;   // optional (PICB200_TRACE): 16 globaltimer stamps per block per Poisson step
    PeerArgs peer;               // kPeer only
    int initial_deposit;         // kPeer only: no summed grid holds this population yet
    Soa park;                    // parking area for step_end_fused's rare case (park_cap electrons)
    int park_cap;
    int adopt_min;               // kAdoptMin, or PICB200_ADOPT_MIN (33: never)
    int compact_policy;          // 0: lazy (default); 1: after every Poisson step (PICB200_COMPACT=eager); 2: lazy, no capacity guard (tests)
    int drain_at;                // staged children from which a warp pushes a batch between two chunks (kFusedDrainAt, PICB200_DRAIN_AT)
    float4* field;               // dense 512^3 per-cell field (field_rows), or nullptr
    int field_ratio;             // a Poisson step computes the field per cell when n >= field_ratio * cells of the occupied box (< 0: never)
    int field_min_n;             // ... and the rank holds at least this many electrons (peer mode: the pass costs a barrier per rank)
    int test_desert;             // PICB200_TEST_DESERT=1 (tests): the last block leaves the kernel at once -> the grid barrier must time out
    PhiloxKeys keys;             // philox_round_keys(seed)
    int warp_fence;              // PICB200_WARP_FENCE=0 switches the per-warp fence at the end of the main phase off (A/B)
    int balance;                 // chunks are dealt to the blocks in proportion to their measured speed (PICB200_BALANCE=0: equal shares)
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)::"memory");      // ("memory": a stamp must not drift across a block barrier)
    return t;
}
#define PIC_TRACE(slot)                                                                             \
    do {                                                                                            \
        if (a.trace && threadIdx.x == 0) a.trace[((size_t)step * gridDim.x + blockIdx.x) * 16 + (slot)] = globaltimer_ns(); \
    } while (0)

// Exact cell index without the fp64 divide of src/grid_operations.cu:21-23: estimate with one fp32
// multiply, then correct against bnd[k] = smallest fp32 x with trunc(double(x)/1e-2) >= k
// (bnd[0] = -inf, bnd[512] = +inf, built on the host; checked for every fp32 in the domain by
// tests/tools/check_cell_index.cpp).  Clamped to the grid like cell_of().
constexpr int kBndBytes = (kGrid + 1) * 4;
// The estimate is taken with a multiplier four ulps below 100, which makes it the true index or one less for every fp32
// in the domain -- never more -- so ONE probe upwards settles it (round 1 estimated with 100.0f and probed both ways:
// three instructions more per coordinate, six coordinates per electron and Poisson step; -DPIC_CELL_TWO_PROBES keeps
// that form).  Both are checked against the reference formula for every fp32 with |x| <= 8 by
// tests/tools/check_cell_index.cpp (-DONE_PROBE: this one).
constexpr float kCellScaleLow = 99.999969482421875f;       // 0x42C7FFFC = 100 - 4 ulp
__device__ __forceinline__ int cell_of_fast(float p, const float* __restrict__ bnd) {
#ifndef PIC_CELL_TWO_PROBES
    int k = __float2int_rz(__fmul_rn(p, kCellScaleLow));
    k = max(0, min(k, kGrid - 1));
    k += (p >= bnd[k + 1]) ? 1 : 0;
    return k;
#else
    int k = __float2int_rz(__fmul_rn(p, 100.0f));
    k = max(0, min(k, kGrid - 1));
    k -= (p < bnd[k]) ? 1 : 0;
    k += (p >= bnd[k + 1]) ? 1 : 0;
    return k;
#endif
}

// Atomics written as the single instruction they are: nvcc wraps an atomicAdd whose address it cannot prove distinct
// per lane in its own warp-aggregation code (vote, leader election, popc, shuffle: ~15 instructions around one
// ATOMS/ATOMG), which is pure overhead where the caller has already elected one lane.  (ptxas still does the same to
// the shared-memory form under a lane predicate -- also to `atom.inc` with an unreachable bound, which it turns back into
// an add; the global `red` is left alone.)
__device__ __forceinline__ int smem_fetch_add(int* p, int v) {
    int old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void red_global_add(int* p, int v) { asm volatile("red.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// NGP deposit of the active lanes with warp-level pre-reduction (lanes that hit the same cell
// are merged into one red.global.add) and a block-level bounding box in shared memory
// (s_box: lo xyz, hi xyz, 16-byte aligned).  The box only grows; nearly every chunk lies inside what the block
// has already seen, so the warp first tests its cells against the box as it stands (a stale read errs on the safe
// side) and reduces its own extent only when some lane falls outside.
__device__ __forceinline__ void deposit_lanes(int* __restrict__ grid, bool active, int ix, int iy, int iz, int lane, int* s_box) {
#ifdef PIC_DEPOSIT_MATCH
    const unsigned cell = active ? (unsigned)cell_index(ix, iy, iz) : (0xffffffffu - (unsigned)lane);
    const unsigned peers = __match_any_sync(kFull, cell);
    if (active && lane == __ffs(peers) - 1) red_global_add(grid + cell, __popc(peers));
#else
    // One red per live lane.  (Round 1 merged lanes that hit the same cell with match.any first; the lanes of a chunk are
    // uncorrelated in space -- 32 electrons among the 2e5 cells of the box share a cell in 0.2 % of the chunks -- so the
    // pre-reduction saved nothing and its latency showed up as 6.5 % of the stall samples of the HBM-side corner.  Lanes
    // that do share a cell are serialised by the L2's atomic unit, which is what happened to the merged red's peers anyway.)
    if (active) red_global_add(grid + cell_index(ix, iy, iz), 1);
#endif
    int4 b0;                                                                 // lo x y z, hi x
    int2 b1;                                                                 // hi y z
    const unsigned box_addr = (unsigned)__cvta_generic_to_shared(s_box);
    asm volatile("ld.volatile.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(b0.x), "=r"(b0.y), "=r"(b0.z), "=r"(b0.w) : "r"(box_addr));
    asm volatile("ld.volatile.shared.v2.s32 {%0,%1}, [%2];" : "=r"(b1.x), "=r"(b1.y) : "r"(box_addr + 16u));
    const bool outside = active & ((ix < b0.x) | (iy < b0.y) | (iz < b0.z) | (ix > b0.w) | (iy > b1.x) | (iz > b1.y));
    if (__any_sync(kFull, outside)) {
        const int lx = __reduce_min_sync(kFull, active ? ix : kGrid), ly = __reduce_min_sync(kFull, active ? iy : kGrid),
                  lz = __reduce_min_sync(kFull, active ? iz : kGrid);
        const int hx = __reduce_max_sync(kFull, active ? ix : -1), hy = __reduce_max_sync(kFull, active ? iy : -1),
                  hz = __reduce_max_sync(kFull, active ? iz : -1);
        if (lane == 0) {
            atomicMin(&s_box[0], lx); atomicMin(&s_box[1], ly); atomicMin(&s_box[2], lz);
            atomicMax(&s_box[3], hx); atomicMax(&s_box[4], hy); atomicMax(&s_box[5], hz);
        }
    }
}

// Grid-wide barrier; the last block to arrive runs `action` (one thread) before everyone is released.
// tr (PICB200_TRACE, barrier 1 only): globaltimer stamps [12] arrived, [13] action done (last block only), [14] released
template <class F, class G>
__device__ __forceinline__ void grid_barrier(DevState* st, unsigned& seq, F action, G fetch, unsigned long long* tr = nullptr) {
    __syncthreads();
    if (threadIdx.x == 0) {
        if (tr) tr[15] = timer_ns();
        __threadfence();
        if (tr) tr[10] = timer_ns();
        const unsigned target = (seq + 1u) * gridDim.x;
        const unsigned prev = atomicAdd(&st->bar_count, 1u);
        if (tr) tr[12] = timer_ns();
        if (prev + 1u == target) {
            __threadfence();
            action();
            if (tr) tr[13] = timer_ns();
            __threadfence();
            atomicExch(&st->bar_release, seq + 1u);
        } else {
            barrier_spin(&st->bar_release, seq + 1u, 2, seq);
        }
        if (tr) tr[14] = timer_ns();
        __threadfence();
        fetch();        // ONE thread per block reads what the next phase needs and leaves it in shared memory
    }
    ++seq;
    __syncthreads();
}
template <class F>
__device__ __forceinline__ void grid_barrier(DevState* st, unsigned& seq, F action) {
    grid_barrier(st, seq, action, []() {});
}

// What one thread per block reads from the control block after a grid barrier, for the whole block.
struct BlockBc {
    int hit_zero, cur, pidx, n;     // start of a Poisson step
    int hw;                         // slots in use (>= n: holes of earlier steps stay in place until a compaction)
    int gen_hi, S, D;               // after barrier 1
    int do_compact, parked;         // after the LAST barrier 1 of a step: does it end with a compaction?
    int H;                          // after barrier 2
    int box_old[6];                 // box of the grid this step empties (lo xyz, hi xyz)
    int box_new[6];                 // peer mode: box of this step's own (local) deposit
    int box_cur[6];                 // box of the grid this step gathers from
    int use_field;                  // this step gathers from the per-cell field (field_rows)
    int peer_box[6], peer_box_valid; // peer mode: union box of the exchange just completed (peer_finish -> fetch_step_start)
};

struct FusedCtx {
    const float* bnd;
    const float4* F;  // per-cell field of this Poisson step (field_rows), or nullptr: difference the counts per electron
    int* A;
    int* B;
    int* s_box;
    double efc;
    bool store_acc;   // last Poisson step of this launch: the frozen acceleration becomes visible to the host
    Soa park;         // finished children whose shifted slot passed the capacity (step_end_fused)
    int park_cap;
};


// updateGrid + gridToParticles for one electron's own cell (src/grid_operations.cu:29-72):
// acc = float(EFC * double(c+ - c-)), absent neighbours at the domain faces count 0.
__device__ __forceinline__ void gather_field_direct(const int* __restrict__ A, double efc, int ix, int iy, int iz, float& ax, float& ay, float& az) {
    const int* q = A + cell_index(ix, iy, iz);
    const int xm = (ix != 0) ? __ldcg(q - 1) : 0, xp = (ix != kGrid - 1) ? __ldcg(q + 1) : 0;
    const int ym = (iy != 0) ? __ldcg(q - kGrid) : 0, yp = (iy != kGrid - 1) ? __ldcg(q + kGrid) : 0;
    const int zm = (iz != 0) ? __ldcg(q - kGrid * kGrid) : 0, zp = (iz != kGrid - 1) ? __ldcg(q + kGrid * kGrid) : 0;
    ax = __double2float_rn(__dmul_rn(efc, (double)(xp - xm)));
    ay = __double2float_rn(__dmul_rn(efc, (double)(yp - ym)));
    az = __double2float_rn(__dmul_rn(efc, (double)(zp - zm)));
}
// Dense plasmas (many electrons per cell): six scattered 32-byte sectors per electron make the L1 -> L2 request
// path the limit of the step (ncu, cfg5 poisson_timestep = 1: l1tex2xbar 81 % busy, 84 % of the load sectors are
// these).  The field is then computed once per CELL of the occupied box -- updateGrid as the reference runs it
// (src/grid_operations.cu:29-56), coalesced row reads, one 16-byte record per cell -- and an electron
// fetches ONE sector.  Same expression, same bits.  Costs one grid barrier per Poisson step, so the host switches
// it on by density (FusedArgs::field_ratio).
__device__ __forceinline__ void field_rows(float4* __restrict__ F, const int* __restrict__ A, const int* box /* lo xyz, hi xyz */,
                                           double efc, int gw, int nw, int lane) {
    const int lx = box[0], ly = box[1], lz = box[2];
    const int wx = box[3] - lx + 1, wy = box[4] - ly + 1, wz = box[5] - lz + 1;
    if (wx <= 0 || wy <= 0 || wz <= 0) return;
    const int rows = wy * wz;
    for (int r = gw; r < rows; r += nw) {
        const int iy = ly + r % wy, iz = lz + r / wy;
        for (int xx = lane; xx < wx; xx += 32) {
            float ax, ay, az;
            gather_field_direct(A, efc, lx + xx, iy, iz, ax, ay, az);
            __stcg(F + cell_index(lx + xx, iy, iz), make_float4(ax, ay, az, 0.f));
        }
    }
}
__device__ __forceinline__ void gather_field(const FusedCtx& f, int ix, int iy, int iz, float& ax, float& ay, float& az) {
    if (f.F) {                                               // uniform over the grid
        const float4 v = __ldcg(f.F + cell_index(ix, iy, iz));
        ax = v.x; ay = v.y; az = v.z;
    } else {
        gather_field_direct(f.A, f.efc, ix, iy, iz, ax, ay, az);
    }
}

// ------------------------------------------------------------------------------------------
// The grid phase as kernels of its own (modes 31/32, the phase-level ABI, the first deposit of a population).
// Round 2: same cell index (single probe of the boundary table, staged in shared memory) and same two gathers as the
// fused kernel -- round 1 ran the reference's double division three times per electron in each of them and always
// took the six-point gather: 190 + 254 us per Poisson step of 1e7 electrons, against 214 us for the whole fused step.
// ------------------------------------------------------------------------------------------
constexpr int kGridPhaseUnroll = 4;
// dense plasma, decided on the device from words every block reads alike: the box of grid `which` and n
__device__ __forceinline__ bool box_is_dense(const DevState* st, int which, int ratio, int* box /* out: lo xyz, hi xyz */) {
    for (int k = 0; k < 3; ++k) { box[k] = ld_state_i32(&st->bbox_lo[which][k]); box[3 + k] = ld_state_i32(&st->bbox_hi[which][k]); }
    const long long cells = (long long)max(0, box[3] - box[0] + 1) * max(0, box[4] - box[1] + 1) * max(0, box[5] - box[2] + 1);
    return ratio >= 0 && cells > 0 && (long long)ld_state_i32(&st->n) >= (long long)ratio * cells;
}
__device__ __forceinline__ void stage_cell_bounds(float* s_bnd, const float* __restrict__ g_bnd) {
    for (int k = threadIdx.x; k <= kGrid; k += blockDim.x) s_bnd[k] = __ldg(g_bnd + k);
    __syncthreads();
}
// K_deposit: particlesToGrid (src/grid_operations.cu:15-26).  Nearest-grid-point count, one red per electron, and a
// block-reduced cell bounding box.
__global__ void k_deposit(Soa s, DevState* st, int* __restrict__ charge, int which, const float* __restrict__ g_bnd) {
    __shared__ int s_lo[3], s_hi[3];
    __shared__ float s_bnd[kGrid + 1];
    const int n = st->n;
    if ((int)(blockIdx.x * blockDim.x) >= n) return;             // small populations: the other blocks stage nothing (whole block, before any barrier)
    if (threadIdx.x < 3) { s_lo[threadIdx.x] = kGrid; s_hi[threadIdx.x] = -1; }
    stage_cell_bounds(s_bnd, g_bnd);
    const int lane = threadIdx.x & 31;
    int lx = kGrid, ly = kGrid, lz = kGrid, hx = -1, hy = -1, hz = -1;
    // four electrons per thread and trip: twelve independent loads in flight (one per thread left the HBM idle half the time)
    const int stride = gridDim.x * blockDim.x;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += kGridPhaseUnroll * stride) {
        float px[kGridPhaseUnroll], py[kGridPhaseUnroll], pz[kGridPhaseUnroll];
#pragma unroll
        for (int u = 0; u < kGridPhaseUnroll; ++u) {
            const int i = i0 + u * stride;
            if (i < n) { px[u] = __ldcg(s.x + i); py[u] = __ldcg(s.y + i); pz[u] = __ldcg(s.z + i); }
        }
#pragma unroll
        for (int u = 0; u < kGridPhaseUnroll; ++u) {
            if (i0 + u * stride < n) {
                const int ix = cell_of_fast(px[u], s_bnd), iy = cell_of_fast(py[u], s_bnd), iz = cell_of_fast(pz[u], s_bnd);
                red_global_add(charge + cell_index(ix, iy, iz), 1);
                lx = min(lx, ix); ly = min(ly, iy); lz = min(lz, iz);
                hx = max(hx, ix); hy = max(hy, iy); hz = max(hz, iz);
            }
        }
    }
    lx = __reduce_min_sync(kFull, lx); ly = __reduce_min_sync(kFull, ly); lz = __reduce_min_sync(kFull, lz);
    hx = __reduce_max_sync(kFull, hx); hy = __reduce_max_sync(kFull, hy); hz = __reduce_max_sync(kFull, hz);
    if (lane == 0 && hx >= 0) {
        atomicMin(&s_lo[0], lx); atomicMin(&s_lo[1], ly); atomicMin(&s_lo[2], lz);
        atomicMax(&s_hi[0], hx); atomicMax(&s_hi[1], hy); atomicMax(&s_hi[2], hz);
    }
    __syncthreads();
    if (threadIdx.x < 3 && s_hi[threadIdx.x] >= 0) {
        atomicMin(&st->bbox_lo[which][threadIdx.x], s_lo[threadIdx.x]);
        atomicMax(&st->bbox_hi[which][threadIdx.x], s_hi[threadIdx.x]);
    }
}
// K_field_box: updateGrid (src/grid_operations.cu:29-56) over the occupied box of grid `which`, when the plasma is dense
__global__ void k_field_box(const DevState* st, const int* __restrict__ charge, float4* __restrict__ F, double efc, int which, int ratio) {
    int box[6];
    if (!box_is_dense(st, which, ratio, box)) return;
    // one thread per cell (field_rows' warp per x-row leaves this launch latency-bound: two dependent trips per warp)
    const int lx = box[0], ly = box[1], lz = box[2];
    const int wx = box[3] - lx + 1, wy = box[4] - ly + 1, wz = box[5] - lz + 1;
    const int cells = wx * wy * wz;                              // <= 512^3 = 2^27
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < cells; k += gridDim.x * blockDim.x) {
        const int x = lx + k % wx, r = k / wx, y = ly + r % wy, z = lz + r / wy;
        float ax, ay, az;
        gather_field_direct(charge, efc, x, y, z, ax, ay, az);
        __stcg(F + cell_index(x, y, z), make_float4(ax, ay, az, 0.f));
    }
}
// K_gather: gridToParticles (src/grid_operations.cu:59-72) from the per-cell field of k_field_box, or -- sparse plasma --
// updateGrid + gridToParticles for the electron's own cell: the six neighbour counts differenced on the spot,
// acc = float(EFC * double(c+ - c-)), absent neighbours at the domain faces count 0.  Same expression, same bits.
// kField: launched with a per-cell field (the host expects a dense plasma; the device decides); without one the six-point
// gather runs with all its loads up front, which costs registers the other instantiation needs for its occupancy.
template <bool kField>
__global__ void k_gather(Soa s, const DevState* st, const int* __restrict__ charge, double efc, const float* __restrict__ g_bnd,
                         const float4* __restrict__ F, int which, int ratio) {
    __shared__ float s_bnd[kGrid + 1];
    const int n = st->n;
    if ((int)(blockIdx.x * blockDim.x) >= n) return;
    stage_cell_bounds(s_bnd, g_bnd);
    int box[6];
    const bool dense = kField && F != nullptr && box_is_dense(st, which, ratio, box);
    const int stride = gridDim.x * blockDim.x;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += kGridPhaseUnroll * stride) {
        float px[kGridPhaseUnroll], py[kGridPhaseUnroll], pz[kGridPhaseUnroll];
#pragma unroll
        for (int u = 0; u < kGridPhaseUnroll; ++u) {
            const int i = i0 + u * stride;
            if (i < n) { px[u] = __ldcg(s.x + i); py[u] = __ldcg(s.y + i); pz[u] = __ldcg(s.z + i); }
        }
        if (dense) {
            float4 v[kGridPhaseUnroll];
#pragma unroll
            for (int u = 0; u < kGridPhaseUnroll; ++u)
                if (i0 + u * stride < n) v[u] = __ldcg(F + cell_index(cell_of_fast(px[u], s_bnd), cell_of_fast(py[u], s_bnd), cell_of_fast(pz[u], s_bnd)));
#pragma unroll
            for (int u = 0; u < kGridPhaseUnroll; ++u) {
                const int i = i0 + u * stride;
                if (i < n) { __stcg(s.ax + i, v[u].x); __stcg(s.ay + i, v[u].y); __stcg(s.az + i, v[u].z); }
            }
        } else if (kField) {
#pragma unroll
            for (int u = 0; u < kGridPhaseUnroll; ++u) {
                const int i = i0 + u * stride;
                if (i < n) {
                    float ax, ay, az;
                    gather_field_direct(charge, efc, cell_of_fast(px[u], s_bnd), cell_of_fast(py[u], s_bnd), cell_of_fast(pz[u], s_bnd), ax, ay, az);
                    __stcg(s.ax + i, ax); __stcg(s.ay + i, ay); __stcg(s.az + i, az);
                }
            }
        } else {
            // all 24 neighbour counts first, then the differences and the stores: taken one electron at a time the stores'
            // memory clobber keeps the next electron's loads behind them -- four L2 round trips in a row per thread
            int d[kGridPhaseUnroll][3];
#pragma unroll
            for (int u = 0; u < kGridPhaseUnroll; ++u) {
                d[u][0] = d[u][1] = d[u][2] = 0;
                if (i0 + u * stride < n) {
                    const int ix = cell_of_fast(px[u], s_bnd), iy = cell_of_fast(py[u], s_bnd), iz = cell_of_fast(pz[u], s_bnd);
                    const int* q = charge + cell_index(ix, iy, iz);
                    const int xm = (ix != 0) ? __ldcg(q - 1) : 0, xp = (ix != kGrid - 1) ? __ldcg(q + 1) : 0;
                    const int ym = (iy != 0) ? __ldcg(q - kGrid) : 0, yp = (iy != kGrid - 1) ? __ldcg(q + kGrid) : 0;
                    const int zm = (iz != 0) ? __ldcg(q - kGrid * kGrid) : 0, zp = (iz != kGrid - 1) ? __ldcg(q + kGrid * kGrid) : 0;
                    d[u][0] = xp - xm; d[u][1] = yp - ym; d[u][2] = zp - zm;
                }
            }
#pragma unroll
            for (int u = 0; u < kGridPhaseUnroll; ++u) {
                const int i = i0 + u * stride;
                if (i < n) {                                     // the expression of gather_field_direct
                    __stcg(s.ax + i, __double2float_rn(__dmul_rn(efc, (double)d[u][0])));
                    __stcg(s.ay + i, __double2float_rn(__dmul_rn(efc, (double)d[u][1])));
                    __stcg(s.az + i, __double2float_rn(__dmul_rn(efc, (double)d[u][2])));
                }
            }
        }
    }
}

// generation-0 chunk of the fused kernel: gather from A, push, deposit into B
// One electron per lane through mobility steps t_first..pt of the current Poisson step, the whole warp in
// lockstep: all lanes run the same step t, so the loop is straight-line code -- the Philox block changes for
// the whole warp at once, lanes that are dead (or, kStaged, not born yet: a child born at step tb takes part
// from tb+1 on) keep computing values nobody stores instead of branching around the body, and the only
// branch is the warp vote "somebody spawned".  updateParticle, src/particle_move.cu:22-87.
//
// kAdopt (chunks of the store): while kAdoptMin or more lanes of the warp are dead, a child is handed to a dead lane
// the moment it is born instead of being staged -- it needs exactly the steps the loop still has to run --
// and later takes over that lane's slot.  In collision-heavy runs (poisson_timestep = 100: every electron is
// replaced 2.8 times per Poisson step) this keeps the lanes busy; the Philox blocks are then refreshed for
// the whole warp, because the new occupants have other ids.  `slot_ok`: the lane has a slot to offer.
constexpr int kAdoptMin = 1;        // measured at cfg5, poisson_timestep = 100: 1 -> 114 ms, 4 -> 122, 8 -> 132, never -> 163
constexpr int kAdoptMinSteps = 24;   // poisson_timestep from which chunks use the adopting loop
constexpr int kStreamMaxSteps = 3;   // poisson_timestep up to which chunks are streamed (k_poisson_fused, kVariant 2)
template <bool kStaged, bool kAdopt, int kSmemRows = kRows>
__device__ __forceinline__ void lockstep_push(const MoverCtx& c, float& x, float& y, float& z, float& vx, float& vy, float& vz,
                                              float& ax, float& ay, float& az, unsigned long long& id, bool& alive,
                                              int tb, const int t_first, int& staged, unsigned& pushes, int& dead,
                                              const bool slot_ok, int& n_adopted) {
    const float dt = (float)1e-12, hdt = (float)(1e-12 / 2);      // src/pic.cu:371
    const float sim = (float)(kGrid * 1e-2);                      // Sim_Size, src/cell.cu:4
    // The Philox block of the first step is worked out in front of the loop: it needs the id only, so its ~50 instructions
    // run while the caller's field gather is still in flight (inside the loop they sat behind the loop header, and the
    // first leapfrog FMA -- i.e. the whole warp -- waited out the gather's latency with nothing to do: 12 % of all stall
    // samples on the HBM-side corner)
    uint4 rnd = philox4x32_10(make_uint4((uint32_t)id, (uint32_t)(id >> 32), (c.g0 + (unsigned)t_first) >> 2, kTagCollide), *c.keys);
    bool refresh = false;
    const bool was_alive = alive;
    int t_end = c.pt;                                             // !kAdopt: the last step this lane's electron is pushed in
#pragma unroll 1
    for (int t = t_first; t <= c.pt; ++t) {
        const unsigned g = c.g0 + (unsigned)t;
        if ((t != t_first && (g & 3u) == 0u) || (kAdopt && refresh)) {       // warp-uniform
            rnd = philox4x32_10(make_uint4((uint32_t)id, (uint32_t)(id >> 32), g >> 2, kTagCollide), *c.keys);
            refresh = false;
        }
        const bool act = kStaged ? (alive & (t > tb)) : alive;
        // leapfrog, src/particle_move.cu:22-39
        float nvx = __fmaf_rn(-ax, hdt, vx), nvy = __fmaf_rn(-ay, hdt, vy), nvz = __fmaf_rn(-az, hdt, vz);
        const float nx = __fmaf_rn(nvx, dt, x), ny = __fmaf_rn(nvy, dt, y), nz = __fmaf_rn(nvz, dt, z);
        nvx = __fmaf_rn(-ax, hdt, nvx); nvy = __fmaf_rn(-ay, hdt, nvy); nvz = __fmaf_rn(-az, hdt, nvz);
        if (!kStaged || act) { x = nx; y = ny; z = nz; vx = nvx; vy = nvy; vz = nvz; }      // (parents: dead lanes just keep computing)
        // checkOutOfBounds, src/particle_move.cu:41-52 (a NaN coordinate is NOT out of bounds)
        const bool oob = (x < 0.f) | (x >= sim) | (y < 0.f) | (y >= sim) | (z < 0.f) | (z >= sim);
        // collider, src/particle_move.cu:55-80
        const uint32_t w01 = (g & 1u) ? rnd.y : rnd.x, w23 = (g & 1u) ? rnd.w : rnd.z;
        const float r = __uint2float_rn((g & 2u) ? w23 : w01);   // the rows hold thresholds on float(w): draw_threshold() on the host
        const float en = __fmaf_rn(vz, vz, __fmaf_rn(vy, vy, __fmul_rn(vx, vx)));
        const float2 th = collision_row<kSmemRows>(c, en);
        const bool coll = act & !oob;
        const bool spawn = coll & (r < th.x);                    // ionisation
        const bool die = act & (oob | (!spawn & (r < th.y)));    // left the domain, or attachment
        if (kAdopt) { if (die) { pushes += (unsigned)(t - tb); ++dead; } }       // it was pushed in steps tb+1 .. t
        else if (die) t_end = t;
        alive = alive & !die;
        unsigned m = __ballot_sync(kFull, spawn);
        if (m) {
            bool stage_it = spawn;
            if (kAdopt) {
                const bool free_lane = slot_ok & !alive;         // dead before this step, or died in it
                const unsigned fm = __ballot_sync(kFull, free_lane);
                if (__popc(fm) >= c.adopt_min) {
                    const int na = min(__popc(m), __popc(fm));   // the first na spawners hand over to the first na free lanes
                    const int my_free = __popc(fm & lanemask_lt());
                    const bool adopt = free_lane && my_free < na;
                    const int src = (int)__fns(m, 0, my_free + 1) & 31;
                    const float cx = __shfl_sync(kFull, x, src), cy = __shfl_sync(kFull, y, src), cz = __shfl_sync(kFull, z, src);
                    const float cvx = __shfl_sync(kFull, vx, src), cvy = __shfl_sync(kFull, vy, src), cvz = __shfl_sync(kFull, vz, src);   // pre-flip
                    const float cax = __shfl_sync(kFull, ax, src), cay = __shfl_sync(kFull, ay, src), caz = __shfl_sync(kFull, az, src);
                    const unsigned long long pid = __shfl_sync(kFull, id, src);
                    if (adopt) {
                        x = cx; y = cy; z = cz; vx = cvx; vy = cvy; vz = cvz; ax = cax; ay = cay; az = caz;
                        id = child_id(pid, (unsigned long long)g);
                        alive = true; tb = t; ++n_adopted;
                    }
                    const int my_spawn = __popc(m & lanemask_lt());
                    stage_it = spawn && my_spawn >= na;
                    refresh = true;                              // new ids in the warp: their Philox blocks, from the next step on
                }
            }
            const unsigned sm = kAdopt ? __ballot_sync(kFull, stage_it) : m;
            if (sm) {
                const int cnt = __popc(sm);
                if (staged + cnt > c.stage_cap) { stage_spill<true>(c, staged); staged = 0; }
                if (stage_it) {                                  // the child copies the pre-flip state
                    uint32_t* q = c.stage + staged + __popc(sm & lanemask_lt());
                    q[0 * kSS] = __float_as_uint(x); q[1 * kSS] = __float_as_uint(y); q[2 * kSS] = __float_as_uint(z);
                    q[3 * kSS] = __float_as_uint(vx); q[4 * kSS] = __float_as_uint(vy); q[5 * kSS] = __float_as_uint(vz);
                    q[6 * kSS] = __float_as_uint(ax); q[7 * kSS] = __float_as_uint(ay); q[8 * kSS] = __float_as_uint(az);
                    q[9 * kSS] = (uint32_t)t;
                    q[10 * kSS] = (uint32_t)id; q[11 * kSS] = (uint32_t)(id >> 32);
                }
                staged += cnt;
            }
            const uint32_t flip = spawn ? 0x80000000u : 0u;      // parent v = -v
            vx = __uint_as_float(__float_as_uint(vx) ^ flip);
            vy = __uint_as_float(__float_as_uint(vy) ^ flip);
            vz = __uint_as_float(__float_as_uint(vz) ^ flip);
        }
    }
    if (kAdopt) {
        if (alive) pushes += (unsigned)(c.pt - tb);
    } else {
        pushes += was_alive ? (unsigned)(t_end - tb) : 0u;
        dead += (was_alive & !alive) ? 1 : 0;
    }
}

template <bool kLaterGen, int kSmemRows = kRows>
__device__ __forceinline__ void fused_chunk(const MoverCtx& c, const FusedCtx& f, int i0, int end, int& staged, unsigned& pushes, int& dead) {
    const int i = i0 + c.lane;
    const bool in = i < end;
    El e{0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int ts = kDead;
    if (in) {
        e.x = __ldcg(c.s.x + i); e.y = __ldcg(c.s.y + i); e.z = __ldcg(c.s.z + i);
        e.vx = __ldcg(c.s.vx + i); e.vy = __ldcg(c.s.vy + i); e.vz = __ldcg(c.s.vz + i);
        ts = __ldcg(c.s.ts + i);
        e.id = __ldcg(c.s.id + i);
        if (kLaterGen) { e.ax = __ldcg(c.s.ax + i); e.ay = __ldcg(c.s.ay + i); e.az = __ldcg(c.s.az + i); }
    }
    const bool was_alive = in && (kLaterGen ? ts >= 0 : ts != kDead);
    if (!kLaterGen && was_alive) {
        // updateGrid + gridToParticles for this electron's own cell (src/grid_operations.cu:29-72)
        const int ix = cell_of_fast(e.x, f.bnd), iy = cell_of_fast(e.y, f.bnd), iz = cell_of_fast(e.z, f.bnd);
        gather_field(f, ix, iy, iz, e.ax, e.ay, e.az);
        if (f.store_acc) { __stcg(c.s.ax + i, e.ax); __stcg(c.s.ay + i, e.ay); __stcg(c.s.az + i, e.az); }
    }
    bool alive = was_alive;
    bool touched = false;
    if (kLaterGen) {                        // spilled children (ts = birth step): the lockstep loop, each from its own step on
        const int t_first = __reduce_min_sync(kFull, was_alive ? ts : 0x3fffffff) + 1;
        int none = 0;
        lockstep_push<true, false, kSmemRows>(c, e.x, e.y, e.z, e.vx, e.vy, e.vz, e.ax, e.ay, e.az, e.id, alive, ts, t_first, staged, pushes, dead, false, none);
        touched = true;
    } else {
        const int t0 = max(1, ts + 1);      // src/pic.cu:218
        advance<true, kSmemRows>(c, e, alive, t0, c.pt, staged, pushes, dead, touched);
    }
    if (was_alive) {
        if (touched) {
            __stcg(c.s.x + i, e.x); __stcg(c.s.y + i, e.y); __stcg(c.s.z + i, e.z);
            __stcg(c.s.vx + i, e.vx); __stcg(c.s.vy + i, e.vy); __stcg(c.s.vz + i, e.vz);
        }
        const int ts1 = alive ? -1 : kDead;
        if (ts1 != ts) __stcg(c.s.ts + i, ts1);
    }
    deposit_lanes(f.B, alive, cell_of_fast(e.x, f.bnd), cell_of_fast(e.y, f.bnd), cell_of_fast(e.z, f.bnd), c.lane, f.s_box);
}


// The common case of the fused kernel: a chunk whose 32 electrons are all "alive since the start of
// the Poisson step" (ts == -1; anything else takes fused_chunk<false>).  All lanes then run the same
// mobility step t, so the step loop is straight-line code: the Philox block changes for the whole
// warp at once, dead lanes keep computing values nobody stores instead of branching around the body,
// and the only branch is the warp vote "somebody spawned".
// cp.async (LDGSTS) copies of one chunk's arrays into a warp-private buffer: no registers are held while the data is in
// flight, so the loads of the NEXT chunk overlap the pushes of the current one.
__device__ __forceinline__ void stream_issue(const MoverCtx& c, int i0, int end, uint32_t* sb) {
    const int i = i0 + c.lane;
    if (i < end) {
        const unsigned d = (unsigned)__cvta_generic_to_shared(sb + c.lane);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(c.s.x + i) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 128u), "l"(c.s.y + i) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 256u), "l"(c.s.z + i) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 384u), "l"(c.s.vx + i) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 512u), "l"(c.s.vy + i) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 640u), "l"(c.s.vz + i) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 768u), "l"(c.s.ts + i) : "memory");
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(sb + 224) + 8u * (unsigned)c.lane), "l"(c.s.id + i) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void stream_wait_older() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void stream_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// kStream: the chunk's eight arrays were copied into the warp's shared-memory buffer `sb` by cp.async while the previous
// chunk was pushed (stream_issue): lane-indexed words x y z vx vy vz ts, then the 64-bit ids.
constexpr int kStreamWordsPerChunk = 9 * 32;
constexpr int kStreamFieldWords = 4 * 32;                       // one float4 per lane: the prefetched field records of a chunk
// kStream, dense plasma (per-cell field): the field records of a chunk whose arrays have landed in `sb` are fetched into the
// warp's field buffer `fb` by cp.async as well -- cells from the buffered positions, one 16-byte record per live lane --
// a whole chunk ahead of their use.  (Fetched when the chunk was pushed, the record's L2/HBM latency was the largest single
// stall of the HBM-side corner: the chain position -> cell -> record -> first leapfrog FMA has nothing to overlap with.)
__device__ __forceinline__ void stream_field_issue(const MoverCtx& c, const FusedCtx& f, int i0, int end, const uint32_t* sb, uint32_t* fb) {
    const int i = i0 + c.lane;
    if (i < end && (int)sb[192 + c.lane] != kDead) {
        const int ix = cell_of_fast(__uint_as_float(sb[c.lane]), f.bnd), iy = cell_of_fast(__uint_as_float(sb[32 + c.lane]), f.bnd),
                  iz = cell_of_fast(__uint_as_float(sb[64 + c.lane]), f.bnd);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(fb + 4 * c.lane)), "l"(f.F + cell_index(ix, iy, iz)) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
struct NoAfterLoad { __device__ __forceinline__ void operator()() const {} };
// after_load (kStream): called once the chunk's buffered words are in registers, i.e. when its buffers may be refilled --
// the caller issues the next prefetches there, in front of the pushes.  fb != nullptr: the field records were prefetched.
template <bool kAdopt, bool kStream, class AfterLoad = NoAfterLoad>
__device__ __forceinline__ void fused_chunk_fast(const MoverCtx& c, const FusedCtx& f, int i0, int end, int& staged, unsigned& pushes, int& dead, int& npool,
                                                 const uint32_t* sb = nullptr, const uint32_t* fb = nullptr, AfterLoad after_load = AfterLoad()) {
    constexpr int kSmemRows = kStream ? kStreamRows : kRows;
    const int i = i0 + c.lane;
    const bool in = i < end;
    float x = 0.f, y = 0.f, z = 0.f, vx = 0.f, vy = 0.f, vz = 0.f;
    unsigned long long id = 0;
    int ts = -1;
    float4 pf = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool prefetched = kStream && fb != nullptr;        // uniform
    if (in) {
        if (kStream) {
            x = __uint_as_float(sb[c.lane]); y = __uint_as_float(sb[32 + c.lane]); z = __uint_as_float(sb[64 + c.lane]);
            vx = __uint_as_float(sb[96 + c.lane]); vy = __uint_as_float(sb[128 + c.lane]); vz = __uint_as_float(sb[160 + c.lane]);
            ts = (int)sb[192 + c.lane];
            id = reinterpret_cast<const unsigned long long*>(sb + 224)[c.lane];
            if (prefetched) pf = reinterpret_cast<const float4*>(fb)[c.lane];      // (a lane without a live electron reads a stale record nobody uses)
        } else {
            x = __ldcg(c.s.x + i); y = __ldcg(c.s.y + i); z = __ldcg(c.s.z + i);
            vx = __ldcg(c.s.vx + i); vy = __ldcg(c.s.vy + i); vz = __ldcg(c.s.vz + i);
            ts = __ldcg(c.s.ts + i);
            id = __ldcg(c.s.id + i);
        }
    }
    if (kStream) { __syncwarp(); after_load(); }
    if (__any_sync(kFull, (ts != -1) & (ts != kDead))) {      // uploaded populations may carry other timestamps
        if (kStream) {
            // the host runs the first launch after an upload with the plain instantiation, which brings every timestamp to -1
            // (this copy of the generic path cost the streaming loop 4.6 %); finding one here is a bug worth a loud failure
            *(volatile int*)&c.st->error = 4;
        } else {
            fused_chunk<false, kSmemRows>(c, f, i0, end, staged, pushes, dead);
            return;
        }
    }
    const bool was_alive = in & (ts != kDead);                // holes of earlier Poisson steps stay in place (lazy compaction)
    // updateGrid + gridToParticles for this electron's own cell (src/grid_operations.cu:29-72)
    float ax = 0.f, ay = 0.f, az = 0.f;
    float x0 = x, y0 = y, z0 = z;
    int ix = 0, iy = 0, iz = 0;
    if (prefetched) {
        ax = pf.x; ay = pf.y; az = pf.z;
        x0 = __int_as_float(0x7fc00000);                      // the cell was worked out a chunk ago and not kept: the deposit works it out again
    } else {
        ix = cell_of_fast(x, f.bnd); iy = cell_of_fast(y, f.bnd); iz = cell_of_fast(z, f.bnd);
        if (was_alive) gather_field(f, ix, iy, iz, ax, ay, az);
    }
    if (was_alive && f.store_acc) { __stcg(c.s.ax + i, ax); __stcg(c.s.ay + i, ay); __stcg(c.s.az + i, az); }
    bool alive = was_alive;
    int n_adopted = 0;                                            // children this lane took over from their parents' lanes
    lockstep_push<false, kAdopt, kSmemRows>(c, x, y, z, vx, vy, vz, ax, ay, az, id, alive, 0, 1, staged, pushes, dead, in, n_adopted);
    if (kAdopt && __any_sync(kFull, n_adopted > 0)) {
        // adopted children were born like any other: the reference's counter n decides whether they exist
        // (src/pic.cu:222).  One atomic per chunk; the lane's LAST adoption has the highest index, the
        // earlier ones have died.  An index >= capacity can only come up in a run that overflows: those
        // children never existed -- their deaths are taken back and a survivor is dropped.
        int incl = n_adopted;
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(kFull, incl, o); if (c.lane >= o) incl += u; }
        int vbase, tbase;
        reserve_fused(c, __shfl_sync(kFull, incl, 31), 0, vbase, tbase);
        const int over = min(n_adopted, max(0, vbase + incl - c.cap));
        if (over > 0) { dead -= over - (alive ? 1 : 0); alive = false; }
    }
    if (in) {
        if (alive) {
            __stcg(c.s.x + i, x); __stcg(c.s.y + i, y); __stcg(c.s.z + i, z);
            __stcg(c.s.vx + i, vx); __stcg(c.s.vy + i, vy); __stcg(c.s.vz + i, vz);
            if (kAdopt && n_adopted > 0) {                        // a child lives in this slot now
                __stcg(c.s.id + i, id);
                if (f.store_acc) { __stcg(c.s.ax + i, ax); __stcg(c.s.ay + i, ay); __stcg(c.s.az + i, az); }
                if (!was_alive) __stcg(c.s.ts + i, -1);
            }
        } else if (was_alive) {
            __stcg(c.s.ts + i, kDead);
        }
    }
    // slots of this chunk whose electron arrived dead or died just now: this warp's children move in (fused_staged)
    {
        const bool free_slot = in & !alive;
        const unsigned fm = __ballot_sync(kFull, free_slot);
        if (fm) {
            const int r = npool + __popc(fm & lanemask_lt());
            if (free_slot && r < kPoolSlots) c.pool[r] = i;
            npool = min(kPoolSlots, npool + __popc(fm));
        }
    }
    // particlesToGrid of the next Poisson step; the cell is recomputed only where the coordinate changed
    if (__any_sync(kFull, alive & ((x != x0) | (y != y0) | (z != z0)))) {
        ix = cell_of_fast(x, f.bnd); iy = cell_of_fast(y, f.bnd); iz = cell_of_fast(z, f.bnd);
    }
    deposit_lanes(f.B, alive, ix, iy, iz, c.lane, f.s_box);
}

// More than 32 children wait in the stage: bring the 32 with the EARLIEST birth steps to the end of the stage
// (entries [k-32, k)) and close the others up at its start, keeping their order.  A batch costs
// poisson_timestep - min(birth step) lockstep steps whatever its lanes hold, so the late-born wait for a
// batch of their own kind.  Warp-uniform; field by field, every entry is read into a register before any is rewritten.
__device__ __forceinline__ void stage_select_earliest(const MoverCtx& c, int k) {
    const int lane = c.lane;
    const bool v1 = lane + 32 < k;                           // (k > 32: entry `lane` always exists)
    const int tb0 = (int)c.stage[9 * kSS + lane], tb1 = v1 ? (int)c.stage[9 * kSS + lane + 32] : 0x7fffffff;
    // T = the 32nd smallest birth step: the smallest T with #(tb <= T) >= 32
    int lo = 0, hi = c.pt;                                   // #(tb <= pt) = k >= 32
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const int cnt = __popc(__ballot_sync(kFull, tb0 <= mid)) + __popc(__ballot_sync(kFull, tb1 <= mid));
        if (cnt >= 32) hi = mid; else lo = mid + 1;
    }
    const int T = lo;
    const unsigned lt = lanemask_lt();
    const unsigned less0 = __ballot_sync(kFull, tb0 < T), less1 = __ballot_sync(kFull, tb1 < T);
    const unsigned eq0 = __ballot_sync(kFull, tb0 == T), eq1 = __ballot_sync(kFull, tb1 == T);
    const int need = 32 - __popc(less0) - __popc(less1);     // entries with tb == T that still fit (>= 1)
    const bool s0 = (tb0 < T) || (tb0 == T && __popc(eq0 & lt) < need);
    const bool s1 = (tb1 < T) || (tb1 == T && __popc(eq0) + __popc(eq1 & lt) < need);
    const unsigned sel0 = __ballot_sync(kFull, s0), sel1 = __ballot_sync(kFull, s1);
    const unsigned val1 = __ballot_sync(kFull, v1);
    const int rest = k - 32;
    // new position: selected entries keep their order behind the others
    const int d0 = s0 ? rest + __popc(sel0 & lt) : __popc(~sel0 & lt);
    const int d1 = s1 ? rest + __popc(sel0) + __popc(sel1 & lt) : __popc(~sel0) + __popc(~sel1 & val1 & lt);
    // the fields are separate arrays: permute them one at a time (two registers instead of twenty-four)
#pragma unroll
    for (int w = 0; w < kStageWords; ++w) {
        uint32_t* fld = c.stage + w * kSS;
        const uint32_t a = fld[lane], b = v1 ? fld[lane + 32] : 0u;
        __syncwarp();
        fld[d0] = a;
        if (v1) fld[d1] = b;
    }
    __syncwarp();
}

// children out of the warp's stage: push, deposit, store once (see process_staged).  At most 32 at a time:
// of more, the earliest-born 32 (stage_select_earliest).
__device__ __forceinline__ void load_stage_entry(const MoverCtx& c, const uint32_t* g, El& e, int& tb) {
    e.x = __uint_as_float(g[0 * kSS]); e.y = __uint_as_float(g[1 * kSS]); e.z = __uint_as_float(g[2 * kSS]);
    e.vx = __uint_as_float(g[3 * kSS]); e.vy = __uint_as_float(g[4 * kSS]); e.vz = __uint_as_float(g[5 * kSS]);
    e.ax = __uint_as_float(g[6 * kSS]); e.ay = __uint_as_float(g[7 * kSS]); e.az = __uint_as_float(g[8 * kSS]);
    tb = (int)g[9 * kSS];
    e.id = staged_child_id(c, g);
}

// k children in registers (lane < k: electron e, born at step tb): slots, remaining pushes, one store, deposit
template <int kSmemRows>
__device__ __forceinline__ void fused_batch(const MoverCtx& c, const FusedCtx& f, El& e, int tb, int k, int& staged, unsigned& pushes, int& dead, int& npool) {
    // the first p children move into free slots this warp holds, the others go above the high-water mark
    const int p = min(k, npool);
    int vslot, tbase;
    reserve_fused(c, k, k - p, vslot, tbase);
    vslot += c.lane;                                         // the reference's index (src/pic.cu:222): the child exists iff it is < capacity
    const int slot = c.lane < p ? c.pool[npool - 1 - c.lane] : tbase + (c.lane - p);
    npool -= p;
    __syncwarp();                                            // pool entries read before anybody adds new ones
    const bool born = c.lane < k && vslot < c.cap;
    const bool in = born && slot < c.cap;
    if (c.lane >= p && c.lane < k && !born && slot < c.cap) __stcg(c.s.ts + slot, kDead);   // (overflow) a slot above the mark nobody was born into: a hole
    bool alive = born;
    const int t_first = __reduce_min_sync(kFull, born ? tb : 0x3fffffff) + 1;
    int none = 0;
    lockstep_push<true, false, kSmemRows>(c, e.x, e.y, e.z, e.vx, e.vy, e.vz, e.ax, e.ay, e.az, e.id, alive, tb, t_first, staged, pushes, dead, false, none);
    if (__any_sync(kFull, born & !in)) {                     // rare: no room above the holes -> parked until this step's compaction
        const bool park = born && !in && alive;
        const unsigned pm = __ballot_sync(kFull, park);
        int pb = 0;
        if (c.lane == 0 && pm) pb = atomicAdd(&c.st->n_parked, __popc(pm));
        pb = __shfl_sync(kFull, pb, 0) + __popc(pm & lanemask_lt());
        if (park) {
            if (pb < f.park_cap) {
                const Soa& q = f.park;
                __stcg(q.x + pb, e.x); __stcg(q.y + pb, e.y); __stcg(q.z + pb, e.z);
                __stcg(q.vx + pb, e.vx); __stcg(q.vy + pb, e.vy); __stcg(q.vz + pb, e.vz);
                __stcg(q.ax + pb, e.ax); __stcg(q.ay + pb, e.ay); __stcg(q.az + pb, e.az);
                __stcg(q.id + pb, e.id);
            } else {
                *(volatile int*)&c.st->error = 2;
            }
        }
    }
    if (in) {
        __stcg(c.s.x + slot, e.x); __stcg(c.s.y + slot, e.y); __stcg(c.s.z + slot, e.z);
        __stcg(c.s.vx + slot, e.vx); __stcg(c.s.vy + slot, e.vy); __stcg(c.s.vz + slot, e.vz);
        if (f.store_acc) { __stcg(c.s.ax + slot, e.ax); __stcg(c.s.ay + slot, e.ay); __stcg(c.s.az + slot, e.az); }
        __stcg(c.s.ts + slot, alive ? -1 : kDead);
        __stcg(c.s.id + slot, e.id);
    }
    deposit_lanes(f.B, alive, cell_of_fast(e.x, f.bnd), cell_of_fast(e.y, f.bnd), cell_of_fast(e.z, f.bnd), c.lane, f.s_box);
}

// the warp's own stage: up to 32 children into registers (of more, the 32 earliest-born); returns their number
__device__ __forceinline__ int staged_load(const MoverCtx& c, int& staged, El& e, int& tb) {
    __syncwarp();
    if (staged > 32) stage_select_earliest(c, staged);
    const int k = min(staged, 32);
    const int rest = staged - k;
    if (c.lane < k) load_stage_entry(c, c.stage + rest + c.lane, e, tb);
    staged = rest;                                           // grandchildren are appended behind the children that still wait
    __syncwarp();
    return k;
}
template <int kSmemRows>
__device__ __forceinline__ void fused_staged(const MoverCtx& c, const FusedCtx& f, int& staged, unsigned& pushes, int& dead, int& npool) {
    El e{0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int tb = 0;
    const int k = staged_load(c, staged, e, tb);
    fused_batch<kSmemRows>(c, f, e, tb, k, staged, pushes, dead, npool);
}

constexpr int kBndSmemBytes = ((kBndBytes + 15) / 16) * 16;
constexpr int kAllTableBytes = kTableSmemBytes + kBndSmemBytes;
__host__ __device__ constexpr int smem_tables_bytes(int R) { return smem_bnd_off(R) + kBndSmemBytes; }      // R == kRows: kAllTableBytes
constexpr int kStreamBytesPerWarp = (2 * kStreamWordsPerChunk + kStreamFieldWords) * 4;  // two chunks in flight per warp + one chunk's field records
// per warp: the children stage (whose first bytes double as the peer exchange's row buffer: that runs between
// barrier 1 and barrier 2, when every stage is empty) and the free-slot pool
static_assert(kPushStageBytesPerWarp <= kStageBytesPerWarp, "the peer row buffer lives inside the children stage");
constexpr int kFusedSmemBytes(int threads, bool stream = false) {
    return smem_tables_bytes(stream ? kStreamRows : kRows) + (threads / 32) * (kStageBytesPerWarp + kPoolSlots * 4 + (stream ? kStreamBytesPerWarp : 0));
}
static_assert(kFusedSmemBytes(1024) + 1024 <= 227 * 1024, "shared memory of the fused kernel");
// both stay inside the 196 KB carve-out (request + static + 1 KB reserved), see "Shared memory against L1" in DESIGN.md
static_assert(kFusedSmemBytes(1024, true) + 512 + 1024 <= 196 * 1024 && kFusedSmemBytes(1024, false) + 512 + 1024 <= 196 * 1024, "196 KB carve-out");
static_assert(smem_tables_bytes(kRows) == kAllTableBytes && smem_tables_bytes(kStreamRows) % 16 == 0, "table windows");
static_assert(kSS >= 32 && kSS <= 64, "stage_select_earliest handles up to 64 entries");

// kAdopt: the chunks use the adopting step loop (lockstep_push); the host picks it for poisson_timestep >= kAdoptMinSteps --
// it carries more state through every step and only pays when lanes die out long before the Poisson step ends.
// Quotas.  Chunks are dealt to the blocks round-robin (chunk k of block b is global chunk k * gridDim + b: the same SM
// sees the same lines every step, which keeps them on its side of the L2), and with equal shares a Poisson step ends
// with its SLOWEST block: on the HBM-side corner the blocks' main phases took 178 / 192 / 206 / 220 us, the same blocks
// every step (correlation 0.99 from step to step: SMs of the larger GPCs share the GPC's port to the L2 with more
// neighbours) -- 13 % of the step spent waiting at the barrier.  So the shares follow the measured speed: every block
// publishes the chunks per microsecond it managed (a running mean); at the start of a step warp 0 of every block
// turns the SAME 148 numbers into the same cumulative shares.  All blocks still take the first `base` = min share
// rounds round-robin; what a block gets beyond that is a contiguous range after those rounds.  Results do not depend
// on who pushes which chunk.  q[0] = base, q[1] = first chunk of this block's extra range, q[2] = its length.
__device__ __forceinline__ void block_quota(const float* __restrict__ rate, int chunks, int lane, int* q) {
    const int G = (int)gridDim.x;
    float r[5];
    bool ok = true;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int b = lane * 5 + j;
        r[j] = b < G ? __ldcg(rate + b) : 0.f;
        ok = ok && (b >= G || r[j] > 0.f);
    }
    if (!__all_sync(kFull, ok)) {                            // no history yet (first step after an upload): equal shares
        if (lane == 0) { q[0] = 0x7fffffff; q[1] = 0; q[2] = 0; }
        return;
    }
    float sum = r[0] + r[1] + r[2] + r[3] + r[4];
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(kFull, sum, o);
    const float mean = sum / (float)G, lo = 0.67f * mean, hi = 1.5f * mean;      // one odd measurement must not starve a block
    float pre[5], acc = 0.f;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int b = lane * 5 + j;
        acc += b < G ? fminf(fmaxf(r[j], lo), hi) : 0.f;
        pre[j] = acc;
    }
    float off = acc;                                         // inclusive scan of the lanes' totals
    for (int o = 1; o < 32; o <<= 1) { const float t = __shfl_up_sync(kFull, off, o); if (lane >= o) off += t; }
    const float total = __shfl_sync(kFull, off, 31);
    off -= acc;                                              // exclusive
    int cum[5], share[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int b = lane * 5 + j;
        cum[j] = b >= G - 1 ? chunks : min(chunks, __float2int_rn((float)chunks * ((off + pre[j]) / total)));
    }
    int prev = __shfl_up_sync(kFull, cum[4], 1);
    if (lane == 0) prev = 0;
    int smin = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int b = lane * 5 + j;
        share[j] = max(0, cum[j] - (j ? cum[j - 1] : prev));
        if (j == 0 && cum[0] < prev) cum[0] = prev;          // (rounding can never run backwards: the prefix is monotone)
        if (b < G) smin = min(smin, share[j]);
    }
    smin = __reduce_min_sync(kFull, smin);
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int b = lane * 5 + j;
        if (b == (int)blockIdx.x) {
            const int before = j ? cum[j - 1] : prev;        // chunks of the blocks before this one
            q[0] = smin; q[1] = smin * G + (before - smin * b); q[2] = share[j] - smin;
        }
    }
}

// kVariant: 0 = plain; 1 = adopt (above); 2 = stream: for small poisson_timestep the step is HBM-side -- a warp then keeps the
// NEXT chunk's arrays in flight (cp.async into a shared-memory double buffer) while it pushes the current one; the room comes
// from a shared-memory window of kStreamRows table rows instead of all of them.
template <bool kPeer, int kVariant>
__global__ void __launch_bounds__(1024, 1) k_poisson_fused(const __grid_constant__ FusedArgs a) {
    constexpr bool kAdopt = kVariant == 1, kStream = kVariant == 2;
    constexpr int kSmemRows = kStream ? kStreamRows : kRows;
    constexpr int kTablesBytes = smem_tables_bytes(kSmemRows);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ __align__(16) int s_box[8];     // lo xyz, hi xyz of this block's deposits (deposit_lanes reads it as int4 + int2)
    __shared__ int s_ticket;       // next chunk of this block (chunks are dealt to blocks round-robin)
    __shared__ int s_quota[4];     // block_quota(): base, first extra chunk, extra chunks; [3] = Poisson step it is valid for, plus one
    __shared__ unsigned long long s_t0;   // start of this block's main phase
    __shared__ float s_rate;              // this block's chunks per microsecond (running mean, see the end of the main phase)
    __shared__ unsigned long long s_pushes;
    __shared__ int s_dead;
    __shared__ BlockBc bc;
    if (threadIdx.x == 0) { s_ticket = 0; s_pushes = 0; s_dead = 0; s_quota[3] = 0; }
    if (!kStream) {
        const float4* src = reinterpret_cast<const float4*>(a.tables);
        float4* dst = reinterpret_cast<float4*>(smem_raw);
        for (int k = threadIdx.x; k < kAllTableBytes / 16; k += blockDim.x) dst[k] = __ldg(src + k);
    } else {        // the first kStreamRows rows of the thresholds and of the edges, then the cell bounds
        const unsigned char* g = reinterpret_cast<const unsigned char*>(a.tables);
        for (int k = threadIdx.x; k < kStreamRows * 8 / 16; k += blockDim.x) reinterpret_cast<float4*>(smem_raw)[k] = __ldg(reinterpret_cast<const float4*>(g) + k);
        for (int k = threadIdx.x; k < kStreamRows * 4 / 16; k += blockDim.x)
            reinterpret_cast<float4*>(smem_raw + smem_edge_off(kSmemRows))[k] = __ldg(reinterpret_cast<const float4*>(g + kThrBytes) + k);
        for (int k = threadIdx.x; k < kBndSmemBytes / 16; k += blockDim.x)
            reinterpret_cast<float4*>(smem_raw + smem_bnd_off(kSmemRows))[k] = __ldg(reinterpret_cast<const float4*>(g + kTableSmemBytes) + k);
    }
    if (threadIdx.x < 6) s_box[threadIdx.x] = threadIdx.x < 3 ? kGrid : -1;
    __syncthreads();

    DevState* st = a.st;
    MoverCtx c;
    c.s = a.s;
    c.st = st;
    c.thr = reinterpret_cast<const float2*>(smem_raw);
    c.edge = reinterpret_cast<const float*>(smem_raw + smem_edge_off(kSmemRows));
    c.g_thr = reinterpret_cast<const float2*>(a.tables);
    c.g_edge = reinterpret_cast<const float*>(reinterpret_cast<const unsigned char*>(a.tables) + kThrBytes);
    c.stage = reinterpret_cast<uint32_t*>(smem_raw + kTablesBytes) + (threadIdx.x >> 5) * (kStageWords * kSS);
    c.cap = a.capacity;
    c.pt = a.pt;
    c.lane = threadIdx.x & 31;
    c.stage_cap = kSS;
    c.stepwise = 0;
    c.adopt_min = a.adopt_min;
    c.seed = a.seed;
    c.keys = &a.keys;
    FusedCtx f;
    f.bnd = reinterpret_cast<const float*>(smem_raw + smem_bnd_off(kSmemRows));
    f.s_box = s_box;
    f.efc = a.efc;
    f.park = a.park;
    f.park_cap = a.park_cap;

    if (a.test_desert && blockIdx.x == gridDim.x - 1) return;    // tests only: a block that never reaches a barrier (barrier_spin's time-out)
    unsigned seq = 0;
    const int gwarp = (int)((threadIdx.x >> 5) * gridDim.x + blockIdx.x), nwarp = (int)(gridDim.x * (blockDim.x >> 5));
    unsigned xe = 0;                                         // kPeer: exchanges completed (same on every rank)
    int* const push_stage = reinterpret_cast<int*>(c.stage);        // peer exchange only, while the stage is empty
    c.pool = reinterpret_cast<int*>(smem_raw + kTablesBytes + (blockDim.x >> 5) * kStageBytesPerWarp) + (threadIdx.x >> 5) * kPoolSlots;
    uint32_t* const sbuf = reinterpret_cast<uint32_t*>(smem_raw + kTablesBytes + (blockDim.x >> 5) * (kStageBytesPerWarp + kPoolSlots * 4)) +
                           (threadIdx.x >> 5) * (kStreamBytesPerWarp / 4);       // kStream only: this warp's two chunk buffers + field buffer
    volatile DevState* const v = st;

    // ---- what ONE thread per block fetches after each grid barrier (BlockBc): every thread asking L2 for
    // the same control words made that one L2 slice the bottleneck of the short phases
    auto load_box = [&](int* dst, const volatile int* lo, const volatile int* hi) {
        for (int k = 0; k < 3; ++k) { dst[k] = lo[k]; dst[3 + k] = hi[k]; }
    };
    bool launch_start = true;                                // every launch starts compacted: hw == n
    auto fetch_step_start = [&](unsigned e_done) {           // e_done: exchanges complete (peer mode)
        bc.hit_zero = v->hit_zero; bc.cur = v->cur; bc.pidx = v->poisson_index; bc.n = v->n;
        bc.hw = launch_start ? bc.n : v->hw;
        s_ticket = 0;
        for (int k = 0; k < 6; ++k) s_box[k] = k < 3 ? kGrid : -1;
        if (kPeer) {                                         // the box of S[e_done&1]: just received, or (start of a launch) kept
            const int p = (int)(e_done & 1u);                //  in PeerLocal, which upload / init do not reset
            const volatile PeerLocal* pl = a.peer.pl;
            for (int k = 0; k < 3; ++k) {
                bc.box_old[k] = bc.peer_box_valid ? bc.peer_box[k] : pl->sbox_lo[p][k];
                bc.box_old[3 + k] = bc.peer_box_valid ? bc.peer_box[3 + k] : pl->sbox_hi[p][k];
            }
            bc.peer_box_valid = 0;
        } else {
            const int old = (bc.cur + 2) % kRot;
            load_box(bc.box_old, v->bbox_lo[old], v->bbox_hi[old]);
        }
        // dense plasma: this step gathers from the per-cell field (field_rows).  The same words in every block: the same
        // decision.  Peer mode: the grid gathered from is S[e_done&1], its box the union box (bc.box_old); the pass and its
        // barrier are paid per rank whatever the slab holds, so the slab must also be large (field_min_n)
        if (!kPeer) load_box(bc.box_cur, v->bbox_lo[bc.cur], v->bbox_hi[bc.cur]);
        else for (int k = 0; k < 6; ++k) bc.box_cur[k] = bc.box_old[k];
        bc.use_field = 0;
        if (a.field && a.field_ratio >= 0) {
            const long long cells = (long long)max(0, bc.box_cur[3] - bc.box_cur[0] + 1) * max(0, bc.box_cur[4] - bc.box_cur[1] + 1) * max(0, bc.box_cur[5] - bc.box_cur[2] + 1);
            bc.use_field = (cells > 0 && (long long)bc.n >= (long long)a.field_ratio * cells && bc.n >= a.field_min_n) ? 1 : 0;
        }
    };
    auto fetch_after_b1 = [&](int nxt) {
        bc.gen_hi = v->gen_hi;
        for (int k = 0; k < 6; ++k) s_box[k] = k < 3 ? kGrid : -1;       // merged into the global box before the barrier
        if (bc.gen_hi >= 0) return;                                     // a fall-back pass comes first
        bc.S = v->scan_S; bc.D = v->scan_D; bc.do_compact = v->do_compact; bc.parked = v->parked;
        if (kPeer) load_box(bc.box_new, v->bbox_lo[nxt], v->bbox_hi[nxt]);     // nxt: the box slot of this step's local deposit
    };
    auto fetch_after_b2 = [&]() { bc.H = min(v->n_holes, v->n_movers); };
    // peer mode.  slot of DevState::bbox that collects the box of the local deposit feeding exchange e (two slots in
    // alternation: the one of the previous exchange is reset at barrier 1, a whole step after everybody has read it)
    auto lslot = [](unsigned e) { return 2 + (int)(e & 1u); };
    // ONE thread per block: wait for exchange e to be complete on this rank, then fetch what the next Poisson step needs.
    // Block 0 also records the exchange for the host and for the next launch.
    auto peer_finish = [&](unsigned e) {
        int ubox[6];
        peer_wait_all(a.peer, e, ubox);                      // all rows of all ranks have been added into S[e&1]
        if (blockIdx.x == 0) {
            const int p = (int)(e & 1u);
            for (int k = 0; k < 3; ++k) {
                a.peer.pl->sbox_lo[p][k] = ubox[k]; a.peer.pl->sbox_hi[p][k] = ubox[3 + k];
                v->bbox_lo[p][k] = ubox[k]; v->bbox_hi[p][k] = ubox[3 + k];
            }
            a.peer.pl->xseq = e;
            v->cur = p;
        }
        bc.peer_box[0] = ubox[0];                            // (fetch_step_start reads S[e&1]'s box from here, not from PeerLocal:
        for (int k = 1; k < 6; ++k) bc.peer_box[k] = ubox[k]; //  only block 0 writes that, for the next launch)
        bc.peer_box_valid = 1;
        fetch_step_start(e);
    };
    if (threadIdx.x == 0) {
        bc.peer_box_valid = 0;
        s_rate = ((int)gridDim.x <= kMaxBlocks) ? __ldcg(&st->rate[ld_volatile_i32(&st->poisson_index) & 1][blockIdx.x]) : 0.f;
    }
    if (kPeer) {
        xe = ld_sys_u32(&a.peer.pl->xseq);
        if (a.initial_deposit) {
            // particlesToGrid of the first Poisson step for a fresh population: local deposit, exchange e
            const unsigned e = xe + 1u;
            const int p = (int)(e & 1u), ls = lslot(e);
            int* mine = a.grid[2];
            const int n = min(ld_volatile_i32(&st->n), a.capacity);
            const int rounded = (n + 31) & ~31;
            for (int i = (gwarp << 5) + c.lane; i < rounded; i += nwarp << 5) {
                const bool in = i < n;
                const float px = in ? __ldcg(c.s.x + i) : 0.f, py = in ? __ldcg(c.s.y + i) : 0.f, pz = in ? __ldcg(c.s.z + i) : 0.f;
                deposit_lanes(mine, in, cell_of_fast(px, f.bnd), cell_of_fast(py, f.bnd), cell_of_fast(pz, f.bnd), c.lane, s_box);
            }
            __syncthreads();
            if (threadIdx.x < 3 && s_box[3 + threadIdx.x] >= 0) {
                atomicMin(&st->bbox_lo[ls][threadIdx.x], s_box[threadIdx.x]);
                atomicMax(&st->bbox_hi[ls][threadIdx.x], s_box[3 + threadIdx.x]);
            }
            grid_barrier(st, seq, []() {}, [&]() { load_box(bc.box_new, v->bbox_lo[ls], v->bbox_hi[ls]); fetch_step_start(xe); });
            if (blockIdx.x == 0 && threadIdx.x == 0) peer_send_box(a.peer, e, bc.box_new);
            peer_add_rows(a.peer, p, a.grid[2], bc.box_new, gwarp, nwarp, c.lane, push_stage);
            // the other summed grid still holds the last exchange of the PREVIOUS population: empty it before
            // this block arrives, the peers add the exchange after this one into it
            clear_box_rows(a.grid[xe & 1u], bc.box_old, gwarp, nwarp, c.lane);
            peer_add_drain(c.lane);
            __syncthreads();
            if (threadIdx.x == 0) {
                peer_arrive(a.peer, e);
                peer_finish(e);
            }
            __syncthreads();
            xe = e;
        }
    }
    if (!(kPeer && a.initial_deposit)) {                     // no barrier has filled the block's copy yet
        if (threadIdx.x == 0) fetch_step_start(xe);
        __syncthreads();
    }
    for (int step = 0; step < a.steps; ++step) {
        if (bc.hit_zero) break;                              // uniform: set before the previous step's last barrier released
        launch_start = false;
        const int cur = bc.cur;
        // single GPU: three grids in rotation (gather / next deposit / being emptied).
        // Peer mode: grid xe&1 holds the sum over the ranks of exchange xe and is gathered from; this rank
        // deposits into its local grid 2, whose rows all ranks add into each other's grid (xe+1)&1 after barrier 1.
        const int nxt = kPeer ? 2 : (cur + 1) % kRot;
        const int nbox = kPeer ? lslot(xe + 1u) : nxt;       // DevState::bbox slot that collects the box of this step's deposit
        const int old = kPeer ? (int)(xe & 1u) : (cur + 2) % kRot;
        f.A = a.grid[kPeer ? old : cur];   // gathered from
        f.B = a.grid[nxt];                 // (local) deposit of the next step
        f.F = nullptr;
        if (bc.use_field) {                // dense plasma: updateGrid over the occupied box first (uniform over the grid)
            field_rows(a.field, f.A, bc.box_cur, a.efc, gwarp, nwarp, c.lane);
            grid_barrier(st, seq, []() {});
            f.F = a.field;
        }
        int* const g_old = a.grid[old];
        // g_old: single GPU: gathered from one step ago, emptied during this step's main phase;
        //        peer mode: the grid gathered from in THIS step, emptied once its gathers are done (barrier 1),
        //        before this rank raises its flag -- the peers add the exchange after next into it
        c.g0 = (unsigned)bc.pidx * (unsigned)a.pt;
        const int hw0 = min(bc.hw, c.cap);                   // slots in use: bc.n live electrons and hw0 - n holes
        const int chunks = (hw0 + 31) >> 5;
        const bool last_step = step == a.steps - 1;
        int staged = 0;
        int npool = 0;                                       // free slots this warp holds (valid within this Poisson step only)
        unsigned pushes = 0;
        int dead = 0;
        PIC_TRACE(0);

        // ---- generation 0 (+ this warp's own children): chunk k of block b is global chunk k*gridDim+b,
        // so every SM gets the same number of chunks to within one; inside the block the warps draw
        // their next chunk from a shared-memory ticket, which evens out the children they pick up
        f.store_acc = last_step;
        // (one copy of each of the two code paths in the instruction stream: the loop runs until the tickets
        // are gone AND the stage is empty)
        // shares by measured speed (block_quota), once the step is long enough for shares to mean anything; until warp 0
        // has worked them out the warps take the rounds every block takes under any quota (shares are clamped to >= 0.44
        // of an equal one)
        const int G = (int)gridDim.x;
        const bool balanced = a.balance != 0 && G <= kMaxBlocks && chunks >= 16 * G;
        const int k_free = balanced ? (chunks / G) * 2 / 5 : 0x7fffffff;
        const int step_tag = step + 1;
        if (threadIdx.x == 0) s_t0 = globaltimer_ns();
        if (balanced && threadIdx.x < 32) {
            block_quota(st->rate[bc.pidx & 1], chunks, c.lane, s_quota);
            __syncwarp();
            if (c.lane == 0) { __threadfence_block(); *(volatile int*)&s_quota[3] = step_tag; }
        }
        auto chunk_of_ticket = [&](int k) {                   // ticket k of this block -> global chunk (0x7fffffff: none left)
            if (k < k_free) return k * G + (int)blockIdx.x;
            while (*(volatile int*)&s_quota[3] != step_tag) __nanosleep(20);      // (ready long before anybody gets here)
            const int base = *(volatile int*)&s_quota[0];
            if (k < base) { const int chunk = k * G + (int)blockIdx.x; return chunk; }
            const int j = k - base;
            return j < *(volatile int*)&s_quota[2] ? *(volatile int*)&s_quota[1] + j : 0x7fffffff;
        };
        auto take_ticket = [&]() {
            int k = 0;
            if (c.lane == 0) k = smem_fetch_add(&s_ticket, 1);
            return chunk_of_ticket(__shfl_sync(kFull, k, 0));
        };
        if (!kStream) {
            for (bool more = true; more || staged;) {
                if (more) {
                    const int chunk = take_ticket();
                    more = chunk < chunks;
                    if (more) fused_chunk_fast<kAdopt, false>(c, f, chunk << 5, hw0, staged, pushes, dead, npool);
                }
                if (staged >= a.drain_at || (!more && staged)) fused_staged<kSmemRows>(c, f, staged, pushes, dead, npool);
            }
        } else {
            // the same loop with the chunk's loads issued one chunk ahead and -- dense plasma -- its field records as soon as
            // its positions have landed (cp.async groups in commit order: arrays k+1, field k | field k+1, arrays k+2 | ...;
            // every iteration starts by waiting for all of them, issued a whole chunk's pushes ago).  Tickets are static
            // here (warp w takes tickets w, w + 32, ...) for most of the step: the shared counter's atomic + shuffle sat in every
            // chunk's dependency chain.
            uint32_t* const fbuf = sbuf + 2 * kStreamWordsPerChunk;
            const bool pf = f.F != nullptr;                      // uniform over the grid
            int tk = (int)(threadIdx.x >> 5);
            const int nwb = (int)(blockDim.x >> 5);
            const int k_static = (((chunks / G) * 3 / 4) / nwb) * nwb;      // ... for the first three quarters of an equal share; the
            auto take_static = [&]() {                                      // rest is drawn from the shared counter, which evens out the tail
                int k = tk;
                if (k < k_static) {
                    tk += nwb;
                } else {
                    if (c.lane == 0) k = k_static + smem_fetch_add(&s_ticket, 1);
                    k = __shfl_sync(kFull, k, 0);
                }
                return chunk_of_ticket(k);
            };
            int cur_chunk = take_static(), next_chunk = 0x7fffffff, b = 0;
            bool have = cur_chunk < chunks, have_next = false;
            if (have) {
                stream_issue(c, cur_chunk << 5, hw0, sbuf);
                next_chunk = take_static();
                have_next = next_chunk < chunks;
                stream_issue(c, have_next ? next_chunk << 5 : hw0, hw0, sbuf + kStreamWordsPerChunk);
                if (pf) {
                    stream_wait_older();
                    __syncwarp();
                    stream_field_issue(c, f, cur_chunk << 5, hw0, sbuf, fbuf);
                }
            }
            while (have || staged) {
                if (have) {
                    stream_wait_all();
                    __syncwarp();
                    const int this_chunk = cur_chunk;
                    uint32_t* const sb_cur = sbuf + b * kStreamWordsPerChunk;
                    uint32_t* const sb_nxt = sbuf + (b ^ 1) * kStreamWordsPerChunk;
                    fused_chunk_fast<false, true>(c, f, this_chunk << 5, hw0, staged, pushes, dead, npool, sb_cur, pf ? fbuf : nullptr, [&]() {
                        // this chunk is in registers: its buffers are free for the chunk after next / the next chunk's field
                        if (pf) stream_field_issue(c, f, have_next ? next_chunk << 5 : hw0, hw0, sb_nxt, fbuf);
                        cur_chunk = next_chunk;
                        have = have_next;
                        if (have_next) {
                            next_chunk = take_static();
                            have_next = next_chunk < chunks;
                        }
                        stream_issue(c, have_next ? next_chunk << 5 : hw0, hw0, sb_cur);
                    });
                    b ^= 1;
                }
                if (staged >= a.drain_at || (!have && staged)) fused_staged<kSmemRows>(c, f, staged, pushes, dead, npool);
            }
        }
        // resetGrid, off the critical path: a warp that has run out of chunks empties its rows of the
        // grid that was gathered from in the PREVIOUS Poisson step (nobody touches it during this one)
        if (!kPeer) clear_box_rows(g_old, bc.box_old, gwarp, nwarp, c.lane);
        // every warp waits for its own stores and deposits to be performed as soon as it is through, while the others are
        // still pushing: the fence in front of the barrier's arrival then has nothing left to wait for (it was 9 us of
        // the last block's -- everybody's -- critical path at cfg2)
        if (a.warp_fence) __threadfence();
        PIC_TRACE(1);          // warp 0 of the block finished its own chunks

        // ---- barrier 1 (+ fall-back passes over spilled children).  The last block to arrive at the LAST
        // barrier 1 of the step does the host's bookkeeping of src/pic.cu:539-559 and decides whether this
        // step ends with remove_dead_particles or leaves its holes in place (step_end_fused)
        int lo = hw0;
        for (;;) {
            for (int o = 16; o > 0; o >>= 1) {
                pushes += __shfl_down_sync(kFull, pushes, o);
                dead += __shfl_down_sync(kFull, dead, o);
            }
            if (c.lane == 0) {                                   // one global atomic per block, not per warp
                if (pushes) atomicAdd(&s_pushes, (unsigned long long)pushes);
                if (dead) atomicAdd(&s_dead, dead);
            }
            pushes = 0; dead = 0;
            __syncthreads();
            PIC_TRACE(2);      // every warp of the block finished
            if (a.balance && threadIdx.x == 0 && lo == hw0 && G <= kMaxBlocks) {
                // chunks per microsecond of this block's main phase for the next step's shares: the plain mean over the first
                // eight Poisson steps of a population, then a running mean with weight 1/8 (one step's measurement carries
                // ~3 % of noise -- children, neighbours -- against a systematic spread of 2 % at cfg2 and 20 % HBM-side)
                const int mine = (balanced && s_quota[0] != 0x7fffffff) ? s_quota[0] + s_quota[2] : (chunks - (int)blockIdx.x + G - 1) / G;
                const float us = (float)(globaltimer_ns() - s_t0) * 1e-3f;
                if (mine >= 8 && us > 0.f) {
                    const float meas = (float)mine / us, w = s_rate > 0.f ? fmaxf(0.125f, 1.0f / (float)(min(bc.pidx, 1 << 20) + 1)) : 1.0f;
                    s_rate += w * (meas - s_rate);
                }
                __stcg(&st->rate[(bc.pidx & 1) ^ 1][blockIdx.x], s_rate);
            }
            if (threadIdx.x == 32) {
                if (s_pushes) atomicAdd(&st->pushes, s_pushes);
                if (s_dead) atomicAdd(&st->n_dead, s_dead);
                s_pushes = 0; s_dead = 0;
            }
            if (threadIdx.x < 3 && s_box[3 + threadIdx.x] >= 0) {
                atomicMin(&st->bbox_lo[nbox][threadIdx.x], s_box[threadIdx.x]);
                atomicMax(&st->bbox_hi[nbox][threadIdx.x], s_box[3 + threadIdx.x]);
            }
            if (a.trace && (threadIdx.x & 31) == 0 && threadIdx.x <= 64)      // (diagnostic stamps 8, 9, 11: warps 0, 1, 2 on their way to the barrier)
                a.trace[((size_t)step * gridDim.x + blockIdx.x) * 16 + (threadIdx.x == 0 ? 8 : threadIdx.x == 32 ? 9 : 11)] = globaltimer_ns();
            grid_barrier(st, seq, [&]() {
                const int unfinished = v->n_unfinished;
                v->n_unfinished = 0;
                if (unfinished > 0) {
                    v->gen_hi = min(v->top_alloc, c.cap);
                    return;
                }
                v->gen_hi = -1;
                step_end_fused(st, c.cap, a.log, a.log_len, a.stop_at_zero, hw0, last_step, a.compact_policy);
                if (!kPeer) {                                    // the grids change roles; the emptied one has no box
                    for (int k = 0; k < 3; ++k) { v->bbox_lo[old][k] = kGrid; v->bbox_hi[old][k] = -1; }
                    v->cur = nxt;
                } else {
                    // reset what the NEXT step will collect into: the other local-box slot (everybody read it a step ago)
                    const int ob = lslot(xe);
                    for (int k = 0; k < 3; ++k) { v->bbox_lo[ob][k] = kGrid; v->bbox_hi[ob][k] = -1; }
                }
            }, [&]() {
                fetch_after_b1(nbox);
                if (!kPeer && bc.gen_hi < 0 && !bc.do_compact) fetch_step_start(0u);     // straight on to the next Poisson step
            }, a.trace ? a.trace + ((size_t)step * gridDim.x + blockIdx.x) * 16 : nullptr);
            const int hi = bc.gen_hi;
            if (hi < 0) break;
            const int rounded = (hi - lo + 31) & ~31;
            for (int k = (blockIdx.x * blockDim.x + threadIdx.x) & ~31;;) {         // (one call site of fused_staged)
                const bool have = k < rounded;
                if (have) { fused_chunk<true, kSmemRows>(c, f, lo + k, hi, staged, pushes, dead); k += gridDim.x * blockDim.x; }
                if (staged >= a.drain_at || (!have && staged)) fused_staged<kSmemRows>(c, f, staged, pushes, dead, npool);
                if (!have && !staged) break;
            }
            lo = hi;
        }
        PIC_TRACE(3);          // barrier 1 passed: every push of this step is done
        const bool compact = bc.do_compact != 0;             // uniform over the grid
        if (!kPeer && !compact) { PIC_TRACE(7); continue; }

        if (kPeer) {
            // every local deposit of this step is in place: hand the rows of this rank's box to the copy engine,
            // which adds them into the summed grid of every rank over NVLink (asynchronous bulk reduce-adds),
            // and empty the grid this step gathered from
            if (blockIdx.x == 0 && threadIdx.x == 0) peer_send_box(a.peer, xe + 1u, bc.box_new);
            peer_add_rows(a.peer, (int)((xe + 1u) & 1u), a.grid[2], bc.box_new, gwarp, nwarp, c.lane, push_stage);
            PIC_TRACE(11);
            clear_box_rows(g_old, bc.box_old, gwarp, nwarp, c.lane);
        }
        PIC_TRACE(8);
        // ---- remove_dead_particles (src/pic.cu:320-357), when this step has one: list the holes below the
        // new n and the live electrons above it
        if (compact) compact_scan_body(c.s, st, bc.S, bc.D, a.holes, a.movers, a.trace ? a.trace + ((size_t)step * gridDim.x + blockIdx.x) * 16 : nullptr);
        // peer mode: the bulk adds were issued a whole scan ago; once they are complete this block arrives at every rank
        if (kPeer) {
            peer_add_drain(c.lane);
            __syncthreads();
            if (threadIdx.x == 0) peer_arrive(a.peer, xe + 1u);
        }
        PIC_TRACE(4);
        if (compact) {
            grid_barrier(st, seq, []() {}, [&]() { fetch_after_b2(); });
            PIC_TRACE(5);
            compact_move_body(c.s, bc.H, a.holes, a.movers);
            PIC_TRACE(6);
            const int parked = bc.parked;
            if (parked > 0) {
                // children that found no room above the holes: the compaction has made it (slots n - parked .. n)
                grid_barrier(st, seq, []() {});
                const int n_new = min(v->n, c.cap) - parked;
                for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < parked; k += gridDim.x * blockDim.x) {
                    const Soa& q = a.park;
                    const int dst = n_new + k;
                    __stcg(c.s.x + dst, __ldcg(q.x + k)); __stcg(c.s.y + dst, __ldcg(q.y + k)); __stcg(c.s.z + dst, __ldcg(q.z + k));
                    __stcg(c.s.vx + dst, __ldcg(q.vx + k)); __stcg(c.s.vy + dst, __ldcg(q.vy + k)); __stcg(c.s.vz + dst, __ldcg(q.vz + k));
                    __stcg(c.s.ax + dst, __ldcg(q.ax + k)); __stcg(c.s.ay + dst, __ldcg(q.ay + k)); __stcg(c.s.az + dst, __ldcg(q.az + k));
                    __stcg(c.s.ts + dst, -1);
                    __stcg(c.s.id + dst, __ldcg(q.id + k));
                }
            }
            // ---- barrier 3: the store is dense again
            grid_barrier(st, seq, [&]() { v->n_holes = 0; v->n_movers = 0; }, [&]() { if (!kPeer) fetch_step_start(0u); });
        }
        if (kPeer) {
            // the one cross-GPU barrier of the step: every block of every rank has added its rows into this rank's S
            if (threadIdx.x == 0) peer_finish(xe + 1u);
            __syncthreads();
            if (!compact) PIC_TRACE(5);
        }
        ++xe;
        PIC_TRACE(7);
    }
}
