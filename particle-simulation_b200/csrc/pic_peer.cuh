// pic_peer.cuh -- multi-GPU exchange of the charge grid INSIDE the persistent kernel, over NVLink
// peer memory (included inside namespace picb200 by pic_mover.cuh).
//
// The reference has no multi-GPU path (device 0 is hard-coded, src/pic.cu:394).  Here every GPU owns
// a slab of the electrons; the only coupling is the charge grid, whose counts are additive integers
// (src/grid_operations.cu:25).  Every rank deposits its survivors into a LOCAL grid.  Once the local
// deposits of a Poisson step are complete (grid barrier 1), each warp stages one x-row of the occupied box in
// shared memory and hands the copy engine one bulk REDUCE-ADD (cp.reduce.async.bulk ... .add.s32, TMA) per
// rank: the row is added, by the L2 atomic units at the destination, into the summed grid of EVERY rank
// (its own included) over NVLink.  There is neither a staging copy per rank nor a summing pass: when every
// block of every rank has arrived, the grid IS the sum.  Integer adds commute, so all ranks hold bit-identical
// fields and the union of the slabs equals the single-GPU run.  No host round trip, no separate collective
// kernel: the whole run stays one cooperative launch per GPU.
//
// Synchronisation (round 2): ONE cross-GPU barrier per exchange, counted by remote atomics -- no system-scope
// grid barrier, no mailbox round trip, and no system-scope fence: a phase trace with the bulk adds switched off
// (PICB200_PEER_DEBUG=3) showed that of the 23 us the exchange cost per Poisson step, 11 us were the
// `fence.sys` in front of the arrival and 8 us the `red.release.sys` + the acquire.sys polling -- not the
// megabyte on the wire.  A block whose lanes have waited for their bulk adds (`cp.async.bulk.wait_group 0`: the
// adds have been PERFORMED at the destination's L2, the point of coherence of that memory, and acknowledged)
// meets at a block barrier; one thread then does a relaxed `red.sys.add` on the arrival counter of every peer and a
// gpu-scope release-add on its own rank's counter.  The arrival cannot overtake data that has already been
// performed.  A block may go on to the next Poisson step as soon as ITS rank's counter shows that all blocks of all
// ranks have arrived (relaxed polling of a word in local memory, one gpu-scope fence after it).  What the peers
// later add into this rank's memory (exchange e+2 into the rows this block has just zeroed) arrives at this rank's
// L2 at least one Poisson step after the gpu-scope release.  PICB200_PEER_FENCE=sys restores system-scope fences on
// both sides.  The ranks' deposit boxes travel as ONE 8-byte word per source rank (box + exchange number: a
// single-copy-atomic store, so a box is never seen without its number), written by one thread right after barrier 1.
//
// Buffers per rank: L = its local deposit (grid 2), S[0], S[1] = summed grids of even / odd exchanges
// (grids 0, 1; peers add into them).  Exchange e fills S[e&1]; the step after it gathers from S[e&1] and, once
// its gathers are done (barrier 1), every block zeroes its rows of it BEFORE arriving for exchange e+1 -- so a
// peer that has seen all arrivals of exchange e+1 knows S[(e+2)&1] is clean and may add its exchange e+2 rows
// at any later time.  The same argument covers the box words (reset at barrier 1 of the step that follows
// their use) and L (a block has read and zeroed its rows of L before it arrives, and nobody deposits into L
// for the next step before all blocks of the own rank have arrived).
#pragma once

constexpr int kMaxRanks = 8;

// One per rank, in that rank's memory, mapped into every peer's address space (CUDA IPC / peer access).
// Everything in it is written by remote atomics of the peers (and reset by the owner at a time when no peer can
// be writing, see above).
struct PeerSync {
    alignas(128) unsigned arrive[2];                   // per exchange parity: arrivals so far = (exchanges of that parity) x ranks x blocks
    alignas(128) unsigned long long box[2][kMaxRanks]; // per exchange parity and SOURCE rank: {lo xyz, hi xyz: 9 bits each, exchange number: 10 bits}
};

// Owned by the rank, survives picb200_upload / init (the exchange counter must never run backwards).
struct PeerLocal {
    unsigned xseq;          // exchanges completed by this rank
    int error;              // (unused since the arrival counters: a time-out traps, see barrier_timeout_abort)
    int sbox_lo[2][3];      // what S[0], S[1] hold: the union box of the last exchange of that parity
    int sbox_hi[2][3];
};

struct PeerArgs {
    int rank, nranks;
    PeerSync* sync[kMaxRanks];      // sync block of rank r as mapped into this GPU's address space (own rank: local pointer)
    int* sum[kMaxRanks][2];         // summed grids S[0], S[1] of rank r, likewise
    int* mc[2];                     // S[0], S[1] of ALL ranks at once: the multicast mapping (NVSwitch reduction, pic_nvls.h), or nullptr
    PeerLocal* pl;
    int strict;                     // PICB200_PEER_FENCE=sys: system-scope fences around the arrival (see above)
    int debug;                      // PICB200_PEER_DEBUG (timing experiments only, results are wrong): 1 = skip the add into the own grid, 2 = skip the adds into the peers' grids
};

__device__ __forceinline__ unsigned ld_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add(unsigned* p, unsigned v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_relaxed_sys_add(unsigned* p, unsigned v) {
    asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_relaxed_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

constexpr int kPushSegCells = 64;                       // cells staged per bulk operation (256 B per warp)
constexpr int kPushStageBytesPerWarp = kPushSegCells * 4;

__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// All warps of the rank, after the local deposits of this step are complete: add the occupied box of the
// local grid L into S[p] of every rank, one warp per x-row, and leave L empty for the next step.
// x-ranges are widened to multiples of 4 cells (bulk operations need 16-byte alignment and size; the extra
// cells of L are zero).  Destinations are visited starting with the next rank, so that at any moment the ranks
// send to different peers.
__device__ __forceinline__ void peer_add_rows(const PeerArgs& P, int p, int* __restrict__ L, const int* box /* lo xyz, hi xyz */, int gw, int nw,
                                              int lane, int* stage /* smem, per warp */) {
    if (box[3] < box[0] || box[4] < box[1] || box[5] < box[2]) return;
    const int lx = box[0] & ~3, hx = box[3] | 3;
    const int wx = hx - lx + 1, wy = box[4] - box[1] + 1, wz = box[5] - box[2] + 1;
    const int rows = wy * wz;
    if (P.mc[p]) {
        // NVSwitch reduction: one multimem.red per pair of cells adds them into the summed grid of EVERY rank (this one
        // included) -- the row leaves this GPU once, the switch fans it out.  Two 32-bit counts travel as one 64-bit add:
        // the low count cannot carry into the high one (a cell never holds 2^32 electrons).  Fire and forget: the
        // block's arrival (peer_arrive) is preceded by a system-scope fence that covers these.
        int* const M = P.mc[p];
        for (int r = gw; r < rows; r += nw) {
            const int y = box[1] + r % wy, z = box[2] + r / wy;
            const size_t base = cell_index(lx, y, z);
            for (int x0 = 2 * lane; x0 < wx; x0 += 64) {
                int2* src = reinterpret_cast<int2*>(L + base + x0);
                const int2 v = __ldcg(src);
                if (v.x | v.y) {
                    __stcg(src, make_int2(0, 0));                            // resetGrid of the local grid, on the way
                    const unsigned long long w = (unsigned long long)(unsigned)v.x | ((unsigned long long)(unsigned)v.y << 32);
                    asm volatile("multimem.red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(M + base + x0), "l"(w) : "memory");
                }
            }
        }
        return;
    }
    const unsigned stage_addr = (unsigned)__cvta_generic_to_shared(stage);
    for (int r = gw; r < rows; r += nw) {
        const int y = box[1] + r % wy, z = box[2] + r / wy;
        const size_t base = cell_index(lx, y, z);
        for (int x0 = 0; x0 < wx; x0 += kPushSegCells) {
            const int cells = min(kPushSegCells, wx - x0);               // multiple of 4
            if (lane == 0) bulk_wait_read();                             // the staging buffer is free again
            __syncwarp();
            if (2 * lane < cells) {
                int2* src = reinterpret_cast<int2*>(L + base + x0 + 2 * lane);
                *reinterpret_cast<int2*>(stage + 2 * lane) = __ldcg(src);
                __stcg(src, make_int2(0, 0));                            // resetGrid of the local grid, on the way
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> async proxy
            __syncwarp();
            if (lane == 0) {
                for (int k = 1; k <= P.nranks; ++k) {
                    const int q = (P.rank + k) % P.nranks;               // (the own grid last)
                    if (P.debug && ((q == P.rank) ? (P.debug & 1) : (P.debug & 2))) continue;
                    int* dst = P.sum[q][p] + base + x0;
                    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.s32 [%0], [%1], %2;" ::"l"(dst), "r"(stage_addr), "r"(cells * 4) : "memory");
                }
                bulk_commit();
            }
        }
    }
}
// Every lane that issued bulk operations (lane 0 of each warp): wait until they are complete.
__device__ __forceinline__ void peer_add_drain(int lane) {
    if (lane == 0) bulk_wait_all();
}

// ONE thread of the rank, right after barrier 1 (the box of the local deposit is final): this rank's box and the exchange
// number as one word into every rank's sync block, its own included.
__device__ __forceinline__ void peer_send_box(const PeerArgs& P, unsigned e, const int* box /* lo xyz, hi xyz of the local deposit */) {
    const bool mine = box[3] >= box[0] && box[4] >= box[1] && box[5] >= box[2];
    unsigned long long w = mine ? ((unsigned long long)(box[0] | (box[1] << 9) | (box[2] << 18)) | ((unsigned long long)(box[3] | (box[4] << 9) | (box[5] << 18)) << 27))
                                : 1ull;                                  // lo x = 1, hi x = 0: empty
    w |= (unsigned long long)(e & 1023u) << 54;
    for (int k = 1; k <= P.nranks; ++k) st_relaxed_sys_u64(&P.sync[(P.rank + k) % P.nranks]->box[e & 1u][P.rank], w);
}
// ONE thread per block, after a __syncthreads() that follows peer_add_drain() and the block's share of the clearing
// of the old summed grid: this block's contribution to exchange `e` is complete.
__device__ __forceinline__ void peer_arrive(const PeerArgs& P, unsigned e) {
    const int p = (int)(e & 1u);
    if (P.strict || P.mc[p]) __threadfence_system();             // (multimem.red has no completion to wait for: the fence is the wait)
    for (int k = 1; k < P.nranks; ++k) red_relaxed_sys_add(&P.sync[(P.rank + k) % P.nranks]->arrive[p], 1u);
    red_release_gpu_add(&P.sync[P.rank]->arrive[p], 1u);        // the block's plain stores (rows of L and of the old grid zeroed) before its own rank goes on
}
// Arrivals this rank's counter of parity e&1 shows once exchange e is complete everywhere: exchanges 1..e of that
// parity, each with one arrival per block of every rank (the ranks run equal grids).
__device__ __forceinline__ unsigned peer_arrival_target(const PeerArgs& P, unsigned e) {
    const unsigned rounds = (e + (e & 1u)) >> 1;        // e odd: (e+1)/2 odd exchanges so far; e even: e/2 even ones
    return rounds * (unsigned)P.nranks * gridDim.x;
}
// ONE thread per block: wait until every block of every rank has arrived for exchange `e`, i.e. until all rows have
// been added into this rank's S[e&1]; returns the union of the ranks' boxes.
__device__ __forceinline__ void peer_wait_all(const PeerArgs& P, unsigned e, int* box /* out: lo xyz, hi xyz */) {
    const int p = (int)(e & 1u);
    const PeerSync* s = P.sync[P.rank];
    const unsigned target = peer_arrival_target(P, e);
    unsigned long long t0 = 0;
    unsigned spins = 0;
    while ((int)(ld_relaxed_sys_u32(&s->arrive[p]) - target) < 0) {
        if (spins == 0) t0 = timer_ns();
        __nanosleep(20);
        if ((++spins & 255u) == 0u && timer_ns() - t0 > g_barrier_timeout_ns) barrier_timeout_abort(3, e);
    }
    for (int a = 0; a < 3; ++a) { box[a] = kGrid; box[3 + a] = -1; }
    for (int q = 0; q < P.nranks; ++q) {
        unsigned long long w = ld_relaxed_sys_u64(&s->box[p][q]);
        while ((unsigned)(w >> 54) != (e & 1023u)) {                 // (sent microseconds before that rank's first arrival: normally there)
            if (spins == 0) t0 = timer_ns();
            if ((++spins & 255u) == 0u && timer_ns() - t0 > g_barrier_timeout_ns) barrier_timeout_abort(3, e);
            w = ld_relaxed_sys_u64(&s->box[p][q]);
        }
        const int lo = (int)(w & 0x7ffffffull), hi = (int)((w >> 27) & 0x7ffffffull);
        if ((lo & 511) <= (hi & 511)) {
            for (int a = 0; a < 3; ++a) { box[a] = min(box[a], (lo >> (9 * a)) & 511); box[3 + a] = max(box[3 + a], (hi >> (9 * a)) & 511); }
        }
    }
    if (P.strict) __threadfence_system(); else __threadfence();     // acquire: what follows reads the summed grid
}
