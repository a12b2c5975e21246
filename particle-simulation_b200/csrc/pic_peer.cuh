// pic_peer.cuh -- multi-GPU exchange of the charge grid INSIDE the persistent kernel, over NVLink
// peer memory (included inside namespace picb200 by pic_mover.cuh).
//
// The reference has no multi-GPU path (device 0 is hard-coded, src/pic.cu:394).  Here every GPU owns
// a slab of the electrons; the only coupling is the charge grid, whose counts are additive integers
// (src/grid_operations.cu:25).  Each rank deposits its survivors into a LOCAL grid; once per Poisson
// step the ranks meet at a mailbox barrier written straight into each other's memory, agree on the
// union bounding box, and every rank PULLS the boxes of all ranks through its peer mappings and
// writes their sum into the grid it gathers from next.  Integer sums are order-independent, so all
// ranks hold bit-identical fields and the union of the slabs equals the single-GPU run.  No host
// round trip, no separate collective kernel: the whole run stays one cooperative launch per GPU.
#pragma once

constexpr int kMaxRanks = 8;

// One message per (exchange parity, source rank), written by the source rank into every peer's mailbox.
struct alignas(16) PeerMsg {
    // {box lo packed 3 x 10 bits, box hi likewise, exchange number, 0}: ONE 16-byte store, so the box can
    // never be seen without its sequence number and no fence separates them (lo > hi on an axis: empty)
    uint4 box;
    unsigned done;         // last exchange whose pull the sender has finished (end-of-launch handshake)
    unsigned pad[3];
};
struct PeerMailbox { PeerMsg msg[2][kMaxRanks]; };

// Survives picb200_upload / init (the exchange counter must never run backwards: peers compare it).
struct PeerLocal {
    unsigned xseq;         // exchanges completed by this rank
    int error;             // 1: a peer did not show up within the time-out
};

struct PeerArgs {
    int rank, nranks;
    PeerMailbox* mail[kMaxRanks];        // mailbox of rank r as mapped into this GPU's address space
    const int* loc[kMaxRanks][2];        // the two local deposit grids of rank r, likewise
    PeerLocal* pl;
};

__device__ __forceinline__ void st_sys_u32(unsigned* p, unsigned v) { asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_sys_i32(int* p, int v) { asm volatile("st.volatile.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int ld_sys_i32(const int* p) {
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long peer_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void st_sys_v4(uint4* p, uint4 v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_sys_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
constexpr unsigned long long kPeerTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;

// Executed by ONE thread of the rank (the last block to arrive at the local grid barrier, i.e. after
// every local deposit of exchange `e` is in place): publish this rank's box to all peers, wait for
// theirs, and leave the union box in st->xbox.
__device__ __forceinline__ void peer_exchange_boxes(DevState* st, const PeerArgs& P, int loc_grid, unsigned e) {
    volatile DevState* v = st;
    int lo[3], hi[3];
    for (int k = 0; k < 3; ++k) { lo[k] = v->bbox_lo[loc_grid][k]; hi[k] = v->bbox_hi[loc_grid][k]; }
    const bool mine = hi[0] >= lo[0];
    // empty box: lo = 1, hi = 0 on every axis
    const uint4 msg = make_uint4(mine ? (unsigned)(lo[0] | (lo[1] << 10) | (lo[2] << 20)) : 0x00100401u,
                                 mine ? (unsigned)(hi[0] | (hi[1] << 10) | (hi[2] << 20)) : 0u, e, 0u);
    __threadfence_system();                                   // the deposits of this GPU before the message
    for (int r = 0; r < P.nranks; ++r)
        if (r != P.rank) st_sys_v4(&P.mail[r]->msg[e & 1u][P.rank].box, msg);
    const unsigned long long t0 = peer_timer_ns();
    for (int r = 0; r < P.nranks; ++r) {
        if (r == P.rank) continue;
        const PeerMsg* m = &P.mail[P.rank]->msg[e & 1u][r];
        uint4 got = ld_sys_v4(&m->box);
        while ((int)(got.z - e) < 0) {
            if (peer_timer_ns() - t0 > kPeerTimeoutNs) { P.pl->error = 1; break; }
            got = ld_sys_v4(&m->box);
        }
        for (int k = 0; k < 3; ++k) {
            const int pl = (int)((got.x >> (10 * k)) & 1023u), ph = (int)((got.y >> (10 * k)) & 1023u);
            if (pl <= ph) { lo[k] = min(lo[k], pl); hi[k] = max(hi[k], ph); }
        }
    }
    __threadfence_system();
    for (int k = 0; k < 3; ++k) { v->xbox_lo[k] = lo[k]; v->xbox_hi[k] = hi[k]; }
}

// End-of-launch handshake (ONE thread): nobody may reuse or free a local grid while a peer still pulls
// from it.  `e` = the last exchange of this launch, whose pull this rank has completed.
__device__ __forceinline__ void peer_finish(const PeerArgs& P, unsigned e) {
    __threadfence_system();
    for (int r = 0; r < P.nranks; ++r)
        if (r != P.rank) st_sys_u32(&P.mail[r]->msg[e & 1u][P.rank].done, e);
    const unsigned long long t0 = peer_timer_ns();
    for (int r = 0; r < P.nranks; ++r) {
        if (r == P.rank) continue;
        const PeerMsg* m = &P.mail[P.rank]->msg[e & 1u][r];
        while ((int)(ld_sys_u32(&m->done) - e) < 0) {
            if (peer_timer_ns() - t0 > kPeerTimeoutNs) { P.pl->error = 1; break; }
            __nanosleep(100);
        }
    }
    __threadfence_system();
}

// All warps of the rank: sum the local grids of every rank over the union box into `sum_grid`, zero
// what is left of the previous contents of `sum_grid` (its old box), one warp per x-row.  Split in two
// so that the NVLink round trip hides behind the compaction scan: peer_pull_issue() starts the loads of
// the warp's first row (two cells per lane per rank), peer_pull_finish() adds them up, stores, and
// handles whatever a larger box leaves over.
struct PeerPull {
    int lo[3], hi[3], hl[3], hh[3];     // union box of this exchange; hull of it and the old box of sum_grid
    bool any, work;
    int v0[kMaxRanks], v1[kMaxRanks];
};
__device__ __forceinline__ bool peer_cell_in(const PeerPull& p, int x, int y, int z) {
    return p.any && x >= p.lo[0] && x <= p.hi[0] && y >= p.lo[1] && y <= p.hi[1] && z >= p.lo[2] && z <= p.hi[2];
}
__device__ __forceinline__ void peer_pull_issue(const DevState* st, const PeerArgs& P, int parity, int sum_idx, int gw, int lane, PeerPull& p) {
    p.any = true;
    bool old_any = true;
    for (int k = 0; k < 3; ++k) {
        p.lo[k] = __ldcg(&st->xbox_lo[k]); p.hi[k] = __ldcg(&st->xbox_hi[k]);
        p.hl[k] = __ldcg(&st->bbox_lo[sum_idx][k]); p.hh[k] = __ldcg(&st->bbox_hi[sum_idx][k]);
    }
    for (int k = 0; k < 3; ++k) { p.any = p.any && p.lo[k] <= p.hi[k]; old_any = old_any && p.hl[k] <= p.hh[k]; }
    p.work = p.any || old_any;
    if (!old_any) { for (int k = 0; k < 3; ++k) { p.hl[k] = p.lo[k]; p.hh[k] = p.hi[k]; } }
    else if (p.any) { for (int k = 0; k < 3; ++k) { p.hl[k] = min(p.hl[k], p.lo[k]); p.hh[k] = max(p.hh[k], p.hi[k]); } }
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q) { p.v0[q] = 0; p.v1[q] = 0; }
    if (!p.work) return;
    const int wx = p.hh[0] - p.hl[0] + 1, wy = p.hh[1] - p.hl[1] + 1, wz = p.hh[2] - p.hl[2] + 1;
    if (gw >= wy * wz) return;
    const int y = p.hl[1] + gw % wy, z = p.hl[2] + gw / wy;
    const size_t base = cell_index(p.hl[0], y, z);
    const bool in0 = lane < wx && peer_cell_in(p, p.hl[0] + lane, y, z);
    const bool in1 = lane + 32 < wx && peer_cell_in(p, p.hl[0] + lane + 32, y, z);
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q) {
        if (q < P.nranks) {
            if (in0) p.v0[q] = __ldcg(P.loc[q][parity] + base + lane);
            if (in1) p.v1[q] = __ldcg(P.loc[q][parity] + base + lane + 32);
        }
    }
}
__device__ __forceinline__ void peer_pull_finish(const PeerArgs& P, int parity, int* __restrict__ sum_grid, int gw, int nw, int lane, const PeerPull& p) {
    if (!p.work) return;
    const int wx = p.hh[0] - p.hl[0] + 1, wy = p.hh[1] - p.hl[1] + 1, wz = p.hh[2] - p.hl[2] + 1;
    const int rows = wy * wz;
    for (int r = gw; r < rows; r += nw) {
        const int y = p.hl[1] + r % wy, z = p.hl[2] + r / wy;
        const size_t base = cell_index(p.hl[0], y, z);
        int xx = lane;
        if (r == gw) {                                       // the cells whose loads were issued early
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int q = 0; q < kMaxRanks; ++q) { s0 += p.v0[q]; s1 += p.v1[q]; }
            if (lane < wx) __stcg(sum_grid + base + lane, s0);
            if (lane + 32 < wx) __stcg(sum_grid + base + lane + 32, s1);
            xx = lane + 64;
        }
        for (; xx < wx; xx += 32) {
            int sum = 0;
            if (peer_cell_in(p, p.hl[0] + xx, y, z))
                for (int q = 0; q < P.nranks; ++q) sum += __ldcg(P.loc[q][parity] + base + xx);
            __stcg(sum_grid + base + xx, sum);
        }
    }
}
