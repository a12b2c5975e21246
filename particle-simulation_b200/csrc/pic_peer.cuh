// pic_peer.cuh -- multi-GPU exchange of the charge grid INSIDE the persistent kernel, over NVLink
// peer memory (included inside namespace picb200 by pic_mover.cuh).
//
// The reference has no multi-GPU path (device 0 is hard-coded, src/pic.cu:394).  Here every GPU owns
// a slab of the electrons; the only coupling is the charge grid, whose counts are additive integers
// (src/grid_operations.cu:25).  Every rank deposits its survivors into a LOCAL grid.  Once the local
// deposits of a Poisson step are complete, each warp stages one x-row of the occupied box in shared
// memory and hands the copy engine one bulk REDUCE-ADD (cp.reduce.async.bulk ... .add.s32, TMA) per
// rank: the row is added, by the L2 atomic units at the destination, into the summed grid of EVERY rank
// (its own included) over NVLink.  The operation is asynchronous -- the warp goes straight on to the
// compaction while the megabyte is in flight (plain peer loads or stores held the issuing warps for
// 13 us per step) -- and there is neither a staging copy per rank nor a summing pass: when the flags of
// all ranks are in, the grid IS the sum.  Integer adds commute, so all ranks hold bit-identical fields
// and the union of the slabs equals the single-GPU run.  No host round trip, no separate collective
// kernel: the whole run stays one cooperative launch per GPU.
//
// Buffers per rank: L = its local deposit (grid 2), S[0], S[1] = summed grids of even / odd exchanges
// (grids 0, 1; peers add into them).  Exchange e fills S[e&1]; the step after it gathers from S[e&1] and,
// once its gathers are done (barrier 1), zeroes it again BEFORE raising the flag of exchange e+1 -- so a
// peer that sees flag e+1 knows S[(e+2)&1] is clean and may add its exchange e+2 rows at any later time.
#pragma once

constexpr int kMaxRanks = 8;

// One message per (exchange parity, source rank), written by the source rank into every peer's mailbox:
// {box lo packed 3 x 10 bits, box hi likewise, exchange number, 0} in ONE 16-byte store, so the box can
// never be seen without its sequence number (lo > hi on an axis: empty box).
struct alignas(16) PeerMsg { uint4 box; };
struct PeerMailbox { PeerMsg msg[2][kMaxRanks]; };

// Survives picb200_upload / init (the exchange counter must never run backwards: peers compare it).
struct PeerLocal {
    unsigned xseq;          // exchanges completed by this rank
    int error;              // 1: a peer did not show up within the time-out
};

struct PeerArgs {
    int rank, nranks;
    PeerMailbox* mail[kMaxRanks];   // mailbox of rank r as mapped into this GPU's address space
    int* sum[kMaxRanks][2];         // summed grids S[0], S[1] of rank r, likewise (own rank: local pointers)
    PeerLocal* pl;
};

__device__ __forceinline__ unsigned ld_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
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

constexpr int kPushSegCells = 64;                       // cells staged per bulk operation (256 B per warp)
constexpr int kPushStageBytesPerWarp = kPushSegCells * 4;

__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// All warps of the rank, after the local deposits of this step are complete: add the occupied box of the
// local grid L into S[p] of every rank, one warp per x-row, and leave L empty for the next step.
// x-ranges are widened to multiples of 4 cells (bulk operations need 16-byte alignment and size; the extra
// cells of L are zero).
__device__ __forceinline__ void peer_add_rows(const PeerArgs& P, int p, int* __restrict__ L, const int* box /* lo xyz, hi xyz */, int gw, int nw,
                                              int lane, int* stage /* smem, per warp */) {
    if (box[3] < box[0] || box[4] < box[1] || box[5] < box[2]) return;
    const int lx = box[0] & ~3, hx = box[3] | 3;
    const int wx = hx - lx + 1, wy = box[4] - box[1] + 1, wz = box[5] - box[2] + 1;
    const int rows = wy * wz;
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
                for (int q = 0; q < P.nranks; ++q) {
                    int* dst = P.sum[q][p] + base + x0;
                    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.s32 [%0], [%1], %2;" ::"l"(dst), "r"(stage_addr), "r"(cells * 4) : "memory");
                }
                bulk_commit();
            }
        }
    }
}
// Every lane that issued bulk operations (lane 0 of each warp): wait until they are complete (the grid
// barrier that follows fences once per block at system scope).
__device__ __forceinline__ void peer_add_drain(int lane) {
    if (lane == 0) bulk_wait_all();
}

// ONE thread, after every block of this rank has drained its bulk adds and the barrier has fenced them at
// system scope: tell the peers that this rank's contribution to exchange `e` is complete, and its box.
__device__ __forceinline__ void peer_signal(const PeerArgs& P, unsigned e, const int* lo, const int* hi) {
    const bool mine = hi[0] >= lo[0];
    const uint4 msg = make_uint4(mine ? (unsigned)(lo[0] | (lo[1] << 10) | (lo[2] << 20)) : 0x00100401u,   // empty: lo = 1, hi = 0
                                 mine ? (unsigned)(hi[0] | (hi[1] << 10) | (hi[2] << 20)) : 0u, e, 0u);
    __threadfence_system();
    for (int r = 0; r < P.nranks; ++r)
        if (r != P.rank) st_sys_v4(&P.mail[r]->msg[e & 1u][P.rank].box, msg);
}
// ONE thread: wait until every peer's contribution to exchange `e` has been added into this rank's grid;
// lo/hi come in as this rank's box and leave as the union over the ranks.
__device__ __forceinline__ void peer_wait(const PeerArgs& P, unsigned e, int* lo, int* hi) {
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
}
