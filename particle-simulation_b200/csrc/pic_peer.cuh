// pic_peer.cuh -- multi-GPU exchange of the charge grid INSIDE the persistent kernel, over NVLink
// peer memory (included inside namespace picb200 by pic_mover.cuh).
//
// The reference has no multi-GPU path (device 0 is hard-coded, src/pic.cu:394).  Here every GPU owns
// a slab of the electrons; the only coupling is the charge grid, whose counts are additive integers
// (src/grid_operations.cu:25).  Every rank keeps one grid PER RANK (and per parity of the exchange):
// its own slot receives its local deposit, the other slots are written by the peers.  After the local
// deposits of a Poisson step are complete, each rank PUSHES the occupied box of its grid into its slot
// on every peer with plain stores over NVLink (posted writes: nobody waits for a round trip; pulling
// the same 1 MB with loads cost 15-20 us per step), raises a flag in the peers' mailboxes, and the next
// step's field gather adds up the N slots on the fly.  Integer sums are order-independent, so all ranks
// see bit-identical fields and the union of the slabs equals the single-GPU run.  No host round trip, no
// separate collective kernel, no summing pass: the whole run stays one cooperative launch per GPU.
#pragma once

constexpr int kMaxRanks = 8;
constexpr int kCellsLog2 = 27;          // 512^3 cells per grid

// One message per (exchange parity, source rank), written by the source rank into every peer's mailbox:
// {box lo packed 3 x 10 bits, box hi likewise, exchange number, 0} in ONE 16-byte store, so the box can
// never be seen without its sequence number (lo > hi on an axis: empty box).
struct alignas(16) PeerMsg { uint4 box; };
struct PeerMailbox { PeerMsg msg[2][kMaxRanks]; };

// Survives picb200_upload / init (the exchange counter must never run backwards: peers compare it).
struct PeerLocal {
    unsigned xseq;          // exchanges completed by this rank
    int error;              // 1: a peer did not show up within the time-out
    int pbox_lo[2][3], pbox_hi[2][3];   // the box this rank last pushed at each parity (what its slot on the peers holds)
};

// recv[r] = base of rank r's block of 2 x nranks grids as mapped into this GPU's address space
// (grid (parity, source) at offset (parity * nranks + source) << kCellsLog2); recv[rank] is local.
struct PeerArgs {
    int rank, nranks;
    PeerMailbox* mail[kMaxRanks];
    int* recv[kMaxRanks];
    PeerLocal* pl;
};
__device__ __forceinline__ int* peer_grid(const PeerArgs& P, int owner, int parity, int source) {
    return P.recv[owner] + ((size_t)(parity * P.nranks + source) << kCellsLog2);
}

__device__ __forceinline__ void st_sys_u32(unsigned* p, unsigned v) { asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
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
constexpr unsigned long long kPeerTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;

// All warps of the rank, after the local deposits of exchange parity `p` are complete: copy the occupied
// box of this rank's own grid into its slot on every peer, one warp per x-row, including what the slot
// still holds of the box pushed two exchanges ago (hull of the old and the new box; this rank's own grid
// is zero there).  The copy is ASYNCHRONOUS: the row is staged in shared memory and lane 0 hands one
// cp.async.bulk (TMA) store per peer to the copy engine, so the warp goes on to the compaction while the
// megabyte crosses NVLink (plain stores, like plain loads, held the issuing warps for 13 us per step).
// x-ranges are widened to multiples of 4 cells: bulk copies need 16-byte alignment and size.
constexpr int kPushSegCells = 64;                       // cells staged per bulk copy (256 B per warp)
constexpr int kPushStageBytesPerWarp = kPushSegCells * 4;

__device__ __forceinline__ void bulk_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void peer_push_rows(const PeerArgs& P, int p, const int* nbox, const int* obox /* lo xyz, hi xyz each */, int gw, int nw,
                                               int lane, int* stage /* smem, per warp */) {
    int lo[3], hi[3], hl[3], hh[3];
    bool any = true, old_any = true;
    for (int k = 0; k < 3; ++k) { lo[k] = nbox[k]; hi[k] = nbox[3 + k]; hl[k] = obox[k]; hh[k] = obox[3 + k]; }
    for (int k = 0; k < 3; ++k) { any = any && lo[k] <= hi[k]; old_any = old_any && hl[k] <= hh[k]; }
    if (!any && !old_any) return;
    if (!old_any) { for (int k = 0; k < 3; ++k) { hl[k] = lo[k]; hh[k] = hi[k]; } }
    else if (any) { for (int k = 0; k < 3; ++k) { hl[k] = min(hl[k], lo[k]); hh[k] = max(hh[k], hi[k]); } }
    hl[0] &= ~3; hh[0] |= 3;                             // 16-byte granules (the grid is 512 = 4 x 128 cells wide)
    const int wx = hh[0] - hl[0] + 1, wy = hh[1] - hl[1] + 1, wz = hh[2] - hl[2] + 1;
    const int rows = wy * wz;
    const int* mine = peer_grid(P, P.rank, p, P.rank);
    const unsigned stage_addr = (unsigned)__cvta_generic_to_shared(stage);
    for (int r = gw; r < rows; r += nw) {
        const int y = hl[1] + r % wy, z = hl[2] + r / wy;
        const size_t base = cell_index(hl[0], y, z);
        for (int x0 = 0; x0 < wx; x0 += kPushSegCells) {
            const int cells = min(kPushSegCells, wx - x0);               // multiple of 4
            if (lane == 0) bulk_store_wait_read();                       // the staging buffer is free again
            __syncwarp();
            if (2 * lane < cells) {
                const int2 v = __ldcg(reinterpret_cast<const int2*>(mine + base + x0 + 2 * lane));
                *reinterpret_cast<int2*>(stage + 2 * lane) = v;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> async proxy
            __syncwarp();
            if (lane == 0) {
                for (int q = 0; q < P.nranks; ++q) {
                    if (q == P.rank) continue;
                    int* dst = peer_grid(P, q, p, P.rank) + base + x0;
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(stage_addr), "r"(cells * 4) : "memory");
                }
                bulk_store_commit();
            }
        }
    }
}
// Every lane that issued bulk stores (lane 0 of each warp): wait until they are complete.
__device__ __forceinline__ void peer_push_drain(int lane) {
    if (lane == 0) bulk_store_wait_all();       // (the grid barrier that follows fences once per block at system scope)
}

__device__ __forceinline__ void st_sys_v4(uint4* p, uint4 v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_sys_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// ONE thread, after every block of this rank has drained its bulk stores and the barrier has fenced them at
// system scope: tell the peers that this rank's slot of exchange `e` is complete, and which box it occupies.
__device__ __forceinline__ void peer_signal(const PeerArgs& P, unsigned e, const int* lo, const int* hi) {
    const bool mine = hi[0] >= lo[0];
    const uint4 msg = make_uint4(mine ? (unsigned)(lo[0] | (lo[1] << 10) | (lo[2] << 20)) : 0x00100401u,   // empty: lo = 1, hi = 0
                                 mine ? (unsigned)(hi[0] | (hi[1] << 10) | (hi[2] << 20)) : 0u, e, 0u);
    __threadfence_system();
    for (int r = 0; r < P.nranks; ++r)
        if (r != P.rank) st_sys_v4(&P.mail[r]->msg[e & 1u][P.rank].box, msg);
}
// ONE thread: wait until every peer's slot of exchange `e` is complete in this rank's memory; lo/hi come in
// as this rank's box and leave as the union over the ranks.
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

// All warps of the rank, once every slot of parity `p` is complete: add the per-rank grids up over the union
// box into `sum_grid` (the one grid the next step gathers from -- six scattered loads per electron from one
// grid, not from one per rank), writing zeros over what is left of its previous box; one warp per x-row.
__device__ __forceinline__ void peer_sum_rows(const PeerArgs& P, int p, int* __restrict__ sum_grid, const int* ubox, const int* obox, int gw, int nw, int lane) {
    int lo[3], hi[3], hl[3], hh[3];
    bool any = true, old_any = true;
    for (int k = 0; k < 3; ++k) { lo[k] = ubox[k]; hi[k] = ubox[3 + k]; hl[k] = obox[k]; hh[k] = obox[3 + k]; }
    for (int k = 0; k < 3; ++k) { any = any && lo[k] <= hi[k]; old_any = old_any && hl[k] <= hh[k]; }
    if (!any && !old_any) return;
    if (!old_any) { for (int k = 0; k < 3; ++k) { hl[k] = lo[k]; hh[k] = hi[k]; } }
    else if (any) { for (int k = 0; k < 3; ++k) { hl[k] = min(hl[k], lo[k]); hh[k] = max(hh[k], hi[k]); } }
    const int wx = hh[0] - hl[0] + 1, wy = hh[1] - hl[1] + 1, wz = hh[2] - hl[2] + 1;
    const int rows = wy * wz;
    const int* slot0 = peer_grid(P, P.rank, p, 0);
    for (int r = gw; r < rows; r += nw) {
        const int y = hl[1] + r % wy, z = hl[2] + r / wy;
        const bool row_in = any && y >= lo[1] && y <= hi[1] && z >= lo[2] && z <= hi[2];
        const size_t base = cell_index(hl[0], y, z);
        for (int xx = lane; xx < wx; xx += 32) {
            const int x = hl[0] + xx;
            int sum = 0;
            if (row_in && x >= lo[0] && x <= hi[0]) {
#pragma unroll
                for (int q = 0; q < kMaxRanks; ++q)
                    if (q < P.nranks) sum += __ldcg(slot0 + ((size_t)q << kCellsLog2) + base + xx);
            }
            __stcg(sum_grid + base + xx, sum);
        }
    }
}
