#!/usr/bin/env python3
"""Summarise a PICB200_TRACE dump: per-phase durations of k_poisson_fused (us), median over Poisson steps."""
import sys
import numpy as np
raw = open(sys.argv[1], "rb").read()
steps, blocks = np.frombuffer(raw[:8], dtype=np.int32)
t = np.frombuffer(raw[8:], dtype=np.uint64).reshape(steps, blocks, 16).astype(np.int64)
names = ["own chunks of warp 0 (0->1)", "wait for block's other warps (1->2)", "barrier 1 (2->3)", "clear A + scan (3->4)",
         "barrier 2 (4->5)", "compact move (5->6)", "barrier 3 (6->7)"]
good = t[:, 0, 7] > 0
t = t[good]
dur = (t[:, :, 7].max(1) - t[:, :, 0].min(1)) / 1e3
comp = t[:, 0, 6] > 0          # steps that ended with a compaction (the others leave their holes in place)
print(f"{t.shape[0]} steps x {blocks} blocks; step time (first start -> last end), us: median {np.median(dur):.1f} mean {dur.mean():.1f}"
      f"; {int(comp.sum())} steps with compaction (mean {dur[comp].mean() if comp.any() else 0:.1f}), {int((~comp).sum())} without (mean {dur[~comp].mean() if (~comp).any() else 0:.1f})")
tl = t[~comp]
if tl.shape[0] and tl[:, :, 12].min() > 0:
    # inside barrier 1: the last block to arrive (the one that stamped slot 13) and everybody's release
    last = tl[:, :, 13].argmax(1)
    ar = np.array([tl[s, last[s], 12] for s in range(tl.shape[0])])
    act = np.array([tl[s, last[s], 13] for s in range(tl.shape[0])])
    pre = np.array([tl[s, last[s], 12] - tl[s, last[s], 2] for s in range(tl.shape[0])])
    if tl[:, :, 15].min() > 0:
        a_ = np.array([[tl[s, last[s], 15] - tl[s, last[s], 2], tl[s, last[s], 10] - tl[s, last[s], 15], tl[s, last[s], 12] - tl[s, last[s], 10]] for s in range(tl.shape[0])])
        print(f"  [barrier 1, last block] counters + box + block barrier {np.median(a_[:,0])/1e3:.2f} us, fence {np.median(a_[:,1])/1e3:.2f} us, arrival atomic {np.median(a_[:,2])/1e3:.2f} us")
    rel = tl[:, :, 14].max(1)
    fin = tl[:, :, 3].max(1)
    print(f"  [barrier 1, last block] syncthreads+fence+arrive {np.median(pre)/1e3:.2f} us, action {np.median(act-ar)/1e3:.2f} us, fence+release seen by all {np.median(rel-act)/1e3:.2f} us, fence+fetch+syncthreads {np.median(fin-rel)/1e3:.2f} us")
if tl.shape[0]:
    rows = [("own chunks of warp 0 (0->1)", 0, 1), ("wait for block's other warps (1->2)", 1, 2), ("barrier 1 + bookkeeping (2->3)", 2, 3)]
    if tl[:, :, 5].min() > 0:      # peer mode: the exchange follows barrier 1 in every step
        rows += [("peer: push rows to the peers (3->11)", 3, 11), ("peer: clear own old grid (11->8)", 11, 8), ("peer: drain bulk adds (8->4)", 8, 4),
                 ("peer: arrive at every rank + wait for all blocks of all ranks (4->5)", 4, 5)]
    for nm, a_, b_ in rows:
        d = (tl[:, :, b_] - tl[:, :, a_]) / 1e3
        print(f"  [no compaction] {nm:50s} mean {d.mean():7.1f}  min-block {np.median(d.min(1)):7.1f}  max-block {np.median(d.max(1)):7.1f}")
t = t[comp]
if not t.shape[0]:
    sys.exit(0)
for k, nm in enumerate(names):
    d = (t[:, :, k + 1] - t[:, :, k]) / 1e3
    print(f"  {nm:40s} mean {d.mean():7.1f}  min-block {np.median(d.min(1)):7.1f}  max-block {np.median(d.max(1)):7.1f}")
for nm, a, b in (("peer: push rows to the peers (3->11)", 3, 11), ("peer: clear own old grid (11->8)", 11, 8), ("scan: counters + ts loads + masks (8->9)", 8, 9), ("scan: block reserve (9->10)", 9, 10), ("scan: list stores (10->4)", 10, 4)):
    if t[:, :, a].min() > 0 and t[:, :, b].min() > 0:
        d = (t[:, :, b] - t[:, :, a]) / 1e3
        print(f"    {nm:40s} mean {d.mean():7.1f}  min-block {np.median(d.min(1)):7.1f}  max-block {np.median(d.max(1)):7.1f}")
end_main = t[:, :, 2] - t[:, :, 0].min(1, keepdims=True)
print(f"  block finishes main phase at: min {np.median(end_main.min(1))/1e3:.1f} median {np.median(end_main)/1e3:.1f} max {np.median(end_main.max(1))/1e3:.1f} us")
