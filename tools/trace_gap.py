#!/usr/bin/env python3
"""Diagnostic: where a block's time goes between the end of its main phase (stamp 2) and its arrival at grid barrier 1
(stamps 8/9/11: warps 0/1/2 in front of the barrier's block barrier, 15: behind it, 12: arrived), by finishing order."""
import sys
import numpy as np
raw = open(sys.argv[1], "rb").read()
steps, blocks = np.frombuffer(raw[:8], dtype=np.int32)
t = np.frombuffer(raw[8:], dtype=np.uint64).reshape(steps, blocks, 16).astype(np.int64)
for s in range(steps):
    if t[s, 0, 6] > 0 or t[s, 0, 7] == 0:
        continue
    t0 = t[s, :, 0].min()
    o = np.argsort(t[s, :, 2])
    q = lambda k: ((t[s, o, k] - t[s, o, 2]) / 1e3)
    f = lambda a: " ".join(f"{x:5.1f}" for x in a)
    sel = [0, 1, 2, 5, 10, 20, 40, 74, 110, 130, 140, 145, 146, 147]
    print(f"step {s}: end of main phase (stamp 2, us after step start), blocks in finishing order #{sel}")
    print("   stamp2     ", f(((t[s, o, 2] - t0) / 1e3)[sel]))
    for k, nm in ((8, "warp0 pre-bar"), (9, "warp1 pre-bar"), (11, "warp2 pre-bar"), (15, "post-bar     "), (12, "arrived      "), (14, "released     "), (3, "b1 done      ")):
        print(f"   {nm}", f(q(k)[sel]))
    print("   step length %.1f" % ((t[s, :, 7].max() - t0) / 1e3 if t[s, :, 7].max() > 0 else (t[s, :, 3].max() - t0) / 1e3))
    if s >= 4:
        break
