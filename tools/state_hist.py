#!/usr/bin/env python3
"""Histogram tool for binary snapshots (picb200_save_state / PICB200_DUMP_DIR / PICB200_DUMP_FINAL): energy |v|^2,
per-cell occupancy and velocity moments of one snapshot, or the total-variation distances between two -- the
1 % density / energy check of north_star without a text dump.
usage: state_hist.py a.bin [b.bin]"""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from golden_stats import state_stats, tvd  # noqa: E402

eng = importlib.import_module("particle-simulation_b200")


def stats(path):
    hdr, a = eng.read_state(path)
    pos = np.stack([a["x"], a["y"], a["z"]], axis=1).astype(np.float64)
    vel = np.stack([a["vx"], a["vy"], a["vz"]], axis=1).astype(np.float64)
    return hdr, state_stats(pos, vel, a["ts"])


def main():
    if len(sys.argv) < 2:
        raise SystemExit(__doc__)
    h1, s1 = stats(sys.argv[1])
    print(json.dumps({k: (v if not isinstance(v, bytes) else v.decode()) for k, v in h1.items()}))
    if len(sys.argv) == 2:
        print(json.dumps({k: (v if not isinstance(v, (list, np.ndarray)) else list(map(float, np.asarray(v).ravel()[:16]))) for k, v in s1.items()})[:2000])
        return
    h2, s2 = stats(sys.argv[2])
    print(f"n: {s1['n']} vs {s2['n']} ({abs(s1['n'] - s2['n']) / max(1, s1['n']):.4%})")
    for k in ("energy_hist", "occupancy_hist"):
        print(f"  TVD {k}: {tvd(s1[k], s2[k]):.4%}")


if __name__ == "__main__":
    main()
