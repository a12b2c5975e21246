#!/usr/bin/env python3
"""Generate tests/golden/*.npz / *.json from the REFERENCE's own CUDA program (oracle/_ref/main =
the unmodified upstream sources compiled for sm_100, see oracle/Makefile).  It needs a GPU, so it
runs on the B200 box (`gpurun -- python tools/make_golden_from_ref.py`) and writes into
gpurun_out/golden/; the files are then copied to tests/golden/ and committed.

  ref_collision_free.npz   `main 31 3 5000 6 256 100000 100 10`: the state dumps the reference prints
      (src/utility.cu:124-137, src/electron.h:19-26) at Poisson steps 0, 3 and 6.  60 pushes at 0.02
      electrons per cell cannot reach table row 1, so no collision happens and the run is a
      deterministic function of the initial positions: it pins cell indexing, deposit, the difference
      stencil, EFC, the gather and the leapfrog against the real reference, independent of the RNG.
  ref_collision_free_dense.npz   `main 31 2 200000 2 256 400000 100 10`: 2e5 electrons in the 62^3-cell box (0.84 per
      cell: cell counts 0..8, every stencil shape), dumps at Poisson steps 0 and 2.  20 pushes cannot reach table row 1
      either (|v|^2 <= 3 x (20 x 2.53e-6 x 8)^2 = 4.9e-7 < 1.005e-6).  Stored: every initial position (fp32, exact: the
      reference draws them as floats, src/random.cu:7-12), every electron's acceleration at step 2 as the integer count
      difference it encodes (acc = float(EFC * dc), src/grid_operations.cu:29-56) and the printed acceleration and
      velocity of every 10th electron (in (x, y, z) order).
  ref_cfg2_stats.json      `main 30 100 1000000 100 256 100000000 100 10` (BASELINE cfg2), FIVE runs: summary
      counts and histograms of the final state (energy |v|^2, per-cell occupancy, velocity moments).
      The reference is not run-to-run deterministic once collisions occur, so
      this is a statistical fixture (north_star: counts/histograms within 1 %); five runs show its own spread.

usage: make_golden_from_ref.py [sparse] [dense] [stats]     (default: dense stats)
"""
import json
import os
import re
import shutil
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
EXE = os.path.join(REF_DIR, "main")
OUT = os.path.join(ROOT, "gpurun_out", "golden")
TRANS = str.maketrans("():,[]", "      ")

sys.path.insert(0, os.path.join(ROOT, "tests"))
from golden_stats import state_stats as _state_stats  # noqa: E402  (one binning for fixture and tests)


def run_ref(args, timeout=1200):
    os.makedirs(os.path.join(REF_DIR, "src"), exist_ok=True)
    os.makedirs(os.path.join(REF_DIR, "out", "visualization"), exist_ok=True)
    shutil.copyfile(os.path.join(ROOT, "src", "cross_section.txt"), os.path.join(REF_DIR, "src", "cross_section.txt"))
    t0 = time.time()
    r = subprocess.run([EXE] + [str(a) for a in args], cwd=REF_DIR, capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise SystemExit(f"reference run failed rc={r.returncode}: {r.stdout[-500:]} {r.stderr[-500:]}")
    return r.stdout, time.time() - t0


def parse_dumps(txt):
    """-> {t: array (n, 11) of index, pos3, vel3, acc3, timestamp}, summary dict"""
    dumps, cur_t, cur = {}, None, []
    summ = {}
    for ln in txt.splitlines():
        if ln.startswith("Time "):
            m = re.match(r"Time (\d+), amount (\d+)", ln)
            cur_t, cur = int(m.group(1)), []
            dumps[cur_t] = cur
        elif cur_t is not None and ln and ln[0].isdigit() and ": (" in ln:
            cur.append(ln.translate(TRANS))
        else:
            for key, pat in (("final_n", r"Final amount of particles: (-?\d+)"), ("added", r"Particles added: (-?\d+)"),
                             ("removed", r"Particles removed: (-?\d+)"), ("gpu_ms", r"GPU time of program: ([0-9.]+) ms")):
                m = re.match(pat, ln)
                if m:
                    summ[key] = float(m.group(1)) if key == "gpu_ms" else int(m.group(1))
    out = {}
    for t, lines in dumps.items():
        a = np.fromstring(" ".join(lines), dtype=np.float64, sep=" ").reshape(-1, 11) if lines else np.zeros((0, 11))
        out[t] = a
    return out, summ


def state_stats(a):
    """Histograms of one state dump (rows: idx, x y z, vx vy vz, ax ay az, ts)."""
    return _state_stats(a[:, 1:4], a[:, 4:7], a[:, 10])


EFC = (1.602176487e-19 * 1.602176487e-19) / (4 * 3.1415926536 * 8.8541878176E-12 * 1e-2 * 1e-2 * 9.1093837015e-31)   # src/cell.cu:5


def by_xyz(pos):
    q = np.round(np.asarray(pos, dtype=np.float64), 9)
    return np.lexsort((q[:, 2], q[:, 1], q[:, 0]))


def make_sparse():
    args = [31, 3, 5000, 6, 256, 100000, 100, 10]
    txt, dt = run_ref(args)
    dumps, summ = parse_dumps(txt)
    assert sorted(dumps) == [0, 3, 6] and all(d.shape[0] == 5000 for d in dumps.values()), {k: v.shape for k, v in dumps.items()}
    assert summ["added"] == 0 and summ["removed"] == 0, summ
    np.savez_compressed(os.path.join(OUT, "ref_collision_free.npz"), args=np.array(args),
                        **{f"t{t}_pos": d[:, 1:4] for t, d in dumps.items()}, **{f"t{t}_vel": d[:, 4:7] for t, d in dumps.items()},
                        **{f"t{t}_acc": d[:, 7:10].astype(np.float32) for t, d in dumps.items()},
                        **{f"t{t}_ts": d[:, 10].astype(np.int32) for t, d in dumps.items()})
    print("collision-free fixture:", summ, f"{dt:.1f}s", "nonzero acc rows:", int((dumps[6][:, 7:10] != 0).any(axis=1).sum()))


def make_dense():
    args = [31, 2, 200000, 2, 256, 400000, 100, 10]
    txt, dt = run_ref(args)
    dumps, summ = parse_dumps(txt)
    del txt
    n = args[2]
    assert sorted(dumps) == [0, 2] and all(d.shape[0] == n for d in dumps.values()), {k: v.shape for k, v in dumps.items()}
    assert summ["added"] == 0 and summ["removed"] == 0 and summ["final_n"] == n, summ
    d0, d2 = dumps[0], dumps[2]
    p0 = d0[:, 1:4]
    # the reference draws the coordinates as floats (src/random.cu:7-12); %.15f prints them to well below half an fp32 ulp
    assert np.abs(p0.astype(np.float32).astype(np.float64) - p0).max() < 1e-15, "initial positions are fp32 values"
    o0, o2 = by_xyz(p0), by_xyz(d2[:, 1:4])
    assert np.abs(p0[o0] - d2[o2, 1:4]).max() < 1e-12 and np.all(d2[:, 10] == -1) and np.all(d0[:, 4:10] == 0)
    acc = d2[o2, 7:10]                                   # in the (x, y, z) order of the initial positions
    dc = np.rint(acc / EFC).astype(np.int64)
    assert np.abs(dc).max() < 127
    # the printed acceleration (%.7f of a float) must be the float the reference's formula gives for that integer difference
    assert np.array_equal(np.round((EFC * dc.astype(np.float64)).astype(np.float32).astype(np.float64), 7), np.round(acc, 7))
    cells = (p0 / 1e-2).astype(np.int64)
    lin = (cells[:, 2] * 512 + cells[:, 1]) * 512 + cells[:, 0]
    _, counts = np.unique(lin, return_counts=True)
    sub = np.arange(0, n, 10)
    np.savez_compressed(os.path.join(OUT, "ref_collision_free_dense.npz"), args=np.array(args), t0_pos=p0[o0].astype(np.float32),
                        t2_dc=dc.astype(np.int8), sub=sub.astype(np.int32), t2_sub_vel=d2[o2][sub, 4:7], t2_sub_acc=acc[sub].astype(np.float32),
                        t2_sub_pos=d2[o2][sub, 1:4], occupancy=np.bincount(counts))
    print("dense collision-free fixture:", summ, f"{dt:.1f}s", "occupied cells", counts.size, "max count", counts.max(), "max |dc|", int(np.abs(dc).max()),
          "rows with a field", int((dc != 0).any(axis=1).sum()), "max |v|^2", float((d2[:, 4:7] ** 2).sum(axis=1).max()))


def make_stats(reps=5):
    runs = []
    for rep in range(reps):
        args2 = [30, 100, 1000000, 100, 256, 100000000, 100, 10]
        txt, dt = run_ref(args2, timeout=2400)
        dumps, summ = parse_dumps(txt)
        del txt
        st = state_stats(dumps[100])
        st0 = state_stats(dumps[0]) if rep == 0 else None
        runs.append({"args": args2, "summary": summ, "final": st, "initial": st0, "wall_s": dt})
        print("cfg2 run", rep, summ, f"{dt:.1f}s", "n parsed", st["n"], flush=True)
        del dumps
    json.dump({"generator": "tools/make_golden_from_ref.py", "reference": "oracle/_ref/main (unmodified upstream sources, sm_100)", "runs": runs},
              open(os.path.join(OUT, "ref_cfg2_stats.json"), "w"))


def main():
    if not os.path.exists(EXE):
        raise SystemExit(f"{EXE} missing: `make -C oracle ref` in the authoring container first")
    os.makedirs(OUT, exist_ok=True)
    what = sys.argv[1:] or ["dense", "stats"]
    if "sparse" in what:
        make_sparse()
    if "dense" in what:
        make_dense()
    if "stats" in what:
        make_stats()
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    sys.exit(main())
