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
  ref_cfg2_stats.json      `main 30 100 1000000 100 256 100000000 100 10` (BASELINE cfg2): summary
      counts and histograms of the final state (energy |v|^2, per-cell occupancy, velocity moments).
      The reference is not run-to-run deterministic once collisions occur, so
      this is a statistical fixture (north_star: counts/histograms within 1 %).
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


def main():
    if not os.path.exists(EXE):
        raise SystemExit(f"{EXE} missing: `make -C oracle ref` in the authoring container first")
    os.makedirs(OUT, exist_ok=True)
    # ---- 1. deterministic, collision-free fixture
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
    # ---- 2. statistical fixture at BASELINE cfg2 (two independent runs: shows the reference's own run-to-run spread)
    runs = []
    for rep in range(2):
        args2 = [30, 100, 1000000, 100, 256, 100000000, 100, 10]
        txt, dt = run_ref(args2, timeout=2400)
        dumps, summ = parse_dumps(txt)
        del txt
        st = state_stats(dumps[100])
        st0 = state_stats(dumps[0]) if rep == 0 else None
        runs.append({"args": args2, "summary": summ, "final": st, "initial": st0, "wall_s": dt})
        print("cfg2 run", rep, summ, f"{dt:.1f}s", "n parsed", st["n"])
        del dumps
    json.dump({"generator": "tools/make_golden_from_ref.py", "reference": "oracle/_ref/main (unmodified upstream sources, sm_100)", "runs": runs},
              open(os.path.join(OUT, "ref_cfg2_stats.json"), "w"))
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    sys.exit(main())
