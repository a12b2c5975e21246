#!/usr/bin/env python3
"""every BASELINE config on one B200, this engine (pic_run) next to the reference's own
CUDA program (oracle/_ref/main), both through their identical 8-argument CLI; parses the summary lines.
usage: measure_configs.py [quick]   -> gpurun_out/configs.json + a markdown table on stdout"""
import json
import os
import re
import shutil
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
OURS = os.path.join(ROOT, "run")
REF = os.path.join(REF_DIR, "main")


def parse(txt):
    out = {}
    for key, pat in (("final_n", r"Final amount of particles: (-?\d+)"), ("added", r"Particles added: (-?\d+)"), ("removed", r"Particles removed: (-?\d+)"),
                     ("gpu_ms", r"GPU time of program: ([0-9.]+) ms"), ("cpu_ms", r"CPU time of program: ([0-9.]+) ms")):
        m = re.search(pat, txt)
        if m:
            out[key] = float(m.group(1)) if key.endswith("_ms") else int(m.group(1))
    out["overflow"] = "OVERFLOW" in txt
    return out


def run(exe, args, cwd, env=None, reps=3, timeout=900):
    best = None
    for _ in range(reps):
        t0 = time.time()
        try:
            r = subprocess.run([exe] + [str(a) for a in args], cwd=cwd, capture_output=True, text=True, timeout=timeout, env=env)
        except subprocess.TimeoutExpired:
            return {"error": "timeout"}
        d = parse(r.stdout)
        d["wall_s"] = time.time() - t0
        if r.returncode != 0 or "gpu_ms" not in d:
            return {"error": f"rc={r.returncode} {r.stdout[-300:]!r} {r.stderr[-300:]!r}"}
        if best is None or d["gpu_ms"] < best["gpu_ms"]:
            best = d
    return best


def main():
    quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
    os.makedirs(os.path.join(REF_DIR, "src"), exist_ok=True)
    os.makedirs(os.path.join(REF_DIR, "out", "visualization"), exist_ok=True)
    shutil.copyfile(os.path.join(ROOT, "src", "cross_section.txt"), os.path.join(REF_DIR, "src", "cross_section.txt"))
    # supercritical table for the avalanche (cfg3): every row "2.0 0.0" -> x1.02 per mobility step
    ava = os.path.join(ROOT, "gpurun_out", "cross_section_avalanche.txt")
    os.makedirs(os.path.dirname(ava), exist_ok=True)
    with open(ava, "w") as f:
        f.write("2.0 0.0\n" * 10000)
    ava_dir = os.path.join(ROOT, "gpurun_out", "ava_ref")
    os.makedirs(os.path.join(ava_dir, "src"), exist_ok=True)
    os.makedirs(os.path.join(ava_dir, "out", "visualization"), exist_ok=True)
    shutil.copyfile(ava, os.path.join(ava_dir, "src", "cross_section.txt"))

    cfgs = []   # (name, args, pushes-model, ours env, reference cwd or None)
    cfgs.append(("cfg1 naive", [32, 0, 10000, 10, 256, 1000000, 100, 10], None, None, REF_DIR))
    for m in (30, 31, 33):
        cfgs.append((f"cfg1 mode {m}", [m, 0, 10000, 10, 256, 1000000, 100, 10], None, None, REF_DIR))
    for m in (30, 31, 32, 33):
        cfgs.append((f"cfg2 mode {m}", [m, 0, 1000000, 100, 256, 100000000, 100, 10], None, None, REF_DIR))
    # cfg3: avalanche 1e5 -> capacity; 50 Poisson x 10 steps = x1.02^500 = 2e4 -> 2e9
    cfgs.append(("cfg3 avalanche cap 2e8 (39 Poisson steps)", [30, 0, 100000, 39, 256, 200000000, 100, 10], None, {"PICB200_TABLE": ava}, ava_dir))
    cfgs.append(("cfg3 avalanche cap 2e9 (50 Poisson steps)", [30, 0, 100000, 50, 256, 2000000000, 100, 10], None, {"PICB200_TABLE": ava}, None))
    if not quick:
        for m in (30, 31, 32):
            for pt in (1, 10, 100):
                cfgs.append((f"cfg5 mode {m} pt {pt}", [m, 0, 10000000, 10, 256, 100000000, 100, pt], None, None, REF_DIR))
    rows = []
    for name, args, _, env_extra, ref_cwd in cfgs:
        env = dict(os.environ)
        if env_extra:
            env.update(env_extra)
        ours = run(OURS, args, ROOT, env=env)
        ref = run(REF, args, ref_cwd, reps=2) if (ref_cwd and os.path.exists(REF)) else {"error": "not run (cannot allocate: 2*capacity overflows int, 176 B/electron)"}
        rows.append({"config": name, "args": args, "ours": ours, "reference": ref})
        print(name, "ours", ours, "ref", ref, flush=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "configs.json"), "w"), indent=1)
    print("\n| config | CLI args | this engine: GPU ms (final n / added / removed) | reference CUDA: GPU ms (final n / added / removed) | speed-up |")
    print("|---|---|---|---|---|")
    for r in rows:
        o, f = r["ours"], r["reference"]
        so = f"{o['gpu_ms']:.2f} ({o['final_n']} / {o['added']} / {o['removed']})" + (" OVERFLOW" if o.get("overflow") else "") if "gpu_ms" in o else o.get("error", "?")[:60]
        sf = f"{f['gpu_ms']:.2f} ({f['final_n']} / {f['added']} / {f['removed']})" + (" OVERFLOW" if f.get("overflow") else "") if "gpu_ms" in f else f.get("error", "?")[:80]
        sp = f"{f['gpu_ms'] / o['gpu_ms']:.1f}x" if "gpu_ms" in o and "gpu_ms" in f else "-"
        print(f"| {r['config']} | `{' '.join(str(a) for a in r['args'])}` | {so} | {sf} | {sp} |")


if __name__ == "__main__":
    main()
