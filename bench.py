#!/usr/bin/env python3
"""bench.py -- electron-pushes/sec of the PIC-MCC hot path on N B200s (BASELINE.json's metric).

Workload (N=1): BASELINE cfg2 = `run 30 0 1000000 100 256 100000000 100 10`: the persistent
scheduler (mode 30), 1e6 initial electrons, 100 Poisson steps x 10 mobility steps, capacity 1e8,
shipped cross-section table.  One bench "step" = one complete run of that configuration (~1e9
pushes).  N>1: every rank runs the same slab size (weak scaling, global init n = N x 1e6); the charge box is
summed across the GPUs INSIDE the persistent kernel -- one multimem.red per pair of cells through the NVSwitch, or
(PICB200_NVLS=0) one bulk reduce-add per peer over NVLink peer memory (csrc/pic_peer.cuh, csrc/pic_nvls.h) -- no NCCL
call on the data path (PICB200_NCCL=1 selects the one-all-reduce-per-Poisson-step fallback).  Before anything is timed, a run with N > 1 checks itself: a short colliding workload on the N slabs
must equal the same population on ONE GPU (counts and an order-independent digest of every electron) -> `parity_check`.

  value     pushes / device time of the Poisson loop (cudaEvent span, as src/pic.cu:476..:563),
            electrons initialised on the device, L2 flushed between runs.
  e2e       the same run driven through the C ABI with HOST buffers: the initial electrons are
            uploaded from pinned host memory and the final population is downloaded to pinned host
            memory inside the timed (wall-clock) region.
  roofline  the dominant kernel (k_poisson_fused: one cooperative launch = the whole run of 100 Poisson
            steps): algorithmic bytes = 68 B per push x pushes of the launch, divided by
            its device time (CUDA events on the launching stream) vs measured HBM copy bandwidth.  The
            kernel keeps an electron in registers for all 10 mobility steps, so the 68-B-accounted
            figure may exceed the HBM peak; `traffic` is the DRAM traffic ncu measured for it (a committed capture,
            labelled with its commit -- not measured in this run).  `issue_frac` is the honest ceiling: pushes/s
            against 4 warp-instructions per SM and cycle / (warp-instructions per push, same capture).
  roofline_hbm_corner  (N = 1) the same kernel where every push really crosses HBM, measured live: cfg5 with one push per
            Poisson step, 1e7 electrons (> L2); there 68 B x pushes / time IS an HBM utilisation.
  cfg4      BASELINE configs[3]: 1e9 electrons in total sharded over the N GPUs, 10 x 10 steps, device-timed
            (strong scaling: the per-N values of this sub-record divide the same work).
  --impl reference   times the reference itself: its own CUDA program (oracle/_ref/main, built from
            the upstream sources where they lie) when it and a GPU are available, else the CPU port.
"""
import argparse
import importlib
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CFG2 = dict(mode=0, init_n=1000000, poisson_steps=100, block_size=256, capacity=100000000, sleep_time_ns=100, poisson_timestep=10)
WORKLOAD = "cfg2: run 30 0 1000000 100 256 100000000 100 10 (mode 30 dynamic, init n 1e6, 100 Poisson x 10 mobility steps, max n 1e8)"
BYTES_PER_PUSH = 68.0
METRIC, UNIT = "electron_pushes_per_sec", "pushes/s"


def measured_hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 50 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        if shutil.which("nvidia-smi"):
            self.p = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)

    def stop(self):
        if self.p:
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except Exception:
                self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for ln in self.f.read().splitlines():
            c = [v.strip() for v in ln.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(n_gpus):
    """Control plane for N>1 (torchrun): gloo for the rendezvous, the barriers, the max over ranks and the all-gather of
    the 256-byte peer handles; the data path is the engine's own NVLink peer exchange (no collective library)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world == 1:
        return None, 0, 1, 0
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    import torch.distributed as dist
    dist.init_process_group(backend="gloo", rank=rank, world_size=world)
    return dist, rank, world, local


def all_max(dist, v):
    if dist is None:
        return v
    import torch
    t = torch.tensor([v], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def all_sum(dist, v):
    if dist is None:
        return v
    import torch
    t = torch.tensor([v], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def ncu_capture():
    """The committed ncu capture of the dominant kernel (profiles/*_ncu_traffic.json), or None."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_traffic.json")))
    return json.load(open(files[-1])) if files else None


def ncu_traffic(pushes_per_launch):
    """DRAM bytes of one k_poisson_fused launch of this bench: dram__bytes_read.sum + dram__bytes_write.sum of the committed
    `ncu --set full` capture (profiles/*_ncu_traffic.json, written by tools/refresh_profiles.sh), scaled from the captured
    launch's push count to this launch's (the capture runs 10 Poisson steps of the same population, the bench 100)."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_traffic.json")))
    if not files:
        return None, "no ncu capture committed"
    t = json.load(open(files[-1]))
    if not t.get("pushes_in_launch"):
        return None, f"{os.path.basename(files[-1])}: push count of the captured launch missing"
    per_push = (t["dram_bytes_read"] + t["dram_bytes_write"]) / t["pushes_in_launch"]
    return per_push * pushes_per_launch, (f"{os.path.basename(files[-1])} (commit {t['commit']}): {per_push:.2f} B of DRAM traffic per push in the captured launch "
                                          f"x {pushes_per_launch:.4g} pushes per launch here; bytes")


def issue_note():
    """What actually limits the kernel (it moves < 1 B per push through DRAM): instruction issue, from the ncu capture."""
    t = ncu_capture()
    if not t or not t.get("warp_instructions") or not t.get("pushes_in_launch"):
        return None
    return {"warp_instructions_per_push": t["warp_instructions"] / t["pushes_in_launch"], "issue_active_pct_whole_launch": t.get("issue_active_pct"),
            "registers_per_thread": t.get("registers_per_thread"), "source": f"profiles ncu capture, commit {t['commit']}"}


def state_digest(d):
    """Order-independent digest of a population: per-electron 64-bit mix of the id and the raw fp32 bits of position and
    velocity; (sum mod 2^64, xor) -- both combine across ranks."""
    import numpy as np
    h = d.id.astype(np.uint64).copy()
    for k in ("x", "y", "z", "vx", "vy", "vz"):
        h = (h * np.uint64(0x9E3779B97F4A7C15)) ^ getattr(d, k).view(np.uint32).astype(np.uint64)
    return int(h.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(h)) if h.size else 0


PEER_TRANSPORT = {"kind": None, "note": None}     # what the last peer_sim() ended up with (for the bench line)


def peer_sim(eng, dist, world, **kw):
    """One rank's Sim with the in-kernel exchange wired up.  The exchange goes through the NVSwitch (a multicast object
    shared between the ranks, csrc/pic_nvls.h) where the box offers it; should that set-up fail on ANY rank -- the ranks
    agree on it over gloo after each phase -- every rank falls back to one bulk reduce-add per peer (PICB200_NVLS=0)
    instead of failing the run."""
    for attempt in (0, 1):
        sim, handle, err = None, None, None
        try:
            sim = eng.Sim(**kw)
            handle = sim.peer_export()
        except eng.EngineError as e:
            err = str(e)
        handles = [None] * world
        dist.all_gather_object(handles, (handle, err))
        bad = [e for _, e in handles if e]
        if not bad:
            try:
                sim.peer_connect([h for h, _ in handles])
            except eng.EngineError as e:
                err = str(e)
            errs = [None] * world
            dist.all_gather_object(errs, err)
            bad = [e for e in errs if e]
        if not bad:
            nvls = os.environ.get("PICB200_NVLS", "1") not in ("0",)
            PEER_TRANSPORT["kind"] = "peer (NVSwitch multicast where available)" if nvls else "peer (one bulk reduce-add per peer, PICB200_NVLS=0)"
            dist.barrier()
            return sim
        if sim is not None:
            sim.close()
        if attempt == 1 or os.environ.get("PICB200_NVLS") == "0":
            raise SystemExit("bench.py: peer set-up failed: " + bad[0])
        os.environ["PICB200_NVLS"] = "0"
        PEER_TRANSPORT["note"] = "NVSwitch set-up failed, fell back to one bulk reduce-add per peer: " + bad[0].strip().splitlines()[0]
    raise SystemExit("bench.py: peer set-up failed")


def parity_check(eng, dist, rank, world, local, uid):
    """N slabs == one GPU (the reference has no multi-GPU path, src/pic.cu:394, so this is the only oracle there is for
    N > 1; tests/test_multi_gpu.py compares the same thing with the CPU oracle).  A hot, dense population (collisions from
    the first push, several electrons per cell) runs 5 + 7 Poisson steps in two launches on the N slabs; counts and the
    digest are combined over the ranks and compared with the same population on rank 0's GPU alone.  Untimed."""
    import numpy as np
    n, pt, steps = 64000 * world, 10, (5, 7)
    cap = 16 * n
    rng = np.random.default_rng(3)
    pos = rng.uniform(2.45, 2.65, size=(3, n)).astype(np.float32)
    vel = (rng.normal(size=(3, n)) * 0.05).astype(np.float32)
    lo, hi = n * rank // world, n * (rank + 1) // world
    kw = dict(mode=0, capacity=cap // world, poisson_timestep=pt, poisson_steps=sum(steps), device=local, rank=rank, nranks=world, nccl_unique_id=uid)
    sim = peer_sim(eng, dist, world, **kw) if uid is None else eng.Sim(**kw)
    sim.upload(*[a[lo:hi] for a in pos], *[a[lo:hi] for a in vel], id=np.arange(lo, hi, dtype=np.uint64))
    dist.barrier()
    for k in steps:
        sim.run_steps(k)
    m = sim.summary()
    mine = (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) + state_digest(sim.download())
    dist.barrier()                       # no rank frees a grid a peer may still map
    sim.close()
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    got = [sum(p[k] for p in parts) for k in range(4)] + [sum(p[4] for p in parts) % (1 << 64)]
    x = 0
    for p in parts:
        x ^= p[5]
    got.append(x)
    res = [None]
    if rank == 0:
        with eng.Sim(mode=0, capacity=cap, poisson_timestep=pt, poisson_steps=sum(steps), device=local) as one:
            one.upload(*pos, *vel, id=np.arange(n, dtype=np.uint64))
            for k in steps:
                one.run_steps(k)
            m1 = one.summary()
            exp = [m1["final_n"], m1["total_added"], m1["total_removed"], m1["pushes"]] + list(state_digest(one.download()))
        res[0] = {"ok": got == exp, "n_initial": n, "final_n": got[0], "added": got[1], "removed": got[2], "pushes": got[3],
                  "digest": "%016x:%016x" % (got[4], got[5]), "transport": "nccl" if uid is not None else "peer", "transport_detail": PEER_TRANSPORT["kind"] if uid is None else None, "transport_note": PEER_TRANSPORT["note"],
                  "what": f"{world} slabs vs the same population on one GPU: counts and (sum, xor) digest of (id, position, velocity) of every electron"}
        if not res[0]["ok"]:
            res[0]["expected"] = exp
            res[0]["got"] = got
    dist.broadcast_object_list(res, src=0)
    return res[0]


CFG4 = dict(mode=0, init_n=1000000000, poisson_steps=10, block_size=256, capacity=1600000000, sleep_time_ns=100, poisson_timestep=10)


def cfg4_record(eng, dist, rank, world, local, uid, runs=3):
    """BASELINE configs[3]: 1e9 electrons in total, sharded over the N GPUs (N = 1: 48 GB on one B200), 10 Poisson x 10
    mobility steps, capacity 1.6e9 in total.  Device-timed, 1 warm-up + `runs` timed runs, max over ranks."""
    kw = dict(CFG4)
    kw["capacity"] = CFG4["capacity"] // world
    kw.update(device=local, rank=rank, nranks=world, nccl_unique_id=uid)
    sim = peer_sim(eng, dist, world, **kw) if (world > 1 and uid is None) else eng.Sim(**kw)
    n = CFG4["init_n"]
    lo, hi = n * rank // world, n * (rank + 1) // world
    ms, pushes, final_n = [], 0, 0
    for it in range(1 + runs):
        sim.init_uniform(hi - lo, lo)
        if dist is not None:
            dist.barrier()
        sim.run_steps(CFG4["poisson_steps"])
        m = sim.summary()
        if it > 0:
            ms.append(all_max(dist, m["gpu_ms"]))
            pushes += m["pushes"]
        final_n = m["final_n"]
    if dist is not None:
        dist.barrier()
    sim.close()
    pushes_all = all_sum(dist, float(pushes))
    final_all = all_sum(dist, float(final_n))
    return {"workload": "cfg4: 1e9 electrons in total sharded over the GPUs, run 30 0 1000000000 10 256 1600000000 100 10", "scaling": "strong",
            "value": pushes_all / (sum(ms) / 1e3), "unit": UNIT, "ms_per_run": sum(ms) / len(ms), "ms_runs": ms, "runs": runs,
            "final_n": int(final_all), "pushes_per_run": int(pushes_all) // runs, "timing": "cudaEvent span of the Poisson loop on the engine's stream, max over ranks"}


def hbm_corner_record(eng, local, peak, runs=3):
    """The memory-side corner of the same kernel, measured live (N = 1): cfg5 with one push per Poisson step -- 1e7 electrons,
    40 Poisson steps, every push really crosses HBM (the 480 MB population does not fit the 126 MB L2), so here the 68-B
    figure IS an HBM utilisation.  Device-timed span of the Poisson loop (the first deposit, two steps of quota warm-up and
    the closing compaction included), 1 warm-up + `runs` timed runs."""
    kw = dict(mode=eng.MODE_DYNAMIC, init_n=10_000_000, poisson_steps=40, capacity=100_000_000, poisson_timestep=1, block_size=256, sleep_time_ns=100)
    ms, pushes = [], 0
    with eng.Sim(device=local, **kw) as sim:
        sim.set_profiling(1)
        for it in range(1 + runs):
            sim.init_uniform(kw["init_n"], 0)
            sim.run_steps(kw["poisson_steps"])
            m = sim.summary()
            if it > 0:
                ms.append(m["gpu_ms"])
                pushes += m["pushes"]
    gbs = pushes * BYTES_PER_PUSH / (sum(ms) / 1e3) / 1e9
    return {"workload": "cfg5-shaped, poisson_timestep = 1: run 30 0 10000000 40 256 100000000 100 1", "bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s",
            "frac": gbs / peak if peak else None, "value": pushes / (sum(ms) / 1e3), "value_unit": UNIT, "ms_per_run": sum(ms) / len(ms), "us_per_poisson_step": 1e3 * sum(ms) / len(ms) / kw["poisson_steps"],
            "runs": runs, "kernel": "k_poisson_fused (streaming instantiation)", "accounting": f"{BYTES_PER_PUSH:.0f} B per push x pushes / device time of the Poisson loop; inputs > L2"}


def cpu_baseline(budget_s=12.0):
    """The CPU port (oracle/, OpenMP over all host cores) on a bounded sample of cfg2: its first
    Poisson steps until ~budget_s of wall time is used."""
    import oracle_lib
    cores = os.cpu_count() or 1
    o = oracle_lib.Oracle(capacity=CFG2["capacity"], poisson_timestep=CFG2["poisson_timestep"], threads=cores)
    o.init_uniform(CFG2["init_n"])
    t0 = time.perf_counter()
    steps = 0
    while steps < CFG2["poisson_steps"] and time.perf_counter() - t0 < budget_s:
        o.poisson_step()
        steps += 1
    dt = time.perf_counter() - t0
    pushes = o.counters()["pushes"]
    o.close()
    return {"value": pushes / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"first {steps} of {CFG2['poisson_steps']} Poisson steps of cfg2 ({pushes} pushes, {dt:.1f} s), oracle/pic_oracle.cpp fp32 + OpenMP"}


def run_ours(args):
    dist, rank, world, local = dist_setup(args.gpus)
    eng = importlib.import_module("particle-simulation_b200")
    if eng.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; libpicb200 has no CPU fallback")
    # N>1: the charge grid is exchanged inside the persistent kernel over NVLink peer memory (handles
    # all-gathered over gloo below); PICB200_NCCL=1 selects the one-NCCL-all-reduce-per-Poisson-step path.
    use_nccl = world > 1 and os.environ.get("PICB200_NCCL", "0") not in ("", "0")
    uid = None
    if use_nccl:
        obj = [eng.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        uid = obj[0]
    # ---- N > 1: the N slabs must equal one GPU before anything is timed (exit code 2 otherwise)
    parity = None
    if world > 1 and not os.environ.get("PICB200_BENCH_NO_PARITY"):
        parity = parity_check(eng, dist, rank, world, local, uid)
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "n_gpus": world, "parity_check": parity, "error": "multi-GPU run differs from the single-GPU run"}), flush=True)
            raise SystemExit(2)
    n_local = CFG2["init_n"]
    kw = dict(CFG2)
    kw["init_n"] = n_local * world
    kw.update(device=local, rank=rank, nranks=world, nccl_unique_id=uid)
    sim = peer_sim(eng, dist, world, **kw) if (world > 1 and not use_nccl) else eng.Sim(**kw)
    id_base = rank * n_local
    P = CFG2["poisson_steps"]

    def barrier():
        if dist is not None:
            dist.barrier()

    # host copy of the initial population in pinned memory (for the e2e leg)
    sim.init_uniform(n_local, id_base)
    st0 = sim.download()
    cap_host = min(CFG2["capacity"], 4 * n_local)
    names_f = ("x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az")
    import numpy as np
    pin_in = eng.State(*[eng.pinned_empty(n_local, np.float32) for _ in names_f], eng.pinned_empty(n_local, np.int32), eng.pinned_empty(n_local, np.uint64))
    for k in list(names_f) + ["ts", "id"]:
        getattr(pin_in, k)[:] = getattr(st0, k)
    pin_out = eng.State(*[eng.pinned_empty(cap_host, np.float32) for _ in names_f], eng.pinned_empty(cap_host, np.int32), eng.pinned_empty(cap_host, np.uint64))

    for _ in range(args.warmup):
        sim.init_uniform(n_local, id_base)
        sim.run_steps(P)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()

    # ---- leg 1: device-resident inputs; device time of the Poisson loop
    sim.set_profiling(1)
    pushes = 0
    gpu_ms = 0.0
    barrier()
    for _ in range(args.steps):
        sim.init_uniform(n_local, id_base)
        sim.flush_l2()                     # after the init: the population it wrote is not L2-resident at t0
        barrier()
        sim.run_steps(P)
        s = sim.summary()
        pushes += s["pushes"]
        gpu_ms += s["gpu_ms"]
    barrier()
    prof = sim.profile()
    final = sim.summary()

    # ---- leg 2: end to end through the C ABI with host buffers (pinned), wall clock
    sim.set_profiling(0)
    e2e_pushes = 0
    d2h_bytes = 0
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sim.upload_state(pin_in)
        sim.run_steps(P)
        sm = sim.summary()
        if sm["final_n"] > cap_host:
            raise SystemExit(f"bench.py: population {sm['final_n']} outgrew the host buffer {cap_host}")
        n_out = sim.download_into(pin_out)
        e2e_pushes += sm["pushes"]
        d2h_bytes += n_out * 48
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None

    # ---- phase breakdown (separate pass; per-phase events perturb the loop slightly)
    sim.set_profiling(2)
    sim.init_uniform(n_local, id_base)
    sim.run_steps(P)
    phases = sim.profile()
    barrier()                              # no rank frees a grid a peer may still map
    sim.close()

    # ---- the drop-in call itself (N = 1): picb200_run(cfg, &summary, electrons_out, cap) -- what the runPIC shim
    # makes (csrc/runpic_shim.cpp) -- allocation, init, the run and the 64-byte AoS population into a pinned host buffer
    dropin = None
    if world == 1 and not os.environ.get("PICB200_BENCH_NO_DROPIN"):
        import ctypes as C
        buf = eng.pinned_empty(cap_host, eng.ELECTRON_DTYPE)
        cfg, keep = eng.make_config(**CFG2)
        walls, dpush, dn = [], 0, 0
        for it in range(6):
            sm = eng.Summary()
            t0 = time.perf_counter()
            rc = eng.lib.picb200_run(C.byref(cfg), C.byref(sm), buf.ctypes.data_as(C.c_void_p), cap_host)
            dt = time.perf_counter() - t0
            if rc != 0:
                raise SystemExit("bench.py: picb200_run failed: " + eng.lib.picb200_last_error().decode(errors="replace"))
            if it > 0:
                walls.append(dt)
                dpush += sm.pushes
                dn = int(sm.final_n)
        walls.sort()
        med = walls[len(walls) // 2]               # (allocation times of a box vary from call to call: median of five)
        dropin = {"value": dpush / len(walls) / med, "unit": UNIT, "wall_ms_per_run": 1e3 * med, "wall_ms_runs": [round(1e3 * w, 2) for w in walls], "gpu_ms_last": sm.gpu_ms, "d2h_bytes_per_run": dn * 64,
                  "what": "picb200_run(): create (all device allocations) + init on the device + 100 Poisson steps + 64-byte AoS export (device transpose) into a pinned host buffer + destroy, wall clock, median of five calls; the reference's `GPU time of program` excludes its allocations"}

    corner = None
    if world == 1 and not os.environ.get("PICB200_BENCH_NO_CORNER"):
        try:
            corner = hbm_corner_record(eng, local, measured_hbm_peak()[0])
        except eng.EngineError as e:
            corner = {"error": str(e)}

    cfg4 = None
    if not os.environ.get("PICB200_BENCH_NO_CFG4"):
        try:
            cfg4 = cfg4_record(eng, dist, rank, world, local, uid)
        except eng.EngineError as e:       # e.g. a smaller GPU: the headline line is still printed
            cfg4 = {"error": str(e)}

    t_max = all_max(dist, gpu_ms / 1e3)
    e2e_max = all_max(dist, e2e_s)
    pushes_all = all_sum(dist, float(pushes))
    e2e_pushes_all = all_sum(dist, float(e2e_pushes))
    if rank != 0:
        return
    peak, peak_src = measured_hbm_peak()
    mover_s = prof["mover_ms"] / 1e3
    achieved = pushes * BYTES_PER_PUSH / mover_s / 1e9 if mover_s > 0 else None
    traffic, traffic_note = ncu_traffic(pushes / max(1, int(prof["mover_launches"])))
    note = issue_note()
    rate1 = pushes / mover_s if mover_s > 0 else None       # rank 0's own kernel rate
    issue = None
    if note and rate1 and clocks and clocks.get("sm_mhz"):
        ceiling = 4.0 * final["sm_count"] * clocks["sm_mhz"] * 1e6 / note["warp_instructions_per_push"]
        issue = {"issue_ceiling_pushes_per_s": ceiling, "issue_frac": rate1 / ceiling,
                 "how": "4 warp-instructions per SM and cycle x SMs x SM clock sampled in this run / warp-instructions per push of the committed ncu capture", "source": note["source"]}
    line = {
        "metric": METRIC, "value": pushes_all / t_max, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_max * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "per_gpu_init_n": n_local, "global_init_n": n_local * world, "final_n_rank0": final["final_n"],
                   "pushes_per_step_rank0": pushes // max(1, args.steps), "parallelism": (f"particle slabs x{world}, " + ("NCCL all-reduce of the charge box per Poisson step" if use_nccl else "charge box summed inside the persistent kernel over NVLink peer memory")) if world > 1 else "single GPU",
                   "collective": ("ncclAllReduce(int32, sum) per Poisson step" if use_nccl else "none: in-kernel adds into every rank's summed grid (multimem.red through the NVSwitch, or PICB200_NVLS=0: TMA bulk reduce-adds into the peers' memory); no NCCL call on the data path") if world > 1 else "none",
                   "l2": "flushed after the population is initialised, before every timed run (512 MiB memset > 126 MB L2); the 48 MB working set then stays L2-resident within a run, as it does for the reference"},
        "e2e": {"value": e2e_pushes_all / e2e_max, "unit": UNIT, "h2d_bytes_per_step": n_local * 48, "d2h_bytes_per_step": d2h_bytes // max(1, args.steps),
                "what": "picb200_upload (pinned host SoA) + picb200_run_steps + picb200_download (pinned host SoA), wall clock"},
        "gpu_launches": int(prof["launches"]),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None, "traffic": traffic, "traffic_note": traffic_note,
                     "traffic_source": "static profile (committed ncu capture, not measured in this run)",
                     "binding_resource": "instruction issue (the kernel moves ~1 B/push through DRAM); `frac` is the 68-B step-synchronous equivalent, see issue_frac",
                     "issue_frac": issue["issue_frac"] if issue else None, "issue": issue,
                     "kernel": "k_poisson_fused", "launches": int(prof["mover_launches"]), "kernel_ms_total": prof["mover_ms"],
                     "kernel_share_of_step": prof["mover_ms"] / gpu_ms if gpu_ms else None,
                     "accounting": f"{BYTES_PER_PUSH:.0f} B per push x pushes per launch (SURVEY.md 8d); peak = {peak_src}; electrons stay in registers for all 10 mobility steps of a launch, so required HBM traffic is ~7.6 B/push and the 68-B-accounted figure may exceed the peak"},
        "issue_bound_note": note,
        "parity_check": parity,
        "e2e_dropin": dropin,
        "cfg4": cfg4,
        "roofline_hbm_corner": corner,
        "phases_ms_per_step": {k: v for k, v in phases.items() if k.endswith("_ms")},
        "clocks": clocks,
    }
    if world == 1 and not os.environ.get("PICB200_BENCH_NO_CPU"):      # (tools/ab.sh: kernel A/B runs skip the 12 s CPU leg)
        try:
            line["cpu_baseline"] = cpu_baseline()
        except Exception as e:  # the bench line must still be printed
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {e}"}
    print(json.dumps(line), flush=True)


def parse_ref_stdout(txt):
    ms = re.search(r"GPU time of program: ([0-9.]+) ms", txt)
    fin = re.search(r"Final amount of particles: (\d+)", txt)
    return (float(ms.group(1)) if ms else None), (int(fin.group(1)) if fin else None)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    exe = os.path.join(ref_dir, "main")
    have_gpu = shutil.which("nvidia-smi") is not None and subprocess.run(["nvidia-smi", "-L"], capture_output=True).returncode == 0
    base = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD}}
    if os.path.exists(exe) and have_gpu and args.ref_kind != "cpu":
        # the reference's own CUDA program, unmodified sources, one B200 (it has no multi-GPU path, src/pic.cu:394)
        os.makedirs(os.path.join(ref_dir, "src"), exist_ok=True)
        os.makedirs(os.path.join(ref_dir, "out", "visualization"), exist_ok=True)
        shutil.copyfile(os.path.join(ROOT, "src", "cross_section.txt"), os.path.join(ref_dir, "src", "cross_section.txt"))
        cmd = [exe, "30", "0", "1000000", "100", "256", "100000000", "100", "10"]
        tot_ms, tot_push, last = 0.0, 0.0, ""
        for it in range(args.warmup + args.steps):
            r = subprocess.run(cmd, cwd=ref_dir, capture_output=True, text=True, timeout=900)
            last = r.stdout[-600:]
            ms, fin = parse_ref_stdout(r.stdout)
            if r.returncode != 0 or ms is None:
                print(json.dumps({"impl": "reference", "unavailable": f"reference CUDA run failed rc={r.returncode}: {last[-200:]!r}"}))
                return
            if it >= args.warmup:
                tot_ms += ms
                # the reference does not count pushes: estimate from the mean population (critical branching, n ~ const)
                tot_push += 0.5 * (CFG2["init_n"] + fin) * CFG2["poisson_steps"] * CFG2["poisson_timestep"]
        v = tot_push / (tot_ms / 1e3)
        base.update({"value": v, "ms_per_step": tot_ms / args.steps,
                     "cpu_baseline": {"value": v, "unit": UNIT, "cores": 0, "kind": "reference",
                                      "sample": "reference CUDA program (oracle/_ref/main, sm_100 rebuild of the unmodified sources) on 1 B200, full cfg2, its own `GPU time of program`; the reference has no CPU implementation of this path; pushes estimated as mean(init n, final n) x 1000 steps"},
                     "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        print(json.dumps(base), flush=True)
        return
    # CPU port of the reference's kernels (the reference ships no CPU mover), all host threads
    tot_p, tot_s, cb = 0.0, 0.0, None
    for it in range(args.warmup + args.steps):
        cb = cpu_baseline(budget_s=8.0)
        if it >= args.warmup:
            m = re.search(r"\((\d+) pushes, ([0-9.]+) s\)", cb["sample"])
            tot_p += float(m.group(1))
            tot_s += float(m.group(2))
    v = tot_p / tot_s
    cb["value"] = v
    base.update({"value": v, "ms_per_step": tot_s * 1e3 / args.steps, "dtype": "f32", "cpu_baseline": cb,
                 "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(base), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-kind", default="auto", choices=["auto", "cuda", "cpu"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
