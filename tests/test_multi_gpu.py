"""Multi-GPU parity (needs >= 2 B200s: `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu`).

Particle slabs: rank r owns initial ids [r*n/k, (r+1)*n/k); the only coupling is the charge grid, summed
over the ranks once per Poisson step -- inside the persistent kernel over NVLink peer memory
(picb200_peer_connect), or by one NCCL all-reduce per step (nccl_unique_id).  Philox is keyed on the
particle id, so the union of the slabs must equal the single-population oracle run bit for bit.
"""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

F32 = ("x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az")


def run_ranks(engine, k, transport, mode, n, cap, pt, steps_list, pos=None, vel=None, launches_per_entry=1):
    """Run k ranks in k host threads of this process; returns per-rank (summary, State) after each entry of steps_list."""
    uid = engine.nccl_unique_id() if transport == "nccl" else None
    handles = [None] * k
    out = [[] for _ in range(k)]
    err = []
    bar = threading.Barrier(k)

    def work(r):
        try:
            sim = engine.Sim(mode=mode, capacity=cap // k, poisson_timestep=pt, poisson_steps=sum(steps_list), device=r, rank=r, nranks=k, nccl_unique_id=uid)
            if transport == "peer":
                handles[r] = sim.peer_export()
                bar.wait()
                sim.peer_connect(handles)
            lo, hi = n * r // k, n * (r + 1) // k
            if pos is None:
                sim.init_uniform(hi - lo, lo)
            else:
                sim.upload(*[a[lo:hi] for a in pos], *[a[lo:hi] for a in vel], id=np.arange(lo, hi, dtype=np.uint64))
            bar.wait()
            for st in steps_list:
                sim.run_steps(st)
                out[r].append((sim.summary(), sim.download()))
            bar.wait()
            sim.close()
        except Exception as e:  # noqa: BLE001
            err.append((r, repr(e)))
            bar.abort()

    th = [threading.Thread(target=work, args=(r,)) for r in range(k)]
    [t.start() for t in th]
    [t.join(timeout=300) for t in th]
    assert not err, err
    return out


def union_sorted(states):
    ids = np.concatenate([s.id for s in states])
    o = np.argsort(ids, kind="stable")
    return {f: np.concatenate([getattr(s, f) for s in states])[o] for f in F32 + ("ts", "id")}


def need_gpus(engine, k):
    if engine.device_count() < k:
        pytest.skip(f"needs {k} CUDA devices, {engine.device_count()} visible")


# "peer": rows added into every rank's summed grid inside the kernel -- through the NVSwitch (one multimem.red per pair of
# cells, csrc/pic_nvls.h) where the box has one, else / "peer-unicast" (PICB200_NVLS=0) by one TMA bulk reduce-add per peer
@pytest.mark.parametrize("transport,pt", [("peer", 10), ("peer-unicast", 10), ("nccl", 10), ("peer", 32)])      # pt 32: the adopting step loop
@pytest.mark.parametrize("k", [2])
def test_slabs_equal_single_population_oracle(engine, oracle, monkeypatch, transport, pt, k):
    need_gpus(engine, k)
    if transport == "peer-unicast":
        monkeypatch.setenv("PICB200_NVLS", "0")
        transport = "peer"
    n, cap = 40000, 800000
    rng = np.random.default_rng(3)
    pos = rng.uniform(2.45, 2.65, size=(3, n)).astype(np.float32)     # dense: ~5 electrons per cell, real fields
    vel = (rng.normal(size=(3, n)) * 0.05).astype(np.float32)          # hot: collisions from the first push
    steps_list = [1, 3, 2]                                             # several launches: the exchange state carries over
    res = run_ranks(engine, k, transport, 0, n, cap, pt, steps_list, pos, vel)
    o = oracle.Oracle(capacity=cap, poisson_timestep=pt, threads=4)
    o.upload(*pos, *vel)
    for j, st in enumerate(steps_list):
        o.run(st)
        c, ref = o.counters(), o.download_sorted()
        sums = [res[r][j][0] for r in range(k)]
        assert sum(s["final_n"] for s in sums) == c["n"]
        assert sum(s["total_added"] for s in sums) == c["added"] and sum(s["total_removed"] for s in sums) == c["removed"]
        assert sum(s["pushes"] for s in sums) == c["pushes"]
        u = union_sorted([res[r][j][1] for r in range(k)])
        assert np.array_equal(u["id"], ref["id"]) and np.array_equal(u["ts"], ref["ts"])
        for f in F32:
            assert np.array_equal(u[f], ref[f].astype(np.float32)), (transport, j, f)
    assert c["added"] > 1000 and c["removed"] > 1000


@pytest.mark.parametrize("k,nvls", [(2, "1"), (4, "1"), (8, "1"), (8, "0")])
def test_peer_mode_equals_single_gpu_run(engine, monkeypatch, k, nvls):
    """cfg1-shaped run (collision free) and a short colliding run: k slabs == 1 GPU, after sorting by id."""
    need_gpus(engine, k)
    monkeypatch.setenv("PICB200_NVLS", nvls)
    n, pt, cap, P = 64000, 10, 1024000, 12
    with engine.Sim(mode=0, capacity=cap, poisson_timestep=pt, poisson_steps=P) as one:
        one.init_uniform(n)
        one.run_steps(P)
        s1, d1 = one.summary(), one.download().sorted_by_id()
    res = run_ranks(engine, k, "peer", 0, n, cap, pt, [P])
    sums = [res[r][0][0] for r in range(k)]
    assert sum(s["final_n"] for s in sums) == s1["final_n"] and sum(s["pushes"] for s in sums) == s1["pushes"]
    u = union_sorted([res[r][0][1] for r in range(k)])
    assert np.array_equal(u["id"], d1.id)
    for f in F32:
        assert np.array_equal(u[f], getattr(d1, f)), f


def test_empty_rank_keeps_stepping(engine):
    """A rank whose slab is empty still takes part in every exchange (no deadlock, right field)."""
    need_gpus(engine, 2)
    n, pt, cap = 1, 5, 1000       # rank 0 gets 0 electrons, rank 1 gets 1
    res = run_ranks(engine, 2, "peer", 0, n, cap, pt, [3])
    assert res[0][0][0]["final_n"] + res[1][0][0]["final_n"] == 1
    assert res[0][0][0]["pushes"] + res[1][0][0]["pushes"] == 15


@pytest.mark.parametrize("k", [2])
def test_second_population_on_the_same_ranks(engine, k):
    """Two populations one after the other on the SAME simulations (what bench.py does between timed runs): the summed
    grid that still holds the last exchange of the first population must be emptied before the second one uses it, and
    the exchange counters carry on.  k slabs == one GPU for both populations."""
    need_gpus(engine, k)
    n, pt, cap = 48000, 10, 960000
    rng = np.random.default_rng(11)
    pops = []
    for shift in (0.0, 0.07):
        pos = (rng.uniform(2.45, 2.65, size=(3, n)) + shift).astype(np.float32)
        vel = (rng.normal(size=(3, n)) * 0.05).astype(np.float32)
        pops.append((pos, vel))
    singles = []
    with engine.Sim(mode=0, capacity=cap, poisson_timestep=pt, poisson_steps=8) as one:
        for pos, vel in pops:
            one.upload(*pos, *vel)
            one.run_steps(3)
            one.run_steps(2)
            singles.append((one.summary(), one.download().sorted_by_id()))
    handles = [None] * k
    out = [[] for _ in range(k)]
    err = []
    bar = threading.Barrier(k)

    def work(r):
        try:
            sim = engine.Sim(mode=0, capacity=cap // k, poisson_timestep=pt, poisson_steps=8, device=r, rank=r, nranks=k)
            handles[r] = sim.peer_export()
            bar.wait()
            sim.peer_connect(handles)
            lo, hi = n * r // k, n * (r + 1) // k
            for pos, vel in pops:
                sim.upload(*[a[lo:hi] for a in pos], *[a[lo:hi] for a in vel], id=np.arange(lo, hi, dtype=np.uint64))
                bar.wait()
                sim.run_steps(3)
                sim.run_steps(2)
                out[r].append((sim.summary(), sim.download()))
                bar.wait()
            sim.close()
        except Exception as e:  # noqa: BLE001
            err.append((r, repr(e)))
            bar.abort()

    th = [threading.Thread(target=work, args=(r,)) for r in range(k)]
    [t.start() for t in th]
    [t.join(timeout=300) for t in th]
    assert not err, err
    for j, (s1, d1) in enumerate(singles):
        sums = [out[r][j][0] for r in range(k)]
        assert sum(s["final_n"] for s in sums) == s1["final_n"] and sum(s["pushes"] for s in sums) == s1["pushes"], j
        assert sum(s["total_added"] for s in sums) == s1["total_added"] and s1["total_added"] > 500
        u = union_sorted([out[r][j][1] for r in range(k)])
        assert np.array_equal(u["id"], d1.id), j
        for f in F32:
            assert np.array_equal(u[f], getattr(d1, f)), (j, f)


def test_large_slabs_use_the_per_cell_field_in_peer_mode(engine, monkeypatch):
    """Peer mode with the per-cell field path forced on (PICB200_FIELD_MIN_N=0, ratio 0) and off: both equal one GPU."""
    need_gpus(engine, 2)
    n, pt, cap, P = 64000, 10, 1024000, 6
    with engine.Sim(mode=0, capacity=cap, poisson_timestep=pt, poisson_steps=P) as one:
        one.init_uniform(n)
        one.run_steps(P)
        s1, d1 = one.summary(), one.download().sorted_by_id()
    for ratio in ("0", "-1"):
        monkeypatch.setenv("PICB200_FIELD_MIN_N", "0")
        monkeypatch.setenv("PICB200_FIELD_RATIO", ratio)
        res = run_ranks(engine, 2, "peer", 0, n, cap, pt, [2, 4])
        sums = [res[r][1][0] for r in range(2)]
        assert sum(s["final_n"] for s in sums) == s1["final_n"] and sum(s["pushes"] for s in sums) == s1["pushes"]
        u = union_sorted([res[r][1][1] for r in range(2)])
        assert np.array_equal(u["id"], d1.id)
        for f in F32:
            assert np.array_equal(u[f], getattr(d1, f)), (ratio, f)
