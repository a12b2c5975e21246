"""GPU parity tests: the CUDA engine (through the C ABI) against the CPU oracle on the same
seeded inputs.  Integer results (counts, ids, timestamps, cell counts, table indices) must be
bit-exact; fp32 state is compared bit-exactly too (both sides evaluate the same fp32 expressions
with explicit fused multiply-adds), which is stricter than north_star's 1e-5 relative bound.
Particle order is an atomic race on the GPU, so states are compared after sorting by particle id
-- the reference's own test sorts before comparing as well (src/test.cu:57, :75).
"""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MODES = [0, 1, 2, 3]            # Dynamic, CPU Sync, Naive, Dynamic Old  (src/main.cu:26-40)
F32 = ("x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az")


def hot_state(n, seed=1, sigma=0.05):
    rng = np.random.default_rng(seed)
    pos = rng.uniform(2.3, 2.8, size=(3, n)).astype(np.float32)
    vel = (rng.normal(size=(3, n)) * sigma).astype(np.float32)
    return pos, vel


def assert_state_equal(dev_state, orc_sorted, tol=None, ts_dead_only=False):
    d = dev_state.sorted_by_id()
    assert d.n == len(orc_sorted["id"])
    assert np.array_equal(d.id, orc_sorted["id"])
    if ts_dead_only:
        # Between the mobility phase and compaction the reference keeps a child's birth step in
        # `timestamp`; the engine already wrote the -1 that remove_dead_particles would write
        # (src/pic.cu:356).  Only dead/alive is comparable at that point.
        assert np.array_equal(d.ts == -2, orc_sorted["ts"] == -2)
        assert np.all((d.ts == -2) | (d.ts == -1))
    else:
        assert np.array_equal(d.ts, orc_sorted["ts"])
    for k in F32:
        a, b = getattr(d, k), orc_sorted[k].astype(np.float32)
        if tol is None:
            assert np.array_equal(a, b), f"{k}: {np.sum(a != b)} of {a.size} differ, max rel {np.max(np.abs(a - b) / (np.abs(b) + 1e-30))}"
        else:
            assert np.allclose(a, b, rtol=tol, atol=0), k


# ------------------------------------------------------------------------------------------
# arithmetic building blocks, evaluated on the device
# ------------------------------------------------------------------------------------------
def test_philox_known_answers_on_device(engine, oracle):
    from test_oracle import PHILOX_KAT
    for ctr, key, exp in PHILOX_KAT:
        assert [int(v) for v in engine.dev_philox(ctr, key)[0]] == exp
    rng = np.random.default_rng(0)
    ctr = rng.integers(0, 2 ** 32, size=(4096, 4), dtype=np.uint64).astype(np.uint32)
    key = np.array([39587, 0], dtype=np.uint32)
    assert np.array_equal(engine.dev_philox(ctr, key), oracle.philox(ctr, key))


def test_energy_index_on_device(engine, oracle):
    rng = np.random.default_rng(1)
    edges = engine.host_energy_edges()
    near = np.concatenate([np.nextafter(edges[1:-1], np.float32(0)), edges[1:-1], np.nextafter(edges[1:-1], np.float32(np.inf))])
    rnd = (10.0 ** rng.uniform(-12, 20, size=200000)).astype(np.float32)
    bits = rng.integers(0, 0x7F800000, size=200000, dtype=np.int64).astype(np.uint32).view(np.float32)
    special = np.array([0.0, 1e-45, 1.17e-38, 1e-6, 1.0, 3.4e38, np.inf, np.nan], dtype=np.float32)
    e = np.concatenate([near, rnd, bits, special]).astype(np.float32)
    got = engine.dev_energy_index(e)
    exp = np.array([oracle.energy_to_index(float(v)) for v in e], dtype=np.int32)
    assert np.array_equal(got, exp)


def test_energy_index_exhaustive_on_device(engine):
    """energy_index() (SFU lg2 estimate + one probe of the exact edge table) against a bisection of the table for EVERY
    non-negative fp32 bit pattern -- 2^31 values incl. zero, subnormals, +inf and all NaNs -- on the device."""
    bad, first = engine.dev_energy_index_sweep(stride=1)
    assert bad == 0, (bad, hex(first))


# ------------------------------------------------------------------------------------------
# kernels, one phase at a time
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 31, 4097, 100000])
def test_init_uniform_matches_oracle(engine, oracle, n):
    with engine.Sim(capacity=max(n, 1) + 5) as sim:
        sim.init_uniform(n, id_base=7)
        o = oracle.Oracle(capacity=max(n, 1) + 5, poisson_timestep=10)
        o.init_uniform(n, 7)
        assert_state_equal(sim.download(), o.download_sorted())


def test_deposit_and_gather_match_oracle(engine, oracle):
    n = 200000
    rng = np.random.default_rng(2)
    # a dense blob (many electrons per cell -> exercises the warp pre-reduction) plus electrons on
    # the domain faces (absent neighbours count as 0, src/grid_operations.cu:41-52)
    pos = rng.uniform(2.50, 2.56, size=(3, n)).astype(np.float32)
    pos[:, :64] = 0.0
    pos[:, 64:128] = np.float32(5.1199994)
    pos[0, 128:192] = 0.001
    with engine.Sim(capacity=n) as sim:
        sim.upload(*pos)
        sim.op_deposit()
        lo, hi = sim.charge_bbox()
        o = oracle.Oracle(capacity=n, poisson_timestep=10)
        o.upload(*pos)
        o.op_deposit()
        assert list(lo) == [0, 0, 0] and list(hi) == [511, 511, 511]
        for blo, bhi in (([248, 248, 248], [258, 258, 258]), ([0, 0, 0], [3, 3, 3]), ([508, 508, 508], [511, 511, 511])):
            got, exp = sim.charge_box(blo, bhi), o.charge_box(blo, bhi)
            assert np.array_equal(got, exp)
        assert sim.charge_box([248] * 3, [258] * 3).sum() + 192 >= n - 64  # nothing lost
        sim.op_gather()
        o.op_gather()
        assert_state_equal(sim.download(), o.download_sorted())
        # clearing restores the all-zero invariant
        sim.op_clear_grid()
        assert sim.charge_box([248] * 3, [258] * 3).sum() == 0 and sim.charge_box([0] * 3, [3] * 3).sum() == 0
        lo, hi = sim.charge_bbox()
        assert np.all(lo > hi)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n,pt", [(1, 10), (33, 1), (5000, 7), (60000, 10)])
def test_mobility_phase_matches_oracle(engine, oracle, mode, n, pt):
    """updateParticle + the scheduler of each mode (src/pic.cu:509-537), collisions switched on."""
    pos, vel = hot_state(n, seed=n)
    acc = (np.random.default_rng(5).normal(size=(3, n)) * 3e6).astype(np.float32)
    cap = 4 * n + 64
    with engine.Sim(capacity=cap, poisson_timestep=pt, mode=mode) as sim:
        sim.upload(*pos, *vel, *acc)
        sim.set_poisson_index(3)
        sim.op_mobility()
        o = oracle.Oracle(capacity=cap, poisson_timestep=pt)
        o.upload(*pos, *vel, *acc)
        o.set_poisson_index(3)
        o.op_mobility(stepwise=(mode == 2))
        assert sim.summary()["pushes"] == o.counters()["pushes"]
        assert_state_equal(sim.download(), o.download_sorted(), ts_dead_only=True)
        sim.op_compact()
        o.op_compact()
        d = sim.download()
        assert np.all(d.ts == -1)
        assert_state_equal(d, o.download_sorted())


def test_out_of_bounds_and_nan(engine, oracle):
    x = np.array([5.1199994, 2.5, 0.0000001, 2.5], dtype=np.float32)
    vx = np.array([1e9, 0.0, -1e9, np.nan], dtype=np.float32)
    y = np.full(4, 2.5, np.float32)
    for mode in MODES:
        with engine.Sim(capacity=16, poisson_timestep=4, mode=mode) as sim:
            sim.upload(x, y, y, vx)
            sim.op_mobility()
            o = oracle.Oracle(capacity=16, poisson_timestep=4)
            o.upload(x, y, y, vx)
            o.op_mobility(stepwise=(mode == 2))
            d = sim.download().sorted_by_id()
            s = o.download_sorted()
            assert list(d.ts == -2) == list(s["ts"] == -2) and d.ts[0] == -2 and d.ts[2] == -2
            assert sim.summary()["pushes"] == o.counters()["pushes"]


# ------------------------------------------------------------------------------------------
# whole Poisson steps and whole runs
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", MODES)
def test_poisson_steps_hot_population(engine, oracle, mode):
    n, pt, P = 20000, 10, 4
    pos, vel = hot_state(n, seed=11)
    with engine.Sim(capacity=200000, poisson_timestep=pt, mode=mode, poisson_steps=P) as sim:
        sim.upload(*pos, *vel)
        o = oracle.Oracle(capacity=200000, poisson_timestep=pt)
        o.upload(*pos, *vel)
        for step in range(P):
            sim.poisson_step()
            o.poisson_step(stepwise=(mode == 2))
            s, c = sim.summary(), o.counters()
            assert (s["final_n"], s["total_added"], s["total_removed"], s["pushes"]) == (c["n"], c["added"], c["removed"], c["pushes"]), step
        assert c["added"] > 1000 and c["removed"] > 1000
        assert_state_equal(sim.download(), o.download_sorted())


@pytest.mark.parametrize("mode", MODES)
def test_cfg1_reference_smallest_run(engine, oracle, mode):
    """BASELINE cfg1: init n 1e4, 10 Poisson steps x 10 mobility steps, capacity 1e6.  Collision
    free (row 0 of the table is `0.0 0.0`), so added = removed = 0 and the state is RNG-independent."""
    summ, el = engine.run(electrons=True, mode=mode, init_n=10000, poisson_steps=10, block_size=256, capacity=1000000, poisson_timestep=10)
    o = oracle.Oracle(capacity=1000000, poisson_timestep=10)
    o.init_uniform(10000)
    o.run(10, stepwise=(mode == 2))
    c = o.counters()
    assert (summ["final_n"], summ["total_added"], summ["total_removed"], summ["pushes"]) == (10000, 0, 0, 1000000)
    assert (c["n"], c["added"], c["removed"], c["pushes"]) == (10000, 0, 0, 1000000)
    s = o.download()          # no spawns, no deaths -> slot order is the id order on both sides
    assert el.shape[0] == 10000
    for k, name in enumerate("xyz"):        # all nine columns of the 64-byte record (src/electron.h:12-18), bit for bit
        assert np.array_equal(el["position"][:, k], s[name].astype(np.float32).astype(np.float64)), f"position.{name}"
        assert np.array_equal(el["velocity"][:, k], s["v" + name].astype(np.float32).astype(np.float64)), f"velocity.{name}"
        assert np.array_equal(el["acceleration"][:, k], s["a" + name].astype(np.float32)), f"acceleration.{name}"
    assert np.all(el["timestamp"] == -1)
    assert np.abs(el["velocity"]).max() > 0 and np.abs(el["acceleration"]).max() > 0


def test_run_with_collisions_all_modes_agree(engine, oracle):
    """1e5 electrons for 30 Poisson steps x 10: collisions start after ~100 mobility steps with the
    shipped table.  All four schedulers and the oracle must produce the same set of electrons."""
    kw = dict(init_n=100000, poisson_steps=30, capacity=2000000, poisson_timestep=10)
    o = oracle.Oracle(capacity=2000000, poisson_timestep=10, threads=4)
    o.init_uniform(100000)
    o.run(30)
    c, s = o.counters(), o.download_sorted()
    assert c["added"] > 100 and c["removed"] > 100
    for mode in MODES:
        with engine.Sim(mode=mode, **kw) as sim:
            sim.init_uniform(100000)
            sim.run_steps(30)
            m = sim.summary()
            assert (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) == (c["n"], c["added"], c["removed"], c["pushes"]), mode
            assert_state_equal(sim.download(), s)


def test_capacity_overflow(engine):
    rows = np.tile(np.array([[30.0, 0.0]], dtype=np.float32), (10000, 1))
    pos, vel = hot_state(1000, seed=5)
    for mode in MODES:
        with engine.Sim(capacity=1500, poisson_timestep=10, mode=mode, table_rows=rows) as sim:
            sim.upload(*pos, *vel)
            sim.poisson_step()
            m = sim.summary()
            assert m["final_n"] == 1500 and m["overflowed"] == 1
            assert m["total_added"] - m["total_removed"] == 500
            d = sim.download()
            assert d.n == 1500 and np.all(d.ts == -1) and len(np.unique(d.id)) == 1500


def test_population_dies_out(engine):
    rows = np.tile(np.array([[0.0, 100.0]], dtype=np.float32), (10000, 1))     # every push removes
    pos, vel = hot_state(500, seed=6)
    for mode in MODES:
        with engine.Sim(capacity=1000, poisson_timestep=3, mode=mode, table_rows=rows, poisson_steps=5) as sim:
            sim.upload(*pos, *vel)
            sim.run_steps(5)
            m = sim.summary()
            assert m["final_n"] == 0 and m["hit_zero"] == 1 and m["total_removed"] == 500 and m["pushes"] == 500
            assert m["poisson_steps_done"] == 1        # the reference breaks out of its loop (src/pic.cu:556-559)


def test_empty_population(engine):
    summ, _ = engine.run(mode=0, init_n=0, poisson_steps=3, capacity=100, poisson_timestep=10)
    assert summ["final_n"] == 0 and summ["pushes"] == 0


# ------------------------------------------------------------------------------------------
# size-independent properties at BASELINE's full cfg2 size
# ------------------------------------------------------------------------------------------
def test_cfg2_full_size_invariants(engine):
    """mode 30, init n 1e6, 100 x 10 steps, capacity 1e8 (BASELINE cfg2)."""
    with engine.Sim(mode=0, init_n=1000000, poisson_steps=100, capacity=100000000, poisson_timestep=10) as sim:
        sim.init_uniform(1000000)
        sim.run_steps(100)
        m = sim.summary()
        assert m["final_n"] == 1000000 + m["total_added"] - m["total_removed"]
        assert m["total_added"] > 10000 and m["total_removed"] > 10000 and m["overflowed"] == 0
        assert 0.9e9 < m["pushes"] < 1.2e9
        d = sim.download()
        assert d.n == m["final_n"] and np.all(d.ts == -1)
        assert len(np.unique(d.id)) == d.n                      # ids stay unique through spawn + compaction
        for k in "xyz":
            a = getattr(d, k)
            assert a.min() > 2.25 and a.max() < 2.89             # nobody left the initial box (dt = 1e-12)
        sim.op_deposit()
        lo, hi = sim.charge_bbox()
        assert sim.charge_box(lo, hi).sum() == d.n               # deposition conserves the count


def _state_digest(d):
    """Order-independent digest of a population: per-electron 64-bit mix of id and the raw fp32 bits, summed."""
    h = d.id.astype(np.uint64).copy()
    for k in ("x", "y", "z", "vx", "vy", "vz"):
        h = (h * np.uint64(0x9E3779B97F4A7C15)) ^ getattr(d, k).view(np.uint32).astype(np.uint64)
    return int(h.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(h))


def test_cfg2_full_size_all_schedulers_identical(engine):
    """BASELINE cfg2 in all four scheduler modes: the schedulers differ in WHEN a child is pushed, never in
    what happens to it, so counts and the multiset of (id, position, velocity) must be identical."""
    ref = None
    for mode in MODES:
        with engine.Sim(mode=mode, init_n=1000000, poisson_steps=100, capacity=100000000, poisson_timestep=10) as sim:
            sim.init_uniform(1000000)
            sim.run_steps(100)
            m = sim.summary()
            got = (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) + _state_digest(sim.download())
        if ref is None:
            ref = got
        assert got == ref, (mode, got, ref)


def test_split_run_equals_single_run(engine):
    """100 Poisson steps as 1 + 7 + 92 calls == one call (the persistent kernel carries no hidden state
    from launch to launch; RNG counters depend on the global step only)."""
    digests = []
    for parts in ((100,), (1, 7, 92)):
        with engine.Sim(mode=0, init_n=300000, poisson_steps=100, capacity=3000000, poisson_timestep=10) as sim:
            sim.init_uniform(300000)
            for k in parts:
                sim.run_steps(k)
            m = sim.summary()
            digests.append((m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) + _state_digest(sim.download()))
    assert digests[0] == digests[1] and digests[0][1] > 1000


def _run_policy(engine, policy, rows, pos, vel, cap, steps, pt=10, parts=None):
    """One fused-kernel run under a compaction policy of the engine (PICB200_COMPACT, read at create)."""
    old = os.environ.get("PICB200_COMPACT")
    if policy is None:
        os.environ.pop("PICB200_COMPACT", None)
    else:
        os.environ["PICB200_COMPACT"] = policy
    try:
        with engine.Sim(capacity=cap, poisson_timestep=pt, mode=0, table_rows=rows, poisson_steps=steps) as sim:
            sim.upload(*pos, *vel)
            for k in (parts or (steps,)):
                sim.run_steps(k)
            m = sim.summary()
            d = sim.download()
            assert np.all(d.ts == -1) and d.n == m["final_n"]
            return (m["final_n"], m["total_added"], m["total_removed"], m["pushes"], m["overflowed"]) + _state_digest(d)
    finally:
        if old is None:
            os.environ.pop("PICB200_COMPACT", None)
        else:
            os.environ["PICB200_COMPACT"] = old


@pytest.mark.parametrize("split,remove", [(0.5, 0.5), (0.2, 1.5), (1.5, 0.2), (0.0, 3.0)])
def test_lazy_compaction_equals_compaction_every_step(engine, oracle, split, remove):
    """The fused kernel leaves dead slots in place until they are worth a compaction; the reference compacts
    after every Poisson step (src/pic.cu:547-551).  Same electrons either way, steady, decaying and growing."""
    rows = np.tile(np.array([[split, remove]], dtype=np.float32), (10000, 1))
    pos, vel = hot_state(20000, seed=21)
    lazy = _run_policy(engine, None, rows, pos, vel, 400000, 12)
    eager = _run_policy(engine, "eager", rows, pos, vel, 400000, 12)
    split_run = _run_policy(engine, None, rows, pos, vel, 400000, 12, parts=(5, 1, 6))
    assert lazy == eager == split_run
    o = oracle.Oracle(capacity=400000, poisson_timestep=10, table_rows=rows)
    o.upload(*pos, *vel)
    o.run(12)
    c = o.counters()
    assert lazy[:4] == (c["n"], c["added"], c["removed"], c["pushes"])


@pytest.mark.parametrize("drain_at", [1, 16, 33, 40, 64])
@pytest.mark.parametrize("pt,split,remove", [(10, 1.8, 1.8), (10, 6.0, 5.0), (7, 12.0, 14.0)])
def test_children_batching_policy_matches_oracle(engine, oracle, drain_at, pt, split, remove):
    """The fused kernel lets children wait in the warp's 64-entry stage and pushes them 32 at a time, earliest-born
    first (stage_select_earliest); PICB200_DRAIN_AT moves the point at which a warp stops to push a batch, from
    "at once" (1) to "only when the stage is full or the chunks are gone" (64: the stage overflows and spills in
    the heavy tables).  Whatever the batches, the electrons are those of the oracle, bit for bit."""
    rows = np.tile(np.array([[split, remove]], dtype=np.float32), (10000, 1))
    n = 640000                                   # ~4 chunks per warp of a 148-SM grid: the stages fill up across chunks
    pos, vel = hot_state(n, seed=33)
    old = os.environ.get("PICB200_DRAIN_AT")
    os.environ["PICB200_DRAIN_AT"] = str(drain_at)
    try:
        with engine.Sim(capacity=4 * n, poisson_timestep=pt, mode=0, table_rows=rows, poisson_steps=3) as sim:
            sim.upload(*pos, *vel)
            sim.run_steps(3)
            m, d = sim.summary(), sim.download()
    finally:
        if old is None:
            os.environ.pop("PICB200_DRAIN_AT", None)
        else:
            os.environ["PICB200_DRAIN_AT"] = old
    o = oracle.Oracle(capacity=4 * n, poisson_timestep=pt, table_rows=rows)
    o.upload(*pos, *vel)
    o.run(3)
    c = o.counters()
    assert (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) == (c["n"], c["added"], c["removed"], c["pushes"])
    assert m["total_added"] > n // 4
    assert_state_equal(d, o.download_sorted())


@pytest.mark.parametrize("ratio", [-1, 0, 1, 16])
@pytest.mark.parametrize("pt,mode", [(1, 0), (10, 0), (4, 1), (4, 2)])
def test_per_cell_field_path_matches_oracle(engine, oracle, monkeypatch, ratio, pt, mode):
    """Dense plasmas: the fused kernel computes the field once per cell of the occupied box (field_rows, the
    reference's updateGrid, src/grid_operations.cu:29-56) and every electron fetches one record, instead of
    differencing six counts per electron.  PICB200_FIELD_RATIO (read at create) moves the density from which it
    does: -1 never, 0 always, 1 and 16 electrons per cell (the population below has ~3).  Same bits.
    Modes 31/32 (mode 1, 2 here) do the same with a kernel of their own (k_field_box) in front of the gather; it is
    part of their grid phase from 5e5 configured electrons, PICB200_FIELD_LAUNCH=1 puts it there for this small one."""
    monkeypatch.setenv("PICB200_FIELD_LAUNCH", "1")
    rows = np.tile(np.array([[3.0, 1.0]], dtype=np.float32), (10000, 1))
    rng = np.random.default_rng(77)
    n = 300000
    pos = rng.uniform(2.4, 2.86, size=(3, n)).astype(np.float32)        # 46^3 cells: ~3 electrons per cell
    vel = (rng.normal(size=(3, n)) * 0.05).astype(np.float32)
    old = os.environ.get("PICB200_FIELD_RATIO")
    os.environ["PICB200_FIELD_RATIO"] = str(ratio)
    try:
        with engine.Sim(capacity=8 * n, poisson_timestep=pt, mode=mode, table_rows=rows, poisson_steps=5) as sim:
            sim.upload(*pos, *vel)
            sim.run_steps(2)
            sim.run_steps(3)
            m, d = sim.summary(), sim.download()
    finally:
        if old is None:
            os.environ.pop("PICB200_FIELD_RATIO", None)
        else:
            os.environ["PICB200_FIELD_RATIO"] = old
    o = oracle.Oracle(capacity=8 * n, poisson_timestep=pt, table_rows=rows)
    o.upload(*pos, *vel)
    o.run(5)
    c = o.counters()
    assert (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) == (c["n"], c["added"], c["removed"], c["pushes"])
    assert_state_equal(d, o.download_sorted())


@pytest.mark.parametrize("stream_steps,pt,field_ratio", [(0, 1, None), (3, 1, None), (3, 1, "-1"), (3, 2, None), (3, 3, None), (3, 3, "-1"), (16, 10, None)])
def test_streaming_step_loop_matches_oracle(engine, oracle, monkeypatch, stream_steps, pt, field_ratio):
    """Small poisson_timestep: the fused kernel streams its chunks (cp.async double buffer, a shared-memory window of the
    first 2048 table rows; k_poisson_fused kVariant 2).  PICB200_STREAM_STEPS moves the threshold (0: never; 16: also at
    poisson_timestep 10).  A third of the electrons is fast enough (|v|^2 > 5.7) to use table rows BEYOND the window,
    which come from the global image.  At 3.7 electrons per cell the per-cell field is on, so the field records are
    prefetched a chunk ahead (stream_field_issue); PICB200_FIELD_RATIO=-1 runs the same loop with the six-point gather.
    Same bits as the oracle; ragged population (not a multiple of 32)."""
    if field_ratio is not None:
        monkeypatch.setenv("PICB200_FIELD_RATIO", field_ratio)
    rows = np.zeros((10000, 2), dtype=np.float32)
    rows[:, 0] = 1.5 + (np.arange(10000) % 7) * 0.25          # every row different from its neighbours: a wrong row shows
    rows[:, 1] = 1.0 + (np.arange(10000) % 5) * 0.2
    rng = np.random.default_rng(5)
    n = 100003
    pos = rng.uniform(2.45, 2.75, size=(3, n)).astype(np.float32)
    vel = (rng.normal(size=(3, n)) * 0.05).astype(np.float32)
    vel[:, ::3] *= 60.0                                       # |v| ~ 3-5 m/s: rows ~3100-3400
    monkeypatch.setenv("PICB200_STREAM_STEPS", str(stream_steps))
    steps = 12 if pt < 10 else 4
    with engine.Sim(capacity=6 * n, poisson_timestep=pt, mode=0, table_rows=rows, poisson_steps=steps) as sim:
        sim.upload(*pos, *vel)
        sim.run_steps(steps // 2)
        sim.run_steps(steps - steps // 2)
        m, d = sim.summary(), sim.download()
    o = oracle.Oracle(capacity=6 * n, poisson_timestep=pt, table_rows=rows, threads=4)
    o.upload(*pos, *vel)
    o.run(steps)
    c = o.counters()
    assert (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) == (c["n"], c["added"], c["removed"], c["pushes"])
    assert c["added"] > 1000 and c["removed"] > 1000
    assert_state_equal(d, o.download_sorted())
    e = np.stack([d.vx, d.vy, d.vz]).astype(np.float64)
    assert ((e ** 2).sum(axis=0) > 5.7).sum() > 1000           # rows beyond the window were in use until the end


@pytest.mark.parametrize("pt", [3, 10])
def test_uploaded_timestamps_first_launch_generic(engine, oracle, pt):
    """A population uploaded WITH timestamps (some electrons "born at step t" of the coming Poisson step, src/pic.cu:218:
    their first push is step t + 1): the fused kernel's generic chunk path.  The streaming instantiation (poisson_timestep
    <= 3) has no copy of that path; the host runs the first launch after such an upload on the plain instantiation, which
    leaves every timestamp at -1, and streams from the second launch on.  Same bits as the oracle after each launch."""
    rows = np.tile(np.array([[2.0, 1.0]], dtype=np.float32), (10000, 1))
    rng = np.random.default_rng(41)
    n = 50001
    pos = rng.uniform(2.45, 2.7, size=(3, n)).astype(np.float32)
    vel = (rng.normal(size=(3, n)) * 0.05).astype(np.float32)
    ts = np.full(n, -1, dtype=np.int32)
    ts[::3] = rng.integers(1, pt, size=len(ts[::3]))            # a third arrives "born" somewhere inside the step
    with engine.Sim(capacity=8 * n, poisson_timestep=pt, mode=0, table_rows=rows, poisson_steps=6) as sim:
        sim.upload(*pos, *vel, ts=ts)
        o = oracle.Oracle(capacity=8 * n, poisson_timestep=pt, table_rows=rows, threads=4)
        o.upload(*pos, *vel, ts=ts)
        for k in (1, 2, 3):
            sim.run_steps(k)
            o.run(k)
            m, c = sim.summary(), o.counters()
            assert (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) == (c["n"], c["added"], c["removed"], c["pushes"]), k
            assert_state_equal(sim.download(), o.download_sorted())
    assert c["added"] > 100 and c["pushes"] < 6 * pt * 2 * n


def test_lazy_compaction_parks_children_near_capacity(engine, oracle):
    """Holes in place shift the slots of new children upwards; a child whose shifted slot passes the capacity
    while the reference's counter does not is parked and placed by the compaction the step then ends with.
    `noguard` switches off the policy that normally compacts before this can happen."""
    rows = np.tile(np.array([[0.5, 0.5]], dtype=np.float32), (10000, 1))
    pos, vel = hot_state(20000, seed=22)
    for cap in (20900, 21500, 23000):       # overflow in step 1 / parking in later steps / mostly lazy
        lazy = _run_policy(engine, "noguard", rows, pos, vel, cap, 8)
        eager = _run_policy(engine, "eager", rows, pos, vel, cap, 8)
        assert bool(lazy[4]) == bool(eager[4]), cap      # (a count of overflowing steps; how many is a race)
        if eager[4]:                         # overflow: WHICH children are dropped is a race (in the reference, too)
            for r in (lazy, eager):
                assert r[0] <= cap and r[0] == 20000 + r[1] - r[2]
        else:                                # nobody is dropped: same electrons
            assert lazy[:4] + lazy[5:] == eager[:4] + eager[5:], cap
            o = oracle.Oracle(capacity=cap, poisson_timestep=10, table_rows=rows)
            o.upload(*pos, *vel)
            o.run(8)
            c = o.counters()
            assert lazy[:4] == (c["n"], c["added"], c["removed"], c["pushes"])


@pytest.mark.parametrize("pt,split,remove", [(64, 2.5, 2.5), (100, 3.0, 2.0), (32, 1.0, 6.0)])
def test_adopting_step_loop_matches_oracle(engine, oracle, pt, split, remove):
    """poisson_timestep >= 24 selects the kernel whose step loop hands a new-born child to a dead lane of the
    warp on the spot (lockstep_push, kAdopt): with several per cent of split / remove per push every lane changes
    its occupant a few times per Poisson step.  Same electrons, bit for bit, as the oracle and as mode 31."""
    rows = np.tile(np.array([[split, remove]], dtype=np.float32), (10000, 1))
    n, P, cap = 20000, 3, 3000000        # (3.0, 2.0) x 100 pushes: the reference's counter n peaks at 30 x the start
    pos, vel = hot_state(n, seed=31)
    o = oracle.Oracle(capacity=cap, poisson_timestep=pt, table_rows=rows, threads=4)
    o.upload(*pos, *vel)
    o.run(P)
    c, s = o.counters(), o.download_sorted()
    assert c["added"] > n // 8 and c["removed"] > n // 2     # lanes die out and are handed children: the adoption path is busy
    assert o.counters()["n"] + c["added"] < cap              # (no overflow: which child is dropped would be a race)
    for mode in (0, 1):
        with engine.Sim(capacity=cap, poisson_timestep=pt, mode=mode, table_rows=rows, poisson_steps=P) as sim:
            sim.upload(*pos, *vel)
            sim.run_steps(P)
            m = sim.summary()
            assert (m["final_n"], m["total_added"], m["total_removed"], m["pushes"]) == (c["n"], c["added"], c["removed"], c["pushes"]), mode
            assert_state_equal(sim.download(), s)


def test_adopting_step_loop_capacity_overflow(engine):
    """Adopted children ask the reference's counter n for their existence only when their chunk ends; in a run
    that overflows those past the capacity are taken back.  Conservation must hold whatever the race."""
    rows = np.tile(np.array([[4.0, 1.0]], dtype=np.float32), (10000, 1))
    pos, vel = hot_state(5000, seed=32)
    for pt, policy in ((64, None), (64, "eager"), (10, None), (10, "noguard")):     # adopting / plain step loop, lazy / eager compaction
        old = os.environ.get("PICB200_COMPACT")
        if policy:
            os.environ["PICB200_COMPACT"] = policy
        try:
            for rep in range(3):                               # which children are dropped is a race: a few rolls of the dice
                with engine.Sim(capacity=9000, poisson_timestep=pt, mode=0, table_rows=rows, poisson_steps=4) as sim:
                    sim.upload(*pos, *vel)
                    sim.run_steps(4)
                    m = sim.summary()
                    assert m["overflowed"] >= 1 and m["final_n"] <= 9000
                    assert m["final_n"] == 5000 + m["total_added"] - m["total_removed"]
                    d = sim.download()
                    assert d.n == m["final_n"] and np.all(d.ts == -1), (pt, policy)
                    assert len(np.unique(d.id)) == d.n, (pt, policy, d.n - len(np.unique(d.id)))    # nobody twice
        finally:
            os.environ.pop("PICB200_COMPACT", None)
            if old is not None:
                os.environ["PICB200_COMPACT"] = old


# ------------------------------------------------------------------------------------------
# the outer boundary: the `run` CLI and its stdout contract
# ------------------------------------------------------------------------------------------
def test_cli_stdout_matches_oracle_cli(engine, oracle, tmp_path):
    """`run 32 1 200 3 256 1000 100 5` prints the reference's lines (src/pic.cu:360, :404, :457-474,
    :573-576, src/utility.cu:124-137, src/electron.h:19-26).  Collision-free, so the engine's and the
    oracle's state dumps must be identical text."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    args = ["30", "1", "200", "3", "256", "1000", "100", "5"]
    a = subprocess.run([os.path.join(root, "run")] + args, capture_output=True, text=True, cwd=root)
    b = subprocess.run([oracle.CLI] + args, capture_output=True, text=True, cwd=root)
    assert a.returncode == 0, a.stdout + a.stderr

    def stable(txt):
        keep = []
        for ln in txt.splitlines():
            if ln.startswith(("GPU time of program", "CPU time of program", "Multiprocessor count", "Blocks per multiprocessor")):
                keep.append(ln.split(":")[0])
            else:
                keep.append(ln)
        return keep

    la, lb = stable(a.stdout), stable(b.stdout)
    assert la[:7] == ["PIC with", "init n: 200", "capacity: 1000", "poisson steps: 3", "poisson_timestep: 5", "block size: 256", "sleep time: 100"]
    assert la == lb
    assert sum(1 for ln in la if ln.startswith("Time ")) == 4 and "Final amount of particles: 200" in la
