"""CPU tests that pin the oracle (oracle/pic_oracle.cpp) -- no GPU needed.

The reference ships no golden vectors for this path, so the oracle is pinned by:
  * the Random123 known-answer vectors of Philox4x32-10,
  * the reference's own energyToIndex formula (src/cross_section.cu:32-35) at hand-computed points,
  * the constants of src/cell.cu:3-5,
  * the collision-free closed form: with zero initial velocity the first steps
    cannot collide, so velocities are an exact function of the frozen accelerations,
  * equivalence of its two scheduling orders (the reference's own `test` mode, src/test.cu:43-100).
"""
import math

import numpy as np
import pytest

PHILOX_KAT = [  # Random123 kat_vectors, philox4x32 10 rounds: counter, key, expected
    ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
    ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
    ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0], [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
]


def test_philox_known_answers(oracle):
    for ctr, key, exp in PHILOX_KAT:
        assert [int(v) for v in oracle.philox(ctr, key)[0]] == exp


def test_uniform_mapping_is_half_open(oracle):
    # curand_uniform: (0, 1]  (src/random.cu:8)
    assert oracle.lib.orc_uniform01(0) == pytest.approx(2.0 ** -33, rel=0, abs=0)
    assert oracle.lib.orc_uniform01(0xFFFFFFFF) == 1.0
    assert oracle.lib.orc_uniform01(0x80000000) == np.float32(0.5 + 2.0 ** -33)


def test_electric_force_constant(oracle):
    # SURVEY.md probe table: 2532637.9954267191 from src/cell.cu:5
    assert oracle.lib.orc_electric_force_constant() == pytest.approx(2532637.9954267191, rel=1e-15)


def test_energy_to_index_reference_points(oracle):
    f = oracle.energy_to_index
    assert f(0.0) == 0                      # log10(0) = -inf -> saturates -> 0
    assert f(1e-7) == 0                     # below 1e-6
    assert f(1e-6) == 0
    assert f(10 ** (-6 + 22 / 10000) * (1 + 1e-12)) == 1
    assert f(1.0) == int(6 * 10000 / 22)    # 2727
    assert f(1e16) == 9999 and f(1e30) == 9999
    assert f(float("inf")) == 9999
    assert f(float("nan")) == 0
    for e in [3.3e-5, 7.7e-3, 12.5, 4.2e9]:
        assert f(e) == min(9999, max(0, math.trunc((math.log10(e) + 6) * 10000 / 22)))


def test_child_ids_are_distinct_and_tagged(oracle):
    ids = {oracle.lib.orc_child_id(p, g) for p in range(200) for g in range(1, 50)}
    assert len(ids) == 200 * 49
    assert all(i >> 63 for i in ids)


def test_table_loader_matches_numpy(oracle):
    r, rows = oracle.load_table()
    assert r == 10000
    ref = np.loadtxt(oracle.DEFAULT_TABLE, dtype=np.float64).astype(np.float32)
    assert np.array_equal(rows, ref)
    assert rows[0, 0] == 0.0 and rows[0, 1] == 0.0


def test_init_positions_inside_reference_box(oracle):
    o = oracle.Oracle(capacity=5000, poisson_timestep=10)
    o.init_uniform(4000)
    s = o.download()
    for k in "xyz":       # (G/2-30)*cs .. (G/2+32)*cs, src/particle_move.cu:13-15
        assert s[k].min() > 2.26 - 1e-6 and s[k].max() <= np.float32(2.88)
    assert np.all(s["ts"] == -1) and np.all(s["vx"] == 0) and np.array_equal(s["id"], np.arange(4000, dtype=np.uint64))
    # fp32 and fp64 oracles start from identical (fp32-representable) positions
    o64 = oracle.Oracle(capacity=5000, poisson_timestep=10, fp64=True)
    o64.init_uniform(4000)
    assert np.array_equal(o64.download()["x"], s["x"])


def test_collision_free_closed_form(oracle):
    """cfg1-like run: no collisions can occur while |v|^2 < 10^(-6+22/10000); velocity after k
    pushes under a frozen field a is -a*dt*k up to fp32 rounding, added = removed = 0."""
    n, pt, P = 3000, 10, 3
    o = oracle.Oracle(capacity=10000, poisson_timestep=pt)
    o.init_uniform(n)
    o64 = oracle.Oracle(capacity=10000, poisson_timestep=pt, fp64=True)
    o64.init_uniform(n)
    for _ in range(P):
        o.poisson_step()
        o64.poisson_step()
    c = o.counters()
    assert (c["added"], c["removed"], c["pushes"], c["n"]) == (0, 0, n * pt * P, n)
    s, d = o.download(), o64.download()
    # positions never move between cells, so the field is the same every Poisson step
    for v, a in (("vx", "ax"), ("vy", "ay"), ("vz", "az")):
        expect = -s[a].astype(np.float64) * 1e-12 * pt * P
        assert np.allclose(s[v], expect, rtol=1e-5, atol=1e-12)
        assert np.allclose(d[v], expect, rtol=1e-9, atol=1e-15)
        assert np.allclose(s[v], d[v], rtol=1e-5, atol=1e-12)      # north_star: 1e-5 relative in fp32
    assert np.array_equal(s["ax"], d["ax"])


def _hot_state(n, seed=1):
    """Electrons fast enough to collide at once (|v|^2 ~ 1e-4..1e-2 -> table rows 900-1800)."""
    rng = np.random.default_rng(seed)
    pos = rng.uniform(2.3, 2.8, size=(3, n)).astype(np.float32)
    vel = (rng.normal(size=(3, n)) * 0.05).astype(np.float32)
    return pos, vel


def test_scheduler_orders_agree(oracle):
    """dynamic/cpuSync semantics (generations) == naive semantics (step by step), as sets."""
    n, pt = 2000, 7
    pos, vel = _hot_state(n)
    res = []
    for stepwise in (False, True):
        o = oracle.Oracle(capacity=20000, poisson_timestep=pt)
        o.upload(*pos, *vel)
        for _ in range(3):
            o.poisson_step(stepwise)
        res.append((o.counters(), o.download_sorted()))
    (c0, s0), (c1, s1) = res
    assert c0 == c1 and c0["added"] > 50 and c0["removed"] > 50
    for k in s0:
        assert np.array_equal(s0[k], s1[k]), k


def test_openmp_oracle_equals_serial(oracle):
    n, pt = 3000, 5
    pos, vel = _hot_state(n, seed=3)
    out = []
    for th in (1, 4):
        o = oracle.Oracle(capacity=30000, poisson_timestep=pt, threads=th)
        o.upload(*pos, *vel)
        o.run(3)
        out.append((o.counters(), o.download_sorted()))
    assert out[0][0] == out[1][0]
    for k in out[0][1]:
        assert np.array_equal(out[0][1][k], out[1][1][k]), k


def test_capacity_overflow_drops_children(oracle):
    n, pt = 1000, 10
    pos, vel = _hot_state(n, seed=5)
    rows = np.tile(np.array([[30.0, 0.0]], dtype=np.float32), (10000, 1))     # 30 % split, no removal
    o = oracle.Oracle(capacity=1500, poisson_timestep=pt, table_rows=rows)
    o.upload(*pos, *vel)
    o.poisson_step()
    c = o.counters()
    assert c["n"] == 1500 and c["n_after_mobility"] == 1500 and c["added"] == 500 and c["removed"] == 0


def test_out_of_bounds_dies_without_draw(oracle):
    # an electron that leaves the domain is marked DEAD before any random draw (src/particle_move.cu:85)
    x = np.array([5.1199994, 2.5], dtype=np.float32)
    vx = np.array([1e9, 0.0], dtype=np.float32)
    o = oracle.Oracle(capacity=10, poisson_timestep=4)
    o.upload(x, np.full(2, 2.5, np.float32), np.full(2, 2.5, np.float32), vx)
    o.op_mobility()
    s = o.download()
    assert list(s["ts"]) == [-2, -1] or list(s["ts"])[0] == -2
    assert o.counters()["pushes"] == 1 + 4
    o.op_compact()
    assert o.n == 1
