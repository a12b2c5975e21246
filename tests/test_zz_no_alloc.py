"""picb200_charge_box_noalloc: the charge box through an entry point that allocates nothing (fixed device scratch array,
include/picb200.h).  No counterpart in the reference; added for callers that cannot
call cudaMalloc.  Collected last on purpose.

The counts are integers, so the comparison with the existing entry point (picb200_charge_box, which cudaMallocs its scratch
per call) is exact.  Two calls in a row share the scratch: the second asks for a DIFFERENT, smaller box of a different
population, so stale contents of the first call would show.
"""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_no_alloc_entry_point_contains_no_allocation_call():
    """Source check (no GPU needed): the body of picb200_charge_box_noalloc calls no cudaMalloc* / cudaFree* / allocating
    container, takes the scratch's device address with cudaGetSymbolAddress and checks its return code."""
    src = open(os.path.join(ROOT, "particle-simulation_b200", "csrc", "pic_engine.cu")).read()
    a = src.index("int picb200_charge_box_noalloc(")
    body = src[a:src.index("\nint picb200_charge_bbox(", a)]
    assert not re.search(r"cudaMalloc|cudaFree|cudaHostAlloc|cudaMallocHost|std::vector|new |malloc\(", body)
    assert "CK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_box_scratch)" in body
    assert re.search(r"static __device__ __align__\(256\) int g_noalloc_box_scratch\[PICB200_NOALLOC_MAX_CELLS\];", src)
    hdr = open(os.path.join(ROOT, "include", "picb200.h")).read()
    assert "#define PICB200_NOALLOC_MAX_CELLS (1 << 20)" in hdr and "int picb200_charge_box_noalloc(" in hdr



def _deposit(sim, n, seed):
    rng = np.random.default_rng(seed)
    pos = rng.uniform(2.40, 2.70, size=(3, n)).astype(np.float32)          # 30^3 cells, several electrons per cell
    sim.upload(*pos)
    sim.op_deposit()
    return sim.charge_bbox()


@pytest.mark.gpu
def test_no_alloc_charge_box_equals_the_allocating_entry_point(engine):
    with engine.Sim(capacity=200000, poisson_timestep=10) as sim:
        # ---- call 1: the whole occupied box
        lo, hi = _deposit(sim, 150000, seed=1)
        ref1 = sim.charge_box(lo, hi)
        got1 = sim.charge_box_noalloc(lo, hi)
        assert got1.shape == ref1.shape and got1.dtype == np.int32
        assert np.array_equal(got1, ref1)
        assert int(got1.sum()) == 150000 and got1.max() >= 3
        # ---- call 2, right after it: another population, a smaller box inside the occupied one
        lo2, hi2 = _deposit(sim, 90000, seed=2)
        lo2, hi2 = lo2 + np.array([3, 5, 7], dtype=np.int32), hi2 - np.array([2, 4, 6], dtype=np.int32)
        ref2 = sim.charge_box(lo2, hi2)
        got2 = sim.charge_box_noalloc(lo2, hi2)
        assert got2.shape == ref2.shape and got2.size < got1.size
        assert np.array_equal(got2, ref2)
        assert 0 < int(got2.sum()) < 90000
        # ---- and the first box again: same answer as the allocating entry point now gives
        assert np.array_equal(sim.charge_box_noalloc(lo, hi), sim.charge_box(lo, hi))


@pytest.mark.gpu
def test_no_alloc_charge_box_rejects_boxes_beyond_its_scratch(engine):
    with engine.Sim(capacity=1000, poisson_timestep=10) as sim:
        sim.init_uniform(100)
        sim.op_deposit()
        lo = np.array([0, 0, 0], dtype=np.int32)
        hi = np.array([127, 127, 63], dtype=np.int32)                   # 128 * 128 * 64 = 2^20 cells: the maximum, accepted
        assert 128 * 128 * 64 == engine.NOALLOC_MAX_CELLS
        assert sim.charge_box_noalloc(lo, hi).shape == (64, 128, 128)
        hi[2] = 64                                                       # one slab more: rejected, clearly
        with pytest.raises(engine.EngineError) as ei:
            sim.charge_box_noalloc(lo, hi)
        assert "exceeds the fixed scratch" in str(ei.value) and "PICB200_NOALLOC_MAX_CELLS" in str(ei.value)
        with pytest.raises(engine.EngineError):
            sim.charge_box_noalloc(np.array([5, 5, 5], dtype=np.int32), np.array([4, 9, 9], dtype=np.int32))      # empty box
        # the engine is still usable after a rejected call
        lo, hi = sim.charge_bbox()
        assert int(sim.charge_box_noalloc(lo, hi).sum()) == 100
