"""Worker of tests/test_dist_cpu.py: runs under torchrun with the gloo backend on CPU (world size 2) and
exercises the host-side multi-rank plumbing of bench.py -- rendezvous, slab partition, the ordered
all-gather of the 256-byte peer handles, max/sum reductions of the timings -- without any GPU call."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

dist, rank, world, local = bench.dist_setup(int(os.environ["WORLD_SIZE"]))
assert dist is not None and world == int(os.environ["WORLD_SIZE"]) and rank == int(os.environ["RANK"])
eng = importlib.import_module("particle-simulation_b200")

# slab partition used by the CLI harness and the tests: ids [r*n/k, (r+1)*n/k)
n = 1000003
lo, hi = n * rank // world, n * (rank + 1) // world
spans = [None] * world
dist.all_gather_object(spans, (lo, hi))
assert spans[0][0] == 0 and spans[-1][1] == n and all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))

# peer handles travel as opaque 256-byte blobs and must arrive in rank order
fake = bytes([rank]) * eng.PEER_HANDLE_BYTES
handles = [None] * world
dist.all_gather_object(handles, fake)
assert [h[0] for h in handles] == list(range(world)) and all(len(h) == eng.PEER_HANDLE_BYTES for h in handles)

# timing reductions: max over ranks for times, sum for work
assert bench.all_max(dist, float(rank + 1)) == float(world)
assert bench.all_sum(dist, float(rank + 1)) == world * (world + 1) / 2

# bench.peer_sim: the multicast set-up of the in-kernel exchange fails on ONE rank (here: a stand-in engine whose rank 1
# cannot connect while PICB200_NVLS is on) -> the ranks agree over gloo and ALL fall back to the unicast exchange
class FakeErr(RuntimeError):
    pass


class FakeSim:
    made = 0

    def __init__(self, **kw):
        FakeSim.made += 1
        self.nvls = os.environ.get("PICB200_NVLS", "1") != "0"
        self.closed = False

    def peer_export(self):
        return bytes([rank]) * eng.PEER_HANDLE_BYTES

    def peer_connect(self, hs):
        assert [h[0] for h in hs] == list(range(world))
        if self.nvls and rank == 1:
            raise FakeErr("CUDA error: multicast set-up failed\nAt: Peer setup")

    def close(self):
        self.closed = True


import types  # noqa: E402
os.environ.pop("PICB200_NVLS", None)
sim = bench.peer_sim(types.SimpleNamespace(Sim=FakeSim, EngineError=FakeErr), dist, world)
assert FakeSim.made == 2 and not sim.nvls and not sim.closed and os.environ["PICB200_NVLS"] == "0"
assert "fell back" in bench.PEER_TRANSPORT["note"] and "PICB200_NVLS=0" in bench.PEER_TRANSPORT["kind"]
os.environ.pop("PICB200_NVLS", None)

# no GPU here: a rank must fail loudly, never fall back to a CPU path
no_gpu = eng.device_count() == 0
if no_gpu:
    try:
        eng.Sim(capacity=100, rank=rank, nranks=world)
        raise SystemExit("engine created a multi-rank sim without a CUDA device")
    except eng.EngineError as e:
        assert "no CUDA device" in str(e)
dist.barrier()
if rank == 0:
    print(json.dumps({"ok": True, "world": world, "no_gpu": no_gpu}))
