"""world_size-2 gloo run of the host-side multi-rank plumbing on CPU (no GPU needed)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_gloo_plumbing():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "tools", "dist_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    assert json.loads(line) == {"ok": True, "world": 2, "no_gpu": True}


def test_reference_arm_only_rank0_prints():
    """bench.py --impl reference under torchrun: rank 0 alone runs and prints, the others exit 0 silently."""
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1", MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_setup_barrier_between_processes(tmp_path):
    """csrc/pic_nvls.h without a GPU: the shared-memory barrier of picb200_peer_connect between forked processes (eight
    ranks meet twice; a rank that reports a failure makes the others give up within a second; an absent rank costs the
    time-out), and a file descriptor duplicated out of another process (pidfd_getfd), as the multicast object's is."""
    exe = tmp_path / "csb"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    inc = "/usr/local/cuda/include"
    subprocess.run([cxx, "-O1", "-std=c++17", "-I", inc, "-o", str(exe), os.path.join(ROOT, "tests", "tools", "check_shm_barrier.cpp"), "-lpthread"], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.startswith("ok"), r.stdout + r.stderr
