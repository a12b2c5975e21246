"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol the
header declares, and fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "picb200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(picb200_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported(engine):
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(engine.lib, n), f"{n} declared in include/picb200.h but not exported"
    assert set(names) == set(engine.SIGNATURES), "binding and header drifted apart"


def test_abi_version_and_struct_layout(engine):
    assert engine.lib.picb200_abi_version() == engine.ABI_VERSION
    cfg = engine.Config()
    engine.lib.picb200_config_default(C.byref(cfg))
    assert (cfg.abi_version, cfg.block_size, cfg.poisson_timestep, cfg.nranks, cfg.device) == (1, 256, 10, 1, -1)
    assert C.sizeof(engine.Config) == 88 and C.sizeof(engine.Summary) == 64
    assert engine.ELECTRON_DTYPE.itemsize == 64
    assert engine.ELECTRON_DTYPE.fields["velocity"][1] == 24 and engine.ELECTRON_DTYPE.fields["timestamp"][1] == 60


def test_library_has_no_oracle_dependency(engine):
    out = subprocess.run(["ldd", engine.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out
    syms = subprocess.run(["nm", "-D", "--defined-only", engine.LIB_PATH], capture_output=True, text=True).stdout
    assert "orc_" not in syms


def test_loader_matches_oracle_loader(engine, oracle):
    r, rows = engine.load_table()
    r2, rows2 = oracle.load_table()
    assert r == r2 == 10000 and np.array_equal(rows, rows2)
    assert engine.load_table("/nonexistent/cross_section.txt")[0] == -1


def test_edge_table_is_the_reference_formula(engine, oracle):
    """edge[j] is the smallest fp32 energy whose reference index (double formula) is >= j."""
    e = engine.host_energy_edges()
    assert e[0] == 0 and np.isinf(e[-1]) and np.all(np.diff(e[:-1]) > 0)
    for j in list(range(1, 40)) + [500, 2727, 5000, 9998, 9999]:
        assert oracle.energy_to_index(float(e[j])) == j
        assert oracle.energy_to_index(float(np.nextafter(e[j], np.float32(0)))) == j - 1


def test_device_energy_index_formula_emulated_on_cpu(engine, tmp_path):
    """The exact bin-edge table the device probes reproduces the reference formula on a strided sweep of
    all non-negative fp32 values plus +-3 ulp around every bin edge (the exhaustive sweep, stride 1, passes
    too: 2 147 553 641 values, 0 mismatches); the device's SFU estimate is swept on the device
    (tests/test_parity_gpu.py::test_energy_index_exhaustive_on_device)."""
    exe = tmp_path / "cei"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([cxx, "-O2", "-march=x86-64-v3", "-ffp-contract=off", "-fopenmp", "-o", str(exe),
                    os.path.join(ROOT, "tests", "tools", "check_energy_index.cpp")], check=True)
    r = subprocess.run([str(exe), "257"], input=engine.host_energy_edges().tobytes(), capture_output=True)
    assert r.returncode == 0, r.stderr.decode()[:2000]
    assert b"mismatches 0" in r.stdout


@pytest.mark.parametrize("variant", [[], ["-DONE_PROBE"]])
def test_device_cell_index_formula_emulated_on_cpu(engine, tmp_path, variant):
    """cell_of_fast() (one fp32 multiply + a probe of exact cell boundaries) equals the reference's
    int(double(pos)/1e-2) (src/grid_operations.cu:21-23) for EVERY fp32 with |x| <= 8 (2.2e9 values);
    -DONE_PROBE: the single-probe form the kernels use (round 1's two-probe form stays behind -DPIC_CELL_TWO_PROBES)."""
    exe = tmp_path / "cci"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([cxx, "-O2", "-march=x86-64-v3", "-ffp-contract=off", "-fopenmp", *variant, "-o", str(exe),
                    os.path.join(ROOT, "tests", "tools", "check_cell_index.cpp")], check=True)
    r = subprocess.run([str(exe)], input=engine.host_cell_bounds().tobytes(), capture_output=True)
    assert r.returncode == 0, r.stderr.decode()[:2000]
    assert b"mismatches 0" in r.stdout


@pytest.mark.skipif(os.environ.get("PICB200_EXPECT_GPU") == "1", reason="GPU present")
def test_no_cpu_fallback(engine):
    if engine.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(engine.EngineError, match="no CUDA device"):
        engine.Sim(capacity=100)
    with pytest.raises(engine.EngineError):
        engine.run(init_n=10, capacity=100)


def test_cli_usage_without_gpu(engine):
    r = subprocess.run([engine.CLI_PATH], capture_output=True, text=True)
    assert r.returncode == 64 and "usage" in r.stderr
    r = subprocess.run([engine.CLI_PATH, "32", "0", "x", "1", "256", "100", "100", "10"], capture_output=True, text=True)
    assert r.returncode == 64
    # unknown mode string: runs nothing, still prints the CPU time (src/main.cu:26-46)
    r = subprocess.run([engine.CLI_PATH, "99", "0", "10", "1", "256", "100", "100", "10"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("CPU time of program: ")


def test_sass_budget_tool_follows_the_source(engine):
    """tools/sass_budget.py finds the step loop of the fused kernel by marker text in csrc/pic_mover.cuh; an edit of the
    loop that moves a marker must show up here, not as a silently dead tool (no GPU needed: SASS of the built library)."""
    import shutil
    import sys
    if not (shutil.which("cuobjdump") and shutil.which("nvdisasm")):
        pytest.skip("CUDA binary utilities not on PATH")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_budget.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-400:] + r.stderr[-400:]
    assert "step loop at items" in r.stdout and "philox" in r.stdout
    m = re.search(r"(\d+) outside the spawn path", r.stdout)
    assert m and 100 <= int(m.group(1)) <= 260, r.stdout[:600]      # (the hot part of a mobility step: ~150 instructions)
