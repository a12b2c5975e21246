"""The reference's two harness modes and its own harness binary on top of the new engine.

* `pic_run test ...`  = runUnitTest (src/test.cu:43-100): base scheduler 1, then every scheduler, sort by the key of
  src/electron.h:28-41, exact compare, `success (n)` / `failure (n)` lines.
* `pic_run bench <csv>` = runBenchmark + printCSV (src/test.cu:4-41, src/utility.cu:87-106): the CSV that
  analyse/plot_pic_*.py read.  PICB200_BENCH_QUICK=1 shrinks the sweep to the cells analyse/plot_pic_mobility.py selects.
* `oracle/_ref_shim/main` = the reference's OWN src/{main,test,utility,cell}.cu, unchanged, linked with
  csrc/runpic_shim.cpp + libpicb200.so (`make -C oracle shim`, built where the reference checkout is present; the
  binary travels to the GPU box): its `test` mode is the reference's only correctness check, now driving this engine.
"""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PIC_RUN = os.path.join(ROOT, "particle-simulation_b200", "pic_run")
SHIM = os.path.join(ROOT, "oracle", "_ref_shim", "main")
# src/utility.cu:87-89 (printTimingHeader), byte for byte
CSV_HEADER = "func,init n,iterations,mobility steps,block size,sleep time,split chance,final n,time\n"
TIMING = ("GPU time of program", "CPU time of program")


def _stable(txt):
    return [ln.split(":")[0] if ln.startswith(TIMING) else ln for ln in txt.splitlines()]


def _verdicts(stdout):
    tail = stdout[stdout.index("Tests done with following results as compared to 1"):]
    return re.findall(r"^(\d): (success|failure) \((\d+)\)$", tail, flags=re.M)


def test_csv_header_is_the_reference_header():
    """The header literal of csrc/pic_harness.cpp equals src/utility.cu:88 (checked against the checkout when present)."""
    src = open(os.path.join(ROOT, "particle-simulation_b200", "csrc", "pic_harness.cpp")).read()
    assert '"' + CSV_HEADER.replace("\n", "\\n") + '"' in src
    ref = "/root/reference/src/utility.cu"
    if os.path.exists(ref):
        assert '"' + CSV_HEADER.replace("\n", "\\n") + '"' in open(ref).read()


def test_shim_binary_is_built_from_the_reference_harness_when_the_checkout_is_present():
    if not os.path.isdir("/root/reference/src"):
        pytest.skip("reference checkout absent (GPU box): the prebuilt binary is used")
    assert os.path.exists(SHIM), "run __graft_entry__.build() (make -C oracle shim)"
    syms = subprocess.run(["nm", "-C", SHIM], capture_output=True, text=True).stdout
    assert re.search(r" T runPIC\(int, int, int, int, int, int, int, int\)", syms)          # the shim defines the operator
    assert re.search(r" T runUnitTest\(", syms) and re.search(r" T runBenchmark\(", syms)     # the reference's harness is in
    assert " U picb200_run_keep" in syms and "updateParticle" not in syms                       # the engine comes from the library


@pytest.mark.gpu
def test_cli_test_mode_all_schedulers_agree(tmp_path):
    """`run test 0 20000 5 256 400000 100 10`: four `success` lines, rc 0.  With collisions (hot table) as well."""
    r = subprocess.run([PIC_RUN, "test", "0", "20000", "5", "256", "400000", "100", "10"], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    v = _verdicts(r.stdout)
    assert [a for a, _, _ in v] == ["0", "1", "3", "2"] and all(b == "success" for _, b, _ in v), v
    assert all(int(c) == 20000 for _, _, c in v)
    assert "Sorting base...\nDone sorting base\n" in r.stdout and "Mismatch" not in r.stdout
    # a run in which electrons are born and die: every row 2.0 % ionisation, 1.5 % attachment
    table = tmp_path / "cs.txt"
    table.write_text("2.0 1.5\n" * 10000)
    r = subprocess.run([PIC_RUN, "test", "0", "20000", "4", "256", "400000", "100", "10"], capture_output=True, text=True, cwd=ROOT, timeout=600,
                       env=dict(os.environ, PICB200_TABLE=str(table)))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    v = _verdicts(r.stdout)
    assert len(v) == 4 and all(b == "success" for _, b, _ in v) and len({c for _, _, c in v}) == 1 and int(v[0][2]) != 20000, v
    added = [int(x) for x in re.findall(r"Particles added: (\d+)", r.stdout)]
    assert len(added) == 5 and min(added) > 1000 and len(set(added)) == 1


@pytest.mark.gpu
def test_cli_bench_mode_writes_the_csv_the_plot_scripts_read(tmp_path):
    """`run bench <csv>` (quick sweep): header byte-equal to src/utility.cu:88, and the pandas filters of
    analyse/plot_pic_mobility.py:8-14 select non-empty frames for every scheduler."""
    pd = pytest.importorskip("pandas")
    csv = tmp_path / "mobility_timesteps.csv"
    r = subprocess.run([PIC_RUN, "bench", str(csv)], capture_output=True, text=True, cwd=ROOT, timeout=900,
                       env=dict(os.environ, PICB200_BENCH_QUICK="1", PICB200_BENCH_INIT_N="200000"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    raw = csv.read_text()
    assert raw.startswith(CSV_HEADER)
    df = pd.read_csv(csv)
    assert list(df.columns) == CSV_HEADER.strip().split(",")
    assert len(df) == 2 * (4 + 3)                                     # 2 mobility-step values x (4 schedulers at 256 + 3 at 1024)
    # analyse/plot_pic_mobility.py:8-14
    crit = ((df["func"] == "Naive") & (df["block size"] == 256)) | ((df["func"] == "Dynamic") & (df["block size"] == 1024)) | \
           ((df["func"] == "CPU Sync") & (df["block size"] == 1024)) | ((df["func"] == "Dynamic Old") & (df["block size"] == 1024))
    sel = df[crit]
    sel = sel[sel["mobility steps"] <= 100]
    assert set(sel["func"]) == {"Naive", "Dynamic", "CPU Sync", "Dynamic Old"} and len(sel) == 8
    assert (sel["time"] > 0).all() and (sel["final n"] > 0).all() and (sel["init n"] == 200000).all() and (sel["iterations"] == 10).all()
    # the same sweep point gives the same population under every scheduler (particle-id keyed RNG)
    for ms, g in df.groupby("mobility steps"):
        assert g["final n"].nunique() == 1, (ms, g[["func", "block size", "final n"]])


@pytest.mark.gpu
def test_reference_harness_runs_on_the_engine(tmp_path):
    """The reference's own `main` (src/main.cu, src/test.cu, src/utility.cu unchanged) linked against libpicb200.so:
    its unit test passes on this engine, and its `30` run prints what pic_run prints."""
    if not os.path.exists(SHIM):
        pytest.skip("oracle/_ref_shim/main not built (needs the reference checkout at build time)")
    r = subprocess.run([SHIM, "test", "0", "20000", "5", "256", "400000", "100", "10"], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    v = _verdicts(r.stdout)
    assert v == [("0", "success", "20000"), ("1", "success", "20000"), ("3", "success", "20000")], v      # src/test.cu:47: {0, 1, 3}
    table = tmp_path / "cs.txt"
    table.write_text("2.0 1.5\n" * 10000)
    r = subprocess.run([SHIM, "test", "0", "20000", "4", "256", "400000", "100", "10"], capture_output=True, text=True, cwd=ROOT, timeout=600,
                       env=dict(os.environ, PICB200_TABLE=str(table)))
    v = _verdicts(r.stdout)
    assert r.returncode == 0 and len(v) == 3 and all(b == "success" for _, b, _ in v) and int(v[0][2]) != 20000, r.stdout[-1500:]
    args = ["30", "1", "200", "3", "256", "1000", "100", "5"]
    a = subprocess.run([SHIM] + args, capture_output=True, text=True, cwd=ROOT, timeout=300)
    b = subprocess.run([PIC_RUN] + args, capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert a.returncode == 0 and b.returncode == 0, a.stdout[-1000:] + a.stderr[-1000:]
    assert _stable(a.stdout) == _stable(b.stdout)
    assert sum(1 for ln in a.stdout.splitlines() if ln.startswith("Time ")) == 4


@pytest.mark.gpu
def test_grid_barrier_times_out_instead_of_spinning_for_ever():
    """A block that never reaches a grid barrier (PICB200_TEST_DESERT=1: the last block leaves the persistent kernel at
    once) must surface as the reference's two error lines (src/utility.cu:108-115) and a non-zero exit code, not as a hang."""
    env = dict(os.environ, PICB200_TEST_DESERT="1", PICB200_BARRIER_TIMEOUT_MS="300")
    r = subprocess.run([PIC_RUN, "30", "0", "5000", "2", "256", "100000", "100", "5"], capture_output=True, text=True, cwd=ROOT, timeout=120, env=env)
    assert r.returncode != 0
    assert "CUDA error: grid barrier" in r.stdout and "timed out" in r.stdout and "\nAt: Mobility steps\n" in r.stdout, r.stdout[-1500:]
    assert "Final amount of particles" not in r.stdout
