"""the PNG snapshot of src/utility.cu:4-74 (host only, checked here by decoding the
file with zlib) and the binary state snapshot with exact restart (GPU)."""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def decode_png(path):
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(raw):
        (ln,), typ = struct.unpack(">I", raw[pos:pos + 4]), raw[pos + 4:pos + 8]
        data = raw[pos + 8:pos + 8 + ln]
        (crc,) = struct.unpack(">I", raw[pos + 8 + ln:pos + 12 + ln])
        assert crc == zlib.crc32(typ + data) & 0xffffffff, typ
        chunks.append((typ, data))
        pos += 12 + ln
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    w, h, depth, colour, comp, filt, lace = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, colour, comp, filt, lace) == (8, 2, 0, 0, 0)
    px = np.frombuffer(zlib.decompress(chunks[1][1]), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert np.all(px[:, 0] == 0)                      # filter type None on every row
    return px[:, 1:].reshape(h, w, 3)


def reference_image(x, y):
    """image()/draw_particle() of src/utility.cu:37-66 restated in numpy (pixels the reference would write
    outside its buffer -- row 500, column 500 -- left out)."""
    img = np.full((500, 500, 3), 255, np.uint8)
    for px, py in zip(np.asarray(x, np.float32), np.asarray(y, np.float32)):
        for i in range(-5, 6):
            for j in range(-5, 6):
                yy, xx = np.float32(py + np.float32(j)), np.float32(px + np.float32(i))
                if yy > 500 or yy < 0 or xx > 500 or xx < 0:
                    continue
                r, c = int(np.float32(500) - yy), int(xx)
                if 0 <= r < 500 and 0 <= c < 500:
                    img[r, c] = 0
    return img


def test_png_snapshot_matches_reference_raster(engine, tmp_path):
    rng = np.random.default_rng(3)
    # electrons of a real run sit in [2.25, 2.89]; also pixel-scale positions, the borders and outside
    x = np.concatenate([rng.uniform(2.25, 2.89, 50), [0.0, 499.5, 500.0, 250.2, 600.0, -3.0, 100.7]]).astype(np.float32)
    y = np.concatenate([rng.uniform(2.25, 2.89, 50), [0.0, 499.5, 500.0, 250.9, 10.0, 4.0, 300.3]]).astype(np.float32)
    path = tmp_path / "test_0000.png"
    engine.snapshot_png(x, y, path)
    got = decode_png(path)
    assert got.shape == (500, 500, 3)
    assert np.array_equal(got, reference_image(x, y))
    assert (got == 0).any() and (got == 255).any()
    engine.snapshot_png(np.zeros(0, np.float32), np.zeros(0, np.float32), path)       # empty population: a white canvas
    assert np.all(decode_png(path) == 255)
    with pytest.raises(engine.EngineError):
        engine.snapshot_png(x, y, tmp_path / "no_such_dir" / "a.png")


@pytest.mark.gpu
def test_state_snapshot_and_exact_restart(engine, tmp_path):
    """12 Poisson steps in one go == 5 steps, snapshot, a NEW sim restarted from the file, 7 more steps."""
    kw = dict(mode=0, init_n=200000, poisson_steps=12, capacity=2000000, poisson_timestep=10)
    with engine.Sim(**kw) as sim:
        sim.init_uniform(200000)
        sim.run_steps(12)
        whole, d_whole = sim.summary(), sim.download().sorted_by_id()
    snap = tmp_path / "state.bin"
    with engine.Sim(**kw) as sim:
        sim.init_uniform(200000)
        sim.run_steps(5)
        sim.save_state(snap)
        mid = sim.summary()
    hdr, arr = engine.read_state(snap)
    assert hdr["n"] == mid["final_n"] and hdr["poisson_index"] == 5 and hdr["pushes"] == mid["pushes"] and hdr["sorted_by_id"] == 1
    assert np.all(np.diff(arr["id"].astype(np.float64)) > 0) and np.all(arr["ts"] == -1)
    with engine.Sim(**kw) as sim:
        sim.load_state(snap)
        sim.run_steps(7)
        resumed, d_res = sim.summary(), sim.download().sorted_by_id()
    for k in ("final_n", "total_added", "total_removed", "pushes", "poisson_steps_done"):
        assert resumed[k] == whole[k], k
    assert whole["total_added"] > 1000
    assert np.array_equal(d_res.id, d_whole.id)
    for k in ("x", "y", "z", "vx", "vy", "vz"):
        assert np.array_equal(getattr(d_res, k), getattr(d_whole, k)), k
    with engine.Sim(**dict(kw, poisson_timestep=7)) as sim:          # a snapshot only continues ITS run
        with pytest.raises(engine.EngineError):
            sim.load_state(snap)


@pytest.mark.gpu
def test_cli_png_dump_and_restart(engine, tmp_path):
    """The `run` CLI: PNG + 'Saved image' line at every logged step when ./out/visualization exists
    (src/utility.cu:133, :34), binary dumps, and PICB200_RESTART continuing a run to the same summary."""
    exe = os.path.join(ROOT, "particle-simulation_b200", "pic_run")
    os.makedirs(tmp_path / "out" / "visualization")
    os.makedirs(tmp_path / "src")
    with open(os.path.join(ROOT, "src", "cross_section.txt"), "rb") as f:
        (tmp_path / "src" / "cross_section.txt").write_bytes(f.read())
    env = dict(os.environ, PICB200_DUMP_DIR=str(tmp_path), PICB200_DUMP_FINAL=str(tmp_path / "final.bin"))
    r = subprocess.run([exe, "30", "2", "300", "4", "256", "10000", "100", "10"], cwd=tmp_path, env=env, capture_output=True, text=True, check=True)
    lines = r.stdout.splitlines()
    for t in (0, 2, 4):
        assert f"Saved image to ./out/visualization/test_{t:04d}.png" in lines
        assert lines[lines.index(f"Saved image to ./out/visualization/test_{t:04d}.png") + 1] == ""      # then log()'s blank line
        assert decode_png(tmp_path / "out" / "visualization" / f"test_{t:04d}.png").shape == (500, 500, 3)
        assert engine.read_state(tmp_path / f"state_{t:04d}.bin")[0]["poisson_index"] == t
    full = [ln for ln in lines if ln.startswith(("Final amount", "Particles added", "Particles removed"))]
    # restart from the step-2 dump, no verbose output: same final summary and the same final state file
    env2 = dict(os.environ, PICB200_RESTART=str(tmp_path / "state_0002.bin"), PICB200_DUMP_FINAL=str(tmp_path / "final2.bin"))
    r2 = subprocess.run([exe, "30", "0", "300", "4", "256", "10000", "100", "10"], cwd=tmp_path, env=env2, capture_output=True, text=True, check=True)
    assert [ln for ln in r2.stdout.splitlines() if ln.startswith(("Final amount", "Particles added", "Particles removed"))] == full
    assert (tmp_path / "final.bin").read_bytes() == (tmp_path / "final2.bin").read_bytes()
