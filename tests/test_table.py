"""The collision table is input data: src/cross_section.txt is the reference's file byte for byte, and its fp32 image
is pinned as well (data/make_cross_section.py regenerates the same fp32 values but 16 of the 10 000 rows differ in the
last printed digit with this libm -- only the fp32 equality is guaranteed there, so the file is shipped, not regenerated)."""
import hashlib
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHA = "f1b1701bbc86b75d041bceb89e4d56902f0de7b4bd189f9da163051c7b8fa6c1"


def test_shipped_table_fp32_image():
    rows = np.loadtxt(os.path.join(ROOT, "src", "cross_section.txt"), dtype=np.float64).astype(np.float32)
    assert rows.shape == (10000, 2)
    assert hashlib.sha256(rows.tobytes()).hexdigest() == SHA
    assert np.array_equal(rows[:, 0], rows[:, 1])        # split == remove in every row: a critical branching process


# sha256 of the file itself: src/cross_section.txt is the reference's input file, byte for byte (it is input data, not source)
FILE_SHA = "dee6b0a82247ce6a57187bc677fa2eb1407c8213bd22dd68acca96f5d40fea49"


def test_shipped_table_is_the_reference_file_byte_for_byte():
    mine = open(os.path.join(ROOT, "src", "cross_section.txt"), "rb").read()
    assert hashlib.sha256(mine).hexdigest() == FILE_SHA
    ref = "/root/reference/src/cross_section.txt"
    if os.path.exists(ref):              # (absent on the GPU box: the pinned hash above stands in)
        assert open(ref, "rb").read() == mine


def _draws():
    """Every value float(w) can take for a 32-bit word w, ascending, and the collider's draw r for each:
    r = fmul(fma(float(w), 2^-32, 2^-33), 100) (src/random.cu:7-12, src/particle_move.cu:61).  The product
    float(w) * 2^-32 is exact, so a float32 multiply followed by a float32 add IS the fused operation."""
    small = np.arange(0, 1 << 24, dtype=np.float32)                                  # every integer below 2^24
    big = [np.arange(1 << 23, 1 << 24, dtype=np.float64) * float(1 << (e - 23)) for e in range(24, 32)]
    f = np.concatenate([small] + [b.astype(np.float32) for b in big] + [np.array([4294967296.0], dtype=np.float32)])
    r = (f * np.float32(2.0 ** -32) + np.float32(2.0 ** -33)) * np.float32(100.0)
    assert r.dtype == np.float32
    return f, r


def test_draw_thresholds_equal_the_literal_collider_test():
    """The engine's tables hold the chances as thresholds on float(w) (draw_threshold, pic_engine.cu), which saves
    the kernels an FMA and a multiply per push.  Checked here against the literal `r < chance` for EVERY value
    float(w) can take (8.4e7), for every distinct chance of the shipped table, the sums split + remove, and the
    edge cases (never, always, NaN, the ends of the draw's range)."""
    import importlib
    import sys
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    eng = importlib.import_module("particle-simulation_b200")
    f, r = _draws()
    assert np.all(np.diff(r) >= 0)                       # the draw never decreases as float(w) grows
    assert r[0] > 0 and r[-1] == np.float32(100.0)       # (0, 100], like curand_uniform * 100
    rows = np.loadtxt(os.path.join(ROOT, "src", "cross_section.txt"), dtype=np.float64).astype(np.float32)
    chances = np.unique(np.concatenate([rows[:, 0], rows[:, 0] + rows[:, 1]]))
    rng = np.random.default_rng(5)
    pick = np.concatenate([chances[:: max(1, len(chances) // 40)], chances[-3:], rng.uniform(0, 100, 8).astype(np.float32),
                           np.array([0.0, -1.0, 1e-9, r[0], np.nextafter(r[0], np.float32(1)), 50.0, 99.99999, 100.0,
                                     np.nextafter(np.float32(100), np.float32(200)), 1e9, np.inf, np.nan], dtype=np.float32)])
    for c in pick:
        t = eng.host_draw_threshold(c)
        assert np.array_equal(f < t, r < c), (c, t)
    # and the cheap boundary form of the same statement for every distinct chance of the table
    for c in chances:
        t = eng.host_draw_threshold(c)
        k = int(np.searchsorted(r, c, side="left"))      # first draw that is not below the chance
        assert (k == len(f) and np.isinf(t)) or (k < len(f) and (f[k] >= t) and (k == 0 or f[k - 1] < t)), (c, t, k)
