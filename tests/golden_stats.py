"""Statistics of an electron population used by the statistical fixture tests/golden/ref_cfg2_stats.json
(generated from the reference's own CUDA program by tools/make_golden_from_ref.py) and by the tests that
compare the engine with it.  One definition, so both sides bin identically."""
import numpy as np

ENERGY_EDGES = np.linspace(-12.0, 0.0, 49)         # log10(|v|^2) bins
OCC_MAX = 40


def state_stats(pos, vel, ts=None):
    """pos, vel: (n, 3) arrays.  Returns the dict stored per run in ref_cfg2_stats.json."""
    pos = np.asarray(pos, dtype=np.float64)
    vel = np.asarray(vel, dtype=np.float64)
    e = (vel ** 2).sum(axis=1)
    with np.errstate(divide="ignore"):
        le = np.where(e > 0, np.log10(e), -np.inf)
    eh, _ = np.histogram(le, bins=ENERGY_EDGES)
    cells = (pos / 1e-2).astype(np.int64)          # src/grid_operations.cu:21-23
    lin = (cells[:, 2] * 512 + cells[:, 1]) * 512 + cells[:, 0]
    _, counts = np.unique(lin, return_counts=True)
    occ = np.bincount(np.minimum(counts, OCC_MAX), minlength=OCC_MAX + 1)
    return {
        "n": int(pos.shape[0]),
        "energy_log10_edges": ENERGY_EDGES.tolist(), "energy_hist": eh.tolist(), "energy_below": int((le < ENERGY_EDGES[0]).sum()),
        "energy_above": int((le >= ENERGY_EDGES[-1]).sum()),
        "occupancy_hist": occ.tolist(), "occupied_cells": int(counts.size),
        "v_mean": vel.mean(axis=0).tolist(), "v_rms": np.sqrt((vel ** 2).mean(axis=0)).tolist(),
        "mean_energy": float(e.mean()), "median_energy": float(np.median(e)),
        "pos_min": pos.min(axis=0).tolist(), "pos_max": pos.max(axis=0).tolist(),
        "ts_all_minus1": bool(ts is None or np.all(np.asarray(ts) == -1)),
    }


def tvd(a, b, merge=1):
    """Total-variation distance of two histograms after merging `merge` adjacent bins."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if merge > 1:
        k = (len(a) // merge) * merge
        a, b = a[:k].reshape(-1, merge).sum(1), b[:k].reshape(-1, merge).sum(1)
    return 0.5 * float(np.abs(a / a.sum() - b / b.sum()).sum())
