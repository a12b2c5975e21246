"""ctypes binding of libpicb200.so -- the B200-native PIC-MCC electron engine.

The product is the CUDA library; this module only mirrors ``include/picb200.h`` so that the
parity tests and ``bench.py`` can drive it the way the reference's ``main``/``test`` drive
``runPIC`` (reference: src/pic.h:9, src/main.cu:26-40, src/test.cu:43-100).  There is no
Python or CPU fallback: if the shared library is missing, importing this module raises.

The package directory name contains a hyphen; import it with
``importlib.import_module("particle-simulation_b200")``.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpicb200.so")
CLI_PATH = os.path.join(_HERE, "pic_run")
REPO_ROOT = os.path.dirname(_HERE)
DEFAULT_TABLE = os.path.join(REPO_ROOT, "src", "cross_section.txt")

MODE_DYNAMIC, MODE_CPU_SYNC, MODE_NAIVE, MODE_DYNAMIC_OLD = 0, 1, 2, 3
DEAD = -2
TABLE_ROWS = 10000
GRID = 512
ABI_VERSION = 1
PEER_HANDLE_BYTES = 256
NOALLOC_MAX_CELLS = 1 << 20

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()). "
        "There is no fallback implementation."
    )
lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)


class Config(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("mode", C.c_int32), ("verbose", C.c_int32), ("init_n", C.c_int32),
        ("poisson_steps", C.c_int32), ("block_size", C.c_int32), ("capacity", C.c_int32),
        ("sleep_time_ns", C.c_int32), ("poisson_timestep", C.c_int32), ("device", C.c_int32),
        ("seed", C.c_uint64), ("table_path", C.c_char_p), ("table_rows", C.POINTER(C.c_float)),
        ("rank", C.c_int32), ("nranks", C.c_int32), ("nccl_unique_id", C.c_void_p),
        ("print_stdout", C.c_int32), ("reserved", C.c_int32),
    ]


class Summary(C.Structure):
    _fields_ = [
        ("final_n", C.c_int64), ("total_added", C.c_int64), ("total_removed", C.c_int64), ("pushes", C.c_int64),
        ("poisson_steps_done", C.c_int64), ("gpu_ms", C.c_double), ("overflowed", C.c_int32), ("hit_zero", C.c_int32),
        ("sm_count", C.c_int32), ("blocks_per_sm", C.c_int32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Profile(C.Structure):
    _fields_ = [("deposit_ms", C.c_double), ("gather_ms", C.c_double), ("clear_ms", C.c_double), ("mover_ms", C.c_double),
                ("compact_ms", C.c_double), ("collective_ms", C.c_double), ("launches", C.c_int64), ("mover_launches", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


ELECTRON_DTYPE = np.dtype(
    [("position", "<f8", 3), ("velocity", "<f8", 3), ("acceleration", "<f4", 3), ("timestamp", "<i4")], align=True
)
assert ELECTRON_DTYPE.itemsize == 64  # src/electron.h:12-18

_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int32)
_up = C.POINTER(C.c_uint64)
_sim = C.c_void_p

# Every symbol include/picb200.h declares, with its signature.  tests/test_abi.py checks this
# list against the header so the two cannot drift apart.
SIGNATURES = {
    "picb200_abi_version": (C.c_int, []),
    "picb200_last_error": (C.c_char_p, []),
    "picb200_device_count": (C.c_int, []),
    "picb200_config_default": (None, [C.POINTER(Config)]),
    "picb200_run": (C.c_int, [C.POINTER(Config), C.POINTER(Summary), C.c_void_p, C.c_int64]),
    "picb200_run_keep": (C.c_int, [C.POINTER(Config), C.POINTER(Summary), C.POINTER(_sim)]),
    "picb200_create": (C.c_int, [C.POINTER(Config), C.POINTER(_sim)]),
    "picb200_destroy": (None, [_sim]),
    "picb200_init_uniform": (C.c_int, [_sim, C.c_int32, C.c_uint64]),
    "picb200_upload": (C.c_int, [_sim, C.c_int32] + [_fp] * 9 + [_ip, _up]),
    "picb200_download": (C.c_int, [_sim, _ip] + [_fp] * 9 + [_ip, _up]),
    "picb200_download_aos": (C.c_int, [_sim, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "picb200_poisson_step": (C.c_int, [_sim]),
    "picb200_run_steps": (C.c_int, [_sim, C.c_int32]),
    "picb200_get_summary": (C.c_int, [_sim, C.POINTER(Summary)]),
    "picb200_set_poisson_index": (C.c_int, [_sim, C.c_int32]),
    "picb200_set_profiling": (C.c_int, [_sim, C.c_int32]),
    "picb200_get_profile": (C.c_int, [_sim, C.POINTER(Profile)]),
    "picb200_flush_l2": (C.c_int, [_sim]),
    "picb200_alloc_pinned": (C.c_void_p, [C.c_size_t]),
    "picb200_free_pinned": (None, [C.c_void_p]),
    "picb200_op_deposit": (C.c_int, [_sim]),
    "picb200_op_gather": (C.c_int, [_sim]),
    "picb200_op_clear_grid": (C.c_int, [_sim]),
    "picb200_op_mobility": (C.c_int, [_sim]),
    "picb200_op_compact": (C.c_int, [_sim]),
    "picb200_charge_box": (C.c_int, [_sim, _ip, _ip, _ip]),
    "picb200_charge_box_noalloc": (C.c_int, [_sim, _ip, _ip, _ip]),
    "picb200_charge_bbox": (C.c_int, [_sim, _ip, _ip]),
    "picb200_dev_philox": (C.c_int, [C.c_int32, C.c_int64, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "picb200_dev_energy_index": (C.c_int, [C.c_int32, C.c_int64, _fp, _ip]),
    "picb200_dev_energy_index_sweep": (C.c_int, [C.c_int32, C.c_uint32, C.POINTER(C.c_int64), C.POINTER(C.c_uint32)]),
    "picb200_load_table": (C.c_int, [C.c_char_p, _fp]),
    "picb200_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "picb200_peer_export": (C.c_int, [_sim, C.c_void_p]),
    "picb200_peer_connect": (C.c_int, [_sim, C.c_void_p]),
    "picb200_snapshot_png": (C.c_int, [_fp, _fp, C.c_int64, C.c_char_p]),
    "picb200_save_state": (C.c_int, [_sim, C.c_char_p, C.c_int32]),
    "picb200_load_state": (C.c_int, [_sim, C.c_char_p]),
}
for _name, (_res, _args) in SIGNATURES.items():
    _f = getattr(lib, _name)  # AttributeError here = the library does not export what the header declares
    _f.restype = _res
    _f.argtypes = _args
# not part of the public header: host copy of the bin-edge table, for the CPU-side exhaustive check
lib.picb200_host_energy_edges.restype = None
lib.picb200_host_energy_edges.argtypes = [_fp]
lib.picb200_host_cell_bounds.restype = None
lib.picb200_host_cell_bounds.argtypes = [_fp]
lib.picb200_host_draw_threshold.restype = C.c_float
lib.picb200_host_draw_threshold.argtypes = [C.c_float]


class EngineError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise EngineError(lib.picb200_last_error().decode(errors="replace"))


def device_count() -> int:
    return int(lib.picb200_device_count())


def _ptr(a, ty):
    return None if a is None else a.ctypes.data_as(ty)


def make_config(mode=MODE_DYNAMIC, verbose=0, init_n=0, poisson_steps=1, block_size=256, capacity=1000000,
                sleep_time_ns=100, poisson_timestep=10, device=-1, seed=0, table_path=None, table_rows=None,
                rank=0, nranks=1, nccl_unique_id=None, print_stdout=0):
    """Build a Config; returns (cfg, keepalive) -- keepalive holds the buffers cfg points into."""
    cfg = Config()
    lib.picb200_config_default(C.byref(cfg))
    keep = []
    cfg.mode, cfg.verbose, cfg.init_n, cfg.poisson_steps = mode, verbose, init_n, poisson_steps
    cfg.block_size, cfg.capacity, cfg.sleep_time_ns, cfg.poisson_timestep = block_size, capacity, sleep_time_ns, poisson_timestep
    cfg.device, cfg.seed, cfg.rank, cfg.nranks, cfg.print_stdout = device, seed, rank, nranks, print_stdout
    if table_rows is not None:
        t = np.ascontiguousarray(table_rows, dtype=np.float32).reshape(TABLE_ROWS, 2)
        keep.append(t)
        cfg.table_rows = t.ctypes.data_as(_fp)
    else:
        p = (table_path or DEFAULT_TABLE).encode()
        keep.append(p)
        cfg.table_path = p
    if nccl_unique_id is not None:
        b = C.create_string_buffer(bytes(nccl_unique_id), 128)
        keep.append(b)
        cfg.nccl_unique_id = C.cast(b, C.c_void_p)
    return cfg, keep


@dataclass
class State:
    """Host copy of the electron store (SoA, fp32 state; src/electron.h:12-18 + a stable id)."""
    x: np.ndarray
    y: np.ndarray
    z: np.ndarray
    vx: np.ndarray
    vy: np.ndarray
    vz: np.ndarray
    ax: np.ndarray
    ay: np.ndarray
    az: np.ndarray
    ts: np.ndarray
    id: np.ndarray

    @property
    def n(self):
        return int(self.x.shape[0])

    def sorted_by_id(self):
        o = np.argsort(self.id, kind="stable")
        return State(*[getattr(self, f)[o] for f in self.__dataclass_fields__])


class Sim:
    """One engine instance = what one call of runPIC owns (src/pic.cu:380-454, :586-596)."""

    def __init__(self, **kw):
        self.cfg, self._keep = make_config(**kw)
        self._h = _sim()
        _check(lib.picb200_create(C.byref(self.cfg), C.byref(self._h)))

    def close(self):
        if self._h:
            lib.picb200_destroy(self._h)
            self._h = _sim()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def init_uniform(self, init_n, id_base=0):
        _check(lib.picb200_init_uniform(self._h, init_n, id_base))

    def upload(self, x, y, z, vx=None, vy=None, vz=None, ax=None, ay=None, az=None, ts=None, id=None):
        n = len(x)
        f = [None if a is None else np.ascontiguousarray(a, dtype=np.float32) for a in (x, y, z, vx, vy, vz, ax, ay, az)]
        ts = None if ts is None else np.ascontiguousarray(ts, dtype=np.int32)
        id = None if id is None else np.ascontiguousarray(id, dtype=np.uint64)
        _check(lib.picb200_upload(self._h, n, *[_ptr(a, _fp) for a in f], _ptr(ts, _ip), _ptr(id, _up)))

    def download(self) -> State:
        cap = self.cfg.capacity
        s = self.summary_raw()
        n_guess = min(cap, max(int(s.final_n), 0) + 0)
        # size buffers from the device's own count: ask once with NULL arrays
        n = C.c_int32(0)
        _check(lib.picb200_download(self._h, C.byref(n), *([None] * 9), None, None))
        n_guess = n.value
        f = [np.empty(n_guess, dtype=np.float32) for _ in range(9)]
        ts = np.empty(n_guess, dtype=np.int32)
        id = np.empty(n_guess, dtype=np.uint64)
        _check(lib.picb200_download(self._h, C.byref(n), *[_ptr(a, _fp) for a in f], _ptr(ts, _ip), _ptr(id, _up)))
        assert n.value == n_guess
        return State(*f, ts, id)

    def download_aos(self, out=None):
        """The population as the reference's 64-byte records (src/electron.h:12-18), transposed on the device."""
        n = C.c_int64(0)
        if out is None:
            _check(lib.picb200_download_aos(self._h, None, 0, C.byref(n)))       # cap 0: only the count
            s = self.summary_raw()
            out = np.zeros(max(int(s.final_n), 0), dtype=ELECTRON_DTYPE)
        _check(lib.picb200_download_aos(self._h, out.ctypes.data_as(C.c_void_p), out.shape[0], C.byref(n)))
        return out[: n.value]

    def poisson_step(self):
        _check(lib.picb200_poisson_step(self._h))

    def run_steps(self, steps):
        _check(lib.picb200_run_steps(self._h, steps))

    def summary_raw(self) -> Summary:
        s = Summary()
        _check(lib.picb200_get_summary(self._h, C.byref(s)))
        return s

    def summary(self) -> dict:
        return self.summary_raw().as_dict()

    def set_profiling(self, level=2):
        _check(lib.picb200_set_profiling(self._h, int(level)))

    def flush_l2(self):
        _check(lib.picb200_flush_l2(self._h))

    def upload_state(self, st):
        """Upload a State (or anything with the same attributes), e.g. pinned host arrays."""
        _check(lib.picb200_upload(self._h, len(st.x), *[_ptr(getattr(st, k), _fp) for k in ("x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az")],
                                  _ptr(st.ts, _ip), _ptr(st.id, _up)))

    def download_into(self, st) -> int:
        """Download into preallocated arrays of a State (must hold at least min(n, capacity) entries)."""
        n = C.c_int32(0)
        _check(lib.picb200_download(self._h, C.byref(n), *[_ptr(getattr(st, k), _fp) for k in ("x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az")],
                                    _ptr(st.ts, _ip), _ptr(st.id, _up)))
        return n.value

    def peer_export(self) -> bytes:
        """This rank's 256-byte peer handle (include/picb200.h: picb200_peer_export)."""
        b = C.create_string_buffer(PEER_HANDLE_BYTES)
        _check(lib.picb200_peer_export(self._h, C.cast(b, C.c_void_p)))
        return b.raw

    def peer_connect(self, handles):
        """handles: the handles of all ranks in rank order (list of bytes)."""
        blob = b"".join(bytes(h) for h in handles)
        assert len(blob) == PEER_HANDLE_BYTES * self.cfg.nranks
        b = C.create_string_buffer(blob, len(blob))
        _check(lib.picb200_peer_connect(self._h, C.cast(b, C.c_void_p)))

    def save_state(self, path, sorted_by_id=True):
        """Binary snapshot of the population and the run counters (picb200_save_state)."""
        _check(lib.picb200_save_state(self._h, os.fsencode(path), 1 if sorted_by_id else 0))

    def load_state(self, path):
        """Replace the population by a snapshot and continue its run (picb200_load_state)."""
        _check(lib.picb200_load_state(self._h, os.fsencode(path)))

    def profile(self) -> dict:
        p = Profile()
        _check(lib.picb200_get_profile(self._h, C.byref(p)))
        return p.as_dict()

    def set_poisson_index(self, idx):
        _check(lib.picb200_set_poisson_index(self._h, idx))

    def op_deposit(self):
        _check(lib.picb200_op_deposit(self._h))

    def op_gather(self):
        _check(lib.picb200_op_gather(self._h))

    def op_clear_grid(self):
        _check(lib.picb200_op_clear_grid(self._h))

    def op_mobility(self):
        _check(lib.picb200_op_mobility(self._h))

    def op_compact(self):
        _check(lib.picb200_op_compact(self._h))

    def charge_bbox(self):
        lo = np.zeros(3, dtype=np.int32)
        hi = np.zeros(3, dtype=np.int32)
        _check(lib.picb200_charge_bbox(self._h, _ptr(lo, _ip), _ptr(hi, _ip)))
        return lo, hi

    def charge_box(self, lo, hi):
        lo = np.ascontiguousarray(lo, dtype=np.int32)
        hi = np.ascontiguousarray(hi, dtype=np.int32)
        shape = tuple(int(v) for v in (hi - lo + 1)[::-1])
        out = np.empty(shape, dtype=np.int32)
        _check(lib.picb200_charge_box(self._h, _ptr(lo, _ip), _ptr(hi, _ip), _ptr(out, _ip)))
        return out  # indexed [z, y, x]

    def charge_box_noalloc(self, lo, hi):
        """charge_box() through the entry point that allocates nothing (fixed device scratch of NOALLOC_MAX_CELLS cells)."""
        lo = np.ascontiguousarray(lo, dtype=np.int32)
        hi = np.ascontiguousarray(hi, dtype=np.int32)
        shape = tuple(int(v) for v in (hi - lo + 1)[::-1])
        out = np.empty(shape, dtype=np.int32)
        _check(lib.picb200_charge_box_noalloc(self._h, _ptr(lo, _ip), _ptr(hi, _ip), _ptr(out, _ip)))
        return out  # indexed [z, y, x]


def run(electrons=False, **kw):
    """runPIC (src/pic.cu:359): returns (summary dict, AoS electrons or None)."""
    cfg, keep = make_config(**kw)
    s = Summary()
    buf = None
    if electrons:
        buf = np.zeros(cfg.capacity, dtype=ELECTRON_DTYPE)
        _check(lib.picb200_run(C.byref(cfg), C.byref(s), buf.ctypes.data_as(C.c_void_p), cfg.capacity))
        buf = buf[: s.final_n]
    else:
        _check(lib.picb200_run(C.byref(cfg), C.byref(s), None, 0))
    return s.as_dict(), buf


def dev_philox(ctr, key, device=-1):
    ctr = np.ascontiguousarray(ctr, dtype=np.uint32).reshape(-1, 4)
    key = np.ascontiguousarray(key, dtype=np.uint32).reshape(2)
    out = np.empty_like(ctr)
    u32 = C.POINTER(C.c_uint32)
    _check(lib.picb200_dev_philox(device, ctr.shape[0], _ptr(ctr, u32), _ptr(key, u32), _ptr(out, u32)))
    return out


def dev_energy_index(energy, device=-1):
    e = np.ascontiguousarray(energy, dtype=np.float32)
    out = np.empty(e.shape, dtype=np.int32)
    _check(lib.picb200_dev_energy_index(device, e.size, _ptr(e, _fp), _ptr(out, _ip)))
    return out


def dev_energy_index_sweep(stride=1, device=-1):
    """(mismatches, bits of one offender) of the device's energy_index over every stride-th non-negative fp32 bit pattern."""
    bad, first = C.c_int64(0), C.c_uint32(0)
    _check(lib.picb200_dev_energy_index_sweep(device, stride, C.byref(bad), C.byref(first)))
    return bad.value, first.value


def load_table(path=None):
    rows = np.zeros((TABLE_ROWS, 2), dtype=np.float32)
    r = lib.picb200_load_table((path or DEFAULT_TABLE).encode(), _ptr(rows, _fp))
    return r, rows


def host_energy_edges():
    e = np.empty(TABLE_ROWS + 1, dtype=np.float32)
    lib.picb200_host_energy_edges(_ptr(e, _fp))
    return e


def snapshot_png(x, y, path):
    """image() of src/utility.cu:52-74 (host only): 500 x 500 PNG, one 11 x 11 black square per electron."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.ascontiguousarray(y, dtype=np.float32)
    assert x.shape == y.shape and x.ndim == 1
    _check(lib.picb200_snapshot_png(_ptr(x, _fp), _ptr(y, _fp), x.size, os.fsencode(path)))


def read_state(path):
    """Parse a picb200_save_state file -> (header dict, dict of arrays)."""
    raw = open(path, "rb").read()
    hdr = np.frombuffer(raw[:72], dtype=np.dtype([("magic", "S8"), ("version", "<u4"), ("header_bytes", "<u4"), ("n", "<i8"),
                                                  ("poisson_index", "<i4"), ("poisson_timestep", "<i4"), ("total_added", "<i8"),
                                                  ("total_removed", "<i8"), ("pushes", "<u8"), ("seed", "<u8"), ("sorted_by_id", "<i4"),
                                                  ("reserved", "<i4")]))[0]
    assert hdr["magic"] == b"PICB200S" and hdr["version"] == 1
    n, off, out = int(hdr["n"]), int(hdr["header_bytes"]), {}
    for k in ("x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az"):
        out[k] = np.frombuffer(raw, dtype="<f4", count=n, offset=off)
        off += 4 * n
    out["ts"] = np.frombuffer(raw, dtype="<i4", count=n, offset=off)
    off += 4 * n
    out["id"] = np.frombuffer(raw, dtype="<u8", count=n, offset=off)
    assert off + 8 * n == len(raw)
    return {k: (hdr[k].item() if k != "magic" else hdr[k]) for k in hdr.dtype.names}, out


def host_draw_threshold(chance):
    """Threshold T on float(w) with (float(w) < T) == (r(w) < chance) for the collider's draw r (draw_threshold, pic_engine.cu)."""
    return np.float32(lib.picb200_host_draw_threshold(C.c_float(float(chance))))


def host_cell_bounds():
    b = np.empty(GRID + 1, dtype=np.float32)
    lib.picb200_host_cell_bounds(_ptr(b, _fp))
    return b


def pinned_empty(n, dtype):
    """numpy array backed by page-locked host memory (freed with the process)."""
    dt = np.dtype(dtype)
    p = lib.picb200_alloc_pinned(max(1, n) * dt.itemsize)
    if not p:
        raise EngineError(lib.picb200_last_error().decode(errors="replace"))
    buf = (C.c_char * (max(1, n) * dt.itemsize)).from_address(p)
    return np.frombuffer(buf, dtype=dt, count=n)


def nccl_unique_id() -> bytes:
    b = C.create_string_buffer(128)
    _check(lib.picb200_nccl_unique_id(C.cast(b, C.c_void_p)))
    return b.raw
