"""ctypes wrapper of oracle/_build/liboracle.so -- the CPU oracle (test infrastructure only)."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
CLI = os.path.join(ORACLE_DIR, "_build", "pic_oracle")
DEFAULT_TABLE = os.path.join(ROOT, "src", "cross_section.txt")
TABLE_ROWS = 10000


def build():
    subprocess.run(["make", "-C", ORACLE_DIR, "-s", "all"], check=True)


if not os.path.exists(LIB):
    build()
lib = C.CDLL(LIB)

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int32)
_up = C.POINTER(C.c_uint64)
_h = C.c_void_p
_u32p = C.POINTER(C.c_uint32)

lib.orc_philox4x32_10.argtypes = [_u32p, _u32p, _u32p]
lib.orc_uniform01.restype = C.c_float
lib.orc_uniform01.argtypes = [C.c_uint32]
lib.orc_child_id.restype = C.c_uint64
lib.orc_child_id.argtypes = [C.c_uint64, C.c_uint64]
lib.orc_energy_to_index.restype = C.c_int
lib.orc_energy_to_index.argtypes = [C.c_double]
lib.orc_electric_force_constant.restype = C.c_double
lib.orc_load_table.restype = C.c_int
lib.orc_load_table.argtypes = [C.c_char_p, _fp]
lib.orc_create.restype = _h
lib.orc_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int]
lib.orc_destroy.argtypes = [_h]
lib.orc_set_table.argtypes = [_h, _fp]
lib.orc_set_table_file.restype = C.c_int
lib.orc_set_table_file.argtypes = [_h, C.c_char_p]
lib.orc_set_poisson_index.argtypes = [_h, C.c_int]
lib.orc_init_uniform.argtypes = [_h, C.c_int, C.c_uint64]
lib.orc_upload.argtypes = [_h, C.c_int] + [_dp] * 6 + [_fp] * 3 + [_ip, _up]
lib.orc_n.restype = C.c_int
lib.orc_n.argtypes = [_h]
lib.orc_download.argtypes = [_h] + [_dp] * 6 + [_fp] * 3 + [_ip, _up]
lib.orc_counters.argtypes = [_h, C.POINTER(C.c_int64)]
lib.orc_poisson_step.argtypes = [_h, C.c_int]
lib.orc_op_deposit.argtypes = [_h]
lib.orc_op_gather.argtypes = [_h]
lib.orc_op_mobility.argtypes = [_h, C.c_int]
lib.orc_op_compact.argtypes = [_h]
lib.orc_charge_at.restype = C.c_int
lib.orc_charge_at.argtypes = [_h, C.c_int, C.c_int, C.c_int]
lib.orc_charge_box.argtypes = [_h, _ip, _ip, _ip]
lib.orc_run.restype = C.c_double
lib.orc_run.argtypes = [_h, C.c_int, C.c_int]
lib.orc_max_threads.restype = C.c_int


def philox(ctr, key):
    ctr = np.ascontiguousarray(ctr, dtype=np.uint32).reshape(-1, 4)
    key = np.ascontiguousarray(key, dtype=np.uint32).reshape(2)
    out = np.empty_like(ctr)
    for i in range(ctr.shape[0]):
        lib.orc_philox4x32_10(ctr[i].ctypes.data_as(_u32p), key.ctypes.data_as(_u32p), out[i].ctypes.data_as(_u32p))
    return out


def energy_to_index(e):
    return int(lib.orc_energy_to_index(float(e)))


def load_table(path=DEFAULT_TABLE):
    rows = np.zeros((TABLE_ROWS, 2), dtype=np.float32)
    r = lib.orc_load_table(path.encode(), rows.ctypes.data_as(_fp))
    return r, rows


class Oracle:
    FIELDS = ("x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az", "ts", "id")

    def __init__(self, capacity, poisson_timestep, fp64=False, seed=39587, threads=1, table_rows=None, table_path=None):
        self.h = lib.orc_create(capacity, poisson_timestep, int(fp64), seed, threads)
        if table_rows is not None:
            t = np.ascontiguousarray(table_rows, dtype=np.float32).reshape(TABLE_ROWS, 2)
            lib.orc_set_table(self.h, t.ctypes.data_as(_fp))
        else:
            r = lib.orc_set_table_file(self.h, (table_path or DEFAULT_TABLE).encode())
            assert r == TABLE_ROWS, f"table read {r} rows"

    def close(self):
        if self.h:
            lib.orc_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def init_uniform(self, n, id_base=0):
        lib.orc_init_uniform(self.h, n, id_base)

    def upload(self, x, y, z, vx=None, vy=None, vz=None, ax=None, ay=None, az=None, ts=None, id=None):
        n = len(x)
        d = [np.ascontiguousarray(np.zeros(n) if a is None else a, dtype=np.float64) for a in (x, y, z, vx, vy, vz)]
        f = [np.ascontiguousarray(np.zeros(n) if a is None else a, dtype=np.float32) for a in (ax, ay, az)]
        ts = np.ascontiguousarray(np.full(n, -1) if ts is None else ts, dtype=np.int32)
        id = np.ascontiguousarray(np.arange(n) if id is None else id, dtype=np.uint64)
        lib.orc_upload(self.h, n, *[a.ctypes.data_as(_dp) for a in d], *[a.ctypes.data_as(_fp) for a in f],
                       ts.ctypes.data_as(_ip), id.ctypes.data_as(_up))

    @property
    def n(self):
        return int(lib.orc_n(self.h))

    def download(self):
        n = self.n
        d = [np.empty(n, dtype=np.float64) for _ in range(6)]
        f = [np.empty(n, dtype=np.float32) for _ in range(3)]
        ts = np.empty(n, dtype=np.int32)
        id = np.empty(n, dtype=np.uint64)
        lib.orc_download(self.h, *[a.ctypes.data_as(_dp) for a in d], *[a.ctypes.data_as(_fp) for a in f],
                         ts.ctypes.data_as(_ip), id.ctypes.data_as(_up))
        out = dict(zip(self.FIELDS, d + f + [ts, id]))
        return out

    def download_sorted(self):
        s = self.download()
        o = np.argsort(s["id"], kind="stable")
        return {k: v[o] for k, v in s.items()}

    def counters(self):
        c = (C.c_int64 * 8)()
        lib.orc_counters(self.h, c)
        return dict(n=c[0], n_after_mobility=c[1], added=c[2], removed=c[3], pushes=c[4], poisson_index=c[5])

    def poisson_step(self, stepwise=False):
        lib.orc_poisson_step(self.h, int(stepwise))

    def run(self, steps, stepwise=False):
        return float(lib.orc_run(self.h, steps, int(stepwise)))

    def set_poisson_index(self, i):
        lib.orc_set_poisson_index(self.h, i)

    def op_deposit(self):
        lib.orc_op_deposit(self.h)

    def op_gather(self):
        lib.orc_op_gather(self.h)

    def op_mobility(self, stepwise=False):
        lib.orc_op_mobility(self.h, int(stepwise))

    def op_compact(self):
        lib.orc_op_compact(self.h)

    def charge_box(self, lo, hi):
        lo = np.ascontiguousarray(lo, dtype=np.int32)
        hi = np.ascontiguousarray(hi, dtype=np.int32)
        shape = tuple(int(v) for v in (hi - lo + 1)[::-1])
        out = np.empty(shape, dtype=np.int32)
        lib.orc_charge_box(self.h, lo.ctypes.data_as(_ip), hi.ctypes.data_as(_ip), out.ctypes.data_as(_ip))
        return out
