"""ctypes binding of the CPU oracle (oracle/psim_oracle.cpp).

TEST INFRASTRUCTURE, NOT PRODUCT: imported only by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs.
PARITY UNPINNED (see psim_oracle.h).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class Params(C.Structure):
    """Mirror of psim_params (include/psim.h)."""

    _fields_ = [
        ("delta_t", C.c_double),
        ("epsilon", C.c_double),
        ("sigma", C.c_double),
        ("mass", C.c_double),
        ("r_cut_factor", C.c_double),
        ("length_of_area", C.c_double * 3),
        ("length_of_cell", C.c_double * 3),
        ("nb_of_threads", C.c_int32),
        ("reserved", C.c_int32),
    ]


def make_params(delta_t=0.001, epsilon=1.0, sigma=1.0, mass=1.0, r_cut=2.5,
                length_of_area=(99.0, 99.0, 99.0), length_of_cell=(3.0, 3.0, 3.0),
                nb_of_threads=1) -> Params:
    """Defaults are the shipped environment_params.nml:2-11 (NB_OF_THREADS aside)."""
    p = Params()
    p.delta_t, p.epsilon, p.sigma, p.mass, p.r_cut_factor = delta_t, epsilon, sigma, mass, r_cut
    p.length_of_area[:] = list(length_of_area)
    p.length_of_cell[:] = list(length_of_cell)
    p.nb_of_threads = nb_of_threads
    p.reserved = 0
    return p


def build(variant: str = "", out_dir: str | None = None, quiet: bool = True) -> str:
    """Compile the oracle with oracle/Makefile; variant '' | 'fast' | 'ofast'."""
    out = os.path.abspath(out_dir or os.path.join(_HERE, "_build"))
    name = "libpsim_oracle%s.so" % ("_" + variant if variant else "")
    target = os.path.join(out, name)
    subprocess.run(["make", "-C", _HERE, "OUT=%s" % out, target], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)
    return target


_libs: dict[str, C.CDLL] = {}


def load(variant: str = "", path: str | None = None) -> C.CDLL:
    key = path or variant
    if key in _libs:
        return _libs[key]
    if path is None:
        path = os.path.join(_HERE, "_build", "libpsim_oracle%s.so" % ("_" + variant if variant else ""))
        if not os.path.exists(path):
            build(variant)
    lib = C.CDLL(path)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    lib.psim_oracle_create.restype = vp
    lib.psim_oracle_create.argtypes = [C.POINTER(Params), i32, i64, vp, vp, vp, i32]
    lib.psim_oracle_destroy.argtypes = [vp]
    lib.psim_oracle_advance.argtypes = [vp, i64]
    lib.psim_oracle_get_frame.argtypes = [vp, vp, vp, vp, vp]
    lib.psim_oracle_get_aux.argtypes = [vp, vp, vp, vp]
    for f in ("rebuilds", "steps", "n_active", "list_entries"):
        getattr(lib, "psim_oracle_" + f).restype = i64
        getattr(lib, "psim_oracle_" + f).argtypes = [vp]
    lib.psim_oracle_loop_seconds.restype = C.c_double
    lib.psim_oracle_loop_seconds.argtypes = [vp]
    lib.psim_oracle_constants.argtypes = [C.POINTER(Params), i32, C.POINTER(C.c_double)]
    lib.psim_oracle_pair_force_x.restype = C.c_double
    lib.psim_oracle_pair_force_x.argtypes = [C.POINTER(Params), i32, C.c_double]
    lib.psim_oracle_bruteforce.restype = i64
    lib.psim_oracle_bruteforce.argtypes = [C.POINTER(Params), i32, i64, vp, vp, vp]
    lib.psim_oracle_energy.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    _libs[key] = lib
    return lib


def _rt(precision):
    return np.float32 if precision == 4 else np.float64


class Oracle:
    """One oracle simulation (the reference's time loop on the CPU)."""

    def __init__(self, params: Params, precision: int, ids, pos, vel, threads: int = 1,
                 variant: str = "", lib_path: str | None = None):
        self.lib = load(variant, lib_path)
        self.precision = precision
        rt = _rt(precision)
        self.n = int(len(ids))
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        pos = np.ascontiguousarray(pos, dtype=rt).reshape(-1)
        vel = np.ascontiguousarray(vel, dtype=rt).reshape(-1)
        assert pos.size == 3 * self.n and vel.size == 3 * self.n
        self.h = self.lib.psim_oracle_create(C.byref(params), precision, self.n, ids.ctypes.data,
                                             pos.ctypes.data, vel.ctypes.data, threads)
        if not self.h:
            raise ValueError("psim_oracle_create failed")

    def close(self):
        if self.h:
            self.lib.psim_oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def advance(self, nsteps: int):
        self.lib.psim_oracle_advance(self.h, nsteps)

    def frame(self):
        rt = _rt(self.precision)
        ids = np.empty(self.n, np.int32)
        pos = np.empty((self.n, 3), rt)
        vel = np.empty((self.n, 3), rt)
        act = np.empty(self.n, np.uint8)
        self.lib.psim_oracle_get_frame(self.h, ids.ctypes.data, pos.ctypes.data, vel.ctypes.data, act.ctypes.data)
        return ids, pos, vel, act

    def aux(self):
        rt = _rt(self.precision)
        d = np.empty((self.n, 3), rt)
        fo = np.empty((self.n, 3), rt)
        fn = np.empty((self.n, 3), rt)
        self.lib.psim_oracle_get_aux(self.h, d.ctypes.data, fo.ctypes.data, fn.ctypes.data)
        return d, fo, fn

    def energy(self):
        k, u = C.c_double(), C.c_double()
        m = (C.c_double * 3)()
        self.lib.psim_oracle_energy(self.h, C.byref(k), C.byref(u), m)
        return k.value, u.value, np.array(list(m))

    rebuilds = property(lambda s: s.lib.psim_oracle_rebuilds(s.h))
    steps = property(lambda s: s.lib.psim_oracle_steps(s.h))
    n_active = property(lambda s: s.lib.psim_oracle_n_active(s.h))
    list_entries = property(lambda s: s.lib.psim_oracle_list_entries(s.h))
    loop_seconds = property(lambda s: s.lib.psim_oracle_loop_seconds(s.h))


def constants(params: Params, precision: int, variant: str = ""):
    out = (C.c_double * 7)()
    load(variant).psim_oracle_constants(C.byref(params), precision, out)
    keys = ("r_min", "r_cut", "min_length_of_cell", "dr_max_tolerable", "F_rcut", "time_mass_factor", "delta_t")
    return dict(zip(keys, list(out)))


def pair_force_x(params: Params, precision: int, r: float) -> float:
    return load().psim_oracle_pair_force_x(C.byref(params), precision, r)


def bruteforce(params: Params, precision: int, pos, active=None):
    rt = _rt(precision)
    pos = np.ascontiguousarray(pos, dtype=rt).reshape(-1)
    n = pos.size // 3
    out = np.empty((n, 3), np.float64)
    act_p = None
    if active is not None:
        active = np.ascontiguousarray(active, dtype=np.uint8)
        act_p = active.ctypes.data
    pairs = load().psim_oracle_bruteforce(C.byref(params), precision, n, pos.ctypes.data, act_p, out.ctypes.data)
    return out, int(pairs)
