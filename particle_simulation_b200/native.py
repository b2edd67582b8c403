"""ctypes binding of libpsim_b200.so (include/psim.h).

The names mirror the reference's integrator interface
(fortran_src/module_stoermer_verlet_method.f08:40-46): a Simulation is created
from the namelist parameters and the particle arrays (set_simulation_parameter +
the particle array), `advance` runs trips of the reference time loop
(program_particle_simulation_openmp.f08:80-117) and `frame` returns what
write_data gathers (:327-365).

There is no CPU path in this package: if the CUDA library is missing or no
device is present, every compute call raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

STATUS_OK = 0
FORCE_ALL_PAIRS, FORCE_CELL_BINNED = 0, 1
OPT_FORCE_ALGO, OPT_FUSED_STEP, OPT_KERNEL_TIMING, OPT_ALLPAIRS_KERNEL = 0, 1, 2, 3
OPT_EXACT_DIFF_FILTER = OPT_ALLPAIRS_KERNEL      # round-1 name
ALLPAIRS_TENSOR, ALLPAIRS_DIFFERENCE_FORM, ALLPAIRS_CUDA_CORE = 0, 1, 5
FRAME_SLOTS = 4


class Params(C.Structure):
    """psim_params: the ENVIRONMENTPARAMS namelist keys the integrator needs
    (program_particle_simulation_openmp.f08:198-200, 219-220)."""

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


class Counters(C.Structure):
    _fields_ = [
        ("steps", C.c_int64),
        ("rebuilds", C.c_int64),
        ("kernel_launches", C.c_int64),
        ("force_launches", C.c_int64),
        ("force_ms", C.c_double),
        ("force_timed", C.c_int64),
        ("n_active", C.c_int64),
        ("filter_pairs", C.c_int64),
        ("filter_tests", C.c_int64),
        ("peer_push_steps", C.c_int64),
    ]


def make_params(delta_t=0.001, epsilon=1.0, sigma=1.0, mass=1.0, r_cut=2.5,
                length_of_area=(99.0, 99.0, 99.0), length_of_cell=(3.0, 3.0, 3.0),
                nb_of_threads=1) -> Params:
    """Defaults are the shipped environment_params.nml:3-9."""
    p = Params()
    p.delta_t, p.epsilon, p.sigma, p.mass, p.r_cut_factor = delta_t, epsilon, sigma, mass, r_cut
    p.length_of_area[:] = [float(v) for v in length_of_area]
    p.length_of_cell[:] = [float(v) for v in length_of_cell]
    p.nb_of_threads = int(nb_of_threads)
    p.reserved = 0
    return p


class PsimError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str = ""):
        self.status = status
        super().__init__("%s failed: status %d (%s)%s" % (where, status, _status_string(status),
                                                         (": " + detail) if detail else ""))


_lib = None

# every symbol include/psim.h declares
EXPORTS = (
    "psim_create", "psim_create_sharded", "psim_comm_unique_id", "psim_destroy", "psim_set_option",
    "psim_advance", "psim_synchronize", "psim_advance_timed", "psim_get_frame", "psim_get_aux",
    "psim_get_counters", "psim_get_shard", "psim_set_positions", "psim_snapshot_async", "psim_snapshot_wait",
    "psim_compute_forces_host", "psim_count_pairs_in_range", "psim_energy", "psim_measure_fma_peak",
    "psim_status_string", "psim_last_error", "psim_version", "psim_compute_locations_static",
    "psim_measure_tensor_peak",
)


def library_path() -> str:
    return _build.LIB


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load the CUDA library; raises if it cannot be found or built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        if not build_if_missing:
            raise FileNotFoundError(path + " is missing: run `python -m particle_simulation_b200.build`")
        _build.build_library()
    lib = C.CDLL(path)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    PP = C.POINTER(Params)
    sig = {
        "psim_create": [PP, i32, i64, vp, vp, vp, i32, C.POINTER(vp)],
        "psim_create_sharded": [PP, i32, i64, vp, vp, vp, i32, i32, i32, vp, C.POINTER(vp)],
        "psim_comm_unique_id": [vp],
        "psim_destroy": [vp],
        "psim_set_option": [vp, i32, i64],
        "psim_advance": [vp, i64],
        "psim_synchronize": [vp],
        "psim_advance_timed": [vp, i64, C.POINTER(C.c_double)],
        "psim_get_frame": [vp, vp, vp, vp, vp],
        "psim_get_aux": [vp, vp, vp, vp],
        "psim_get_counters": [vp, C.POINTER(Counters)],
        "psim_get_shard": [vp, C.POINTER(i64), C.POINTER(i64)],
        "psim_set_positions": [vp, vp, vp],
        "psim_snapshot_async": [vp, i32],
        "psim_snapshot_wait": [vp, i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)],
        "psim_compute_forces_host": [PP, i32, i64, vp, vp, i32, i32, vp, C.POINTER(i64)],
        "psim_count_pairs_in_range": [vp, C.POINTER(i64)],
        "psim_energy": [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)],
        "psim_measure_fma_peak": [i32, i32, C.POINTER(C.c_double), C.POINTER(C.c_double)],
        "psim_measure_tensor_peak": [i32, i32, C.POINTER(C.c_double), C.POINTER(C.c_double)],
        "psim_compute_locations_static": [PP, i32, i64, vp, vp, vp, vp, vp, vp, vp, vp],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    lib.psim_status_string.restype = C.c_char_p
    lib.psim_status_string.argtypes = [C.c_int]
    lib.psim_last_error.restype = C.c_char_p
    lib.psim_version.restype = C.c_char_p
    _lib = lib
    return lib


def _status_string(status: int) -> str:
    try:
        return load(False).psim_status_string(status).decode()
    except Exception:
        return "?"


def _check(status: int, where: str):
    if status != STATUS_OK:
        detail = ""
        try:
            detail = load(False).psim_last_error().decode()
        except Exception:
            pass
        raise PsimError(status, where, detail)


def rt(precision: int):
    if precision == 4:
        return np.float32
    if precision == 8:
        return np.float64
    raise ValueError("precision must be 4 or 8")


def _ptr(a):
    return None if a is None else a.ctypes.data


class Simulation:
    """One particle system resident on one GPU (or one i-shard of it)."""

    def __init__(self, params: Params, ids, pos, vel, precision: int = 4, device: int = -1,
                 rank: int = 0, world: int = 1, nccl_unique_id: bytes | None = None):
        self.lib = load()
        self.precision = precision
        self.params = params
        r = rt(precision)
        pos = np.ascontiguousarray(pos, dtype=r).reshape(-1)
        self.n = pos.size // 3
        vel = np.ascontiguousarray(vel, dtype=r).reshape(-1)
        ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(-1)
        if vel.size != 3 * self.n or ids.size != self.n:
            raise ValueError("ids[n], pos[n,3], vel[n,3] expected")
        h = C.c_void_p()
        if world > 1:
            uid = C.create_string_buffer(nccl_unique_id, 128)
            st = self.lib.psim_create_sharded(C.byref(params), precision, self.n, _ptr(ids), _ptr(pos), _ptr(vel),
                                              device, rank, world, C.cast(uid, C.c_void_p), C.byref(h))
        else:
            st = self.lib.psim_create(C.byref(params), precision, self.n, _ptr(ids), _ptr(pos), _ptr(vel),
                                      device, C.byref(h))
        _check(st, "psim_create")
        self.h = h
        lo, hi = C.c_int64(), C.c_int64()
        _check(self.lib.psim_get_shard(self.h, C.byref(lo), C.byref(hi)), "psim_get_shard")
        self.i_lo, self.i_hi = lo.value, hi.value

    # -- life cycle
    def close(self):
        if getattr(self, "h", None):
            self.lib.psim_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_option(self, option: int, value: int):
        _check(self.lib.psim_set_option(self.h, option, value), "psim_set_option")

    # -- time loop
    def advance(self, nsteps: int, sync: bool = True):
        _check(self.lib.psim_advance(self.h, nsteps), "psim_advance")
        if sync:
            self.synchronize()

    def advance_timed(self, nsteps: int) -> float:
        ms = C.c_double()
        _check(self.lib.psim_advance_timed(self.h, nsteps, C.byref(ms)), "psim_advance_timed")
        return ms.value

    def synchronize(self):
        _check(self.lib.psim_synchronize(self.h), "psim_synchronize")

    # -- state
    def frame(self):
        r = rt(self.precision)
        ids = np.empty(self.n, np.int32)
        pos = np.zeros((self.n, 3), r)
        vel = np.zeros((self.n, 3), r)
        act = np.empty(self.n, np.uint8)
        _check(self.lib.psim_get_frame(self.h, _ptr(ids), _ptr(pos), _ptr(vel), _ptr(act)), "psim_get_frame")
        return ids, pos, vel, act

    def aux(self):
        r = rt(self.precision)
        d = np.zeros((self.n, 3), r)
        fo = np.zeros((self.n, 3), r)
        fn = np.zeros((self.n, 3), r)
        _check(self.lib.psim_get_aux(self.h, _ptr(d), _ptr(fo), _ptr(fn)), "psim_get_aux")
        return d, fo, fn

    def forces(self, out=None):
        r = rt(self.precision)
        fn = np.zeros((self.n, 3), r) if out is None else out
        _check(self.lib.psim_get_aux(self.h, None, None, _ptr(fn)), "psim_get_aux")
        return fn

    def set_positions(self, pos, active=None):
        pos = np.ascontiguousarray(pos, dtype=rt(self.precision)).reshape(-1)
        if active is not None:
            active = np.ascontiguousarray(active, dtype=np.uint8)
        _check(self.lib.psim_set_positions(self.h, _ptr(pos), _ptr(active)), "psim_set_positions")

    def counters(self) -> Counters:
        c = Counters()
        _check(self.lib.psim_get_counters(self.h, C.byref(c)), "psim_get_counters")
        return c

    def pairs_in_range(self) -> int:
        v = C.c_int64()
        _check(self.lib.psim_count_pairs_in_range(self.h, C.byref(v)), "psim_count_pairs_in_range")
        return v.value

    def energy(self):
        k, u = C.c_double(), C.c_double()
        m = (C.c_double * 3)()
        _check(self.lib.psim_energy(self.h, C.byref(k), C.byref(u), m), "psim_energy")
        return k.value, u.value, np.array(list(m))

    # -- frame ring
    def snapshot_async(self, slot: int):
        _check(self.lib.psim_snapshot_async(self.h, slot), "psim_snapshot_async")

    def snapshot_wait(self, slot: int):
        """Views into the slot's pinned buffers (valid until the slot is reused)."""
        pid, pp, pv, pa = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        _check(self.lib.psim_snapshot_wait(self.h, slot, C.byref(pid), C.byref(pp), C.byref(pv), C.byref(pa)),
               "psim_snapshot_wait")
        r = rt(self.precision)
        ct = C.c_float if self.precision == 4 else C.c_double
        ids = np.ctypeslib.as_array(C.cast(pid, C.POINTER(C.c_int32)), shape=(self.n,))
        pos = np.ctypeslib.as_array(C.cast(pp, C.POINTER(ct)), shape=(self.n, 3)).view(r)
        vel = np.ctypeslib.as_array(C.cast(pv, C.POINTER(ct)), shape=(self.n, 3)).view(r)
        act = np.ctypeslib.as_array(C.cast(pa, C.POINTER(C.c_uint8)), shape=(self.n,))
        return ids, pos, vel, act


def compute_forces(params: Params, pos, precision: int = 4, active=None, algo: int = FORCE_ALL_PAIRS,
                   device: int = -1):
    """compute_forces (module_stoermer_verlet_method.f08:87-132) as an operator
    over host arrays.  Returns (force[n,3], ordered pairs inside the cutoff)."""
    lib = load()
    r = rt(precision)
    pos = np.ascontiguousarray(pos, dtype=r).reshape(-1)
    n = pos.size // 3
    out = np.zeros((n, 3), r)
    if active is not None:
        active = np.ascontiguousarray(active, dtype=np.uint8)
    pairs = C.c_int64()
    _check(lib.psim_compute_forces_host(C.byref(params), precision, n, _ptr(pos), _ptr(active), algo, device,
                                        _ptr(out), C.byref(pairs)), "psim_compute_forces_host")
    return out, pairs.value


STATIC_MAX_PARTICLES = 1048576


def compute_locations_static(params: Params, precision: int, n: int, pos_dev: int, vel_dev: int, force_old_dev: int,
                             force_new_dev: int, displacement_dev: int, active_dev: int | None,
                             rebuild_flag_dev: int, cuda_stream: int | None = None):
    """psim_compute_locations_static: one drift step, in place, on device arrays
    the caller owns (raw device addresses), for hosts that cannot let the
    library allocate GPU memory.  Limited to STATIC_MAX_PARTICLES; calls that
    can overlap on one GPU must be ordered by the caller (shared fixed scratch)."""
    _check(load().psim_compute_locations_static(C.byref(params), precision, n, pos_dev, vel_dev, force_old_dev,
                                                force_new_dev, displacement_dev, active_dev, rebuild_flag_dev,
                                                cuda_stream), "psim_compute_locations_static")


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    _check(load().psim_comm_unique_id(C.cast(buf, C.c_void_p)), "psim_comm_unique_id")
    return buf.raw


def measure_fma_peak(precision: int, device: int = 0):
    tf, mhz = C.c_double(), C.c_double()
    _check(load().psim_measure_fma_peak(precision, device, C.byref(tf), C.byref(mhz)), "psim_measure_fma_peak")
    return tf.value, mhz.value


def measure_tensor_peak(kind: int = 0, device: int = 0):
    """Dense tcgen05.mma rate (TFLOP/s) and SM clock (MHz): kind 0 = kind::tf32, 1 = kind::f16."""
    tf, mhz = C.c_double(), C.c_double()
    _check(load().psim_measure_tensor_peak(kind, device, C.byref(tf), C.byref(mhz)), "psim_measure_tensor_peak")
    return tf.value, mhz.value
