"""The reference's three file formats, restated for the Python host and tests.

* parameter file: Fortran namelist group ENVIRONMENTPARAMS
  (program_particle_simulation_openmp.f08:198-209; environment_params.nml)
* particle file: free-format text `id x y z vx vy vz` per line (:252-267)
* output file: direct-access stream of 4-byte records (:54-56, :327-365); the
  reader of record is python_src/play_simulation.py:45-101.

The C++ host (csrc/host/) implements the same formats; these functions exist so
that tests and tools need neither pygame nor OpenGL.
"""
from __future__ import annotations

import re
import struct

import numpy as np

NAMELIST_GROUP = "ENVIRONMENTPARAMS"
# key -> (type, length); program_...:198-200
NAMELIST_KEYS = {
    "nb_of_threads": (int, 1),
    "sigma": (float, 1),
    "epsilon": (float, 1),
    "mass": (float, 1),
    "r_cut": (float, 1),
    "length_of_area": (float, 3),
    "length_of_cell": (float, 3),
    "delta_t": (float, 1),
    "write_time_step": (int, 1),
    "max_time_step": (int, 1),
}


class NamelistError(ValueError):
    pass


def _fortran_float(tok: str) -> float:
    t = tok.strip().lower().replace("d", "e")
    return float(t)


def parse_namelist(text: str) -> dict:
    """Parse the ENVIRONMENTPARAMS group: case-insensitive keys, `=`, comma or
    blank separated values, `r*c` repeats, `!` comments, `/` terminator.  Keys
    that are absent keep no value (the reference leaves them undefined)."""
    # strip comments
    lines = [ln.split("!", 1)[0] for ln in text.splitlines()]
    body = "\n".join(lines)
    m = re.search(r"&\s*" + NAMELIST_GROUP + r"\b(.*?)(^|[\s,])/", body, re.S | re.I | re.M)
    if not m:
        m2 = re.search(r"&\s*" + NAMELIST_GROUP + r"\b(.*)", body, re.S | re.I)
        if not m2:
            raise NamelistError("namelist group &%s not found" % NAMELIST_GROUP)
        content = m2.group(1)   # no terminator: gfortran reports EOF (istat < 0), which the reference ignores (:210)
    else:
        content = m.group(1)
    out: dict = {}
    # split on `name =`
    parts = re.split(r"([A-Za-z_][A-Za-z0-9_]*)\s*(?:\(\s*\d+\s*(?::\s*\d*\s*)?\))?\s*=", content)
    # parts = [pre, key1, vals1, key2, vals2, ...]
    if parts[0].strip(" \t\r\n,"):
        raise NamelistError("unexpected text before first key: %r" % parts[0].strip())
    for k in range(1, len(parts), 2):
        key = parts[k].lower()
        if key not in NAMELIST_KEYS:
            raise NamelistError("Cannot match namelist object name %s" % parts[k])
        typ, length = NAMELIST_KEYS[key]
        toks = [t for t in re.split(r"[,\s]+", parts[k + 1].strip()) if t]
        vals = []
        for t in toks:
            if "*" in t:
                rep, v = t.split("*", 1)
                vals.extend([v] * int(rep))
            else:
                vals.append(t)
        if len(vals) > length:
            raise NamelistError("too many values for %s" % parts[k])
        try:
            conv = [int(v) if typ is int else _fortran_float(v) for v in vals]
        except ValueError as e:
            raise NamelistError("bad value for %s: %s" % (parts[k], e))
        if length == 1:
            if conv:
                out[key] = conv[0]
        else:
            prev = list(out.get(key, [None] * length))
            prev[: len(conv)] = conv
            out[key] = prev
    return out


def read_namelist(path: str) -> dict:
    with open(path, "r") as f:
        return parse_namelist(f.read())


def read_particles(path: str):
    """Return (ids int32[n], pos float64[n,3], vel float64[n,3]).  List-directed
    read of `id, location(3), velocity(3)` until end of file
    (program_...:252-258); zero particles is an error (:269-270)."""
    ids, rows = [], []
    with open(path, "r") as f:
        toks = f.read().replace(",", " ").split()
    if len(toks) % 7 != 0:
        # a short last record ends the read in the reference too
        toks = toks[: len(toks) - len(toks) % 7]
    if not toks:
        raise ValueError("No particles in particle file")
    a = np.array([_fortran_float(t) for t in toks], dtype=np.float64).reshape(-1, 7)
    ids = a[:, 0].astype(np.int32)
    return ids, np.ascontiguousarray(a[:, 1:4]), np.ascontiguousarray(a[:, 4:7])


def write_particles(path: str, ids, pos, vel, append: bool = False):
    """`id x y z vx vy vz ` with %.4f, as write_particle_file.py:72-75 does."""
    with open(path, "a" if append else "w") as f:
        for i, p, v in zip(ids, pos, vel):
            f.write("%d " % i)
            f.write("".join("%.4f " % x for x in p))
            f.write("".join("%.4f " % x for x in v))
            f.write("\n")


class OutputWriter:
    """The `.output` stream (program_...:54-56, 327-365): records of `nbytes`
    (4 as shipped; 8 when rt => real64, integers then occupy the low 4 bytes of
    an 8-byte record).  Record 1-3 = length_of_area, record 4 = frames so far
    (written by write_data only, so a run that never reaches WRITE_TIME_STEP
    leaves a 3-record file), then per frame one count record (7 * n_active) and
    n_active * 7 records."""

    def __init__(self, path: str, length_of_area, nbytes: int = 4):
        assert nbytes in (4, 8)
        self.nbytes = nbytes
        self.f = open(path, "wb")
        self.rt = np.dtype("<f4") if nbytes == 4 else np.dtype("<f8")
        self.f.write(np.asarray(length_of_area, dtype=self.rt).tobytes())
        self.frames = 0          # record 4 (frames so far) first appears with the first frame (program:336-337)

    def _int_record(self, v: int) -> bytes:
        return struct.pack("<i", v) + (b"\0" * (self.nbytes - 4))

    def _write_count(self):
        self.f.seek(3 * self.nbytes)
        self.f.write(self._int_record(self.frames))
        self.f.seek(0, 2)

    def write_frame(self, ids, pos, vel, active=None):
        ids = np.asarray(ids, dtype=np.int32)
        pos = np.asarray(pos, dtype=self.rt).reshape(-1, 3)
        vel = np.asarray(vel, dtype=self.rt).reshape(-1, 3)
        if active is not None:
            keep = np.asarray(active).astype(bool)
            ids, pos, vel = ids[keep], pos[keep], vel[keep]
        n = len(ids)
        self.frames += 1
        self._write_count()
        raw = np.zeros((n, 7, self.nbytes), dtype=np.uint8)
        raw[:, 0, :4] = ids.astype("<i4").view(np.uint8).reshape(n, 4)
        raw[:, 1:4, :] = pos.view(np.uint8).reshape(n, 3, self.nbytes)
        raw[:, 4:7, :] = vel.view(np.uint8).reshape(n, 3, self.nbytes)
        self.f.write(self._int_record(7 * n))
        self.f.write(raw.tobytes())

    def close(self):
        if self.f:
            self.f.close()
            self.f = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def read_output(path: str, nbytes: int = 4):
    """Headless restatement of play_simulation.py:45-101 (`get_data`), keeping
    ids and velocities instead of discarding them.  Returns
    (length_of_area[3], declared_frames, [(ids, pos, vel), ...])."""
    ff, fi = ("<f", "<i") if nbytes == 4 else ("<d", "<q")
    with open(path, "rb") as f:
        data = f.read()
    off = 0

    def rd(fmt):
        nonlocal off
        if off + nbytes > len(data):
            return None
        v = struct.unpack_from(fmt, data, off)[0]
        off += nbytes
        return v

    area = [rd(ff) for _ in range(3)]
    declared = rd(fi)
    if declared is None:         # no frame was ever written: the file ends after record 3
        declared = 0
    frames = []
    while True:
        cnt = rd(fi)
        if cnt is None:
            break
        npart = int(cnt / 7)
        need = npart * 7 * nbytes
        if off + need > len(data):
            raise ValueError("truncated frame")
        block = np.frombuffer(data, dtype=np.uint8, count=need, offset=off).reshape(npart, 7, nbytes)
        off += need
        ids = block[:, 0, :4].copy().view("<i4").reshape(npart)
        rt_ = "<f4" if nbytes == 4 else "<f8"
        pos = block[:, 1:4, :].copy().view(rt_).reshape(npart, 3)
        vel = block[:, 4:7, :].copy().view(rt_).reshape(npart, 3)
        frames.append((ids, pos, vel))
    return area, declared, frames


def params_from_namelist(nml: dict):
    """namelist dict -> native.Params, mapping SIGMA -> particle_radius, R_CUT ->
    r_cut_factor, EPSILON -> input_epsilon_value (program_...:219-220)."""
    from .native import make_params

    missing = [k for k in ("sigma", "epsilon", "mass", "r_cut", "length_of_area", "length_of_cell", "delta_t")
               if k not in nml]
    if missing:
        raise NamelistError("missing namelist keys: " + ", ".join(missing))
    return make_params(delta_t=nml["delta_t"], epsilon=nml["epsilon"], sigma=nml["sigma"], mass=nml["mass"],
                       r_cut=nml["r_cut"], length_of_area=nml["length_of_area"],
                       length_of_cell=nml["length_of_cell"], nb_of_threads=nml.get("nb_of_threads", 1))
