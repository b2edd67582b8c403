"""The C-ABI library loads and exports every symbol include/psim.h declares, and
fails loudly (never silently on the CPU) when there is no CUDA device."""
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "psim.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(psim_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(psim):
    lib = psim.native.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "libpsim_b200.so does not export " + n
    assert set(names) == set(psim.native.EXPORTS)
    assert b"sm_100a" in lib.psim_version()


def test_params_struct_layout_matches_header(psim, oracle):
    import ctypes as C
    assert C.sizeof(psim.native.Params) == 5 * 8 + 2 * 3 * 8 + 2 * 4 == C.sizeof(oracle.Params)
    assert C.sizeof(psim.native.Counters) == 10 * 8


def test_no_device_is_a_loud_error_not_a_cpu_fallback(psim):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    pos = np.zeros((4, 3))
    with pytest.raises(psim.PsimError) as e:
        psim.Simulation(psim.make_params(), [1, 2, 3, 4], pos + 5, pos, 8)
    assert e.value.status == 2 and "no CPU fallback" in str(e.value)
    with pytest.raises(psim.PsimError):
        psim.compute_forces(psim.make_params(), pos + 5, 8)
    with pytest.raises(psim.PsimError):
        psim.native.measure_fma_peak(8)


def test_argument_validation_needs_no_device(psim):
    import ctypes as C
    lib = psim.native.load()
    h = C.c_void_p()
    p = psim.make_params()
    assert lib.psim_create(C.byref(p), 8, 0, None, None, None, -1, C.byref(h)) != 0
    assert lib.psim_create(C.byref(p), 3, 4, None, None, None, -1, C.byref(h)) == 1      # bad precision
    assert lib.psim_advance(None, 1) == 1 and lib.psim_get_frame(None, None, None, None, None) == 1
    assert lib.psim_status_string(4) == b"No particles in particle file"


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import, link,
    load or include it (comments may mention it)."""
    pkg = os.path.join(ROOT, "particle_simulation_b200")
    bad = re.compile(r"(^\s*(from|import)\s+oracle\b|oracle_py|psim_oracle|libpsim_oracle|#include\s+[\"<][^\">]*oracle)", re.M)
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp")):
                src = open(os.path.join(base, f), errors="ignore").read()
                assert not bad.search(src), os.path.join(base, f) + " uses the oracle"


def test_fortran_interface_source_binds_every_entry_point(psim):
    """fortran/psim_binding.f08 cannot be compiled on this project's machines (no Fortran compiler), but it can
    be kept complete: every symbol include/psim.h declares has a bind(C, name=...) interface in it, and the two
    interoperable structs have the C structs' field counts."""
    import os
    import re
    src = open(os.path.join(os.path.dirname(psim.__file__), "fortran", "psim_binding.f08")).read()
    bound = set(re.findall(r'name="(psim_\w+)"', src))
    assert set(psim.native.EXPORTS) <= bound, sorted(set(psim.native.EXPORTS) - bound)
    counters = re.search(r"type, bind\(C\) :: psim_counters(.*?)end type psim_counters", src, re.S).group(1)
    fields = [f for line in counters.strip().splitlines() for f in line.split("::")[1].split(",")]
    assert len(fields) == len(psim.native.Counters._fields_)
