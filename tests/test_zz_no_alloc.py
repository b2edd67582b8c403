"""psim_compute_locations_static — the drift step for hosts that cannot let a
library allocate GPU memory (a driver whose own memory plan already owns the
device, or a loop that cannot afford cudaMalloc's device-wide synchronisation).
Its reduction scratch is a fixed array inside the library, so this test checks
two things the ordinary path gets for free: that the result equals the
ordinary path's, and that a second call, which finds the scratch as the first
one left it, is still right.

The entry point works on device arrays the CALLER owns; here they are torch
CUDA tensors, passed as raw addresses through the C ABI.  The reference for
comparison is the existing handle-based path: psim_advance(1) moves positions
by exactly one drift (force and kick do not touch positions), leaves
force_old = previous force_new, and counts a rebuild when the trigger fired.

This file is named so that pytest collects it last.
"""
import numpy as np
import pytest

from conftest import two_cubes

pytestmark = pytest.mark.gpu

# same tolerances as tests/test_gpu_parity.py (the arithmetic is in fact identical)
XTOL = {4: 2e-4, 8: 1e-9}
FTOL = {4: 1e-5, 8: 1e-12}


def _dev(a, torch):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("precision", [8, 4])
def test_two_calls_in_a_row_match_the_handle_path(psim, precision):
    import torch
    from particle_simulation_b200 import native

    # fast cubes and a long time step: the trigger stays off on the first call and fires on the second
    ids, pos, vel = two_cubes(side=5, speed=25.0)
    params = psim.make_params(delta_t=0.004)
    rt = native.rt(precision)
    n = len(ids)

    sim = psim.Simulation(params, ids, pos, vel, precision)
    # caller-owned device state, seeded from the handle's start-up state
    d_pos = _dev(pos.astype(rt), torch)
    d_vel = _dev(vel.astype(rt), torch)
    d_fnew = _dev(sim.forces(), torch)
    d_fold = torch.zeros_like(d_fnew)
    d_disp = torch.zeros_like(d_fnew)
    d_flag = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()

    flags = []
    for call in range(2):
        f_before = d_fnew.cpu().numpy().copy()
        rebuilds_before = sim.counters().rebuilds
        d_flag.fill_(-1)
        native.compute_locations_static(params, precision, n, d_pos.data_ptr(), d_vel.data_ptr(),
                                        d_fold.data_ptr(), d_fnew.data_ptr(), d_disp.data_ptr(), None,
                                        d_flag.data_ptr(), stream)
        torch.cuda.synchronize()
        sim.advance(1)
        _, x_ref, v_ref, act = sim.frame()
        disp_ref, fold_ref, fnew_ref = sim.aux()
        fired = sim.counters().rebuilds - rebuilds_before
        flag = int(d_flag.item())
        flags.append(flag)

        assert act.all()
        assert flag == fired, "rebuild verdict differs on call %d" % (call + 1)
        assert np.abs(d_pos.cpu().numpy() - x_ref).max() <= XTOL[precision]
        scale = max(np.abs(fold_ref).max(), 1e-30)
        assert np.abs(d_fold.cpu().numpy() - fold_ref).max() <= FTOL[precision] * scale
        assert np.abs(d_fold.cpu().numpy() - f_before).max() == 0      # force_old <- force_new
        assert np.abs(d_fnew.cpu().numpy()).max() == 0                  # force_new reset
        if not fired:   # a rebuild zeroes the handle's displacement; the entry point only reports the verdict
            assert np.abs(d_disp.cpu().numpy() - disp_ref).max() <= XTOL[precision]
        # hand the next call the state the time loop would have: new forces and kicked velocities
        d_fnew.copy_(_dev(fnew_ref, torch))
        d_vel.copy_(_dev(v_ref, torch))
        if fired:
            d_disp.zero_()
    assert flags == [0, 1], "test lost its point: expected the trigger off, then on (got %r)" % (flags,)
    sim.close()


def test_sizes_beyond_the_fixed_scratch_are_refused(psim):
    import torch
    from particle_simulation_b200 import native

    params = psim.make_params()
    small = torch.zeros(3, dtype=torch.float64, device="cuda")
    flag = torch.full((1,), 7, dtype=torch.int32, device="cuda")
    with pytest.raises(psim.PsimError) as err:
        native.compute_locations_static(params, 8, native.STATIC_MAX_PARTICLES + 1, small.data_ptr(), small.data_ptr(),
                                        small.data_ptr(), small.data_ptr(), small.data_ptr(), None,
                                        flag.data_ptr(), None)
    assert err.value.status == 1          # PSIM_ERR_INVALID_ARGUMENT, and nothing ran:
    torch.cuda.synchronize()
    assert int(flag.item()) == 7
