"""psim_compute_locations_static — the drift step for hosts that cannot let a
library allocate GPU memory (a driver whose own memory plan already owns the
device, or a loop that cannot afford cudaMalloc's device-wide synchronisation).
Its reduction scratch is a fixed array inside the library, so this file checks
what the ordinary path gets for free: that the result equals the ordinary
path's (and the CPU oracle's), that a second call, which finds the scratch as
the first one left it, is still right, and that the per-block partial results
and the hand-over between blocks work when thousands of blocks take part.

The entry point works on device arrays the CALLER owns; here they are torch
CUDA tensors, passed as raw addresses through the C ABI.  The references for
comparison: the handle-based path (psim_advance(1) moves positions by exactly
one drift — force and kick do not touch positions —, leaves force_old =
previous force_new, and counts a rebuild when the trigger fired) and
Oracle.advance(1), the CPU restatement of compute_locations
(module_stoermer_verlet_method.f08:136-187).

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
def test_two_calls_in_a_row_match_the_handle_path_and_the_oracle(psim, oracle, precision):
    import torch
    from particle_simulation_b200 import native

    # fast cubes and a long time step: the trigger stays off on the first call and fires on the second
    ids, pos, vel = two_cubes(side=5, speed=25.0)
    params = psim.make_params(delta_t=0.004)
    rt = native.rt(precision)
    n = len(ids)

    sim = psim.Simulation(params, ids, pos, vel, precision)
    orc = oracle.Oracle(oracle.make_params(delta_t=0.004), precision, ids, pos, vel)
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
        orc_rebuilds_before = orc.rebuilds
        d_flag.fill_(-1)
        native.compute_locations_static(params, precision, n, d_pos.data_ptr(), d_vel.data_ptr(),
                                        d_fold.data_ptr(), d_fnew.data_ptr(), d_disp.data_ptr(), None,
                                        d_flag.data_ptr(), stream)
        torch.cuda.synchronize()
        sim.advance(1)
        orc.advance(1)
        _, x_ref, v_ref, act = sim.frame()
        disp_ref, fold_ref, fnew_ref = sim.aux()
        _, x_orc, _, act_orc = orc.frame()
        disp_orc, fold_orc, _ = orc.aux()
        fired = sim.counters().rebuilds - rebuilds_before
        fired_orc = orc.rebuilds - orc_rebuilds_before
        flag = int(d_flag.item())
        flags.append(flag)

        assert act.all() and act_orc.all()
        assert flag == fired == fired_orc, "rebuild verdict differs on call %d" % (call + 1)
        got_x, got_fold, got_disp = d_pos.cpu().numpy(), d_fold.cpu().numpy(), d_disp.cpu().numpy()
        for x_want, fold_want, disp_want, who in ((x_ref, fold_ref, disp_ref, "handle path"), (x_orc, fold_orc, disp_orc, "oracle")):
            assert np.abs(got_x - x_want).max() <= XTOL[precision], who
            scale = max(np.abs(fold_want).max(), 1e-30)
            assert np.abs(got_fold - fold_want).max() <= FTOL[precision] * scale, who
            if not fired:   # a rebuild zeroes the displacement there; the entry point only reports the verdict
                assert np.abs(got_disp - disp_want).max() <= XTOL[precision], who
        assert np.abs(got_fold - f_before).max() == 0                   # force_old <- force_new
        assert np.abs(d_fnew.cpu().numpy()).max() == 0                  # force_new reset
        # hand the next call the state the time loop would have: new forces and kicked velocities
        d_fnew.copy_(_dev(fnew_ref, torch))
        d_vel.copy_(_dev(v_ref, torch))
        if fired:
            d_disp.zero_()
    assert flags == [0, 1], "test lost its point: expected the trigger off, then on (got %r)" % (flags,)
    sim.close()
    orc.close()


def _numpy_drift(rt, pos, vel, fnew, disp, active, delta_t, mass, tol):
    """compute_locations (stoermer:148-185) with one rounding per operation in precision rt."""
    dt, tmf = rt(delta_t), rt(0.5) * rt(delta_t) / rt(mass)
    act = active.astype(bool)
    dx = (dt * (vel + tmf * fnew).astype(rt)).astype(rt)
    x = pos.copy(); d = disp.copy(); fo = np.zeros_like(fnew); fn = fnew.copy()
    x[act] = (pos[act] + dx[act]).astype(rt)
    d[act] = (disp[act] + dx[act]).astype(rt)
    fo[act] = fnew[act]; fn[act] = 0
    sq = (d * d).astype(rt)
    dr = np.sqrt(((sq[:, 0] + sq[:, 1]).astype(rt) + sq[:, 2]).astype(rt)).astype(rt)
    top = np.sort(dr[act])[-2:] if act.sum() >= 2 else np.array([0, 0], rt)
    return x, d, fo, fn, int(rt(top[0] + top[1]) > rt(tol))


@pytest.mark.parametrize("precision", [8, 4])
def test_many_blocks_a_mask_and_a_smaller_second_call(psim, precision):
    """n = 1,048,576 (the stated maximum: 4,096 blocks) with the two largest displacements in different blocks
    and an inactive particle that would otherwise be the largest; then a call with far fewer particles, so
    that stale partial results of the first call sit beyond the second call's grid."""
    import torch
    from particle_simulation_b200 import native

    rt = native.rt(precision)
    rng = np.random.default_rng(5)
    n_big, n_small = native.STATIC_MAX_PARTICLES, 70001
    params = psim.make_params(delta_t=0.001)         # dr_max_tolerable = 3.0 - 2.5 = 0.5
    for n, movers in ((n_big, (17, 900001, 500000)), (n_small, (3, 65000, 40000))):
        pos = rng.uniform(10.0, 80.0, (n, 3)).astype(rt)
        vel = rng.normal(0.0, 1.0, (n, 3)).astype(rt)
        fnew = rng.normal(0.0, 1.0, (n, 3)).astype(rt)
        disp = (rng.uniform(-1.0, 1.0, (n, 3)) * 0.01).astype(rt)
        active = np.ones(n, np.uint8)
        a, b, hidden = movers
        disp[a] = (0.26, 0.0, 0.0); disp[b] = (0.0, 0.25, 0.0)       # 0.26 + 0.25 > 0.5: fires, from two blocks
        disp[hidden] = (3.0, 0.0, 0.0); active[hidden] = 0           # masked out: must not count
        for second_largest, expect in ((0.25, 1), (0.20, 0)):
            disp[b] = (0.0, second_largest, 0.0)
            d_pos, d_vel, d_fnew, d_disp = (_dev(v, torch) for v in (pos, vel, fnew, disp))
            d_fold = torch.zeros_like(d_fnew)
            d_act = _dev(active, torch)
            d_flag = torch.full((1,), -1, dtype=torch.int32, device="cuda")
            native.compute_locations_static(params, precision, n, d_pos.data_ptr(), d_vel.data_ptr(), d_fold.data_ptr(),
                                            d_fnew.data_ptr(), d_disp.data_ptr(), d_act.data_ptr(), d_flag.data_ptr(), None)
            torch.cuda.synchronize()
            x, d, fo, fn, flag = _numpy_drift(rt, pos, vel, fnew, disp, active, 0.001, 1.0, 0.5)
            assert flag == expect
            assert int(d_flag.item()) == flag
            assert np.array_equal(d_pos.cpu().numpy(), x) and np.array_equal(d_disp.cpu().numpy(), d)
            assert np.array_equal(d_fold.cpu().numpy(), fo) and np.array_equal(d_fnew.cpu().numpy(), fn)


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
