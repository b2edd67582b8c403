"""GPU tests of the boundary: the committed golden vectors, the C++ `simu_b200`
command line end to end, the pinned frame ring, the force operator over host
buffers, and (with two or more GPUs) the sharded path."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, two_cubes

pytestmark = pytest.mark.gpu


def _golden(psim):
    from particle_simulation_b200 import formats
    ids, pos, vel = formats.read_particles(os.path.join(GOLDEN, "two_cubes.dat"))
    nml = formats.read_namelist(os.path.join(GOLDEN, "two_cubes.nml"))
    return ids, pos, vel, nml, formats.params_from_namelist(nml), np.load(os.path.join(GOLDEN, "two_cubes_golden.npz"))


@pytest.mark.parametrize("algo", ["allpairs", "cells"])
def test_cuda_path_matches_committed_golden_vectors(psim, algo):
    """Golden = independent numpy brute force in float64 (tests/golden/make_golden.py)."""
    ids, pos, vel, nml, params, g = _golden(psim)
    sim = psim.Simulation(params, ids, pos, vel, 8)
    sim.set_option(psim.OPT_FORCE_ALGO, psim.FORCE_CELL_BINNED if algo == "cells" else psim.FORCE_ALL_PAIRS)
    sim.set_positions(pos)
    f0 = sim.forces()
    assert np.abs(f0 - g["f0"]).max() <= 1e-12 * np.abs(g["f0"]).max()
    assert sim.pairs_in_range() == int(g["pairs0"])
    sim.advance(int(g["k_steps"]))
    _, x, v, a = sim.frame()
    d, fo, fn = sim.aux()
    assert a.all()
    assert np.abs(x - g["x"]).max() <= 1e-9 and np.abs(v - g["v"]).max() <= 1e-7
    assert np.abs(fn - g["f"]).max() <= 1e-9 * np.abs(g["f"]).max()
    assert np.abs(fo - g["f_old"]).max() <= 1e-9 * np.abs(g["f_old"]).max()
    k, u, mom = sim.energy()
    assert k == pytest.approx(g["e1"][0], rel=1e-10) and u == pytest.approx(g["e1"][1], rel=1e-9)
    assert np.abs(mom - g["e1"][2:]).max() <= 1e-9 * np.abs(g["e1"][2:]).max()
    sim.close()


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("algo", ["cells", "allpairs"])
def test_simu_cli_end_to_end_against_oracle_frames(psim, oracle, tmp_path, precision, algo):
    """`simu_b200 nml dat out` (the reference's command line, README.md:8) writes the frames the
    oracle's trajectory predicts, in the reference's byte layout."""
    from particle_simulation_b200 import build, formats
    ids, pos, vel, nml, params, g = _golden(psim)
    out = str(tmp_path / "run.output")
    r = subprocess.run([build.SIMU, os.path.join(GOLDEN, "two_cubes.nml"), os.path.join(GOLDEN, "two_cubes.dat"), out,
                        "--precision", str(precision), "--algo", algo], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "[INFO] write for time step: 20" in r.stdout and "[INFO] write for time step: 40" in r.stdout
    area, declared, frames = formats.read_output(out, nbytes=precision)
    assert area == nml["length_of_area"] and declared == 2 and len(frames) == 2
    o = oracle.Oracle(oracle.make_params(length_of_area=nml["length_of_area"]), precision, ids, pos, vel)
    tol = 1e-9 if precision == 8 else 2e-4
    for fr in frames:
        o.advance(nml["write_time_step"])
        _, x_ref, v_ref, a_ref = o.frame()
        assert np.array_equal(fr[0], ids[a_ref.astype(bool)])
        assert np.abs(fr[1] - x_ref).max() <= tol and np.abs(fr[2] - v_ref).max() <= 100 * tol
    size = os.path.getsize(out)
    assert size == precision * (4 + 2 * (1 + 7 * len(ids)))


def test_simu_cli_writer_thread_file_is_byte_identical_to_sync_io(psim, tmp_path):
    """Frame pipeline: with a frame every 2 steps (20 frames through 4 pinned slots) the
    writer thread produces exactly the file the single-threaded path writes."""
    from particle_simulation_b200 import build
    nml = tmp_path / "fast.nml"
    text = open(os.path.join(GOLDEN, "two_cubes.nml")).read()
    import re
    text = re.sub(r"(?i)(WRITE_TIME_STEP\s*=\s*)\d+", r"\g<1>2", text)
    nml.write_text(text)
    outs = []
    for extra in ([], ["--sync-io"]):
        out = str(tmp_path / ("run%d.output" % len(outs)))
        r = subprocess.run([build.SIMU, str(nml), os.path.join(GOLDEN, "two_cubes.dat"), out, "--precision", "4", *extra],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        steps = [int(l.rsplit(":", 1)[1]) for l in r.stdout.splitlines() if "write for time step" in l]
        assert steps == sorted(steps) and len(steps) == 20
        outs.append(open(out, "rb").read())
    assert outs[0] == outs[1] and len(outs[0]) > 4 * 4


@pytest.mark.parametrize("precision", [4, 8])
def test_simu_cli_run_without_frames_leaves_three_records(psim, tmp_path, precision):
    """MAX_TIME_STEP < WRITE_TIME_STEP: write_data never runs, so record 4 is never written
    (program:336-337) and the file is length_of_area alone."""
    import re
    from particle_simulation_b200 import build, formats
    text = open(os.path.join(GOLDEN, "two_cubes.nml")).read()
    text = re.sub(r"(?i)(WRITE_TIME_STEP\s*=\s*)\d+", r"\g<1>50", text)
    text = re.sub(r"(?i)(MAX_TIME_STEP\s*=\s*)\d+", r"\g<1>7", text)
    nml = tmp_path / "short.nml"
    nml.write_text(text)
    out = str(tmp_path / "short.output")
    r = subprocess.run([build.SIMU, str(nml), os.path.join(GOLDEN, "two_cubes.dat"), out, "--precision", str(precision)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert os.path.getsize(out) == 3 * precision
    area, declared, frames = formats.read_output(out, nbytes=precision)
    assert declared == 0 and frames == []


def test_frame_ring_matches_synchronous_frames(psim):
    ids, pos, vel = two_cubes()
    sim = psim.Simulation(psim.make_params(), ids, pos, vel, 4)
    sim.set_option(psim.OPT_FORCE_ALGO, psim.FORCE_CELL_BINNED)
    got = []
    for f in range(6):                      # more frames than slots: slots are reused
        sim.advance(10, sync=False)
        slot = f % psim.native.FRAME_SLOTS
        sim.snapshot_async(slot)
        sim.advance(3, sync=False)          # the loop keeps running while the frame drains
        i_, x, v, a = sim.snapshot_wait(slot)
        got.append((x.copy(), v.copy(), a.copy()))
    sim.close()
    ref = psim.Simulation(psim.make_params(), ids, pos, vel, 4)
    ref.set_option(psim.OPT_FORCE_ALGO, psim.FORCE_CELL_BINNED)
    for x, v, a in got:
        ref.advance(10)
        _, xr, vr, ar = ref.frame()
        assert np.array_equal(x, xr) and np.array_equal(v, vr) and np.array_equal(a, ar)   # same device arithmetic
        ref.advance(3)
    ref.close()


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("algo", [0, 1])
def test_force_operator_on_host_buffers(psim, oracle, precision, algo):
    """psim_compute_forces_host = compute_forces (stoermer:87-132) as an operator, with an
    explicit active mask, against the brute-force double evaluation."""
    rng = np.random.Generator(np.random.PCG64(11))
    ids, pos, vel = two_cubes(side=5)
    active = (rng.uniform(size=len(ids)) > 0.15).astype(np.uint8)
    f, pairs = psim.compute_forces(psim.make_params(), pos, precision, active=active, algo=algo)
    bf, pairs_ref = oracle.bruteforce(oracle.make_params(), precision, pos, active)
    tol = 4e-12 if precision == 8 else 4e-5
    assert pairs == pairs_ref
    assert np.abs(f - bf).max() <= tol * np.abs(bf).max()
    assert np.all(f[active == 0] == 0)


def test_edge_cases(psim, oracle):
    p = psim.make_params()
    # a single particle; a lone pair; more particles than one tile with ragged padding
    sim = psim.Simulation(p, [1], [[10.0, 10.0, 10.0]], [[1.0, 0.0, 0.0]], 8)
    sim.advance(10)
    assert sim.frame()[1][0, 0] == pytest.approx(10.01, abs=1e-12) and np.all(sim.forces() == 0)
    sim.close()
    f, pairs = psim.compute_forces(p, [[11.0, 10.0, 10.0], [10.0, 10.0, 10.0]], 8)
    assert pairs == 2 and f[0, 0] == pytest.approx(24.0155998, rel=1e-8) and f[1, 0] == -f[0, 0]
    ids, pos, vel = two_cubes(side=5)
    ids, pos, vel = ids[:131], pos[:131], vel[:131]          # 131 = 128 + 3
    for algo in (0, 1):
        f, _ = psim.compute_forces(p, pos, 8, algo=algo)
        bf, _ = oracle.bruteforce(oracle.make_params(), 8, pos)
        assert np.abs(f - bf).max() <= 4e-12 * np.abs(bf).max()
    # everything outside the box: all inactive, nothing moves
    sim = psim.Simulation(p, [1, 2], [[-5.0, 1.0, 1.0], [200.0, 1.0, 1.0]], np.ones((2, 3)), 4)
    sim.advance(5)
    _, x, _, a = sim.frame()
    assert not a.any() and x[0, 0] == -5.0
    sim.close()
    # cell narrower than the cutoff is rejected loudly, not computed wrongly
    with pytest.raises(psim.PsimError) as e:
        psim.Simulation(psim.make_params(length_of_cell=(2.0, 3.0, 3.0)), [1], [[1.0, 1.0, 1.0]], [[0.0, 0.0, 0.0]], 8)
    assert e.value.status == 5


@pytest.mark.parametrize("gather", ["peer_stores", "nccl"])
def test_sharded_run_matches_single_gpu(gather):
    """Two ranks against the oracle (needs 2 GPUs), with both ways of exchanging the position shards: peer-memory
    stores from inside the fused force kernel (the default between processes of one node) and the NCCL all-gather
    (PSIM_GATHER=nccl; also the fallback)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ)
    if gather == "nccl":
        env["PSIM_GATHER"] = "nccl"
    else:
        env.pop("PSIM_GATHER", None)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533" if gather == "nccl" else "29534",
                        os.path.join(ROOT, "tools", "sharded_check.py")], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "sharded check ok" in r.stdout
    pushed = "57 of 60 steps gathered by peer stores" in r.stdout
    if gather == "nccl":
        assert not pushed, r.stdout[-1500:]     # (the other mode may legitimately fall back when peers cannot be mapped)
