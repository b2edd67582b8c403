"""The reference's file formats (the drop-in boundary):
namelist in, particle text in, 4-byte-record stream out — Python host
(particle_simulation_b200/formats.py) and C++ host (csrc/host/) must agree byte
for byte."""
import os
import struct
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN

# same keys, order and style as the reference's environment_params.nml
SHIPPED_STYLE = """&ENVIRONMENTPARAMS
 NB_OF_THREADS   = 3,
 SIGMA           = 1.0,
 EPSILON         = 1.0,
 MASS            = 1.0,
 R_CUT           = 2.5,
 LENGTH_OF_AREA  = 99.0, 99.0, 99.0,
 LENGTH_OF_CELL  = 3.0, 3.0, 3.0,
 DELTA_T         = 0.001,
 WRITE_TIME_STEP = 100,
 MAX_TIME_STEP   = 5001
 /
"""


def test_namelist_shipped_style():
    from particle_simulation_b200 import formats
    n = formats.parse_namelist(SHIPPED_STYLE)
    assert n == {"nb_of_threads": 3, "sigma": 1.0, "epsilon": 1.0, "mass": 1.0, "r_cut": 2.5,
                 "length_of_area": [99.0, 99.0, 99.0], "length_of_cell": [3.0, 3.0, 3.0], "delta_t": 0.001,
                 "write_time_step": 100, "max_time_step": 5001}
    p = formats.params_from_namelist(n)
    assert p.r_cut_factor == 2.5 and p.sigma == 1.0 and list(p.length_of_area) == [99.0] * 3 and p.nb_of_threads == 3


def test_namelist_variants():
    from particle_simulation_b200 import formats
    txt = "! a comment\n&environmentparams sigma=1.5d0 epsilon = 2. , length_of_area = 3*30.0\n length_of_cell(1:3)=3.0,4.0,5.0 ! trailing\n delta_t=1e-3 mass=.5 r_cut=2.5 /\n junk after"
    n = formats.parse_namelist(txt)
    assert n["sigma"] == 1.5 and n["epsilon"] == 2.0 and n["length_of_area"] == [30.0] * 3
    assert n["length_of_cell"] == [3.0, 4.0, 5.0] and n["mass"] == 0.5
    with pytest.raises(formats.NamelistError):
        formats.parse_namelist("&ENVIRONMENTPARAMS bogus = 1 /")
    with pytest.raises(formats.NamelistError):
        formats.parse_namelist("&OTHER sigma = 1 /")
    with pytest.raises(formats.NamelistError):
        formats.parse_namelist("&ENVIRONMENTPARAMS sigma = abc /")
    with pytest.raises(formats.NamelistError):
        formats.params_from_namelist(formats.parse_namelist("&ENVIRONMENTPARAMS sigma = 1 /"))


def test_particle_file_round_trip(tmp_path):
    from particle_simulation_b200 import formats
    ids = np.array([1, 2, 1], np.int32)                     # ids need not be unique
    pos = np.array([[10.0, 10.0, 10.0], [11.1225, 10.0, 10.0], [40.0, 40.0, 40.0]])
    vel = np.array([[5.1218, 5.0229, 5.0969], [-4.9, 0.0, 0.1], [0.0, 0.0, 0.0]])
    path = str(tmp_path / "p.dat")
    formats.write_particles(path, ids, pos, vel)
    first = open(path).readline()
    assert first == "1 10.0000 10.0000 10.0000 5.1218 5.0229 5.0969 \n"      # write_particle_file.py:72-75
    i2, p2, v2 = formats.read_particles(path)
    assert np.array_equal(i2, ids) and np.array_equal(p2, pos) and np.array_equal(v2, vel)
    open(path, "w").close()
    with pytest.raises(ValueError, match="No particles"):
        formats.read_particles(path)


def test_output_stream_bytes_hand_checked(tmp_path):
    """3 particles, one inactive, two frames: every record checked by hand against
    program_...:54-56 and :327-365."""
    from particle_simulation_b200 import formats
    path = str(tmp_path / "t.output")
    ids = [7, 8, 9]
    pos = np.array([[1, 2, 3], [4, 5, 6], [7, 8, 9]], np.float32)
    vel = -pos
    with formats.OutputWriter(path, (99.0, 98.0, 97.0)) as w:
        w.write_frame(ids, pos, vel, active=[1, 0, 1])
        w.write_frame(ids, pos + 1, vel, active=[1, 1, 1])
    raw = open(path, "rb").read()
    f = lambda off: struct.unpack_from("<f", raw, 4 * off)[0]
    i = lambda off: struct.unpack_from("<i", raw, 4 * off)[0]
    assert (f(0), f(1), f(2)) == (99.0, 98.0, 97.0)          # records 1-3
    assert i(3) == 2                                          # record 4: frames written, rewritten per frame
    assert i(4) == 14                                         # frame 1: 7 * n_active
    assert i(5) == 7 and [f(6), f(7), f(8)] == [1, 2, 3] and [f(9), f(10), f(11)] == [-1, -2, -3]
    assert i(12) == 9 and f(13) == 7.0                        # particle 8 (inactive) skipped
    assert i(19) == 21 and i(20) == 7 and f(21) == 2.0        # frame 2 follows contiguously
    assert len(raw) == 4 * (4 + 1 + 14 + 1 + 21)
    area, declared, frames = formats.read_output(path)
    assert area == [99.0, 98.0, 97.0] and declared == 2 and len(frames) == 2
    assert list(frames[0][0]) == [7, 9] and np.array_equal(frames[1][1], pos + 1)


def test_output_stream_8_byte_records(tmp_path):
    from particle_simulation_b200 import formats
    path = str(tmp_path / "t8.output")
    pos = np.array([[1.5, 2.5, 3.5]]); vel = np.array([[0.1, 0.2, 0.3]])
    with formats.OutputWriter(path, (99.0, 99.0, 99.0), nbytes=8) as w:
        w.write_frame([5], pos, vel)
    raw = open(path, "rb").read()
    assert len(raw) == 8 * (4 + 1 + 7)
    assert struct.unpack_from("<q", raw, 8 * 3)[0] == 1 and struct.unpack_from("<q", raw, 8 * 4)[0] == 7
    area, declared, frames = formats.read_output(path, nbytes=8)       # play_simulation.py:50-52 branch
    assert declared == 1 and frames[0][0][0] == 5 and np.array_equal(frames[0][1], pos)


@pytest.mark.parametrize("precision", [4, 8])
def test_cpp_host_io_agrees_with_python(tmp_path, precision):
    """simu_b200 --io-selftest parses both inputs with the C++ readers and writes the
    initial state as one frame (no GPU needed): must equal the Python writer's bytes."""
    from particle_simulation_b200 import build, formats
    simu = build.build_host() or build.SIMU
    if not os.path.exists(build.LIB):
        build.build_library()
        simu = build.build_host()
    nml, dat = os.path.join(GOLDEN, "two_cubes.nml"), os.path.join(GOLDEN, "two_cubes.dat")
    out_cpp = str(tmp_path / "cpp.output")
    r = subprocess.run([simu, nml, dat, out_cpp, "--io-selftest", "--quiet", "--precision", str(precision)],
                       capture_output=True, text=True, check=True)
    n = formats.read_namelist(nml)
    fields = r.stdout.split()
    assert [int(x) for x in fields[:3]] == [n["nb_of_threads"], n["write_time_step"], n["max_time_step"]]
    assert float(fields[3]) == n["delta_t"] and float(fields[4]) == n["r_cut"]
    ids, pos, vel = formats.read_particles(dat)
    out_py = str(tmp_path / "py.output")
    with formats.OutputWriter(out_py, n["length_of_area"], nbytes=precision) as w:
        w.write_frame(ids, pos, vel)
    assert open(out_cpp, "rb").read() == open(out_py, "rb").read()


REFERENCE = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REFERENCE, "particle_files")),
                    reason="the reference checkout only exists in the build container")
@pytest.mark.parametrize("name,count", [("test_particles.dat", 2000), ("more_test_particles.dat", 39440)])
def test_both_hosts_read_the_reference_s_own_files(tmp_path, name, count):
    """BASELINE.json configs[0..1]: the shipped environment_params.nml and particle files go through the C++
    readers (simu_b200 --io-selftest) and the Python readers to the same .output bytes, in the reference's
    4-byte-record layout (program_...:54-56, 327-365).  particle_files/big_body_test.dat is not in the checkout
    (.MISSING_LARGE_BLOBS).  Nothing is copied: the files are read where they lie."""
    from particle_simulation_b200 import build, formats
    simu = build.build_host() or build.SIMU
    nml, dat = os.path.join(REFERENCE, "environment_params.nml"), os.path.join(REFERENCE, "particle_files", name)
    n = formats.read_namelist(nml)
    assert (n["nb_of_threads"], n["write_time_step"], n["max_time_step"]) == (3, 100, 5001)
    assert n["length_of_area"] == [99.0, 99.0, 99.0] and n["length_of_cell"] == [3.0, 3.0, 3.0]
    assert (n["sigma"], n["epsilon"], n["mass"], n["r_cut"], n["delta_t"]) == (1.0, 1.0, 1.0, 2.5, 0.001)
    ids, pos, vel = formats.read_particles(dat)
    assert len(ids) == count and pos.shape == vel.shape == (count, 3)
    out_cpp, out_py = str(tmp_path / "cpp.output"), str(tmp_path / "py.output")
    r = subprocess.run([simu, nml, dat, out_cpp, "--io-selftest", "--precision", "4"], capture_output=True, text=True, check=True)
    assert "Number of particles in file: %d" % count in r.stdout
    with formats.OutputWriter(out_py, n["length_of_area"], nbytes=4) as w:
        w.write_frame(ids, pos, vel)
    data = open(out_cpp, "rb").read()
    assert data == open(out_py, "rb").read()
    assert len(data) == 4 * (4 + 1 + 7 * count)                     # area[3], frame count, then 7 n + 1 records
    area, declared, frames = formats.read_output(out_cpp, nbytes=4)
    assert area == [99.0, 99.0, 99.0] and declared == 1 and np.array_equal(frames[0][0], ids)
    assert np.array_equal(frames[0][1], pos.astype(np.float32)) and np.array_equal(frames[0][2], vel.astype(np.float32))


def test_cpp_host_reference_error_behaviour(tmp_path):
    """Wrong argument count / missing file / empty particle file: message on stderr and
    exit status 0, as stop_program's bare `stop` does (aux:24-49)."""
    from particle_simulation_b200 import build
    simu = build.build_host() or build.SIMU
    r = subprocess.run([simu, "a", "b"], capture_output=True, text=True)
    assert r.returncode == 0 and "Not enough arguments!" in r.stderr and "Usage is:" in r.stderr
    r = subprocess.run([simu, "a", "b", "c", "d"], capture_output=True, text=True)
    assert r.returncode == 0 and "To many arguments!" in r.stderr
    r = subprocess.run([simu, "/nonexistent.nml", "/nonexistent.dat", str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode == 0 and "does not exist" in r.stderr
    empty = tmp_path / "empty.dat"; empty.write_text("")
    r = subprocess.run([simu, os.path.join(GOLDEN, "two_cubes.nml"), str(empty), str(tmp_path / "o"), "--io-selftest"],
                       capture_output=True, text=True)
    assert r.returncode == 0 and "No particles in particle file" in r.stderr


def test_generator_matches_reference_script_semantics():
    from particle_simulation_b200 import generate as G
    ids, pos, vel = G.lattice_block((3, 2, 2), start_point=(10.0, 10.0, 10.0), velocity=(1.0, 0.0, 0.0), thermal=0.0)
    r_min = 2.0 ** (1.0 / 6.0)
    assert list(ids) == list(range(1, 13))
    assert pos[0] == pytest.approx(np.round([10 - 1.5 * r_min, 10 - r_min, 10 - r_min], 4))
    assert pos[1, 0] - pos[0, 0] == pytest.approx(r_min, abs=1e-4)              # x runs fastest
    assert pos[3, 1] - pos[0, 1] == pytest.approx(r_min, abs=1e-4)
    assert np.all(vel == [1.0, 0.0, 0.0])
    a = G.lattice_block((4, 4, 4), thermal=0.1, seed=5)[2]
    b = G.lattice_block((4, 4, 4), thermal=0.1, seed=5)[2]
    assert np.array_equal(a, b) and (a >= 0).all()                               # seeded; noise never negative
    for n in (65536, 262144):
        ids, pos, vel, area = G.synthetic(n)
        assert len(ids) == n and pos.min() > 0 and (pos.max(0) < np.array(area)).all()


def test_cli_tools_round_trip(tmp_path):
    """tools/write_particle_file.py (seeded restatement of the reference generator) and
    tools/check_output.py (headless play_simulation.py reader)."""
    import sys
    from conftest import ROOT
    from particle_simulation_b200 import formats
    dat = str(tmp_path / "b.dat")
    cmd = [sys.executable, os.path.join(ROOT, "tools", "write_particle_file.py"), "-o", dat, "-n", "3", "2", "2",
           "-s", "10", "10", "10", "-vt", "0.1", "--seed", "7"]
    subprocess.run(cmd, check=True, capture_output=True)
    ids, pos, vel = formats.read_particles(dat)
    assert len(ids) == 12 and (vel >= 0).all()
    assert subprocess.run(cmd, capture_output=True).returncode != 0            # mode 'x': no overwrite
    subprocess.run(cmd + ["-a"], check=True, capture_output=True)              # append restarts the ids
    ids2, _, _ = formats.read_particles(dat)
    assert list(ids2) == list(range(1, 13)) * 2
    out = str(tmp_path / "t.output")
    with formats.OutputWriter(out, (99.0, 99.0, 99.0)) as w:
        w.write_frame(ids, pos, vel)
        w.write_frame(ids, pos + 0.5, vel)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check_output.py"), "-i", out, "--compare", out],
                       capture_output=True, text=True)
    assert r.returncode == 0 and "Number of blocks: 2 (header says 2)" in r.stdout and "max |dx| = 0" in r.stdout


@pytest.mark.parametrize("nbytes", [4, 8])
def test_a_run_without_frames_leaves_three_records(tmp_path, nbytes):
    """The reference writes record 4 (frames so far) inside write_data only (program:336-337): a run whose
    MAX_TIME_STEP never reaches WRITE_TIME_STEP leaves length_of_area and nothing else."""
    from particle_simulation_b200 import formats
    path = str(tmp_path / "empty.output")
    with formats.OutputWriter(path, (99.0, 98.0, 97.0), nbytes=nbytes):
        pass
    assert os.path.getsize(path) == 3 * nbytes
    area, declared, frames = formats.read_output(path, nbytes=nbytes)
    assert area == [99.0, 98.0, 97.0] and declared == 0 and frames == []


BIG_BODY_STANDIN = dict(nb=(40, 40, 40), start=(49.5, 49.5, 49.5), velocity=(0.0, 0.0, 0.0), thermal=0.1, seed=12345)


@pytest.mark.parametrize("precision", [4, 8])
def test_big_body_stand_in_bit_for_bit_through_both_hosts(tmp_path, precision):
    """particle_files/big_body_test.dat is not in the reference checkout (.MISSING_LARGE_BLOBS:1; BASELINE.json
    configs[1]).  Its documented stand-in (one 40 x 40 x 40 body, N = 64,000, made by
    tools/write_particle_file.py -n 40 40 40 -s 49.5 49.5 49.5 -vt 0.1 --seed 12345) goes through the C++ host's
    readers and frame writer and through the Python host's: the .output bytes must be identical."""
    from particle_simulation_b200 import build, formats, generate
    b = BIG_BODY_STANDIN
    ids, pos, vel = generate.lattice_block(b["nb"], b["start"], b["velocity"], b["thermal"], seed=b["seed"])
    assert len(ids) == 64000
    dat = str(tmp_path / "big_body_standin.dat")
    formats.write_particles(dat, ids, pos, vel)
    simu = build.build_host() or build.SIMU
    nml = os.path.join(GOLDEN, "two_cubes.nml")
    out_cpp = str(tmp_path / "cpp.output")
    subprocess.run([simu, nml, dat, out_cpp, "--io-selftest", "--quiet", "--precision", str(precision)],
                   capture_output=True, text=True, check=True)
    ids2, pos2, vel2 = formats.read_particles(dat)
    assert np.array_equal(ids2, ids) and np.array_equal(pos2, pos) and np.array_equal(vel2, vel)
    out_py = str(tmp_path / "py.output")
    with formats.OutputWriter(out_py, formats.read_namelist(nml)["length_of_area"], nbytes=precision) as w:
        w.write_frame(ids2, pos2, vel2)
    assert open(out_cpp, "rb").read() == open(out_py, "rb").read()
    assert os.path.getsize(out_py) == precision * (4 + 1 + 7 * 64000)
