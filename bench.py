#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--particles N] [--precision 8|4]
                  [--algo allpairs|cells] [--impl reference]

One "step" is one time step of the reference loop
(program_particle_simulation_openmp.f08:80-117): drift, rebuild side effects,
force pass, kick — over one synthetic lattice block generated with the
reference generator's semantics (particle_simulation_b200/generate.py).

Workload: N = 262,144 (BASELINE.json configs[3]) on 1, 2 and 4 GPUs; N = 1,048,576
(configs[4], box 198 x 198 x 99) on 8 GPUs; --particles overrides.

metric  = pair-interactions/s = N (N-1) ordered pair tests x steps / time,
          whole job over all ranks; timesteps/s is reported
          beside it.
value   : state resident in HBM, K steps in one psim_advance (fused kernels),
          CUDA events on the library's stream, max over ranks.  On N > 1 GPUs the
          positions of a step reach the other ranks as peer-memory stores from
          inside the force kernel (config.parallelism says how many of the timed
          steps went that way; PSIM_GATHER=nccl keeps the NCCL all-gather).
e2e     : the force operator through the C ABI with HOST buffers: every step
          uploads all positions from pinned memory (psim_set_positions),
          evaluates the forces and reads them back (psim_get_aux).
roofline: the force kernel against the unit that bounds it (tensor pipe for the
          default all-pairs kernel, FMA pipes otherwise), per-launch duration
          from CUDA events on the library's stream.
parity  : after the timed regions, every rank compares its slice of the forces at
          the initial positions, and the state after K steps, with the CPU oracle
          (the reference's algorithm, single thread).  A failed comparison makes
          the process exit non-zero.
cell_binned: the same job, measured the same way, through the cell-binned path
          (SURVEY.md 8(f)-1), plus a second workload with a drift velocity so that
          the reference's rebuild trigger fires and the re-binning is timed.
cpu_baseline / --impl reference: the CPU oracle (a port of the reference's
          cell-list algorithm; the Fortran original cannot be compiled here)
          on the host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_STRONG = 262144        # BASELINE.json configs[3]: the configuration listed for 1/2/4 GPUs
N_EIGHT = 1048576        # BASELINE.json configs[4]: 8 x B200
FTOL = {4: 1e-5, 8: 1e-12}      # one force evaluation against the oracle, relative to max|f|
XTOL = {4: 2e-4, 8: 1e-9}       # positions after K steps, absolute


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--particles", "--n", dest="n", type=int, default=0,
                    help="particles (65536, 262144, 1048576 are SURVEY.md 8(d) shapes; default 262144, or 1048576 "
                         "on 8 GPUs); under torchrun spell it --particles (torchrun's own parser trips over --n)")
    ap.add_argument("--precision", type=int, default=8, choices=(4, 8))
    ap.add_argument("--algo", default="allpairs", choices=("allpairs", "cells"))
    ap.add_argument("--impl", default="b200", choices=("b200", "reference"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cells", action="store_true", help="skip the cell-binned companion measurement")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison (profiling runs)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--exact-diff", action="store_true", help="difference-form distance test (no pre-filter)")
    ap.add_argument("--cuda-core", "--ffma2", dest="cuda_core", action="store_true",
                    help="all-pairs on the CUDA cores only (packed-FP32 filter kernel) instead of the default "
                         "tensor-core filter kernel")
    args = ap.parse_args()
    if args.n <= 0:
        args.n = N_EIGHT if args.gpus >= 8 else N_STRONG
    return args


def workload(n, drift=(0.0, 0.0, 0.0)):
    from particle_simulation_b200 import generate as G
    ids, pos, vel, area = G.synthetic(n)
    if any(drift):
        vel = vel + np.asarray(drift, dtype=np.float64)[None, :]
    return ids, pos, vel, area


def workload_name(n, area, drift=None):
    s = ("synthetic lattice block N=%d (write_particle_file.py semantics, seed %d, vt=0.1), box %gx%gx%g, "
         "shipped environment_params.nml otherwise" % (n, 12345, area[0], area[1], area[2]))
    if drift is not None and any(drift):
        s += ", drift velocity (%g, %g, %g)" % tuple(drift)
    return s


# --------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            pass
        sm, mx, reasons = [], [], set()
        try:
            for ln in open(self.path):
                f = [x.strip() for x in ln.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


# --------------------------------------------------------------------------
def _oracle_lib(variant):
    """Compile the oracle on the box that times it (-march=native)."""
    from oracle import oracle_py as O
    build_dir = os.path.join(tempfile.gettempdir(), "psim_oracle_%d" % os.getuid())
    return O.build(variant, out_dir=build_dir)


def cpu_reference_run(n, precision, seconds, threads=None, variant="fast", drift=(0.0, 0.0, 0.0)):
    """Time the oracle port (reference algorithm: cells + half lists + Newton's
    third law) on the host cores over a bounded number of steps."""
    from oracle import oracle_py as O
    threads = threads or (os.cpu_count() or 1)
    lib_path = _oracle_lib(variant)
    ids, pos, vel, area = workload(n, drift)
    p = O.make_params(length_of_area=area)
    o = O.Oracle(p, precision, ids, pos, vel, threads=threads, lib_path=lib_path)
    entries = o.list_entries
    o.advance(1)                                        # warm caches / first-touch
    t0 = o.loop_seconds
    steps = 0
    chunk = 1
    wall0 = time.time()
    while True:
        o.advance(chunk)
        steps += chunk
        el = o.loop_seconds - t0
        if el >= seconds or time.time() - wall0 > 3 * seconds + 30:
            break
        per = el / steps
        chunk = int(max(1, min(200, (seconds - el) / max(per, 1e-9) / 2)))
    el = o.loop_seconds - t0
    res = dict(steps=steps, seconds=el, timesteps_per_s=steps / el, list_entries=entries,
               list_pair_visits_per_s=entries * steps / el, rebuilds=o.rebuilds, threads=threads)
    o.close()
    return res


def reference_arm(args, rank, world):
    """--impl reference: the reference's own algorithm on the host cores
    (oracle port; the Fortran build is unavailable on this box)."""
    if rank != 0:
        return
    n = args.n
    ids, pos, vel, area = workload(n)
    from oracle import oracle_py as O
    threads = os.cpu_count() or 1
    o = O.Oracle(O.make_params(length_of_area=area), args.precision, ids, pos, vel, threads=threads,
                 lib_path=_oracle_lib("fast"))
    o.advance(args.warmup)
    t0 = o.loop_seconds
    o.advance(args.steps)
    el = o.loop_seconds - t0
    pairs = float(n) * (n - 1)
    value = pairs * args.steps / el
    sample = "%d steps of the full N=%d workload, time loop only (list build at start-up excluded)" % (args.steps, n)
    line = {
        "impl": "reference", "metric": "pair_interactions_per_s", "value": value, "unit": "pair-interactions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": el / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64" if args.precision == 8 else "f32", "data": "synthetic",
        "timesteps_per_s": args.steps / el,
        "config": {"workload": workload_name(n, area), "n_particles": n,
                   "note": "CPU path visits O(N) half-list pairs per step; value counts the N(N-1) ordered pairs "
                           "whose interactions one step resolves, the unit of the GPU arm"},
        "cpu_baseline": {"value": value, "unit": "pair-interactions/s", "cores": threads, "kind": "port",
                         "sample": sample, "list_pair_visits_per_s": o.list_entries * args.steps / el,
                         "rebuilds": o.rebuilds},
        "e2e": {"value": value, "unit": "pair-interactions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    o.close()
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------
class OracleAnswers:
    """What the CPU oracle (reference algorithm, single thread = the reference's summation order) says about
    one workload: forces at the initial positions and the state after k steps.  Computed once per workload
    and compared with every GPU path."""

    def __init__(self, ids, pos, vel, area, precision, k):
        from oracle import oracle_py as O
        t0 = time.time()
        o = O.Oracle(O.make_params(length_of_area=area), precision, ids, pos, vel, threads=1)
        self.f0 = o.aux()[2].copy()
        o.advance(k)
        _, x, v, a = o.frame()
        self.x, self.v, self.a = x.copy(), v.copy(), a.copy()
        self.fk = o.aux()[2].copy()
        self.rebuilds = int(o.rebuilds)
        self.k = k
        o.close()
        self.seconds = time.time() - t0


# --------------------------------------------------------------------------
class Ctx:
    """Process-wide plumbing of the GPU arm: rank layout, barriers, reductions over ranks."""

    def __init__(self, args):
        import torch
        self.torch = torch
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
        torch.cuda.set_device(self.local_rank)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _reduce(self, x, op):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, x):
        return self._reduce(x, self.dist.ReduceOp.MAX) if self.dist else x

    def min(self, x):
        return self._reduce(x, self.dist.ReduceOp.MIN) if self.dist else x

    def sum(self, x):
        return self._reduce(x, self.dist.ReduceOp.SUM) if self.dist else x

    def fresh_uid(self):
        """One ncclUniqueId per sharded handle, made on rank 0 and broadcast."""
        if self.world == 1:
            return None
        from particle_simulation_b200 import native
        torch = self.torch
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if self.rank == 0:
            buf.copy_(torch.frombuffer(bytearray(native.comm_unique_id()), dtype=torch.uint8))
        self.dist.broadcast(buf, 0)
        return bytes(buf.cpu().numpy().tobytes())

    def simulation(self, params, ids, pos, vel, prec, kernel):
        """A handle on this rank's GPU with the force path `kernel` selected:
        'allpairs' (tensor-core default), 'cuda_core', 'exact_diff' or 'cells'."""
        import particle_simulation_b200 as P
        from particle_simulation_b200 import native
        sim = P.Simulation(params, ids, pos, vel, prec, device=self.local_rank, rank=self.rank, world=self.world,
                           nccl_unique_id=self.fresh_uid())
        if kernel == "cells":
            sim.set_option(P.OPT_FORCE_ALGO, P.FORCE_CELL_BINNED)
        elif kernel == "cuda_core":
            sim.set_option(P.OPT_ALLPAIRS_KERNEL, P.ALLPAIRS_CUDA_CORE)
        elif kernel == "exact_diff":
            sim.set_option(P.OPT_ALLPAIRS_KERNEL, P.ALLPAIRS_DIFFERENCE_FORM)
        if kernel != "allpairs":
            sim.set_positions(pos.astype(native.rt(prec)))     # initial forces through the selected path
        return sim


def parity_check(ctx, params, ids, pos, vel, prec, kernel, ans):
    """Forces at the initial positions and the state after ans.k steps against the oracle, on this rank's
    slice (positions and flags: all particles, every rank holds them).  Returns the parity object."""
    sim = ctx.simulation(params, ids, pos, vel, prec, kernel)
    lo, hi = sim.i_lo, sim.i_hi
    f = sim.forces()
    scale = float(np.abs(ans.f0).max())
    f_err = float(np.abs(f[lo:hi] - ans.f0[lo:hi]).max()) / scale if hi > lo else 0.0
    sim.advance(ans.k)
    _, x, v, a = sim.frame()
    fk = sim.forces()
    c = sim.counters()
    flags_equal = bool(np.array_equal(a, ans.a))
    keep = ans.a.astype(bool)
    x_err = float(np.abs(x[keep] - ans.x[keep]).max()) if keep.any() else 0.0
    own = np.zeros(len(ids), bool); own[lo:hi] = True
    sel = keep & own
    v_err = float(np.abs(v[sel] - ans.v[sel]).max()) if sel.any() else 0.0
    fk_scale = float(np.abs(ans.fk).max())
    fk_err = float(np.abs(fk[sel] - ans.fk[sel]).max()) / fk_scale if sel.any() else 0.0
    rebuilds = int(c.rebuilds)
    sim.close()
    ok = (f_err <= FTOL[prec] and flags_equal and rebuilds == ans.rebuilds and x_err <= XTOL[prec]
          and v_err <= 100 * XTOL[prec] and fk_err <= 1e3 * FTOL[prec])
    # over ranks: worst error, and everybody must agree
    return {
        "against": "CPU oracle (reference algorithm, single thread), same inputs", "n": int(len(ids)),
        "ranks": ctx.world, "k_steps": ans.k,
        "force_max_rel": ctx.max(f_err), "force_tol": FTOL[prec],
        "x_max_abs": ctx.max(x_err), "x_tol": XTOL[prec], "v_max_abs": ctx.max(v_err),
        "force_after_k_max_rel": ctx.max(fk_err),
        "active_flags_equal": bool(ctx.min(1.0 if flags_equal else 0.0) == 1.0),
        "rebuilds": [rebuilds, ans.rebuilds],
        "ok": bool(ctx.min(1.0 if ok else 0.0) == 1.0),
    }


def tensor_peak(ctx, f16=True):
    """Dense tcgen05.mma rate of this GPU for the kind the kernel issues (kind::tf32, or kind::f16), measured
    live by the library (one issuing thread per SM, M128 N256 back to back); falls back to MEASURED_PEAKS.json's
    cuBLAS bf16 burst figure (/ 2 for kind::tf32)."""
    from particle_simulation_b200 import native
    kind = "kind::f16 M128 N256 K16" if f16 else "kind::tf32 M128 N256 K8"
    scale = 1.0 if f16 else 0.5
    try:
        tf, mhz = native.measure_tensor_peak(1 if f16 else 0, ctx.local_rank)
        if tf > 0:
            return tf, "psim_measure_tensor_peak: dense tcgen05.mma %s on all SMs, this run, %.0f MHz (of measured)" % (kind, mhz)
    except Exception:
        pass
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return scale * json.load(f)["bf16_tflops"], "MEASURED_PEAKS.json bf16_tflops (burst) x %g for %s" % (scale, kind)
    except Exception:
        return scale * 1590.0, "fallback of B200_PROFILING.md (1590 TFLOP/s dense bf16) x %g for %s" % (scale, kind)


def recorded_traffic(kernel, n, prec, world):
    """dram__bytes_read.sum + dram__bytes_write.sum of the force kernel from the ncu capture of this very
    configuration (profiles/roofline_traffic.json, keyed by kernel, N, precision and GPU count); null when no
    capture of this configuration is on file — bench.py cannot run under ncu itself."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get("%s_n%d_f%d_g%d" % (kernel, n, prec * 8, world))
    except Exception:
        return None


def measure_path(ctx, params, ids, pos, vel, prec, kernel, steps, warmup, sample_clocks=False):
    """value, roofline and e2e of one force path on one workload (all ranks take part)."""
    import particle_simulation_b200 as P
    from particle_simulation_b200 import native
    torch = ctx.torch
    n = len(ids)
    pairs_total = float(n) * (n - 1)
    sim = ctx.simulation(params, ids, pos, vel, prec, kernel)
    n_own = sim.i_hi - sim.i_lo

    # ---- value: device-resident stepping loop ----
    sim.advance(warmup)
    c0 = sim.counters()
    sampler = ClockSampler(ctx.local_rank)
    if sample_clocks and ctx.rank == 0:
        sampler.start()
    ctx.barrier()
    ms = sim.advance_timed(steps)
    ctx.barrier()
    clocks = sampler.stop() if (sample_clocks and ctx.rank == 0) else None
    c1 = sim.counters()
    ms = ctx.max(ms)
    launches = int(ctx.sum(c1.kernel_launches - c0.kernel_launches - 1))   # -1: the counters' own kernel
    rebuilds = int(c1.rebuilds - c0.rebuilds)
    out = {"value": pairs_total * steps / (ms * 1e-3), "unit": "pair-interactions/s", "ms_per_step": ms / steps,
           "timesteps_per_s": steps / (ms * 1e-3), "steps": steps, "gpu_launches": launches,
           "rebuilds_in_timed_steps": rebuilds, "clocks": clocks,
           "peer_push_steps": int(c1.peer_push_steps - c0.peer_push_steps)}

    # ---- roofline: per-launch duration of the force kernel, events on the library's stream ----
    pin_own = sim.pairs_in_range()
    sim.set_option(P.OPT_KERNEL_TIMING, 1)
    k0 = sim.counters()
    sim.advance(max(3, min(steps, 10)))
    k1 = sim.counters()
    pin_now = sim.pairs_in_range()          # exact O(N^2) count at the positions of the last timed force launch
    sim.set_option(P.OPT_KERNEL_TIMING, 0)
    n_timed = k1.force_timed - k0.force_timed
    kernel_ms = (k1.force_ms - k0.force_ms) / max(1, n_timed)
    sec = kernel_ms * 1e-3
    tests_own = float(k1.filter_tests) if k1.filter_tests > 0 else float(n_own) * (n - 1)
    # 8 flops per ordered pair tested + 15 per ordered pair in range
    flops_survey = 8.0 * (float(n_own) * (n - 1) if kernel != "cells" else tests_own) + 15.0 * pin_own
    survey_tflops = flops_survey / sec / 1e12 if sec > 0 else 0.0
    peak64, _ = native.measure_fma_peak(8, ctx.local_rank)
    peak32, _ = native.measure_fma_peak(4, ctx.local_rank)
    bytes_launch = n_own * (4 * prec) * 6.0 + n * 4 * prec    # owned records r/w + one pass over all positions
    common = {
        "kernel_ms": kernel_ms, "launches_timed": int(n_timed),
        "share_of_step": kernel_ms / (ms / steps) if ms > 0 else None,
        "pairs_in_range_per_step": int(ctx.sum(pin_own)),
        "hbm_bytes_per_launch_algorithmic": bytes_launch,
        "hbm_gbs_algorithmic": bytes_launch / sec / 1e9 if sec > 0 else None,
        "fp64_fma_peak": peak64, "fp32_fma_peak": peak32,
        "traffic": recorded_traffic(kernel, n, prec, ctx.world),
        "traffic_note": "dram bytes per launch from the ncu capture of this configuration under profiles/ (null: none on file)",
    }
    if kernel == "allpairs":
        f16 = True                          # the kernel issues kind::f16 MMAs
        peak, peak_src = tensor_peak(ctx, f16)
        n_pad_cols = ((n + 1024 * ctx.world - 1) // (1024 * ctx.world)) * 1024 * ctx.world
        tensor_flops = 32.0 * n_own * n_pad_cols          # 2 * K flops per accumulator entry, K = 16
        achieved = tensor_flops / sec / 1e12 if sec > 0 else 0.0
        roofline = dict(common, **{
            "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
            "frac": achieved / peak if peak else None, "peak_source": peak_src,
            "kernel": "tc::allpairs_tc_kernel", "flops_per_launch": tensor_flops,
            "flops_accounting": "tensor pipe: 2*K = 32 flops per ordered pair tested (split-%s Gram matrix, K = 16, "
                                "of which %d columns carry data); SURVEY.md 8(d)'s CUDA-core "
                                "accounting (8 per pair tested + 15 per pair in range) is reported as survey_tflops"
                                % (("FP16", 15) if f16 else ("TF32", 12)),
            "survey_flops_per_launch": flops_survey, "survey_tflops": survey_tflops,
            "survey_frac_of_fp32_fma_peak": survey_tflops / peak32 if peak32 else None,
            "survey_frac_of_fp64_fma_peak": survey_tflops / peak64 if peak64 else None,
            "filter_pairs_last_launch": int(ctx.sum(k1.filter_pairs)) if k1.filter_pairs >= 0 else None,
            "pairs_in_range_last_launch": int(ctx.sum(pin_now)),    # equal: the filter dropped no pair
            "note": "north_star asks for the fraction of the FP64 FMA peak: survey_frac_of_fp64_fma_peak; it exceeds 1 "
                    "because the pair tests run on the tensor pipe, not on the FP64 pipe (DESIGN.md 3.4)",
        })
    else:
        fp32_bound = (prec == 4) or kernel in ("cuda_core", "cells")     # both filters run in packed FP32
        peak = peak32 if fp32_bound else peak64
        roofline = dict(common, **{
            "bound": "fp32_fma" if fp32_bound else "fp64_fma", "achieved": survey_tflops, "peak": peak, "unit": "TFLOP/s",
            "frac": survey_tflops / peak if peak else None,
            "peak_source": "psim_measure_fma_peak: dependent-free FMA loop on all SMs, this run (of measured; "
                           "MEASURED_PEAKS.json has no vector-FMA figure)",
            "kernel": {"cuda_core": "allpairs_warp_kernel", "exact_diff": "allpairs_step_kernel",
                       "cells": "cell_force_kernel"}[kernel],
            "flops_per_launch": flops_survey,
            "flops_accounting": "SURVEY.md 8(d): 8 per ordered pair tested + 15 per ordered pair in range"
                                + ("; pairs tested = the distance tests the kernel executed over the 27-cell "
                                   "neighbourhoods (psim_counters.filter_tests)" if kernel == "cells" else ""),
            "pair_tests_per_launch": tests_own,
            "frac_of_fp64_fma_peak": survey_tflops / peak64 if peak64 else None,
        })
    out["roofline"] = roofline

    # ---- e2e: the force operator with host buffers through the C ABI ----
    rt = native.rt(prec)
    host_pos = torch.empty((n, 3), dtype=torch.float32 if prec == 4 else torch.float64).pin_memory()
    host_pos.numpy()[:] = pos.astype(rt)
    host_f = torch.empty((n, 3), dtype=host_pos.dtype).pin_memory()
    hp, hf = host_pos.numpy(), host_f.numpy()
    e2e_steps = max(3, min(steps, 10))
    for _ in range(2):
        sim.set_positions(hp); sim.forces(out=hf)
    ctx.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        sim.set_positions(hp)
        sim.forces(out=hf)
    ctx.barrier()
    e2e_s = ctx.max(time.perf_counter() - t0)
    out["e2e"] = {"value": pairs_total * e2e_steps / e2e_s, "unit": "pair-interactions/s",
                  "h2d_bytes_per_step": int(n * 3 * prec), "d2h_bytes_per_step": int(n_own * 3 * prec),
                  "ms_per_step": e2e_s / e2e_steps * 1e3, "steps": e2e_steps,
                  "call": "psim_set_positions (pinned H2D of all positions + force pass) + psim_get_aux (D2H of the owned forces)"}
    sim.close()
    return out


# --------------------------------------------------------------------------
def main():
    args = parse_args()
    if args.impl == "reference":
        reference_arm(args, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return

    import particle_simulation_b200 as P

    ctx = Ctx(args)
    rank, world = ctx.rank, ctx.world
    prec, n = args.precision, args.n
    ids, pos, vel, area = workload(n)
    params = P.make_params(length_of_area=area)
    kernel = "cells" if args.algo == "cells" else "exact_diff" if args.exact_diff else "cuda_core" if args.cuda_core else "allpairs"

    main_m = measure_path(ctx, params, ids, pos, vel, prec, kernel, args.steps, args.warmup, sample_clocks=True)

    k_par = max(1, min(args.steps, 20))
    ans = None
    failed = []
    if not args.no_parity:
        ans = OracleAnswers(ids, pos, vel, area, prec, k_par)
        main_m["parity"] = parity_check(ctx, params, ids, pos, vel, prec, kernel, ans)
        main_m["parity"]["oracle_seconds"] = ans.seconds
        if not main_m["parity"]["ok"]:
            failed.append(kernel)

    line = {
        "metric": "pair_interactions_per_s", "value": main_m["value"], "unit": "pair-interactions/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": main_m["ms_per_step"],
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64" if prec == 8 else "f32", "data": "synthetic",
        "timesteps_per_s": main_m["timesteps_per_s"],
        "config": {"workload": workload_name(n, area), "n_particles": n, "algo": args.algo, "kernel": kernel,
                   "parallelism": ("i-sharded x%d; positions reach the other ranks as peer-memory stores from inside the "
                                   "fused force kernel (NVLink), one flag per rank per step (%d of %d timed steps; the "
                                   "rest: NCCL all-gather)" % (world, main_m["peer_push_steps"], args.steps)
                                   if main_m["peer_push_steps"] > 0 else
                                   "i-sharded x%d, NCCL all-gather of the position shards per step" % world)
                                  if world > 1 else "single GPU",
                   "sizes": "N=262,144 (BASELINE configs[3]) on 1/2/4 GPUs, N=1,048,576 (configs[4]) on 8; "
                            "strong_n262144 carries the fixed-size line on 8 GPUs",
                   "l2": "not flushed between steps: every step reads the positions the previous step wrote "
                         "(state is carried on the device, %.0f MB working set); the force kernel is compute bound "
                         "(see roofline.bound)" % (n * 4 * prec * 6 / 1e6)},
        "roofline": main_m["roofline"], "e2e": main_m["e2e"], "gpu_launches": main_m["gpu_launches"],
        "clocks": main_m["clocks"],
    }
    if "parity" in main_m:
        line["parity"] = main_m["parity"]

    if kernel == "allpairs" and not args.no_cells:
        # The same job through the cell-binned path: what a user who wants timesteps/s
        # selects.  Same measurement: value, roofline, e2e, parity.
        csteps = max(args.steps, 200)
        cm = measure_path(ctx, params, ids, pos, vel, prec, "cells", csteps, 30)
        if ans is not None:
            cm["parity"] = parity_check(ctx, params, ids, pos, vel, prec, "cells", ans)
            if not cm["parity"]["ok"]:
                failed.append("cells")
        cm.pop("clocks", None)
        cm["n_gpus"] = world
        cm["note"] = ("value is the equivalent rate N(N-1) x timesteps/s; the path tests only the 27-cell "
                      "neighbourhood of each particle (roofline.pair_tests_per_launch)")
        # Second workload: a drift velocity, so that the reference's displacement trigger fires (every ~50 steps
        # at |v| dt = 0.005 against dr_max_tolerable = 0.5) and the device-side re-binning is inside the timed steps.
        drift = (5.0, 0.0, 0.0)
        ids_d, pos_d, vel_d, _ = workload(n, drift)
        sim = ctx.simulation(params, ids_d, pos_d, vel_d, prec, "cells")
        sim.advance(30)
        r0 = sim.counters().rebuilds
        ctx.barrier()
        dms = ctx.max(sim.advance_timed(csteps))
        ctx.barrier()
        r1 = sim.counters().rebuilds
        sim.close()
        nreb = int(r1 - r0)
        per_rebuild = (dms - cm["ms_per_step"] * csteps) / nreb if nreb else None
        cm["with_rebuilds"] = {"workload": workload_name(n, area, drift), "steps": csteps, "rebuilds": nreb,
                               "ms_per_step": dms / csteps, "timesteps_per_s": csteps / (dms * 1e-3),
                               "ms_per_rebuild": per_rebuild,
                               "note": "ms_per_rebuild = (time of these steps - time of the same number of steps of "
                                       "the drift-free workload) / rebuilds"}
        if ans is not None:
            # long enough to cross rebuilds (one every ~50 steps); the single-thread oracle sets the length at N = 1M
            ans_d = OracleAnswers(ids_d, pos_d, vel_d, area, prec, 120 if n <= N_STRONG else 60)
            pd = parity_check(ctx, params, ids_d, pos_d, vel_d, prec, "cells", ans_d)
            cm["with_rebuilds"]["parity"] = pd
            if not pd["ok"]:
                failed.append("cells+drift")
        line["cell_binned"] = cm

    if kernel == "allpairs" and world >= 8 and n != N_STRONG:
        # the fixed-size strong-scaling line of configs[3] on this many GPUs, beside the configs[4] headline
        ids_s, pos_s, vel_s, area_s = workload(N_STRONG)
        params_s = P.make_params(length_of_area=area_s)
        sm = measure_path(ctx, params_s, ids_s, pos_s, vel_s, prec, kernel, args.steps, args.warmup)
        if not args.no_parity:
            ans_s = OracleAnswers(ids_s, pos_s, vel_s, area_s, prec, k_par)
            sm["parity"] = parity_check(ctx, params_s, ids_s, pos_s, vel_s, prec, kernel, ans_s)
            if not sm["parity"]["ok"]:
                failed.append("strong_n262144")
        sm.pop("clocks", None)
        sm["workload"] = workload_name(N_STRONG, area_s)
        line["strong_n262144"] = sm

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            pairs_total = float(n) * (n - 1)
            r = cpu_reference_run(n, prec, args.cpu_seconds)
            line["cpu_baseline"] = {
                "value": pairs_total * r["timesteps_per_s"], "unit": "pair-interactions/s", "cores": r["threads"],
                "kind": "port",
                "sample": "%d steps of the same N=%d workload in %.1f s (time loop only)" % (r["steps"], n, r["seconds"]),
                "timesteps_per_s": r["timesteps_per_s"], "list_pair_visits_per_s": r["list_pair_visits_per_s"],
                "rebuilds": r["rebuilds"],
                "note": "oracle port of the reference's O(N) cell-list algorithm; value = N(N-1) x its timesteps/s",
            }
            # the reference builds with -Ofast (Makefile:5 there): the same port with -ffast-math, reported separately
            rf = cpu_reference_run(n, prec, 0.5 * args.cpu_seconds, variant="ofast")
            line["cpu_baseline"]["ofast"] = {"value": pairs_total * rf["timesteps_per_s"],
                                             "timesteps_per_s": rf["timesteps_per_s"], "steps": rf["steps"],
                                             "seconds": rf["seconds"], "flags": "-O3 -march=native -ffast-math -fopenmp"}
            # and with a drift velocity, so that the sample pays list rebuilds like the cell_binned.with_rebuilds leg
            rd = cpu_reference_run(n, prec, 0.5 * args.cpu_seconds, drift=(5.0, 0.0, 0.0))
            line["cpu_baseline"]["with_rebuilds"] = {"timesteps_per_s": rd["timesteps_per_s"], "steps": rd["steps"],
                                                     "rebuilds": rd["rebuilds"], "seconds": rd["seconds"]}
        except Exception as e:   # the baseline must never take the GPU line down
            line["cpu_baseline"] = {"error": repr(e)}
    if ctx.dist is not None:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()
    if rank == 0:
        sys.stderr.flush()
        print(json.dumps(line), flush=True)      # the JSON line is the last thing this job prints
    if failed:
        sys.stderr.write("bench.py: PARITY FAILED for %s\n" % ", ".join(failed))
        sys.exit(3)


if __name__ == "__main__":
    main()
