#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n PARTICLES] [--precision 8|4]
                  [--algo allpairs|cells] [--impl reference]

One "step" is one time step of the reference loop
(program_particle_simulation_openmp.f08:80-117): drift, rebuild side effects,
all-pairs force pass, kick — over one synthetic lattice block generated with
the reference generator's semantics (particle_simulation_b200/generate.py).

metric  = pair-interactions/s = N (N-1) ordered pair tests x steps / time,
          whole job over all ranks; timesteps/s is reported
          beside it.
value   : state resident in HBM, K steps in one psim_advance (fused kernels),
          CUDA events on the library's stream, max over ranks.
e2e     : the force operator through the C ABI with HOST buffers: every step
          uploads all positions from pinned memory (psim_set_positions),
          evaluates the forces and reads them back (psim_get_aux).
roofline: the force kernel against the FP64 (or FP32) FMA pipe, peak measured
          live by psim_measure_fma_peak (MEASURED_PEAKS.json has no vector-FMA
          figure); flops per launch as fixed in SURVEY.md 8(d).
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
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DEFAULT_N = 262144       # BASELINE.json configs[3]: the configuration listed for 1/2/4 GPUs


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--particles", "--n", dest="n", type=int, default=DEFAULT_N,
                    help="particles (65536, 262144, 1048576 are SURVEY.md 8(d) shapes); under torchrun spell it "
                         "--particles (torchrun's own parser trips over --n)")
    ap.add_argument("--precision", type=int, default=8, choices=(4, 8))
    ap.add_argument("--algo", default="allpairs", choices=("allpairs", "cells"))
    ap.add_argument("--impl", default="b200", choices=("b200", "reference"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cells", action="store_true", help="skip the cell-binned companion measurement")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--exact-diff", action="store_true", help="difference-form distance test (no pre-filter)")
    ap.add_argument("--tc-pipelined", action="store_true", help="EXPERIMENTAL all-pairs kernel (option value 6, "
                    "csrc/psim_allpairs_tc3.cuh): faster, but an intermittent launch failure is unresolved")
    ap.add_argument("--ffma2", action="store_true", help="all-pairs on the CUDA cores only (packed-FP32 filter kernel) "
                    "instead of the default tensor-core filter kernel")
    return ap.parse_args()


def workload(n):
    from particle_simulation_b200 import generate as G
    ids, pos, vel, area = G.synthetic(n)
    return ids, pos, vel, area


def workload_name(n, area):
    return ("synthetic lattice block N=%d (write_particle_file.py semantics, seed %d, vt=0.1), box %gx%gx%g, "
            "shipped environment_params.nml otherwise" % (n, 12345, area[0], area[1], area[2]))


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
def cpu_reference_run(n, precision, seconds, threads=None, variant="fast"):
    """Time the oracle port (reference algorithm: cells + half lists + Newton's
    third law) on the host cores over a bounded number of steps."""
    from oracle import oracle_py as O
    threads = threads or (os.cpu_count() or 1)
    build_dir = os.path.join(tempfile.gettempdir(), "psim_oracle_%d" % os.getuid())
    lib_path = O.build(variant, out_dir=build_dir)      # -march=native: compile on the box that times it
    ids, pos, vel, area = workload(n)
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
    build_dir = os.path.join(tempfile.gettempdir(), "psim_oracle_%d" % os.getuid())
    lib_path = O.build("fast", out_dir=build_dir)
    o = O.Oracle(O.make_params(length_of_area=area), args.precision, ids, pos, vel, threads=threads, lib_path=lib_path)
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
def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import particle_simulation_b200 as P
    from particle_simulation_b200 import native

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    prec = args.precision
    n = args.n
    ids, pos, vel, area = workload(n)
    params = P.make_params(length_of_area=area)

    uid = None
    if world > 1:
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf.copy_(torch.frombuffer(bytearray(native.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        uid = bytes(buf.cpu().numpy().tobytes())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    sim = P.Simulation(params, ids, pos, vel, prec, device=local_rank, rank=rank, world=world, nccl_unique_id=uid)
    if args.algo == "cells":
        sim.set_option(P.OPT_FORCE_ALGO, P.FORCE_CELL_BINNED)
    if args.exact_diff or args.ffma2 or args.tc_pipelined:
        sim.set_option(P.OPT_EXACT_DIFF_FILTER, 1 if args.exact_diff else 5 if args.ffma2 else 6)
        sim.set_positions(pos.astype(native.rt(prec)))
    pairs_total = float(n) * (n - 1)
    n_own = sim.i_hi - sim.i_lo

    # ---- value: device-resident stepping loop ----
    sim.advance(args.warmup)
    c0 = sim.counters()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ms = sim.advance_timed(args.steps)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    c1 = sim.counters()
    ms = max_over_ranks(ms)
    launches = int(sum_over_ranks(c1.kernel_launches - c0.kernel_launches - 1))   # -1: the counters' own kernel
    value = pairs_total * args.steps / (ms * 1e-3)

    # ---- roofline: per-launch duration of the force kernel, events on the library's stream ----
    pin_own = sim.pairs_in_range()
    sim.set_option(P.OPT_KERNEL_TIMING, 1)
    k0 = sim.counters()
    sim.advance(max(3, min(args.steps, 10)))
    k1 = sim.counters()
    pin_now = sim.pairs_in_range()          # exact O(N^2) count at the positions of the last timed force launch
    sim.set_option(P.OPT_KERNEL_TIMING, 0)
    n_timed = k1.force_timed - k0.force_timed
    kernel_ms = (k1.force_ms - k0.force_ms) / max(1, n_timed)
    flops_launch = 8.0 * n_own * (n - 1) + 15.0 * pin_own
    achieved = flops_launch / (kernel_ms * 1e-3) / 1e12 if kernel_ms > 0 else 0.0
    peak64, _ = native.measure_fma_peak(8, local_rank)
    peak32, _ = native.measure_fma_peak(4, local_rank)
    # Which unit bounds the kernel.
    #  * default all-pairs kernel (allpairs_tcws_kernel): the distance test of every ordered pair is one entry of a
    #    Gram matrix computed by tcgen05.mma kind::tf32 on split operands (K = 16: 32 tensor flops per pair); the
    #    CUDA cores only take minima over 32-column chunks and evaluate the few pairs that pass -> bound "tensor".
    #    Peak: MEASURED_PEAKS.json's dense bf16 figure / 2 (kind::tf32 runs at half the bf16 rate).
    #  * --ffma2: the same filter as packed FP32 FMAs for both working precisions -> FP32 FMA pipe.
    #  * --exact-diff and the cell-binned kernel in f64 -> FP64 FMA pipe.
    tensor = args.algo == "allpairs" and not args.exact_diff and not args.ffma2
    fp32_bound = (prec == 4) or (args.algo == "allpairs" and not args.exact_diff)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            traffic = json.load(f).get("%s_n%d_f%d" % (args.algo, n, prec * 8))
    except Exception:
        pass
    bytes_launch = n_own * (4 * prec) * 6.0 + n * 4 * prec    # owned records r/w + one pass over all positions
    survey_tflops = achieved
    if tensor:
        peaks, peak_src = None, "fallback of B200_PROFILING.md (1590 TFLOP/s dense bf16) / 2 for kind::tf32"
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
            peak_src = "MEASURED_PEAKS.json bf16_tflops (burst; the kernel is timed alone) / 2 for kind::tf32"
        except Exception:
            pass
        peak = 0.5 * (peaks["bf16_tflops"] if peaks else 1590.0)
        n_pad_cols = ((n + 1023) // 1024) * 1024
        tensor_flops = 32.0 * n_own * n_pad_cols          # 2 * K flops per accumulator entry, K = 16
        achieved = tensor_flops / (kernel_ms * 1e-3) / 1e12 if kernel_ms > 0 else 0.0
        roofline = {
            "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
            "frac": achieved / peak if peak else None, "traffic": traffic, "peak_source": peak_src,
            "kernel": "tcp::allpairs_tcws_kernel (experimental)" if args.tc_pipelined else "tc::allpairs_tcws_kernel", "kernel_ms": kernel_ms, "launches_timed": int(n_timed),
            "flops_per_launch": tensor_flops,
            "flops_accounting": "tensor pipe: 2*K = 32 flops per ordered pair tested (split-TF32 Gram matrix, K = 16); "
                                "SURVEY.md 8(d)'s CUDA-core accounting (8 per pair tested + 15 per pair in range) is "
                                "reported as survey_tflops",
            "survey_flops_per_launch": flops_launch, "survey_tflops": survey_tflops,
            "survey_frac_of_fp32_fma_peak": survey_tflops / peak32 if peak32 else None,
            "survey_frac_of_fp64_fma_peak": survey_tflops / peak64 if peak64 else None,
            "pairs_in_range_per_step": int(sum_over_ranks(pin_own)),
            "filter_pairs_last_launch": int(sum_over_ranks(k1.filter_pairs)) if k1.filter_pairs >= 0 else None,
            "pairs_in_range_last_launch": int(sum_over_ranks(pin_now)),    # equal: the filter dropped no pair
            "share_of_step": kernel_ms / (ms / args.steps) if ms > 0 else None,
            "hbm_bytes_per_launch_algorithmic": bytes_launch,
            "fp64_fma_peak": peak64, "fp32_fma_peak": peak32,
            "note": "the survey-accounting rate can exceed the FP32 FMA peak because the pair tests run on the tensor pipe",
        }
    else:
        peak = peak32 if fp32_bound else peak64
        executed = 6.0 * n_own * (n - 1) / (kernel_ms * 1e-3) / 1e12 if kernel_ms > 0 and args.algo == "allpairs" else None
        roofline = {
            "bound": "fp32_fma" if fp32_bound else "fp64_fma", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
            "frac": achieved / peak if peak else None, "traffic": traffic,
            "peak_source": "psim_measure_fma_peak: dependent-free FMA loop on all SMs, this run (of measured; "
                           "MEASURED_PEAKS.json has no vector-FMA figure)",
            "kernel": ("allpairs_warp_kernel" if not args.exact_diff else "allpairs_step_kernel") if args.algo == "allpairs" else "cell_force_kernel",
            "kernel_ms": kernel_ms, "launches_timed": int(n_timed), "flops_per_launch": flops_launch,
            "flops_accounting": "SURVEY.md 8(d): 8 per ordered pair tested + 15 per ordered pair in range",
            "pairs_in_range_per_step": int(sum_over_ranks(pin_own)),
            "share_of_step": kernel_ms / (ms / args.steps) if ms > 0 else None,
            "hbm_bytes_per_launch_algorithmic": bytes_launch,
            "fp64_fma_peak": peak64, "fp32_fma_peak": peak32,
            "frac_of_fp64_fma_peak": achieved / peak64 if peak64 else None,
            "executed_filter_tflops": executed,
            "executed_frac": (executed / peak) if (executed and peak) else None,
            "note": "frac uses the survey's 8-flop difference-form count per pair; the kernel's norm-expansion filter "
                    "executes 3 FMAs = 6 flops per pair (executed_*), on the FP32 pipe, so frac_of_fp64_fma_peak can exceed 1",
        }

    # ---- e2e: the force operator with host buffers through the C ABI ----
    rt = native.rt(prec)
    host_pos = torch.empty((n, 3), dtype=torch.float32 if prec == 4 else torch.float64).pin_memory()
    host_pos.numpy()[:] = pos.astype(rt)
    host_f = torch.empty((n, 3), dtype=host_pos.dtype).pin_memory()
    hp, hf = host_pos.numpy(), host_f.numpy()
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        sim.set_positions(hp); sim.forces(out=hf)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        sim.set_positions(hp)
        sim.forces(out=hf)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": pairs_total * e2e_steps / e2e_s, "unit": "pair-interactions/s",
           "h2d_bytes_per_step": int(n * 3 * prec), "d2h_bytes_per_step": int(n_own * 3 * prec),
           "ms_per_step": e2e_s / e2e_steps * 1e3,
           "call": "psim_set_positions (pinned H2D of all positions + force pass) + psim_get_aux (D2H of the owned forces)"}

    line = {
        "metric": "pair_interactions_per_s", "value": value, "unit": "pair-interactions/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64" if prec == 8 else "f32", "data": "synthetic",
        "timesteps_per_s": args.steps / (ms * 1e-3),
        "config": {"workload": workload_name(n, area), "n_particles": n, "algo": args.algo,
                   "parallelism": "i-sharded x%d, position all-gather per step" % world if world > 1 else "single GPU",
                   "l2": "not flushed between steps: every step reads the positions the previous step wrote "
                         "(state is carried on the device, %.0f MB working set); the force kernel is compute bound (see roofline.bound)" %
                         (n * 4 * prec * 6 / 1e6)},
        "roofline": roofline, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
    }

    if args.algo == "allpairs" and not args.no_cells:
        # the same job through the cell-binned path: what a user who wants
        # timesteps/s selects.  Sharded runs keep ownership by particle index and all-gather the
        # positions per step exactly like the all-pairs path.
        uid2 = None
        if world > 1:
            buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                buf.copy_(torch.frombuffer(bytearray(native.comm_unique_id()), dtype=torch.uint8))
            dist.broadcast(buf, 0)
            uid2 = bytes(buf.cpu().numpy().tobytes())
        cs = P.Simulation(params, ids, pos, vel, prec, device=local_rank, rank=rank, world=world, nccl_unique_id=uid2)
        cs.set_option(P.OPT_FORCE_ALGO, P.FORCE_CELL_BINNED)
        cs.advance(30)
        csteps = max(args.steps, 200)
        barrier()
        cms = max_over_ranks(cs.advance_timed(csteps))
        barrier()
        cc = cs.counters()
        line["cell_binned"] = {"timesteps_per_s": csteps / (cms * 1e-3), "ms_per_step": cms / csteps,
                               "value": pairs_total * csteps / (cms * 1e-3), "unit": "pair-interactions/s",
                               "steps": csteps, "rebuilds": int(cc.rebuilds), "n_gpus": world,
                               "note": "equivalent rate: N(N-1) x timesteps/s; the path tests only the 27-cell "
                                       "neighbourhood of each particle"}
        cs.close()

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            r = cpu_reference_run(n, prec, args.cpu_seconds)
            line["cpu_baseline"] = {
                "value": pairs_total * r["timesteps_per_s"], "unit": "pair-interactions/s", "cores": r["threads"],
                "kind": "port",
                "sample": "%d steps of the same N=%d workload in %.1f s (time loop only)" % (r["steps"], n, r["seconds"]),
                "timesteps_per_s": r["timesteps_per_s"], "list_pair_visits_per_s": r["list_pair_visits_per_s"],
                "rebuilds": r["rebuilds"],
                "note": "oracle port of the reference's O(N) cell-list algorithm; value = N(N-1) x its timesteps/s",
            }
        except Exception as e:   # the baseline must never take the GPU line down
            line["cpu_baseline"] = {"error": repr(e)}
    sim.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        sys.stderr.flush()
        print(json.dumps(line), flush=True)      # the JSON line is the last thing this job prints


if __name__ == "__main__":
    main()
