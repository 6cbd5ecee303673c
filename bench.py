#!/usr/bin/env python
"""Benchmark of the direct-sum hot path (BASELINE.json metric: pairwise interactions/sec).

    python bench.py [--gpus N --steps K --warmup W]                 # this repo's CUDA path
    python bench.py --impl reference [--gpus N --steps K --warmup W] # the reference's CPU path
    torchrun ... bench.py --gpus N ...                               # one rank per GPU

Default workload = BASELINE.json configs[3]: N = 1,048,576 Plummer sphere (the configuration the
1/2/4/8-GPU metric is quoted on; it fits one GPU), STRONG scaling: the same N on every GPU count.
A "step" is one pass of the hot path: force accumulation over all N sources for the rank's
targets + the semi-implicit Euler update (+ the position exchange when N > 1).  value =
N^2 * K / seconds (the GPU convention: both directions of each pair).  One JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pairwise interactions/sec"
UNIT = "interactions/s"
WORKLOADS = {
    # name: (N, generator, BASELINE.json config it mirrors)
    "c2": (65536, "plummer", "configs[1]: N=65536, 1 B200"),
    "c3": (262144, "plummer", "configs[2]: N=262144 roofline capture"),
    "c4": (1048576, "plummer", "configs[3]: N=1,048,576 Plummer strong scaling 1/2/4/8 B200"),
    "c5": (4194304, "plummer", "configs[4]: N=4,194,304, 8 B200"),
}
SM_LANES = 128          # FP32 lanes per SM
FLOP_PER_INTERACTION = 20


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--particles", type=int, default=0,
                    help="override the particle count of the workload (not --n: torchrun would read that as one of its own options)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--mode", default="fast", choices=["fast", "strict", "fp64"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-general", action="store_true", help="skip the general-loop (disk input) roofline measurement")
    ap.add_argument("--ref-n", type=int, default=0, help="sample size of the reference arm (default: chosen to bound its run time)")
    return ap.parse_args()


def workload_config(args):
    n, gen, mirrors = WORKLOADS[args.workload]
    if args.particles:
        n = args.particles
    return n, gen, mirrors


GENERATOR_SOURCE = ("nbody_b200.datasets (NumPy restatement of the reference's generatePlummerSphere / generateGalaxyDisk, "
                    "src/Simulation.cpp:88-136,173-267: same distributions and softening rule, not the same mt19937 stream)")


def shared_config(args, n, gen, mirrors, eps):
    """The workload description -- the SAME dictionary in both arms, so the two lines compare like for like."""
    return {"workload": f"{args.workload}: {mirrors}", "n": n, "generator": gen, "generator_source": GENERATOR_SOURCE,
            "G": 1.0, "dt": 0.5, "min_r2": eps,
            "l2": "GPU arm: flushed between steps (160 MiB memset, inside the timed region); CPU arm: not applicable"}


def make_particles(nb, n, gen, seed=12345):
    if gen == "plummer":
        return nb.datasets.plummer_sphere(n, 100.0, float(n), seed=seed)   # headless_main.cpp:191-195
    return nb.datasets.galaxy_disk(n, 300.0, 50.0, seed=seed)              # headless_main.cpp:200-205


# ------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            pass
        sm, mx, pw, reasons = [], [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), power_w_max=max(pw), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------------------------------------
# reference arm / CPU baseline (the ONLY places bench.py touches oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_reference_rate(p, eps, steps, threads, kind_hint=None):
    """Times the reference's CPU implementation on the given particles.  threads == 1: sequential
    NaiveSimulation (src/NaiveSimulation.cpp); threads > 1: MpiNaiveSimulation with one thread-rank
    per host core (src/MpiNaiveSimulation.cpp through oracle/mpi_shim).  Returns (int/s, seconds, kind)."""
    from oracle import Oracle, Reference

    n = p.shape[0]
    if Reference.available():
        ref = Reference()
        if threads == 1:
            _, secs = ref.naive_steps(p, 1.0, 0.5, eps, steps, timed=True)
        else:
            _, secs = ref.mpi_naive_steps(p, 1.0, 0.5, eps, steps, nranks=threads, timed=True)
        kind = "reference"
    else:
        orc = Oracle()
        t0 = time.perf_counter()
        if threads == 1:
            orc.step_pairs(p, 1.0, 0.5, eps, steps)
        else:
            orc.step_rows(p, 1.0, 0.5, eps, steps)      # OpenMP rows: the port of the sharded path
        secs = time.perf_counter() - t0
        kind = "port"
    return float(n) * n * steps / secs, secs, kind


def reference_sample_size(args, n, cores):
    # bounded sample: ~1.5 s of wall time per step on `cores` threads at ~7 ns per interaction per core
    return args.ref_n or int(min(n, max(4096, (1.5 * cores / 7e-9) ** 0.5 // 1024 * 1024)))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import nbody_b200 as nb

    n, gen, mirrors = workload_config(args)
    cores = os.cpu_count() or 1
    n_s = reference_sample_size(args, n, cores)
    p, eps = make_particles(nb, n, gen)
    ps = p[:n_s].copy()
    if args.warmup:
        cpu_reference_rate(ps, eps, 1, cores)
    rate, secs, kind = cpu_reference_rate(ps, eps, args.steps, cores)
    sample = (f"first {n_s} of {n} particles, {args.steps} steps, MpiNaiveSimulation on {cores} thread-ranks "
              f"(oracle/mpi_shim), compute-only time; rate is size-independent (N^2 law)" if kind == "reference"
              else f"first {n_s} of {n} particles, {args.steps} steps, row-order port on {cores} OpenMP threads")
    # the reference's OWN clock (src/apps/headless_main.cpp:344-421): sim.step() + per-step gather to rank 0
    own = None
    if kind == "reference":
        from oracle import Reference

        _, secs_own = Reference().mpi_headless_loop(ps, 1.0, 0.5, eps, args.steps, nranks=cores)
        own = {"value": float(n_s) * n_s * args.steps / secs_own, "unit": UNIT, "ms_per_step": secs_own / args.steps * 1e3,
               "definition": "wall clock of the headless step loop as the reference measures it: NBodySimulation::step "
                             "(computeStep + render-buffer copy) + per-step MPI_Gatherv of all particles to rank 0 "
                             "(src/apps/headless_main.cpp:344-421, --disable-output); same sample"}
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": shared_config(args, n, gen, mirrors, eps),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_own_clock": own,
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def rel_err_rows(a, b):
    import numpy as np

    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    d = np.linalg.norm(b, axis=1)
    return np.linalg.norm(a - b, axis=1) / np.where(d == 0, 1.0, d)


def sharded_parity_check(nb, dist, local_rank, group, exchange, n=65536, steps=2):
    """Before anything is timed at N > 1: the sharded path (i-slices by the reference's rule,
    src/apps/headless_main.cpp:249-259, positions republished by the selected exchange) against the
    UNSHARDED run of the same system on this rank's own device.  Strict mode must agree bit for bit in
    every field (it reproduces the sequential NaiveSimulation whatever the rank count); fast mode within
    rounding.  Unequal masses, so the general loop and the mass flag are exercised.  Returns the verdict
    of all ranks (MIN / MAX reduced)."""
    import numpy as np
    import torch

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    p, eps = nb.datasets.plummer_sphere(n, 100.0, float(n), seed=777)
    p["mass"] = np.random.default_rng(11).uniform(0.5, 2.0, n).astype(np.float32)
    params = nb.SimParams(G=1.0, dt=0.5, min_r2=eps)
    counts, displs = nb.slice_counts(n, world)
    lo, cnt = displs[rank], counts[rank]
    out = {"n": n, "steps": steps, "exchange": exchange}
    strict_ok, fast_pos, fast_acc = 1, 0.0, 0.0
    for mode in ("strict", "fast"):
        single = nb.CudaNaiveSimulation(device=local_rank, mode=mode)
        want = p.copy()
        single.upload(want); single.step(params, steps); single.download(want)
        single.close()
        sim = nb.CudaNaiveSimulation(device=local_rank, mode=mode, group=group, exchange=exchange)
        got = p[lo:lo + cnt].copy()
        sim.upload(got, n_total=n); sim.step(params, steps); sim.download(got)
        out["exchange"] = sim.exchange                    # "nccl" if peer mapping was refused on some rank
        sim.close()
        ref = want[lo:lo + cnt]
        if mode == "strict":
            strict_ok = int(got.tobytes() == ref.tobytes())
        else:
            fast_pos = float(rel_err_rows(got["position"], ref["position"]).max())
            fast_acc = float(rel_err_rows(got["acceleration"], ref["acceleration"]).max())
    t = torch.tensor([float(strict_ok), -fast_pos, -fast_acc], dtype=torch.float64, device=f"cuda:{local_rank}")
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    out.update(strict_bitwise_vs_1gpu=bool(t[0].item() == 1.0), fast_max_rel_pos_vs_1gpu=float(-t[1].item()),
               fast_max_rel_acc_vs_1gpu=float(-t[2].item()))
    out["ok"] = out["strict_bitwise_vs_1gpu"] and out["fast_max_rel_pos_vs_1gpu"] < 1e-5 and out["fast_max_rel_acc_vs_1gpu"] < 1e-4
    return out


def general_loop_roofline(nb, sim, n, lo, cnt, nominal_peak, steps=3):
    """The force kernel's general (unequal-mass) loop on the same N: the galaxy disk the reference's
    sequential --random run uses (src/apps/headless_main.cpp:200-205; masses U[0.5,2) and one 1e6 body,
    src/Simulation.cpp:92-98,121).  Same per-launch CUDA-event method as the headline roofline; outside
    the headline's timed region."""
    p, eps = make_particles(nb, n, "disk")
    params = nb.SimParams(G=1.0, dt=0.5, min_r2=eps)
    sim.upload(p[lo:lo + cnt].copy(), n_total=n)
    sim.step(params, 1)                       # warm-up; the synchronising call below also teaches the host the mass flag
    sim.synchronize()
    sim.profile_force(True, reserve=steps + 1)
    sim.force_time(reset=True)
    for _ in range(steps):
        sim.step(params, 1)
    ms, launches = sim.force_time(reset=True)
    sim.profile_force(False)
    if not launches:
        return None
    flop = FLOP_PER_INTERACTION * float(cnt) * float(n)
    achieved = flop / (ms / launches * 1e-3) / 1e12
    return {"bound": "fp32_fma", "kernel": "force_fast_kernel (general loop: m_j multiplied in)", "input": "galaxy disk, unequal masses",
            "achieved": achieved, "peak": nominal_peak, "unit": "TFLOP/s", "frac": achieved / nominal_peak,
            "avg_launch_ms": ms / launches, "launches_timed": launches, "issue_ceiling_frac": 20.0 / 24.0}


def run_ours(args):
    import torch

    import nbody_b200 as nb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun --nproc-per-node N for --gpus N (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    group = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
        group = dist.group.WORLD

    n, gen, mirrors = workload_config(args)
    p, eps = make_particles(nb, n, gen)          # same seed on every rank -> identical global set
    params = nb.SimParams(G=1.0, dt=0.5, min_r2=eps)   # headless_main.cpp:33-36 + generator's softening
    counts, displs = nb.slice_counts(n, world)
    lo, cnt = displs[rank], counts[rank]

    parity = None
    if world > 1:
        parity = sharded_parity_check(nb, dist, local_rank, group, args.exchange)
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"error": "sharded run disagrees with the single-GPU run", "parity": parity}))
            dist.barrier()
            dist.destroy_process_group()
            return 3

    sim = nb.CudaNaiveSimulation(device=local_rank, mode=args.mode, group=group, exchange=args.exchange)
    lib = nb.load_library()
    ext = torch.cuda.ExternalStream(lib.nbody_cuda_stream(sim._h), device=local_rank)
    flush_buf = torch.empty(160 << 20, dtype=torch.uint8, device=f"cuda:{local_rank}")   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_region(fn_step, steps):
        """K steps bracketed by barrier + synchronize, device-timed on the handle's stream, max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record(ext)
            for s in range(steps):
                fn_step(s)
                flush_buf.zero_()               # L2 flush between timed iterations (its cost is inside the region)
            e1.record(ext)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local_rank}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- device-resident throughput (value) ----
    local = p[lo:lo + cnt].copy()
    sim.upload(local, n_total=n)
    sim.synchronize()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()          # before the warm-up: nvidia-smi's own start-up must not land in the timed region
    for s in range(args.warmup):
        sim.step(params, 1)
        with torch.cuda.stream(ext):
            flush_buf.zero_()    # the warm-up also loads the flush kernel: not in the timed region
    sim.synchronize()
    # short steps (small N): W steps are over in milliseconds, before the SM clock has left idle; keep
    # warming up (untimed) until half a second of device work has been done
    extra_warmup = 0
    per_step_ms = max(sim.last_step_ms(), 1e-3)
    if per_step_ms * args.warmup < 500.0:
        extra_warmup = int(min(20000, (500.0 - per_step_ms * args.warmup) / per_step_ms))
        if extra_warmup > 0:
            sim.step(params, extra_warmup)
            sim.synchronize()
    sim.profile_force(True, reserve=2 * args.steps + 2)     # timing events exist before the clock starts
    launches0 = sim.launch_count()
    sim.force_time(reset=True)
    ms = timed_region(lambda s: sim.step(params, 1), args.steps)
    force_ms, force_launches = sim.force_time(reset=True)
    sim.profile_force(False)
    gpu_launches = sim.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else {}
    value = float(n) * n * args.steps / (ms * 1e-3)

    # ---- end to end through the reference-facing call: computeStep(host slice) ----
    e2e = None
    if not args.no_e2e:
        pinned = nb.PinnedParticles(cnt)
        pinned.array[:] = local
        for s in range(min(2, args.warmup)):
            sim.computeStep(pinned.array, params, n_total=n)
        barrier()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record(ext)
        for s in range(args.steps):
            sim.computeStep(pinned.array, params, n_total=n)     # H2D of the slice, step (+exchange), D2H of the slice
        with torch.cuda.stream(ext):
            e1.record(ext)
        barrier()
        ms_e2e = e0.elapsed_time(e1)
        wall = time.perf_counter() - t0
        ms_e2e = max(ms_e2e, 0.0)
        if world > 1:
            t = torch.tensor([ms_e2e], dtype=torch.float64, device=f"cuda:{local_rank}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_e2e = float(t.item())
        e2e = {"value": float(n) * n * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(cnt * 40), "d2h_bytes_per_step": int(cnt * 40),
               "ms_per_step": ms_e2e / args.steps, "host_wall_ms_per_step": wall / args.steps * 1e3,
               "call": "CudaNaiveSimulation.computeStep(pinned host slice, params, n_total) -> nbody_cuda_upload_aos/[_exchange]/_step/_download_aos"}
        pinned.close()

    # ---- roofline of the dominant kernel (force accumulation) ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    props = torch.cuda.get_device_properties(local_rank)
    sm_max_mhz = float(peaks.get("sm_max_mhz") or clocks.get("sm_max_mhz") or 1965.0)
    lanes = 64 if args.mode == "fp64" else SM_LANES          # B200: 64 FP64 lanes per SM
    nominal_peak = props.multi_processor_count * lanes * 2 * sm_max_mhz * 1e6 / 1e12
    measured_fma = sim.measure_fma_peak(True) / 1e12
    flop_per_launch = FLOP_PER_INTERACTION * float(cnt) * float(n)
    achieved = flop_per_launch / (force_ms / max(force_launches, 1) * 1e-3) / 1e12 if force_launches else None
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "force_kernel_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(f"n{n}_gpus{world}")
        except Exception:
            traffic = None
    roofline = {
        "bound": "fp64_fma" if args.mode == "fp64" else "fp32_fma", "kernel": f"force_{args.mode}_kernel", "achieved": achieved, "peak": nominal_peak,
        "unit": "TFLOP/s", "frac": (achieved / nominal_peak) if achieved else None, "traffic": traffic,
        "peak_source": f"{props.multi_processor_count} SMs x {lanes} {'FP64' if args.mode == 'fp64' else 'FP32'} lanes x 2 flop x {sm_max_mhz:.0f} MHz "
                       "(sm_max_mhz of MEASURED_PEAKS.json, which holds no FP32 figure); tensor cores unused by design",
        "measured_ffma2_chain_tflops": measured_fma,
        "frac_of_measured_ffma2_chain": (achieved / measured_fma) if achieved else None,
        "flop_per_interaction": FLOP_PER_INTERACTION, "interactions_per_launch": float(cnt) * float(n),
        "avg_launch_ms": force_ms / max(force_launches, 1), "launches_timed": force_launches,
        "issue_ceiling_frac": 20.0 / 24.0,
    }

    # ---- the general (unequal-mass) loop of the same kernel, same N, disk input ----
    roofline_general = None
    if args.mode == "fast" and not args.no_general:
        roofline_general = general_loop_roofline(nb, sim, n, lo, cnt, nominal_peak)

    # ---- CPU baseline: the reference's sequential path on a bounded sample (rank 0) ----
    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        n_s = min(n, 32768)                                 # ~5e8 pair evaluations per step: 5-10 s on one core
        ps = p[:n_s].copy()
        rate, secs, kind = cpu_reference_rate(ps, eps, 2, 1)
        cpu = {"value": rate, "unit": UNIT, "cores": 1, "kind": kind,
               "sample": f"first {n_s} of {n} particles, 2 steps of sequential NaiveSimulation::computeStep "
                         f"({secs:.1f} s); N^2-convention rate (it evaluates N(N-1)/2 pairs)"}

    if rank == 0:
        plan = sim.plan() if args.mode == "fast" else {}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64" if args.mode == "fp64" else "f32", "data": "synthetic",
            "config": shared_config(args, n, gen, mirrors, eps),
            "details": {"mode": args.mode, "exchange": sim.exchange if world > 1 else "none",
                        "parallelism": f"i-slice x{world}", "extra_warmup_steps": extra_warmup, "plan": plan},
            "parity": parity if world > 1 else "single GPU: checked against the oracle by tests/ and smoke()",
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(gpu_launches), "roofline": roofline,
            "roofline_general": roofline_general, "cpu_baseline": cpu,
        }
        print(json.dumps(line))
    sim.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
