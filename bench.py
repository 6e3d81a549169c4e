#!/usr/bin/env python
"""bench.py -- headline measurement of the nbody hot path on B200.

Metric (BASELINE.json): pair-interactions per second (and steps/s) of the all-pairs 2D gravity +
integrator step at N = 1,048,576 galaxy-initialised particles, on 1/2/4/8 B200, with the fraction
of the FP32 roofline (14 FLOP per pair).  One "step" = one Jacobi update of all N targets against
all mass_len sources = total_len * mass_len pair interactions.

  python bench.py [--gpus N] [--steps K] [--warmup W]           our CUDA path (one rank per GPU)
  python bench.py --impl reference ...                          the reference's own CPU path
                                                                (oracle/_ref, all host threads,
                                                                bounded sample of the same workload)

Prints ONE JSON line on rank 0.  For N > 1 launch with
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port P bench.py --gpus N ...
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_PAIR = 14  # 2 sub, 5 mul, 4 add, 1 sqrt, 1 div (ref src/lib/sim_cpu.c:169-188)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    # "--particles" because torchrun's own parser rejects "--n" as an ambiguous abbreviation of its options
    ap.add_argument("--n", "--particles", dest="n", type=int, default=1 << 20,
                    help="particles (default: N=1M, the metric's config)")
    ap.add_argument("--galaxies", type=int, default=2)
    ap.add_argument("--seed", type=int, default=11037)
    ap.add_argument("--dt", type=float, default=1.0, help="ref src/bench.c:21 UPDATE_STEP")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU work per cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_name(n, galaxies, seed=11037):
    return f"N={n} galaxy init ({galaxies} galaxies, srand({seed})), all-pairs step, dt per ref bench.c"


def make_config(args, total_len, mass_len, gpus):
    """The `config` object of the JSON line.  Both arms print the SAME object (the reference arm runs "on
    your arm's config"): it names the workload and how the GPU arm is run; what the CPU arm samples of that
    workload per step is stated in its `cpu_baseline.sample`, not here."""
    return {"workload": workload_name(args.n, args.galaxies, args.seed), "total_len": int(total_len),
            "mass_len": int(mass_len), "dt": args.dt, "parallelism": f"targets sharded over {gpus} GPU(s)",
            "l2": "flushed between timed steps (256 MiB memset, untimed)"}


# ---- clocks sampled while the timed region runs -------------------------------------------------

class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s, p in zip(sm, pw) if p >= 0.5 * max(pw)] or sm
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- the reference's CPU path, timed on a bounded sample -----------------------------------------

def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm is entitled to all host threads."""
    import ctypes as C
    n = len(os.sched_getaffinity(0))
    try:
        C.CDLL("libgomp.so.1").omp_set_num_threads(n)
    except OSError:
        pass
    return n


def cpu_reference_runner():
    """Returns (kind, threads, step_range(arr, mass_len, first, last, dt)) for the CPU arm."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import Oracle, RefLib
    use_all_host_threads()
    if RefLib.available("avx"):
        r = RefLib("avx")
        return "reference", r.threads(), r.lib.refglue_step_range, "oracle/_ref/libnbody_ref.so (gcc -O3 -DUSE_AVX -mavx -fopenmp)"
    o = Oracle()

    def step(ptr, m, first, last, dt):
        o.lib.oracle_step_lanes_range(ptr, m, first, last, dt, 8)
    return "port", o.threads(), step, "oracle/_build/liboracle.so (C restatement, 8-lane order)"


def time_cpu_sample(step_fn, arr, mass_len, targets, dt):
    import ctypes as C
    work = arr.copy()
    ptr = work.ctypes.data_as(C.c_void_p)
    t0 = time.perf_counter()
    step_fn(ptr, mass_len, 0, targets, dt)
    return time.perf_counter() - t0


def calibrate_cpu_sample(step_fn, arr, mass_len, dt, seconds):
    """Pick how many leading targets make about `seconds` of CPU work for one step."""
    n = len(arr)
    probe = min(n, 2048)
    t = time_cpu_sample(step_fn, arr, mass_len, probe, dt)
    t = min(t, time_cpu_sample(step_fn, arr, mass_len, probe, dt))
    rate = probe * mass_len / max(t, 1e-6)
    targets = int(min(n, max(probe, seconds * rate / max(mass_len, 1))))
    return max(20, targets // 20 * 20)


def run_reference(args):
    """--impl reference: the reference's own CPU update on the host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import Oracle, RefLib, mass_len_of
    kind, threads, step_fn, how = cpu_reference_runner()
    if RefLib.available("avx"):
        # inputs and World order from the reference's own MakeGalaxies / CreateWorld
        r = RefLib("avx")
        arr = r.world_state(r.make_galaxies(args.n, args.galaxies, seed=args.seed))
        m = mass_len_of(arr)
    else:
        import nbody_b200 as nb
        arr, m = Oracle().partition(nb.make_galaxies(args.n, args.galaxies, seed=args.seed))
    # size the per-step sample so that (warmup + steps) * sample stays within a few minutes
    budget = min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup))
    targets = calibrate_cpu_sample(step_fn, arr, m, args.dt, budget)
    for _ in range(args.warmup):
        time_cpu_sample(step_fn, arr, m, targets, args.dt)
    times = [time_cpu_sample(step_fn, arr, m, targets, args.dt) for _ in range(args.steps)]
    total = sum(times)
    pairs = float(targets) * m * args.steps
    value = pairs / total * 1e-9
    sample = f"first {targets} of {len(arr)} targets x all {m} sources per step ({how})"
    line = {
        "impl": "reference",
        "metric": "pair_interactions_per_second", "value": value, "unit": "Gpairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        # a step of this arm is the bounded sample: ms_per_step is what was timed, not an extrapolation
        "ms_per_step": total / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(args, len(arr), m, args.gpus),
        "cpu_baseline": {"value": value, "unit": "Gpairs/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Gpairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "ms_per_full_step_extrapolated": total / args.steps * 1e3 * (len(arr) / targets),
        "steps_per_s_extrapolated": value * 1e9 / (float(len(arr)) * m),
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def fp32_peak_tflops(info):
    """SMs x 128 lanes x 2 (FMA) x max SM clock.  MEASURED_PEAKS.json records no FP32 figure, so the
    roofline denominator is derived from the device attributes."""
    return info["sm_count"] * 128 * 2 * info["sm_clock_khz"] * 1e3 * 1e-12


def run_ours(args):
    import torch  # plumbing only: device selection, L2 flush buffer, torch.distributed rendezvous
    import torch.distributed as dist
    import nbody_b200 as nb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
        raise SystemExit(f"WORLD_SIZE={world} but --gpus {args.gpus}")
    if not torch.cuda.is_available() or nb.device_count() == 0:
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # identical input on every rank: the generator is deterministic (srand(seed); MakeGalaxies)
    ps = nb.make_galaxies(args.n, args.galaxies, seed=args.seed)
    w0 = nb.World(ps)                      # partition through the product's CreateWorld
    arr = w0.particles()
    m = w0.mass_len
    w0.close()
    pairs_per_step = float(len(arr)) * m

    uid = None
    if world > 1:
        box = [nb.get_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]
    sim = nb.SimPipeline(len(arr), m, device=local_rank, local_gpus=1, rank=rank, world=world, unique_id=uid)
    sim.set_data(arr)
    info = sim.info()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def one_step():
        flush.zero_()                      # evict L2 between timed steps (untimed)
        barrier()                          # ranks start the step together: without it the skew of the untimed
                                           # flush is counted as time spent waiting in the position exchange
        sim.step_async(1, args.dt)
        return sim.last_step_ms()          # CUDA events on the launching stream; synchronises

    for _ in range(args.warmup):
        one_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = sim.info()["kernels_launched"]
    t_wall0 = time.perf_counter()
    step_ms = [one_step() for _ in range(args.steps)]
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = sim.info()["kernels_launched"] - launches0
    clocks = sampler.stop() if rank == 0 else None

    total_ms = float(sum(step_ms))
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = pairs_per_step / (ms_per_step * 1e-3) * 1e-9          # G pairs/s, whole job

    # ---- e2e: host buffers, H2D + step + D2H inside the timed region, through the C ABI ----------
    e2e = None
    if not args.no_e2e:
        n_e2e = max(2, min(args.steps, 5))
        if world == 1:
            wld = nb.World(arr)            # World API: pinned host array inside the World

            def e2e_step():
                wld.update_cpu(args.dt, 0)         # marks the host copy newer -> forces the upload
                wld.update_gpu(args.dt, 1)         # H2D (if stale) + step, synchronous
                wld.particles(copy=False)          # D2H into the World's array
        else:
            pinned = nb.PinnedParticles(len(arr))   # page-locked host buffer, like the World's own array
            host = pinned.array
            host[:] = arr

            def e2e_step():
                sim.set_data(host)                 # sim_gpu.h seam: H2D of the particles this rank owns,
                                                   # source data replicated GPU to GPU (NCCL)
                sim.update(1, args.dt)
                sim.get_data_local(host)           # D2H of the particles this rank owns (sharded read-back)
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        dt_e2e = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt_e2e], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt_e2e = float(t.item())
        e2e = {"value": pairs_per_step * n_e2e / dt_e2e * 1e-9, "unit": "Gpairs/s",
               "h2d_bytes_per_step": int(len(arr)) * 32, "d2h_bytes_per_step": int(len(arr)) * 32,
               "ms_per_step": dt_e2e / n_e2e * 1e3, "steps": n_e2e,
               "api": "CreateWorld/UpdateWorld_GPU/GetWorldParticles" if world == 1
                      else "SetSimulationData/PerformSimUpdate/NbSimGetDataLocal per rank (each rank moves its own shard)"}
        if world == 1:
            wld.close()

    if rank != 0:
        sim.close()
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak = fp32_peak_tflops(info) * world
    achieved = value * 1e9 * FLOP_PER_PAIR * 1e-12
    # DRAM bytes of the dominant kernel per launch come from a committed ncu capture of the same
    # configuration (profiles/ncu_traffic.json); null when this configuration was never captured.
    traffic = None
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        key = f"{info['total_len']},{info['mass_len']},{info['threads']}x{info['tile_targets'] // info['threads']},{info['slabs']}"
        if world == 1 and key in table:
            traffic = table[key]["bytes"]
    except (OSError, ValueError):
        pass
    roofline = {
        "bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
        "traffic": traffic,
        "peak_source": "derived: SMs x 128 x 2 x max SM clock (MEASURED_PEAKS.json has no FP32 figure)",
        "kernel": f"nb::step_kernel {info['variant']} ({info['threads']} threads, {info['tile_targets']} targets per CTA, "
                  f"{info['slabs']} slabs of {info['slab_len']} sources)",
        "flop_per_pair": FLOP_PER_PAIR,
        "ceiling_frac": 14.0 / 18.0,
        # dispatch bound of the shipped loop: per 32 pairs 144 f32x2 instructions hold the
        # SM sub-partition's dispatch port for 2 cycles each, 32 MUFU + 6 LDS + 5 loop instructions for 1
        "dispatch_bound_frac": 288.0 / 331.0 * 14.0 / 18.0,
    }
    if clocks and clocks.get("sm_mhz"):
        roofline["frac_at_measured_clock"] = achieved / (peak * clocks["sm_mhz"] * 1e3 / info["sm_clock_khz"])

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        kind, threads, step_fn, how = cpu_reference_runner()
        targets = calibrate_cpu_sample(step_fn, arr, m, args.dt, args.cpu_seconds)
        t = time_cpu_sample(step_fn, arr, m, targets, args.dt)
        cpu_baseline = {"value": targets * float(m) / t * 1e-9, "unit": "Gpairs/s", "cores": threads, "kind": kind,
                        "sample": f"one step over the first {targets} of {len(arr)} targets x all {m} sources, "
                                  f"{t:.1f} s ({how})"}

    line = {
        "metric": "pair_interactions_per_second", "value": value, "unit": "Gpairs/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(args, len(arr), m, world),
        "steps_per_s": 1e3 / ms_per_step,
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        "roofline": roofline, "cpu_baseline": cpu_baseline,
        "wall_s_timed_region": t_wall, "device": info["device_name"],
    }
    print(json.dumps(line), flush=True)
    sim.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
