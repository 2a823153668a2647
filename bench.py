#!/usr/bin/env python
"""bench.py — headline benchmark of the escape-time path (BASELINE.json: G iter/s and ms/frame, 3000x3000 Mandelbrot
@1e6x zoom, 1/2/4/8 B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload C2|C2b|C1|C3|C4]
                    [--scaling weak|strong]

One "step" = one pass of the hot path over one frame (per rank).  Prints ONE JSON line (rank 0).

  value      G iter/s, device-resident: the frame is rendered into HBM through the C ABI's device entry point
             (frb_render_device_async) on torch's current stream; K steps timed with CUDA events, max over ranks.
  e2e        the same metric through the reference-facing call get_image (frb_render) with a HOST (pinned) buffer:
             config in (104 B, travels as kernel arguments), finished RGB24 frame copied device->host every step.
  roofline   the iterate kernel against the FP64 pipe: 10 flop per iteration as the reference writes it
             (6 mul + 4 add, calc/src/lib.rs:88-104) over the kernel's own CUDA-event time, against the DFMA peak
             measured on this GPU in this run (MEASURED_PEAKS.json has no FP64 entry).
  cpu_baseline  the CPU oracle (C restatement of the reference path, OpenMP over rows like rayon) on this box's cores,
             on a stated row-strided sample of the same frame.

--impl reference times that CPU path alone (the Rust reference cannot be built on this image: no rustc/cargo).
N > 1 (torchrun): weak scaling = one whole frame per rank and step (frames of a sequence are independent);
strong = one frame split into interleaved 16-row bands (rank r owns bands k % N == r).  No collective on the data path.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_ITER = 10  # 6 mul + 4 add/sub per loop iteration as written in the reference
FP64_NOMINAL_TFLOPS = 37.2  # 148 SM x 64 lanes x 2 x 1.965 GHz


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2", "C2b", "C3", "C4"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel", default="auto", choices=["auto", "stream6", "stream7", "simple", "dense6", "defer6", "tile"])
    return ap.parse_args()


WORKLOAD_TEXT = {
    "C1": "C1 Mandelbrot 1920x1080 it256 full view",
    "C2": "C2 Mandelbrot 3000x3000 it4000 scale 5e5 (README 1e6x-zoom screenshot view), RGB24",
    "C2b": "C2b Mandelbrot 3000x3000 it4000 scale 1e6",
    "C3": "C3 Julia 4096x4096 it1000 c=(-0.7269,0.1889)",
    "C4": "C4 Mandelbrot 16384x16384 it10000 scale 1e12",
}


class ClockSampler:
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        mhz, mx, power, reasons = [], None, [], set()
        rows = [l for (t, l) in self.lines if t0 <= t <= t1 + 0.2] or [l for (_, l) in self.lines]
        for l in rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                mhz.append(float(f[1]))
                mx = float(f[2])
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(mhz) if mhz else None, "sm_max_mhz": mx, "samples": len(mhz),
                "power_w_max": max(power) if power else None, "reasons": sorted(reasons)}


def cpu_sample_rows(height: int, stride: int):
    import numpy as np

    return np.arange(stride // 2, height, stride, dtype=np.uint32)


def host_threads() -> int:
    """All host cores this process may use — NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def time_cpu(cfg, stride: int, threads: int = 0):
    """Oracle (OpenMP rows, like rayon) on a row-strided sample of the frame.  Returns (G iter/s, seconds, iterations)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    threads = threads or host_threads()

    rows = cpu_sample_rows(cfg.height, stride)
    t = time.perf_counter()
    r = ob.render_rows(cfg, rows=rows, want_iters=False, want_z=False, want_rgb=True, threads=threads)
    dt = time.perf_counter() - t
    return r["total"] / dt / 1e9, dt, r["total"], len(rows)


def cpu_baseline(cfg) -> dict:
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    cores = host_threads()
    stride = 16
    g, dt, iters, nrows = time_cpu(cfg, stride)
    # grow the sample until it is ~10 s of CPU work (bounded: never more than the whole frame)
    while dt < 4.0 and stride > 1:
        stride = max(1, stride // 2)
        g, dt, iters, nrows = time_cpu(cfg, stride)
    return {"value": round(g, 4), "unit": "G iter/s", "cores": cores, "kind": "port",
            "sample": f"every {stride}th row of the frame ({nrows} rows x {cfg.width} px, {iters} iterations) in {dt:.2f} s; "
                      "oracle/calc_oracle.c (C restatement; the Rust reference cannot be built here), OpenMP dynamic rows",
            "ms_per_frame_extrapolated": None}


def run_reference(args, cfg, total_iterations):
    """--impl reference: the reference's CPU path (oracle port) on the host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    cores = host_threads()
    # size the per-step sample to ~1-2 s so that K+W steps finish within a few minutes on any core count
    stride = 16
    g, dt, iters, nrows = time_cpu(cfg, stride)
    while dt < 0.75 and stride > 1:
        stride //= 2
        g, dt, iters, nrows = time_cpu(cfg, stride)
    for _ in range(args.warmup):
        time_cpu(cfg, stride)
    t0 = time.perf_counter()
    tot = 0
    for _ in range(args.steps):
        _, _, it, _ = time_cpu(cfg, stride)
        tot += it
    dt = time.perf_counter() - t0
    value = tot / dt / 1e9
    sample = (f"each step = every {stride}th row of the frame ({nrows} rows x {cfg.width} px, {iters} iterations); oracle/calc_oracle.c "
              "(C restatement of calc/src/lib.rs; the Rust reference cannot be built on this image), OpenMP dynamic rows = rayon rows")
    line = {
        "impl": "reference", "metric": "G iter/s", "value": round(value, 4), "unit": "G iter/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic (the fractal itself; no RNG)",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "iterations_per_frame": total_iterations},
        "ms_per_frame": round(total_iterations / (value * 1e9) * 1e3, 2),
        "cpu_baseline": {"value": round(value, 4), "unit": "G iter/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 4), "unit": "G iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    from fractal_renderer_b200 import configs

    cfg = configs.BY_NAME[args.workload]()
    total_iterations = configs.TOTAL_ITERATIONS[args.workload]

    if args.impl == "reference":
        run_reference(args, cfg, total_iterations)
        return

    import numpy as np
    import torch

    import fractal_renderer_b200 as frb

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available() or frb.device_count() < 1:
        raise SystemExit("bench.py needs a B200: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        # NCCL may print its version banner on stdout (NCCL_DEBUG=VERSION on some boxes): keep stdout for the ONE JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist_mod.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
        dist = dist_mod
    frb.init(1 << local)
    variant = {"auto": frb.KERNEL_AUTO, "stream6": frb.KERNEL_STREAM6, "stream7": frb.KERNEL_STREAM7, "simple": frb.KERNEL_SIMPLE, "dense6": frb.KERNEL_DENSE6, "defer6": frb.KERNEL_DEFER6, "tile": frb.KERNEL_TILE}[args.kernel]
    frb.set_kernel_variant(variant)

    shard = frb.Shard(rank, world, 16) if (world > 1 and args.scaling == "strong") else frb.Shard()
    rows = shard.rows(cfg.height)
    nbytes = rows * cfg.width * 3
    dev_out = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
    host_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    host_np = host_out.numpy()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # FP64 peak of this GPU, measured now (rank-local; ~0.3 s)
    peak = frb.measure_fp64_peak(local)

    # a real (non-default) stream: the library launches on it and the CUDA events below are recorded on it
    stream = torch.cuda.Stream(device=local)
    assert stream.cuda_stream != 0
    c_cfg, c_shard = cfg.to_c(), shard.to_c()

    def step_device():
        frb.render_device_async(cfg, dev_out.data_ptr(), nbytes, frb.FRB_RGB24, device=local, shard=shard, stream=stream.cuda_stream)

    def step_e2e():
        frb._lib.check(frb._lib.load().frb_render_shard(ctypes.byref(c_cfg), ctypes.byref(c_shard), host_np.ctypes.data,
                                                        host_np.nbytes, frb.FRB_RGB24, local))

    # ---------------- device-resident arm (`value`) ----------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tw0 = time.perf_counter()
    e0.record(stream)
    iter_ms, shade_ms, launches, my_iters, lane_iters = [], [], 0, 0, 0
    def harvest(back):
        nonlocal launches, my_iters, lane_iters
        st = frb.async_stats(back)
        iter_ms.append(st["iterate_ms"])
        shade_ms.append(st["shade_ms"])
        launches += st["kernel_launches"]
        my_iters += st["iterations"]
        lane_iters += st["lane_iterations"]
        return st

    for k in range(args.steps):
        step_device()
        # keep one frame enqueued ahead while harvesting the previous one's statistics (host-side event wait only)
        if k > 0:
            harvest(1)
    e1.record(stream)
    st = harvest(0)
    barrier()
    tw1 = time.perf_counter()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop(tw0, tw1) if rank == 0 else None
    all_iters = sum_over_ranks(float(my_iters))
    value = all_iters / (dev_ms * 1e-3) / 1e9
    kernel_variant = st["kernel_variant"]

    # ---------------- end-to-end arm (`e2e`): public host-buffer call, D2H inside the timed region ----------------
    for _ in range(max(args.warmup, 3)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    e2e_iters = 0
    e2e_launches = 0
    for _ in range(args.steps):
        step_e2e()
        s2 = frb.last_stats()
        e2e_iters += s2["iterations"]
        e2e_launches += s2["kernel_launches"]
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = sum_over_ranks(float(e2e_iters)) / e2e_s / 1e9

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the dominant kernel (iterate) ----------------
    it_ms = statistics.mean(iter_ms)
    per_launch_iters = my_iters / args.steps
    achieved_tflops = per_launch_iters * FLOP_PER_ITER / (it_ms * 1e-3) / 1e12
    peak_tflops = peak["dfma_per_s"] * 2 / 1e12
    form = {frb.KERNEL_TILE: 6, frb.KERNEL_DEFER6: 6, frb.KERNEL_STREAM6: 6, frb.KERNEL_DENSE6: 6, frb.KERNEL_STREAM7: 7}.get(kernel_variant, 11)
    kname = {frb.KERNEL_TILE: "frb_iterate_tile_kernel<6> (iterate + shade)", frb.KERNEL_DEFER6: "frb_iterate_deferred_kernel<6>", frb.KERNEL_STREAM6: "frb_iterate_sparse_kernel<6>", frb.KERNEL_DENSE6: "frb_iterate_stream_kernel<6>",
             frb.KERNEL_STREAM7: "frb_iterate_stream_kernel<7>"}.get(kernel_variant, "frb_iterate_simple_kernel")
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json"))).get(args.workload)
        if tj and tj["kernel"] == kname and world == 1:
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
    except (OSError, ValueError, KeyError):
        pass
    roofline = {
        "bound": "fp64",
        "achieved": round(achieved_tflops, 3), "peak": round(peak_tflops, 3), "unit": "TFLOP/s",
        "frac": round(achieved_tflops / peak_tflops, 4), "traffic": traffic,
        "traffic_note": "DRAM bytes read+written by this kernel per launch from one ncu --set full capture (profiles/r1_traffic.json); algorithmic = 20 B/px raw result written, 0 read",
        "kernel": kname,
        "kernel_ms": round(it_ms, 4), "shade_kernel_ms": round(statistics.mean(shade_ms), 4),
        "shade_kernel_note": "frb_shade_kernel; with the defer6 iterate kernel it also resolves each pixel's provisional record (<= block+1 iterations with the reference's test)",
        "kernel_share_of_step": round(it_ms * args.steps / dev_ms, 4) if world == 1 else None,
        "peak_source": "measured in this run: frb_measure_fp64_peak DFMA stream x 2 flop (MEASURED_PEAKS.json has no FP64 entry)",
        "peak_nominal": FP64_NOMINAL_TFLOPS, "frac_of_nominal": round(achieved_tflops / FP64_NOMINAL_TFLOPS, 4),
        "fp64_instr_per_iter": form,
        "pipe_frac": round(per_launch_iters * form / (it_ms * 1e-3) / peak["dfma_per_s"], 4),
        "lane_utilisation": round(my_iters / lane_iters, 4) if lane_iters else None,
        "peak_detail": {k: round(v / 1e12, 3) for k, v in peak.items() if k.endswith("_per_s")},
        "note": "bound is the FP64 pipe (scalar recurrence, 3 B written per ~1400 iterations): HBM and tensor rooflines do not apply; "
                "achieved counts the reference's 10 flop/iteration, the kernel issues fp64_instr_per_iter FP64 instructions per iteration "
                "(pipe_frac = FP64 instructions issued for useful iterations / measured DFMA instruction peak)",
    }

    cpu = None if args.no_cpu_baseline else cpu_baseline(cfg)
    if cpu is not None:
        cpu["ms_per_frame_extrapolated"] = round(total_iterations / (cpu["value"] * 1e9) * 1e3, 1)

    frames_per_step = world if args.scaling == "weak" else 1
    line = {
        "metric": "G iter/s", "value": round(value, 2), "unit": "G iter/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": round(dev_ms / args.steps, 4), "higher_is_better": True, "scaling": args.scaling if world > 1 else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic (the fractal itself: deterministic, no RNG, no dataset)",
        "config": {
            "workload": WORKLOAD_TEXT[args.workload], "iterations_per_frame": total_iterations, "frames_per_step": frames_per_step,
            "partition": ("one whole frame per rank" if args.scaling == "weak" else "one frame, interleaved 16-row bands per rank") if world > 1 else "single GPU",
            "kernel": frb.KERNEL_NAMES[kernel_variant], "format": "RGB24",
            "l2": "not flushed: no reusable input (48 KB of coordinate tables are recomputed every step); each step writes 207 MB (> 126 MB L2)",
        },
        "ms_per_frame": round(dev_ms / args.steps, 4),
        "e2e": {"value": round(e2e_value, 2), "unit": "G iter/s", "h2d_bytes_per_step": 104, "d2h_bytes_per_step": nbytes + 16,
                "ms_per_step": round(e2e_s / args.steps * 1e3, 4),
                "api": "frb_render_shard (what get_image binds to): config struct in, finished RGB24 frame into the caller's pinned host memory (written over PCIe by the streaming finisher kernel while the iterate kernel runs)"},
        "gpu_launches": launches,
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
