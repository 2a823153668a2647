#!/usr/bin/env python
"""bench.py — headline benchmark of the escape-time path (BASELINE.json: G iter/s and ms/frame, 3000x3000 Mandelbrot
@1e6x zoom, 1/2/4/8 B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload C2|C2b|C1|C3|C4|C5]
                    [--scaling auto|strong|weak]

A "step" is `frames_per_step` passes of the hot path over one frame (the inner repeat count is chosen so that the timed region
lasts >= 1 s whatever K is; both are reported).  Prints ONE JSON line (rank 0).

  value      G iter/s, device-resident: the frame (N > 1: this rank's interleaved 16-row bands of it) is rendered into HBM
             through the C ABI's device entry point (frb_render_device_async) on a torch stream; timed with CUDA events, max
             over ranks.
  e2e        the same metric through the call the reference's get_image binds to, host buffers, copies inside the timed region.
             N = 1: frb_render(cfg, out, ...) into a FRESH pageable allocation per frame — exactly what rust/ffi.rs does
             (Vec::with_capacity, reference src/lib.rs:253); `e2e_pinned` beside it is the same call into a reused page-locked
             buffer.  N > 1: every rank's frb_render_bands writes its bands into ONE shared host frame (frb_shared_frame_*) and a
             barrier inside the timed region ends each frame when the whole image is on the host.
  roofline   the iterate kernel against the FP64 pipe: 10 flop per iteration as the reference writes it (6 mul + 4 add,
             calc/src/lib.rs:88-104) over the kernel's own CUDA-event time, against the DFMA peak measured on this GPU in this
             run (MEASURED_PEAKS.json has no FP64 entry); step_frac / e2e_frac put the whole step / the e2e call in the denominator.
  cpu_baseline  the CPU oracle (C restatement of the reference path, OpenMP over rows like rayon) on this box's cores, on a
             stated row-strided sample of the same frame (rank 0, N = 1 only).
  extra      c4_strong (16384^2 deep zoom cut over the N ranks, with this run's own N = 1 time for the efficiency), at N > 1
             n1_same_run (the headline frame on one GPU of this very run), inprocess (frb_render(device = -1) from ONE process
             over the N GPUs) and multidevice_check (bytes of the N-GPU image == bytes of the 1-GPU image).

N > 1 (torchrun) defaults to STRONG scaling: one frame cut into interleaved 16-row bands (rank r owns bands k % N == r), the
partition north_star names; `--scaling weak` renders one whole frame per rank (replicas).  No collective on the data path:
torch.distributed (NCCL) only carries the barrier and the max-over-ranks of the timing.
--impl reference times the reference's CPU path alone (the C port of it: the Rust reference cannot be built on this image).
"""
from __future__ import annotations

import argparse
import ctypes
import json
import mmap
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
MIN_TIMED_S = 1.0  # the timed region of every arm lasts at least this long


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2", "C2b", "C3", "C4", "C5"])
    ap.add_argument("--scaling", default="auto", choices=["auto", "weak", "strong"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip extra.c4_strong / n1_same_run / inprocess")
    ap.add_argument("--kernel", default="auto", choices=["auto", "stream7", "simple", "dense6", "defer6", "tile"])
    return ap.parse_args()


WORKLOAD_TEXT = {
    "C1": "C1 Mandelbrot 1920x1080 it256 full view",
    "C2": "C2 Mandelbrot 3000x3000 it4000 scale 5e5 (README 1e6x-zoom screenshot view), RGB24",
    "C2b": "C2b Mandelbrot 3000x3000 it4000 scale 1e6",
    "C3": "C3 Julia 4096x4096 it1000 c=(-0.7269,0.1889)",
    "C4": "C4 Mandelbrot 16384x16384 it10000 scale 1e12",
    "C5": "C5 240-frame zoom 1e0->1e12 at 1920x1080 it10000 (a step = the whole animation)",
}


class ClockSampler:
    """Samples SM clock and throttle reasons with nvidia-smi while the timed regions run."""

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

    def window(self, t0: float, t1: float) -> dict:
        """Statistics of the samples taken between t0 and t1 (perf_counter)."""
        mhz, mx, power, reasons = [], None, [], set()
        for (t, l) in list(self.lines):
            if not (t0 <= t <= t1 + 0.1):
                continue
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

    def stop(self):
        if self.proc is None:
            return
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()


def host_threads() -> int:
    """All host cores this process may use — NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


# ---------------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (test infrastructure) as the reference's CPU implementation of the path
# ---------------------------------------------------------------------------------------------------------------------------

def time_cpu(cfg, stride: int, threads: int = 0):
    """Oracle (OpenMP rows, like rayon) on a row-strided sample of the frame.  Returns (G iter/s, seconds, iterations, rows)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_binding as ob

    threads = threads or host_threads()
    rows = np.arange(stride // 2, cfg.height, stride, dtype=np.uint32)
    t = time.perf_counter()
    r = ob.render_rows(cfg, rows=rows, want_iters=False, want_z=False, want_rgb=True, threads=threads)
    dt = time.perf_counter() - t
    return r["total"] / dt / 1e9, dt, r["total"], len(rows)


def cpu_sample_config(args, cfg_or_list):
    """The frame the CPU arm samples: the workload's frame, or for C5 a frame from the expensive half of the animation."""
    if args.workload == "C5":
        return cfg_or_list[160], "frame 160 of 240"
    return cfg_or_list, "the frame"


def server_latency(device):
    """Median round trip of the headless frame server on 1920x1080 views, a zoom step per frame (scripts/server_latency.py)."""
    import numpy as np
    from fractal_renderer_b200 import configs
    from fractal_renderer_b200.interactive import FrameServer, ViewState
    views = {"C1 view it256": configs.c1(), "C2 view it4000": configs.c2().with_(width=1920, height=1080)}
    out = {"note": "request -> poll of the RGBA frame, ms; poll copies the frame into a fresh array, acquire hands out the server's own buffer"}
    with FrameServer(device=device) as srv:
        for name, cfg in views.items():
            view = ViewState(cfg)
            for _ in range(3):
                srv.request(view.config)
                srv.poll(10000)
            for copy in (True, False):
                lat = []
                for k in range(30):
                    view.apply_scroll(1.0 if k % 2 else -1.0)
                    t = time.perf_counter()
                    srv.request(view.config)
                    frame = srv.poll(10000, copy=copy)
                    lat.append((time.perf_counter() - t) * 1e3)
                    if frame is None:
                        raise RuntimeError("frame server timed out")
                out[f"{name} [{'poll' if copy else 'acquire'}]"] = {"median_ms": round(float(np.median(lat)), 3), "p95_ms": round(float(np.percentile(lat, 95)), 3)}
    return out


def cpu_baseline(args, cfg) -> dict:
    cores = host_threads()
    cfg1, what = cpu_sample_config(args, cfg)
    stride = 64
    g, dt, iters, nrows = time_cpu(cfg1, stride)
    # grow the sample until it is several seconds of CPU work (bounded: never more than the whole frame)
    while dt < 4.0 and stride > 1:
        stride = max(1, stride // 2)
        g, dt, iters, nrows = time_cpu(cfg1, stride)
    return {"value": round(g, 4), "unit": "G iter/s", "cores": cores, "kind": "port",
            "sample": f"every {stride}th row of {what} ({nrows} rows x {cfg1.width} px, {iters} iterations) in {dt:.2f} s; "
                      "oracle/calc_oracle.c (C restatement; the Rust reference cannot be built here), OpenMP dynamic rows",
            "ms_per_frame_extrapolated": None}


def run_reference(args, cfg, total_iterations):
    """--impl reference: the reference's CPU path (oracle port) on the host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_threads()
    cfg1, what = cpu_sample_config(args, cfg)
    # size the per-step sample to ~1-2 s so that K+W steps finish within a few minutes on any core count
    stride = 64
    g, dt, iters, nrows = time_cpu(cfg1, stride)
    while dt < 0.75 and stride > 1:
        stride //= 2
        g, dt, iters, nrows = time_cpu(cfg1, stride)
    for _ in range(args.warmup):
        time_cpu(cfg1, stride)
    t0 = time.perf_counter()
    tot = 0
    for _ in range(args.steps):
        _, _, it, _ = time_cpu(cfg1, stride)
        tot += it
    dt = time.perf_counter() - t0
    value = tot / dt / 1e9
    sample = (f"each step = every {stride}th row of {what} ({nrows} rows x {cfg1.width} px, {iters} iterations); oracle/calc_oracle.c "
              "(C restatement of calc/src/lib.rs; the Rust reference cannot be built on this image), OpenMP dynamic rows = rayon rows")
    line = {
        "impl": "reference", "metric": "G iter/s", "value": round(value, 4), "unit": "G iter/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 and args.scaling != "weak" else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (the fractal itself; no RNG)",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "iterations_per_frame": total_iterations},
        "ms_per_frame": round(total_iterations / (value * 1e9) * 1e3, 2),
        "cpu_baseline": {"value": round(value, 4), "unit": "G iter/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 4), "unit": "G iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------------------

class Bench:
    def __init__(self, args):
        import numpy as np
        import torch

        import fractal_renderer_b200 as frb

        self.np, self.torch, self.frb, self.args = np, torch, frb, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}")
        if not torch.cuda.is_available() or frb.device_count() < 1:
            raise SystemExit("bench.py needs a B200: the product has no CPU path (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist_mod

            # NCCL may print its version banner on stdout (NCCL_DEBUG=VERSION on some boxes): keep stdout for the ONE JSON line
            sys.stdout.flush()
            saved_stdout = os.dup(1)
            os.dup2(2, 1)
            try:
                dist_mod.init_process_group("nccl", device_id=torch.device("cuda", self.local))
                dist_mod.barrier()
                torch.cuda.synchronize()
            finally:
                sys.stdout.flush()
                os.dup2(saved_stdout, 1)
                os.close(saved_stdout)
            self.dist = dist_mod
        frb.init(1 << self.local)
        variant = {"auto": frb.KERNEL_AUTO, "stream7": frb.KERNEL_STREAM7, "simple": frb.KERNEL_SIMPLE, "dense6": frb.KERNEL_DENSE6,
                   "defer6": frb.KERNEL_DEFER6, "tile": frb.KERNEL_TILE}[args.kernel]
        frb.set_kernel_variant(variant)
        self.stream = torch.cuda.Stream(device=self.local)  # a real (non-default) stream: launches and events go on it
        self.stream2 = torch.cuda.Stream(device=self.local)  # device-resident arm: frames alternate between the two (two in flight)
        assert self.stream.cuda_stream != 0 and self.stream2.cuda_stream != 0
        self.job = os.environ.get("MASTER_PORT", "0") + "-" + os.environ.get("TORCHELASTIC_RUN_ID", str(os.getppid()))
        self._frames = {}

    # ---- plumbing ----
    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x: float, op: str) -> float:
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def shared_frame(self, tag: str, nbytes: int):
        """One host image for all ranks (POSIX shared memory, pinned by every rank); a plain pinned buffer at N = 1."""
        if tag not in self._frames:
            self._frames[tag] = self.frb.SharedFrame(f"/frb-bench-{self.job}-{tag}", nbytes, self.world, self.rank)
            self._frames[tag].barrier()  # everybody has attached before anybody may close it
        return self._frames[tag]

    def close(self):
        for f in self._frames.values():
            f.barrier()
            f.close()
        if self.dist is not None:
            self.dist.destroy_process_group()

    # ---- one workload, device-resident: K steps of R frames, CUDA events, max over ranks ----
    def time_device(self, cfg, shard, steps, warmup, min_s=MIN_TIMED_S, pipelined=True):
        """Device-resident throughput: frames are independent, so — as frb_render_frames does for an animation — they are issued
        on TWO streams into two output buffers, two in flight per GPU: the next frame's persistent CTAs take over every SM the
        moment the previous frame's CTA leaves it, and its shade kernel runs beside them.  Timed with CUDA events on the first
        stream, which waits for the second before the closing event.  pipelined=False: one stream, strictly one after another."""
        frb, torch = self.frb, self.torch
        rows = shard.rows(cfg.height)
        nbytes = rows * cfg.width * 3
        streams = [self.stream, self.stream2] if pipelined else [self.stream]
        outs = [torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda") for _ in streams]
        issued = [0]

        def one():
            k = issued[0] % len(streams)
            issued[0] += 1
            frb.render_device_async(cfg, outs[k].data_ptr(), nbytes, frb.FRB_RGB24, device=self.local, shard=shard, stream=streams[k].cuda_stream)

        def open_region(ev):
            ev.record(self.stream)
            for st in streams[1:]:
                st.wait_event(ev)

        def close_region(ev):
            for st in streams[1:]:
                done = torch.cuda.Event()
                done.record(st)
                self.stream.wait_event(done)
            ev.record(self.stream)

        for _ in range(max(warmup, 3) + 1):
            one()
        torch.cuda.synchronize()
        # frames per step: from a short probe, so that K x R frames last >= min_s on the slowest rank
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        open_region(e0)
        for _ in range(4):
            one()
        close_region(e1)
        torch.cuda.synchronize()
        est_ms = self.reduce(e0.elapsed_time(e1) / 4.0, "max")
        reps = max(1, int(-(-min_s * 1e3 // (steps * est_ms))))
        self.barrier()
        tw0 = time.perf_counter()
        open_region(e0)
        iter_ms, shade_ms, acc = [], [], {"launches": 0, "iters": 0, "lane": 0}

        def harvest(back):
            st = frb.async_stats(back)
            iter_ms.append(st["iterate_ms"])
            shade_ms.append(st["shade_ms"])
            acc["launches"] += st["kernel_launches"]
            acc["iters"] += st["iterations"]
            acc["lane"] += st["lane_iterations"]
            return st

        n = steps * reps
        ahead = len(streams)  # frames kept enqueued ahead while the statistics of an earlier one are harvested (host-side event wait only)
        for k in range(n):
            one()
            if k >= ahead:
                harvest(ahead)
        close_region(e1)
        st = None
        for back in range(min(ahead, n) - 1, -1, -1):
            st = harvest(back)
        e1.synchronize()
        torch.cuda.synchronize()
        self.barrier()
        tw1 = time.perf_counter()
        ms = self.reduce(e0.elapsed_time(e1), "max")
        iters_all = self.reduce(float(acc["iters"]), "sum")
        return {"ms_total": ms, "frames": n, "reps": reps, "ms_per_frame": ms / n, "g_iter_s": iters_all / (ms * 1e-3) / 1e9,
                "iterate_ms": statistics.mean(iter_ms), "shade_ms": statistics.mean(shade_ms), "launches": acc["launches"],
                "my_iters": acc["iters"], "lane_iters": acc["lane"], "variant": st["kernel_variant"], "window": (tw0, tw1), "nbytes": nbytes,
                "in_flight": len(streams)}

    # ---- one workload, end to end through a host-buffer call: K steps of R frames, wall clock, max over ranks ----
    def time_e2e(self, call, steps, warmup, min_s=MIN_TIMED_S, after=None):
        frb = self.frb
        for _ in range(max(warmup, 3)):
            call()
            if after:
                after()
        self.barrier()
        t = time.perf_counter()
        for _ in range(3):
            call()
            if after:
                after()
        est = self.reduce((time.perf_counter() - t) / 3.0, "max")
        reps = max(1, int(-(-min_s // (steps * est))))
        n = steps * reps
        self.barrier()
        t0 = time.perf_counter()
        iters = launches = 0
        per_call = []
        tc = t0
        for _ in range(n):
            call()
            s = frb.last_stats()
            iters += s["iterations"]
            launches += s["kernel_launches"]
            if after:
                after()  # N > 1: the frame ends when every rank's bands are on the host
            tn = time.perf_counter()
            per_call.append(tn - tc)
            tc = tn
        t1 = time.perf_counter()
        self.barrier()
        secs = self.reduce(t1 - t0, "max")
        iters_all = self.reduce(float(iters), "sum")
        # median and min per frame beside the mean (this rank's calls; with a barrier per frame every rank sees the same)
        return {"s_total": secs, "frames": n, "reps": reps, "ms_per_frame": secs / n * 1e3, "g_iter_s": iters_all / secs / 1e9,
                "ms_median": statistics.median(per_call) * 1e3, "ms_min": min(per_call) * 1e3,
                "launches": launches, "window": (t0, t1)}


def run_c5(b: Bench, args, cfgs):
    """C5: the whole 240-frame animation per step; frames round-robin over the ranks, each rank through frb_render_frames into
    its slice of one shared host buffer (frame k at offset k * frame_bytes)."""
    frb, np = b.frb, b.np
    fbytes = cfgs[0].width * cfgs[0].height * 3
    frame = b.shared_frame("c5", fbytes * len(cfgs))

    def call():
        # this rank's frames (k % N == rank), each straight into its place in the one shared host buffer
        frb.render_frames(cfgs, frb.FRB_RGB24, device=b.local, out=frame.array, first=b.rank, stride=b.world)

    steps = max(1, min(args.steps, 5))
    res = b.time_e2e(call, steps, 1, min_s=0.0, after=frame.barrier if b.world > 1 else None)
    return res, steps


def workload_record(b: Bench, name: str, cfg, total_iterations: int, strong: bool, peak: dict, sampler) -> dict:
    """One more configuration measured the way the headline is: device-resident (two frames in flight), the iterate kernel alone
    (one frame at a time), and end to end — N = 1: frb_render into a malloc'd buffer per frame, a new mapping per frame and a pinned buffer; N > 1: frb_render_bands
    per rank into one shared host frame with a barrier per frame.  >= 0.3 s per timed region."""
    frb, np = b.frb, b.np
    L = frb._lib.load()
    shard = frb.Shard(b.rank, b.world, 16) if strong else frb.Shard()
    dev = b.time_device(cfg, shard, 5, 3, min_s=0.3)
    one = b.time_device(cfg, shard, 5, 3, min_s=0.3, pipelined=False)
    c_cfg = cfg.to_c()
    nbytes = cfg.width * cfg.height * 3
    peak_tflops = peak["dfma_per_s"] * 2 / 1e12
    flop = (total_iterations / b.world if strong else total_iterations) * FLOP_PER_ITER
    rec = {"workload": WORKLOAD_TEXT[name], "iterations_per_frame": total_iterations, "kernel": frb.KERNEL_NAMES[dev["variant"]],
           "device_ms": round(dev["ms_per_frame"], 4), "g_iter_s": round(dev["g_iter_s"], 1),
           "single_stream_ms": round(one["ms_per_frame"], 4), "kernel_ms": round(one["iterate_ms"], 4), "shade_kernel_ms": round(one["shade_ms"], 4),
           "frac": round(one["my_iters"] / one["frames"] * FLOP_PER_ITER / (one["iterate_ms"] * 1e-3) / 1e12 / peak_tflops, 4),
           "step_frac": round(flop / (dev["ms_per_frame"] * 1e-3) / 1e12 / peak_tflops, 4),
           "lane_utilisation": round(one["my_iters"] / one["lane_iters"], 4) if one["lane_iters"] else None}
    if strong:
        frame = b.shared_frame("w" + name, nbytes)
        c_shard = shard.to_c()
        e = b.time_e2e(lambda: frb._lib.check(L.frb_render_bands(ctypes.byref(c_cfg), ctypes.byref(c_shard), frame.array.ctypes.data, nbytes,
                                                                    frb.FRB_RGB24, b.local)), 5, 3, min_s=0.3, after=frame.barrier)
        rec.update({"e2e_bands_ms": round(e["ms_per_frame"], 4), "e2e_bands_ms_median": round(e["ms_median"], 4), "e2e_bands_ms_min": round(e["ms_min"], 4),
                    "e2e_g_iter_s": round(e["g_iter_s"], 1), "partition": "one frame, interleaved 16-row bands into one shared host frame"})
        window = e["window"]
    else:
        keep = []

        def fresh():
            out = np.frombuffer(mmap.mmap(-1, nbytes), dtype=np.uint8)
            keep[:] = [out]
            frb._lib.check(L.frb_render(ctypes.byref(c_cfg), out.ctypes.data, nbytes, frb.FRB_RGB24, b.local))

        def recycled():
            out = np.empty(nbytes, np.uint8)
            keep[:] = [out]
            frb._lib.check(L.frb_render(ctypes.byref(c_cfg), out.ctypes.data, nbytes, frb.FRB_RGB24, b.local))

        pinned = frb.pinned_empty(nbytes)
        er = b.time_e2e(recycled, 5, 3, min_s=0.3)
        ef = b.time_e2e(fresh, 5, 3, min_s=0.3)
        ep = b.time_e2e(lambda: frb._lib.check(L.frb_render(ctypes.byref(c_cfg), pinned.ctypes.data, nbytes, frb.FRB_RGB24, b.local)), 5, 3, min_s=0.3)
        # e2e_ms: malloc per frame (the drop-in's steady state: recycled resident blocks; "e2e_fresh_ms" of rounds 1-2);
        # e2e_fresh_map_ms: a new anonymous mapping per frame; e2e_pinned_ms: a reused page-locked buffer
        rec.update({"e2e_ms": round(er["ms_per_frame"], 4), "e2e_ms_median": round(er["ms_median"], 4), "e2e_ms_min": round(er["ms_min"], 4),
                    "e2e_fresh_map_ms": round(ef["ms_per_frame"], 4), "e2e_fresh_map_ms_median": round(ef["ms_median"], 4), "e2e_fresh_map_ms_min": round(ef["ms_min"], 4),
                    "e2e_pinned_ms": round(ep["ms_per_frame"], 4), "e2e_pinned_ms_median": round(ep["ms_median"], 4), "e2e_pinned_ms_min": round(ep["ms_min"], 4),
                    "e2e_g_iter_s": round(er["g_iter_s"], 1)})
        window = er["window"]
    if b.rank == 0:
        rec["clocks"] = sampler.window(*window)
    return rec


def c5_record(b: Bench, cfgs, sampler) -> dict:
    """C5 in the headline line: the 240-frame zoom through frb_render_frames_strided, frames k % N == rank per rank, every frame into its
    place in one shared host buffer, a barrier per animation; >= 3 animations.  N > 1: rank 0 also runs the whole animation alone
    (this run's N = 1 denominator)."""
    frb = b.frb
    fbytes = cfgs[0].width * cfgs[0].height * 3
    frame = b.shared_frame("x5", fbytes * len(cfgs))

    def call():
        frb.render_frames(cfgs, frb.FRB_RGB24, device=b.local, out=frame.array, first=b.rank, stride=b.world)

    res = b.time_e2e(call, 3, 1, min_s=0.0, after=frame.barrier if b.world > 1 else None)
    rec = {"workload": WORKLOAD_TEXT["C5"], "frames": len(cfgs), "animations_timed": res["frames"],
           "ms_per_animation": round(res["ms_per_frame"], 2), "ms_median": round(res["ms_median"], 2), "ms_min": round(res["ms_min"], 2),
           "frames_per_s": round(len(cfgs) * res["frames"] / res["s_total"], 1), "g_iter_s": round(res["g_iter_s"], 1),
           "frames_per_gpu": [len(range(r, len(cfgs), b.world)) for r in range(b.world)],
           "api": "frb_render_frames_strided per rank (two frames in flight per GPU), each frame copied to its place in ONE shared pinned host buffer"}
    if b.rank == 0:
        rec["clocks"] = sampler.window(*res["window"])
    if b.world > 1:
        if b.rank == 0:
            t = []
            for _ in range(2):
                t0 = time.perf_counter()
                frb.render_frames(cfgs, frb.FRB_RGB24, device=b.local, out=frame.array)
                t.append((time.perf_counter() - t0) * 1e3)
            rec["n1_ms_per_animation_same_run"] = round(min(t), 2)
            rec["efficiency"] = round(min(t) / (b.world * res["ms_per_frame"]), 4)
        frame.barrier()
    return rec


def main():
    args = parse()
    from fractal_renderer_b200 import configs

    if args.workload == "C5":
        cfg = configs.c5()
        total_iterations = None
    else:
        cfg = configs.BY_NAME[args.workload]()
        total_iterations = configs.TOTAL_ITERATIONS[args.workload]

    if args.impl == "reference":
        run_reference(args, cfg, total_iterations or 0)
        return

    b = Bench(args)
    frb, np, torch = b.frb, b.np, b.torch
    rank, world, local = b.rank, b.world, b.local
    strong = world > 1 and args.scaling != "weak"
    peak = frb.measure_fp64_peak(local)  # FP64 peak of this GPU, measured now (rank-local; ~0.3 s)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    extra = {}

    if args.workload == "C5":
        res, steps = run_c5(b, args, cfg)
        if rank == 0:
            sampler.stop()
            clocks = sampler.window(*res["window"])
            line = {"metric": "G iter/s", "value": round(res["g_iter_s"], 2), "unit": "G iter/s", "n_gpus": world, "steps": steps, "warmup": 1,
                    "ms_per_step": round(res["s_total"] / steps * 1e3, 3), "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
                    "vs_baseline": None, "dtype": "f64", "data": "synthetic (the fractal itself: deterministic, no RNG, no dataset)",
                    "config": {"workload": WORKLOAD_TEXT["C5"], "frames": len(cfg), "partition": "frames round-robin over the ranks" if world > 1 else "single GPU",
                               "format": "RGB24"},
                    "frames_per_s": round(len(cfg) * steps / res["s_total"], 2),
                    "e2e": {"value": round(res["g_iter_s"], 2), "unit": "G iter/s", "h2d_bytes_per_step": 104 * len(cfg),
                            "d2h_bytes_per_step": cfg[0].width * cfg[0].height * 3 * len(cfg),
                            "api": "frb_render_frames_strided per rank (its frames k % N == rank, two in flight), every frame copied device->host into its place in ONE shared pinned host buffer; barrier per animation"},
                    "gpu_launches": res["launches"], "clocks": clocks, "roofline": None, "cpu_baseline": None if args.no_cpu_baseline else cpu_baseline(args, cfg)}
            print(json.dumps(line), flush=True)
        b.close()
        return

    # ---------------- device-resident arm (`value`) ----------------
    shard = frb.Shard(rank, world, 16) if strong else frb.Shard()
    dev = b.time_device(cfg, shard, args.steps, args.warmup)
    # the same frames strictly one after another on one stream: what a caller that never overlaps frames gets, and the run the
    # kernel's own duration is taken from (with two frames in flight a kernel's event bracket includes its wait for SMs)
    one_stream = b.time_device(cfg, shard, max(2, args.steps // 4), 3, min_s=0.3, pipelined=False)
    extra["single_stream"] = {"ms_per_frame": round(one_stream["ms_per_frame"], 4), "g_iter_s": round(one_stream["g_iter_s"], 1),
                              "note": "device-resident arm with ONE frame in flight; the headline `value` keeps two in flight on two streams; "
                                      "roofline.kernel_ms comes from this run"}

    # ---------------- end-to-end arms ----------------
    c_cfg = cfg.to_c()
    L = frb._lib.load()
    full_bytes = cfg.width * cfg.height * 3
    e2e_pinned = e2e_fresh_map = None
    if strong:
        frame = b.shared_frame("main", full_bytes)
        c_shard = shard.to_c()

        def call_bands():
            frb._lib.check(L.frb_render_bands(ctypes.byref(c_cfg), ctypes.byref(c_shard), frame.array.ctypes.data, full_bytes, frb.FRB_RGB24, local))

        e2e = b.time_e2e(call_bands, args.steps, args.warmup, after=frame.barrier)
        e2e_api = ("frb_render_bands per rank into ONE shared host frame (frb_shared_frame: POSIX shared memory pinned by every rank; the streaming "
                   "finisher writes each rank's interleaved 16-row bands over PCIe while it iterates) + a shared-memory barrier per frame: "
                   "the frame ends when the whole image is on the host")
        d2h = full_bytes
    else:
        keep = []

        def call_fresh():
            # A destination NO ONE has touched: an anonymous mapping of its own per frame (what malloc hands a Rust
            # Vec::with_capacity of this size the first time, or whenever the caller keeps the previous frames alive); it is
            # unmapped when the next frame replaces it.  np.empty would not do: after the first free glibc raises its mmap
            # threshold (up to 32 MB) and serves the same two, already resident, heap blocks from then on.
            out = np.frombuffer(mmap.mmap(-1, full_bytes), dtype=np.uint8)
            keep[:] = [out]
            frb._lib.check(L.frb_render(ctypes.byref(c_cfg), out.ctypes.data, full_bytes, frb.FRB_RGB24, local))

        def call_recycled():
            # what a caller that drops its Vec every frame gets from glibc after the first few frames: pageable, but resident
            out = np.empty(full_bytes, np.uint8)
            keep[:] = [out]
            frb._lib.check(L.frb_render(ctypes.byref(c_cfg), out.ctypes.data, full_bytes, frb.FRB_RGB24, local))

        pinned = frb.pinned_empty(full_bytes)

        def call_pinned():
            frb._lib.check(L.frb_render(ctypes.byref(c_cfg), pinned.ctypes.data, full_bytes, frb.FRB_RGB24, local))

        e2e = b.time_e2e(call_recycled, args.steps, args.warmup)
        e2e_fresh_map = b.time_e2e(call_fresh, args.steps, args.warmup, min_s=0.5)
        e2e_pinned = b.time_e2e(call_pinned, args.steps, args.warmup)
        e2e_api = ("frb_render(cfg, out, len, RGB24, device) into a pageable buffer allocated per frame with malloc (np.empty = Vec::with_capacity, as "
                   "rust/ffi.rs get_image_b200 does) and dropped when the next frame arrives: after the first frames glibc serves these allocations "
                   "from the same resident heap blocks (its mmap threshold rises on the first free), so this is the steady state of the drop-in, NOT "
                   "untouched memory — rounds 1-2 called this arm 'fresh'; the cost of a destination nobody has touched (a new anonymous mapping "
                   "per frame: the first call of a process, or a caller that keeps every frame) is e2e_fresh_map; weak N > 1: one such frame per rank")
        d2h = full_bytes

    # ---------------- extras: evidence the driver can only get from this run ----------------
    if not args.no_extras and args.workload in ("C2", "C2b"):
        # (1) N > 1: the headline frame on ONE GPU of this very run (the efficiency denominator), rank 0 alone
        if strong:
            n1_dev = n1_e2e = None
            if rank == 0:
                saved = b.dist
                b.dist = None
                n1_dev = b.time_device(cfg, frb.Shard(), max(5, args.steps // 2), 3, min_s=0.5)
                pin1 = frb.pinned_empty(full_bytes)
                n1_e2e = b.time_e2e(lambda: frb._lib.check(L.frb_render(ctypes.byref(c_cfg), pin1.ctypes.data, full_bytes, frb.FRB_RGB24, local)),
                                    max(5, args.steps // 2), 3, min_s=0.5)
                b.dist = saved
                extra["n1_same_run"] = {"arm": "two frames in flight on two streams, the same arm as `value`",
                                        "ms_per_frame": round(n1_dev["ms_per_frame"], 4), "g_iter_s": round(n1_dev["g_iter_s"], 1),
                                        "e2e_pinned_ms_per_frame": round(n1_e2e["ms_per_frame"], 4)}
                extra["strong_efficiency"] = {"device_resident": round(n1_dev["ms_per_frame"] / (world * dev["ms_per_frame"]), 4),
                                              "e2e": round(n1_e2e["ms_per_frame"] / (world * e2e["ms_per_frame"]), 4),
                                              "note": "T1 / (N x TN), T1 measured on rank 0's GPU in this run; limiter at small shards: the tail of the frame "
                                                      "(the last pixels handed out are full-length orbits: a fixed ~0.1 ms whatever the shard size)"}
            frame.barrier()
        # (2) C4 (16384^2 deep zoom): the configuration north_star's >= 85 % target is quoted on
        c4 = configs.c4()
        c4_bytes = c4.width * c4.height * 3
        c4_shard = frb.Shard(rank, world, 16) if world > 1 else frb.Shard()
        c4_frame = b.shared_frame("c4", c4_bytes)
        c4_c, c4_s = c4.to_c(), c4_shard.to_c()

        def call_c4():
            frb._lib.check(L.frb_render_bands(ctypes.byref(c4_c), ctypes.byref(c4_s), c4_frame.array.ctypes.data, c4_bytes, frb.FRB_RGB24, local))

        c4_res = b.time_e2e(call_c4, 5, 1, min_s=1.0, after=c4_frame.barrier if world > 1 else None)
        rec = {"workload": WORKLOAD_TEXT["C4"], "partition": "one frame, interleaved 16-row bands" if world > 1 else "single GPU",
               "ms_per_frame": round(c4_res["ms_per_frame"], 3), "ms_median": round(c4_res["ms_median"], 3), "ms_min": round(c4_res["ms_min"], 3),
               "g_iter_s": round(c4_res["g_iter_s"], 1), "frames_timed": c4_res["frames"], "timed_region_s": round(c4_res["s_total"], 3),
               "clocks": sampler.window(*c4_res["window"]) if rank == 0 else None,
               "measured": "end to end: frb_render_bands per rank into one shared pinned host frame + barrier per frame (gather to host included)"}
        if world > 1:
            n1 = None
            if rank == 0:
                whole = frb.Shard().to_c()
                t = []
                for k in range(3):
                    t0 = time.perf_counter()
                    frb._lib.check(L.frb_render_bands(ctypes.byref(c4_c), ctypes.byref(whole), c4_frame.array.ctypes.data, c4_bytes, frb.FRB_RGB24, local))
                    t.append((time.perf_counter() - t0) * 1e3)
                n1 = min(t[1:])
                rec["n1_ms_per_frame_same_run"] = round(n1, 3)
                rec["efficiency"] = round(n1 / (world * c4_res["ms_per_frame"]), 4)
            c4_frame.barrier()
        extra["c4_strong"] = rec
        # (3) N > 1: the in-process path — ONE process (rank 0) drives all N GPUs through frb_render(device = -1), the call a
        # Rust get_image makes on a multi-GPU box — timed while the other ranks wait on a host-side barrier; and the check that
        # its bytes equal the one-GPU image (C1-, C3- and C2-shaped frames).
        if strong:
            if rank == 0:
                try:
                    frb.init((1 << world) - 1)
                    chk = {}
                    for name, c in (("C3", configs.c3().with_(width=2048, height=1024)), ("C1", configs.c1()), ("C2", cfg)):
                        nb = c.width * c.height * 3
                        buf = frb.pinned_empty(nb)
                        frb.get_image_into(c, buf, frb.FRB_RGB24, device=-1)
                        ndev = frb.last_stats()["devices"]
                        one = frb.render_shard(c, frb.Shard(), frb.FRB_RGB24, device=local).reshape(-1)
                        chk[name] = bool(np.array_equal(buf, one)) and ndev == world
                    extra["multidevice_check"] = "ok" if all(chk.values()) else f"MISMATCH {chk}"
                    buf = frb.pinned_empty(full_bytes)
                    for _ in range(5):
                        frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
                    t0 = time.perf_counter()
                    nrep = 200
                    for _ in range(nrep):
                        frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
                    ms = (time.perf_counter() - t0) / nrep * 1e3
                    extra["inprocess"] = {"api": "frb_render(device = -1) from one process, one host worker per GPU, pinned destination",
                                          "ms_per_frame": round(ms, 4), "devices": world,
                                          "efficiency_vs_n1_same_run": round(extra["n1_same_run"]["e2e_pinned_ms_per_frame"] / (world * ms), 4)}
                except frb.FrbError as e:  # evidence, not a reason to lose the headline line
                    extra["multidevice_check"] = f"FAILED: {e}"
                finally:
                    frb.init(1 << local)
            frame.barrier()
            # the per-rank path of this bench, checked too: the shared frame every rank just wrote == rank 0's one-GPU image
            call_bands()
            frame.barrier()
            if rank == 0:
                one = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24, device=local).reshape(-1)
                extra["shared_frame_check"] = "ok" if np.array_equal(frame.array[:full_bytes], one) else "MISMATCH"
            frame.barrier()

    # (4) the other configurations of BASELINE.json in the same driver-run line: C1 and C3 (device-resident, kernel, end to end; at
    # N > 1 the same band partition into a shared host frame) and C5 (the zoom animation sharded by frame, examples.md:20-29)
    if not args.no_extras and args.workload in ("C2", "C2b"):
        extra["workloads"] = {}
        for wname in ("C1", "C3"):
            try:
                rec = workload_record(b, wname, configs.BY_NAME[wname](), configs.TOTAL_ITERATIONS[wname], strong, peak, sampler)
            except Exception as e:  # evidence, not a reason to lose the headline line — but only where no other rank waits on a barrier
                if world > 1:
                    raise
                rec = f"FAILED: {e}"
            if rank == 0:
                extra["workloads"][wname] = rec
        try:
            rec = c5_record(b, configs.c5(), sampler)
        except Exception as e:
            if world > 1:
                raise
            rec = f"FAILED: {e}"
        if rank == 0:
            extra["c5"] = rec

    if rank != 0:
        b.close()
        return
    sampler.stop()
    # (f)1 frame server round trips (request -> render thread -> RGBA frame), the GUI loop of src/gui.rs:37-117 without the window
    if not args.no_extras and world == 1:
        try:
            extra["server_latency"] = server_latency(local)
        except Exception as e:  # evidence, not the measurement
            extra["server_latency"] = f"FAILED: {e}"

    # ---------------- roofline of the dominant kernel (iterate) ----------------
    it_ms = one_stream["iterate_ms"]
    per_launch_iters = one_stream["my_iters"] / one_stream["frames"]
    achieved_tflops = per_launch_iters * FLOP_PER_ITER / (it_ms * 1e-3) / 1e12
    peak_tflops = peak["dfma_per_s"] * 2 / 1e12
    kv = dev["variant"]
    form = {frb.KERNEL_TILE: 6, frb.KERNEL_DEFER6: 6, frb.KERNEL_DENSE6: 6, frb.KERNEL_STREAM7: 7}.get(kv, 11)
    kname = {frb.KERNEL_TILE: "frb_iterate_tile_kernel<6, 8, 4> (iterate + shade)", frb.KERNEL_DEFER6: "frb_iterate_deferred_kernel<6>",
             frb.KERNEL_DENSE6: "frb_iterate_stream_kernel<6>", frb.KERNEL_STREAM7: "frb_iterate_stream_kernel<7>"}.get(kv, "frb_iterate_simple_kernel")
    traffic, traffic_note = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json"))).get(args.workload)
        if tj and tj["kernel"] == kname and world == 1:
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
            traffic_note = tj.get("note")
    except (OSError, ValueError, KeyError):
        pass
    frame_flop = (total_iterations / world if strong else total_iterations) * FLOP_PER_ITER
    roofline = {
        "bound": "fp64",
        "achieved": round(achieved_tflops, 3), "peak": round(peak_tflops, 3), "unit": "TFLOP/s",
        "frac": round(achieved_tflops / peak_tflops, 4),
        "step_frac": round(frame_flop / (dev["ms_per_frame"] * 1e-3) / 1e12 / peak_tflops, 4),
        "e2e_frac": round(frame_flop / (e2e["ms_per_frame"] * 1e-3) / 1e12 / peak_tflops, 4),
        "traffic": traffic,
        "traffic_note": traffic_note or "DRAM bytes read+written by this kernel per launch from one ncu --set full capture (profiles/r2_traffic.json)",
        "algorithmic_bytes": cfg.width * cfg.height * 3,
        "algorithmic_bytes_note": "3 B/px written, 0 read (SURVEY §8d); the kernel is FP64-bound: bytes say what is wasted, not what bounds it",
        "kernel": kname,
        "kernel_ms": round(it_ms, 4), "shade_kernel_ms": round(one_stream["shade_ms"], 4),
        "kernel_share_of_step": round(it_ms / one_stream["ms_per_frame"], 4),
        "kernel_ms_note": "CUDA events around the iterate kernel alone, frames issued one at a time on one stream (extra.single_stream); "
                          "step_frac is over the headline arm (two frames in flight)",
        "peak_source": "measured in this run: frb_measure_fp64_peak DFMA stream x 2 flop (MEASURED_PEAKS.json has no FP64 entry; the same "
                       "measurement with its clock sample is kept in profiles/r3_fp64_peak.json)",
        # one DFMA per FP64 lane per cycle: the rate itself says what clock the pipe ran at (nvidia-smi's sample is in `clocks`)
        "peak_implied_sm_clock_mhz": round(peak["dfma_per_s"] / (peak["sm_count"] * 64) / 1e6, 1),
        "peak_nominal": FP64_NOMINAL_TFLOPS, "frac_of_nominal": round(achieved_tflops / FP64_NOMINAL_TFLOPS, 4),
        "fp64_instr_per_iter": form,
        "pipe_frac": round(per_launch_iters * form / (it_ms * 1e-3) / peak["dfma_per_s"], 4),
        "lane_utilisation": round(one_stream["my_iters"] / one_stream["lane_iters"], 4) if one_stream["lane_iters"] else None,
        "peak_detail": {k: round(v / 1e12, 3) for k, v in peak.items() if k.endswith("_per_s")},
        "note": "bound is the FP64 pipe (scalar recurrence, 3 B written per ~1400 iterations): HBM and tensor rooflines do not apply; "
                "achieved counts the reference's 10 flop/iteration, the kernel issues fp64_instr_per_iter FP64 instructions per iteration "
                "(pipe_frac = FP64 instructions issued for useful iterations / measured DFMA instruction peak); "
                "frac = over the iterate kernel's own time, step_frac = over the device-resident step, e2e_frac = over the e2e call",
    }

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cpu = cpu_baseline(args, cfg)
        cpu["ms_per_frame_extrapolated"] = round(total_iterations / (cpu["value"] * 1e9) * 1e3, 1)

    w_dev, w_e2e = dev["window"], e2e["window"]
    clocks = sampler.window(w_dev[0], w_dev[1])
    clocks["e2e_arm"] = sampler.window(w_e2e[0], w_e2e[1])
    frames_per_step = dev["reps"] * (world if (world > 1 and not strong) else 1)
    line = {
        "metric": "G iter/s", "value": round(dev["g_iter_s"], 2), "unit": "G iter/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": round(dev["ms_total"] / args.steps, 4), "higher_is_better": True, "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic (the fractal itself: deterministic, no RNG, no dataset)",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "iterations_per_frame": total_iterations},  # the same two keys as --impl reference
        "run": {
            "frames_per_step": frames_per_step,
            "timed_region_s": round(dev["ms_total"] * 1e-3, 3),
            "partition": ("one frame, interleaved 16-row bands" if strong else "one whole frame per rank") if world > 1 else "single GPU",
            "kernel": frb.KERNEL_NAMES[kv], "format": "RGB24",
            "issue": "device-resident arm: independent frames on two streams, two in flight per GPU (the next frame's CTAs fill the SMs the "
                     "previous frame has left, as frb_render_frames does); e2e arm: one blocking call at a time"
                     + ("; at N > 1 `value` leaves every rank's bands in its own HBM and gathers nothing: the partition's number, with the "
                        "whole image on the host, is `e2e`" if world > 1 else ""),
            "l2": "not flushed: no reusable input (%d KB of coordinate tables are recomputed every frame); each frame writes %.0f MB of new data (%s)" % (
                (cfg.width + cfg.height) * 8 // 1024, cfg.width * cfg.height * (3 if kv == frb.KERNEL_TILE else 23) / 1e6,
                "3 B/px image; smaller than the 126 MB L2, nothing is read back" if kv == frb.KERNEL_TILE else "20 B/px provisional records + 3 B/px image"),
        },
        "ms_per_frame": round(dev["ms_per_frame"], 4),
        "e2e": {"value": round(e2e["g_iter_s"], 2), "unit": "G iter/s", "h2d_bytes_per_step": 104 * e2e["reps"], "d2h_bytes_per_step": (d2h + 16 * (world if strong else 1)) * e2e["reps"],
                "ms_per_frame": round(e2e["ms_per_frame"], 4), "ms_median": round(e2e["ms_median"], 4), "ms_min": round(e2e["ms_min"], 4),
                "frames_per_step": e2e["reps"], "timed_region_s": round(e2e["s_total"], 3), "api": e2e_api},
        "gpu_launches": dev["launches"] + one_stream["launches"] + e2e["launches"],
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "extra": extra,
    }
    if e2e_pinned is not None:
        line["e2e_fresh_map"] = {"value": round(e2e_fresh_map["g_iter_s"], 2), "unit": "G iter/s", "ms_per_frame": round(e2e_fresh_map["ms_per_frame"], 4),
                                 "ms_median": round(e2e_fresh_map["ms_median"], 4), "ms_min": round(e2e_fresh_map["ms_min"], 4),
                                 "api": "the same frb_render call into a NEW anonymous mapping per frame (mmap and the previous frame's munmap inside the "
                                        "timed region): every page of the destination is first-touched during the frame — ~0.2 ms per MB of kernel page "
                                        "faults on this box, which the staging pipeline hides only as far as the iterations last"}
        line["gpu_launches"] += e2e_fresh_map["launches"]
        line["e2e_pinned"] = {"value": round(e2e_pinned["g_iter_s"], 2), "unit": "G iter/s", "ms_per_frame": round(e2e_pinned["ms_per_frame"], 4),
                              "ms_median": round(e2e_pinned["ms_median"], 4), "ms_min": round(e2e_pinned["ms_min"], 4),
                              "api": "the same frb_render call into a reused page-locked buffer (frb_host_alloc): written by the streaming finisher over PCIe while the iterate kernel runs"}
        line["gpu_launches"] += e2e_pinned["launches"]
    print(json.dumps(line), flush=True)
    b.close()


if __name__ == "__main__":
    main()
