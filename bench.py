#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the b200 fused-region hot path.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME]

Metric (BASELINE.json): fused-region HBM GB/s.  A "step" is one forward + one backward pass of a
fused region over one batch of synthetic input; the default workload is BASELINE.json configs[1],
RMSNorm + residual add, bf16, [8192 tokens x 4096 hidden] per GPU (weak scaling: every rank runs the
region on its own micro-batch, no data-path collective — SURVEY §8(e)).

``value``      algorithmic bytes of all ranks / device time, inputs resident in HBM.
``e2e``        same metric through the public launcher API with HOST (pinned) buffers: H2D of the
               step's inputs and D2H of its outputs inside the timed region.
``roofline``   dominant kernel: algorithmic bytes per launch / CUDA-event time per launch, against the
               measured HBM copy bandwidth in MEASURED_PEAKS.json.
``cpu_baseline``  the reference's torch-executor path (``kind: reference``, thunder imported from
               baseline/_ref) or the oracle port, timed on this box's host cores.  Reported, not a target.

Timing hygiene: >= 3 warm-up steps, CUDA events on the launch stream, barrier + synchronize on both
sides, max over ranks; every step rotates through 4 input/output buffer sets (1.07 GB working set,
> 126 MB L2) so no step re-reads data the previous one left in L2.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "lightning-thunder_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
sys.dont_write_bytecode = True

import torch  # noqa: E402

NSETS = 4
METRIC = "fused_region_hbm_gbps"
UNIT = "GB/s"


# --------------------------------------------------------------------------------------------------
# Workloads
# --------------------------------------------------------------------------------------------------
class RmsnormResidual:
    """BASELINE.json configs[1]; algorithmic bytes from SURVEY §8(d) row 2."""

    name = "rmsnorm_residual_fwd_bwd"
    rows, hidden, eps = 8192, 4096, 1e-5
    dtype = torch.bfloat16
    launches_per_step = 3  # rmsnorm_fwd_kernel, rmsnorm_bwd_kernel, rmsnorm_dw_reduce_kernel

    def __init__(self):
        e = 2
        t = self.rows * self.hidden * e
        self.fwd_bytes = 4 * t + self.hidden * e + self.rows * 4  # read x, r, w; write h, y, rstd
        self.bwd_bytes = 4 * t + self.hidden * e + self.rows * 4 + self.hidden * e  # read dy, h, dh_in, w, rstd; write dx, dw
        self.bytes_per_step = self.fwd_bytes + self.bwd_bytes
        self.h2d_bytes = 4 * t + self.hidden * e  # x, r, dy, dh_in, w
        self.d2h_bytes = 3 * t + self.rows * 4 + self.hidden * e  # h, y, dx, rstd, dw

    def describe(self):
        return f"{self.name} bf16 [{self.rows} tokens x {self.hidden} hidden] per GPU"

    def host_inputs(self, seed, pin):
        g = torch.Generator().manual_seed(seed)
        mk = lambda *s: torch.randn(*s, generator=g).to(self.dtype)
        ts = dict(x=mk(self.rows, self.hidden), r=mk(self.rows, self.hidden), w=mk(self.hidden),
                  dy=mk(self.rows, self.hidden), dh=mk(self.rows, self.hidden))
        return {k: (v.pin_memory() if pin else v) for k, v in ts.items()}

    def alloc(self, device):
        self.sets = []
        for s in range(NSETS):
            g = torch.Generator(device=device).manual_seed(1234 + s)
            mk = lambda *sh: torch.randn(*sh, generator=g, device=device, dtype=torch.float32).to(self.dtype)
            self.sets.append(dict(x=mk(self.rows, self.hidden), r=mk(self.rows, self.hidden), w=mk(self.hidden),
                                  dy=mk(self.rows, self.hidden), dh=mk(self.rows, self.hidden)))
        # saved-for-backward of each set, produced once, so the bwd-only kernel loop has its inputs
        from b200ex import ops

        for s in self.sets:
            s["h"], _, s["rstd"] = ops.rmsnorm_fwd(s["x"], s["r"], s["w"], self.eps)

    def fwd(self, i):
        from b200ex import ops

        s = self.sets[i % NSETS]
        return ops.rmsnorm_fwd(s["x"], s["r"], s["w"], self.eps)

    def bwd(self, i):
        from b200ex import ops

        s = self.sets[i % NSETS]
        return ops.rmsnorm_bwd(s["dy"], s["h"], s["w"], s["rstd"], s["dh"])

    def step(self, i):
        from b200ex import ops

        s = self.sets[i % NSETS]
        h, y, rstd = ops.rmsnorm_fwd(s["x"], s["r"], s["w"], self.eps)
        return ops.rmsnorm_bwd(s["dy"], h, s["w"], rstd, s["dh"])

    def kernels(self):
        return [("rmsnorm_fwd_kernel", self.fwd, self.fwd_bytes), ("rmsnorm_bwd_kernel(+dw_reduce)", self.bwd, self.bwd_bytes)]

    def e2e_step(self, host, device, out_host):
        from b200ex import ops

        d = {k: v.to(device, non_blocking=True) for k, v in host.items()}
        h, y, rstd = ops.rmsnorm_fwd(d["x"], d["r"], d["w"], self.eps)
        dx, dw = ops.rmsnorm_bwd(d["dy"], h, d["w"], rstd, d["dh"])
        for k, v in (("h", h), ("y", y), ("rstd", rstd), ("dx", dx), ("dw", dw)):
            out_host[k].copy_(v, non_blocking=True)

    def e2e_out_host(self):
        bf, f32 = self.dtype, torch.float32
        return dict(h=torch.empty(self.rows, self.hidden, dtype=bf).pin_memory(), y=torch.empty(self.rows, self.hidden, dtype=bf).pin_memory(),
                    rstd=torch.empty(self.rows, 1, dtype=f32).pin_memory(), dx=torch.empty(self.rows, self.hidden, dtype=bf).pin_memory(),
                    dw=torch.empty(self.hidden, dtype=bf).pin_memory())

    # ---- CPU legs -------------------------------------------------------------------------------
    def cpu_step_port(self):
        from oracle import regions as O

        hi = self.host_inputs(0, pin=False)

        def run():
            h, y, rstd = O.rmsnorm_residual_fwd(hi["x"], hi["r"], hi["w"], self.eps)
            return O.rmsnorm_residual_bwd(hi["dy"], hi["dh"], h, hi["w"], rstd)

        return run

    def cpu_step_reference(self, thunder):
        import torch.nn.functional as F

        hi = self.host_inputs(0, pin=False)
        hidden, eps = self.hidden, self.eps

        def fn(x, r, w):
            h = x + r
            return h, F.rms_norm(h, (hidden,), w, eps)

        jf = thunder.jit(fn, executors=["torch"])
        x, r, w = (hi[k].clone().requires_grad_(True) for k in ("x", "r", "w"))

        def run():
            h, y = jf(x, r, w)
            return torch.autograd.grad([h, y], [x, r, w], grad_outputs=[hi["dh"], hi["dy"]])

        return run


WORKLOADS = {"rmsnorm": RmsnormResidual}


# --------------------------------------------------------------------------------------------------
# Helpers
# --------------------------------------------------------------------------------------------------
def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    FIELDS = "timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None
        self.t_load0 = self.t_load1 = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def mark_load(self, begin):
        if begin:
            self.t_load0 = time.time()
        else:
            self.t_load1 = time.time()

    def stop(self):
        """Median SM clock and any throttle reason seen while the GPU was under this benchmark's load
        (timed steps + per-kernel loops + e2e steps)."""
        import datetime

        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "window": "timed steps + per-kernel loops + e2e steps"}
        if self.proc is None:
            return out
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [c.strip() for c in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                    clk, cmax = float(f[1]), float(f[2])
                except ValueError:
                    continue
                if self.t_load0 is not None and (ts < self.t_load0 - 0.02 or ts > (self.t_load1 or time.time()) + 0.02):
                    continue
                sm.append(clk), smax.append(cmax)
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(smax), reasons=sorted(reasons), samples=len(sm))
        return out


def import_reference():
    """Import the unmodified reference from baseline/_ref (travels with the repo) with the dependency shims."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "thunder")):
        return None
    for p in (ref, os.path.join(ROOT, "baseline", "shims")):
        if p not in sys.path:
            sys.path.insert(0, p)
    try:
        import warnings

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            import thunder
        return thunder
    except Exception:
        return None


def time_cpu(wl, budget_s=12.0, max_iters=10):
    """Time the reference's CPU path for one full step of the workload on all host cores."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    thunder = import_reference()
    kind = "reference" if thunder is not None else "port"
    run = wl.cpu_step_reference(thunder) if thunder is not None else wl.cpu_step_port()
    run()  # warm-up (includes thunder.jit compilation)
    times, t_end = [], time.perf_counter() + budget_s
    while len(times) < max_iters and (time.perf_counter() < t_end or not times):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    best = min(times)
    return {
        "value": wl.bytes_per_step / best / 1e9, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
        "sample": f"full workload ({wl.describe()}), 1 warm-up + best of {len(times)} steps, {best * 1e3:.1f} ms/step",
        "ms_per_step": best * 1e3,
    }


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    thunder = import_reference()
    kind = "reference" if thunder is not None else "port"
    run = wl.cpu_step_reference(thunder) if thunder is not None else wl.cpu_step_port()
    for _ in range(max(1, args.warmup)):
        run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run()
    dt = time.perf_counter() - t0
    val = wl.bytes_per_step * args.steps / dt / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic", "config": {"workload": wl.describe(), "device": "host CPU", "executor": "thunder.jit(executors=['torch'])" if kind == "reference" else "oracle port"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
                         "sample": f"full workload, {args.steps} steps after {max(1, args.warmup)} warm-up on rank 0 only"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# Main
# --------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="rmsnorm", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    wl = WORKLOADS[args.workload]()

    if args.impl == "reference":
        run_reference_arm(args, wl)
        return

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 path has no CPU fallback (use --impl reference for the CPU arm)")
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(device)

    from b200ex import _lib

    _lib.lib()  # fail loudly before timing anything if the CUDA library is missing
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()  # nvidia-smi needs ~1 s to start printing; allocation + warm-up cover that
    wl.alloc(device)
    stream = torch.cuda.current_stream(device)

    # ---- whole-step throughput, inputs resident in HBM ------------------------------------------
    for i in range(args.warmup):
        wl.step(i)
    barrier()
    sampler.mark_load(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for i in range(args.steps):
        wl.step(i)
    ev1.record(stream)
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=device, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = ms.item()

    # ---- per-kernel timing for the roofline (same process, CUDA events on the launch stream) ------
    kernels = []
    kiters = max(200, min(args.steps, 2000))
    for name, fn, nbytes in wl.kernels():
        for i in range(3):
            fn(i)
        torch.cuda.synchronize(device)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for i in range(kiters):
            fn(i)
        b.record(stream)
        torch.cuda.synchronize(device)
        us = a.elapsed_time(b) * 1e3 / kiters
        kernels.append({"kernel": name, "us_per_launch": us, "algorithmic_bytes": nbytes, "achieved_gbps": nbytes / us / 1e3})

    # ---- end to end with host buffers ---------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        host = wl.host_inputs(99 + rank, pin=True)
        out_host = wl.e2e_out_host()
        e2e_steps = max(3, min(args.steps, 20))
        for _ in range(2):
            wl.e2e_step(host, device, out_host)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(e2e_steps):
            wl.e2e_step(host, device, out_host)
            stream.synchronize()  # the caller reads the step's result on the host
        b.record(stream)
        barrier()
        ems = torch.tensor([a.elapsed_time(b)], device=device, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(ems, op=dist.ReduceOp.MAX)
        e2e = {"value": world * wl.bytes_per_step * e2e_steps / (ems.item() / 1e3) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": wl.h2d_bytes, "d2h_bytes_per_step": wl.d2h_bytes,
               "ms_per_step": ems.item() / e2e_steps, "steps": e2e_steps,
               "api": "b200ex.ops.rmsnorm_fwd/rmsnorm_bwd on pinned host tensors (H2D + launch + D2H per step)"}

    sampler.mark_load(False)
    clocks = sampler.stop() if rank == 0 else None
    if rank == 0:
        peak, peak_src = load_peaks()
        dom = max(kernels, key=lambda k: k["us_per_launch"])
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = time_cpu(wl)
        line = {
            "metric": METRIC, "value": world * wl.bytes_per_step * args.steps / (total_ms / 1e3) / 1e9, "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": wl.describe(), "l2": f"inputs larger than L2: rotating {NSETS} buffer sets (1.07 GB working set)",
                       "sharding": "each rank runs the region on its own micro-batch; no data-path collective"},
            "roofline": {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved_gbps"], "peak": peak, "unit": UNIT,
                         "frac": dom["achieved_gbps"] / peak, "frac_of_nominal_8TBps": dom["achieved_gbps"] / 8000.0,
                         "traffic": None, "peak_source": peak_src},
            "kernels": kernels,
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": wl.launches_per_step * args.steps,
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
