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
# Workloads: one fused region pair (forward + backward) each.  Algorithmic bytes follow SURVEY §8(d):
# every region input and output tensor counted once.
# --------------------------------------------------------------------------------------------------
def _nbytes(*ts):
    return sum(t.numel() * t.element_size() for t in ts if t is not None)


class Workload:
    """Subclasses define make(gen, device) -> dict of inputs (incl. cotangents), fwd(s) -> dict, bwd(s, f) -> dict,
    and the CPU legs.  `sets` are NSETS independent input sets the timed loop rotates through."""

    dtype = torch.bfloat16
    launches_per_step = 2

    def host_inputs(self, seed, pin):
        ts = self.make(torch.Generator().manual_seed(seed), torch.device("cpu"))
        return {k: (v.pin_memory() if pin else v) for k, v in ts.items()}

    def set_bytes(self, inputs):
        """Algorithmic bytes per step from one input set and the analytic output sizes (out_nbytes)."""
        fo, bo = self.out_nbytes()
        saved = self.saved_nbytes()
        self.fwd_bytes = _nbytes(*[inputs[k] for k in self.fwd_inputs]) + fo
        self.bwd_bytes = _nbytes(*[inputs[k] for k in self.bwd_inputs]) + saved + bo
        self.bytes_per_step = self.fwd_bytes + self.bwd_bytes
        self.h2d_bytes = _nbytes(*inputs.values())
        self.d2h_bytes = fo + bo

    def saved_nbytes(self):
        return 0

    def alloc(self, device):
        self.sets = [self.make(torch.Generator(device=device).manual_seed(1234 + s), device) for s in range(NSETS)]
        self.saved = [self.fwd(s) for s in self.sets]
        self.set_bytes(self.sets[0])
        f, b = self.saved[0], self.bwd(self.sets[0], self.saved[0])
        assert _nbytes(*f.values()) + _nbytes(*b.values()) == self.d2h_bytes, "out_nbytes() disagrees with the launcher outputs"

    def step(self, i):
        s = self.sets[i % NSETS]
        return self.bwd(s, self.fwd(s))

    def kernels(self):
        return [(self.name + " fwd", lambda i: self.fwd(self.sets[i % NSETS]), self.fwd_bytes),
                (self.name + " bwd", lambda i: self.bwd(self.sets[i % NSETS], self.saved[i % NSETS]), self.bwd_bytes)]

    # ---- end to end through the plugin boundary: thunder.jit(fn, executors=[b200]) ---------------------------------
    def jit_setup(self, thunder, b200):
        """The workload's defining PyTorch function (reference_fn) compiled by the reference with the b200 executor:
        the call a user of the reference makes.  Forward = the jitted callable, backward = autograd through it."""
        fn, self._diff, self._cot = self.reference_fn()
        self._jf = thunder.jit(fn, executors=[b200.ex])

    def jit_fwd(self, d):
        args = [d[k].requires_grad_(True) if (k in self._diff and d[k].is_floating_point()) else d[k] for k in self.fwd_inputs]
        outs = self._jf(*args)
        outs = list(outs) if isinstance(outs, (tuple, list)) else [outs]
        f = {f"out{i}": o for i, o in enumerate(outs)}
        f["_args"] = args
        return f

    def jit_bwd(self, d, f):
        outs = [f[k] for k in sorted(k for k in f if k.startswith("out"))]
        dargs = [a for a in f["_args"] if a.requires_grad]
        grads = torch.autograd.grad(outs, dargs, grad_outputs=[d[k] for k in self._cot])
        return {f"grad{i}": g for i, g in enumerate(grads)}

    def e2e_out_host(self, use_jit=False):
        if use_jit:
            d = {k: v.detach().clone() for k, v in self.sets[0].items()}
            f = self.jit_fwd(d)
            b = self.jit_bwd(d, f)
            f = {k: v.detach() for k, v in f.items() if k != "_args"}
        else:
            f = self.saved[0]
            b = self.bwd(self.sets[0], f)
        return {k: torch.empty(v.shape, dtype=v.dtype).pin_memory() for k, v in {**f, **b}.items()}

    def e2e_run(self, host, device, out_host, steps: int, stream, use_jit=False):
        """`steps` steps from pinned host buffers.  Every step copies its inputs host -> device, launches forward and
        backward, copies every result device -> host and waits for them (the caller reads the step's results on the
        host).  Three streams: a copy-in stream runs one step ahead of the compute stream, a copy-out stream drains
        results behind it, so both PCIe directions stay busy; nothing is reused between steps."""
        if not hasattr(self, "_h2d"):
            self._h2d, self._d2h = torch.cuda.Stream(device), torch.cuda.Stream(device)
        h2d, d2h = self._h2d, self._d2h
        bwd_only = [k for k in host if k not in self.fwd_inputs]
        fwd_fn, bwd_fn = (self.jit_fwd, self.jit_bwd) if use_jit else (self.fwd, self.bwd)

        def load(keys):
            with torch.cuda.stream(h2d):
                d = {k: host[k].to(device, non_blocking=True) for k in keys}
                ev = torch.cuda.Event()
                ev.record(h2d)
            return d, ev

        def drain(res):
            ev = torch.cuda.Event()
            ev.record(stream)
            d2h.wait_event(ev)
            with torch.cuda.stream(d2h):
                for k, v in res.items():
                    if k in out_host:
                        out_host[k].copy_(v.detach(), non_blocking=True)

        h2d.wait_stream(stream)
        nxt = load(self.fwd_inputs)
        for i in range(steps):
            d, ev_f = nxt
            late, ev_b = load(bwd_only)
            if i + 1 < steps:
                nxt = load(self.fwd_inputs)  # next step's inputs queue up behind this step's
            stream.wait_event(ev_f)
            f = fwd_fn(d)
            drain(f)
            stream.wait_event(ev_b)
            d.update(late)
            b = bwd_fn(d, f)
            drain(b)
            d2h.synchronize()  # this step's results are on the host
        stream.wait_stream(d2h)

    def cpu_step_port(self):
        hi = self.host_inputs(0, pin=False)
        self.set_bytes(hi)
        return lambda: self.oracle(hi)

    def cpu_step_reference(self, thunder):
        hi = self.host_inputs(0, pin=False)
        self.set_bytes(hi)
        fn, diff, cot = self.reference_fn()
        jf = thunder.jit(fn, executors=["torch"])
        args = [hi[k].clone().requires_grad_(k in diff) if hi[k].is_floating_point() else hi[k] for k in self.fwd_inputs]
        dargs = [a for a, k in zip(args, self.fwd_inputs) if k in diff]

        def run():
            outs = jf(*args)
            outs = list(outs) if isinstance(outs, (tuple, list)) else [outs]
            return torch.autograd.grad(outs, dargs, grad_outputs=[hi[k] for k in cot])

        return run


def _randn(gen, device, *shape, dtype=torch.bfloat16, scale=1.0):
    return (torch.randn(*shape, generator=gen, device=device, dtype=torch.float32) * scale).to(dtype)


class RmsnormResidual(Workload):
    """BASELINE.json configs[1]: fwd reads x, r, w and writes h, y, rstd (268 476 416 B); bwd reads dy, h, dh_in, w, rstd
    and writes dx (= dr, written once) and dw (268 484 608 B)."""

    name = "rmsnorm_residual"
    rows, hidden, eps = 8192, 4096, 1e-5
    launches_per_step = 3  # rmsnorm_fwd_kernel, rmsnorm_bwd_tma_kernel, colsum_finalize_kernel
    fwd_inputs, bwd_inputs, bwd_saved = ("x", "r", "w"), ("dy", "dh", "w"), ("h", "rstd")

    def describe(self):
        return f"{self.name}_fwd_bwd bf16 [{self.rows} tokens x {self.hidden} hidden] per GPU"

    def out_nbytes(self):
        t = self.rows * self.hidden * 2
        return 2 * t + self.rows * 4, t + self.hidden * 2  # (h, y, rstd), (dx, dw)

    def saved_nbytes(self):
        return self.rows * self.hidden * 2 + self.rows * 4  # h, rstd

    def make(self, g, dev):
        R, H = self.rows, self.hidden
        return dict(x=_randn(g, dev, R, H), r=_randn(g, dev, R, H), w=_randn(g, dev, H), dy=_randn(g, dev, R, H), dh=_randn(g, dev, R, H))

    def fwd(self, s):
        from b200ex import ops

        h, y, rstd = ops.rmsnorm_fwd(s["x"], s["r"], s["w"], self.eps)
        return dict(h=h, y=y, rstd=rstd)

    def bwd(self, s, f):
        from b200ex import ops

        dx, dw = ops.rmsnorm_bwd(s["dy"], f["h"], s["w"], f["rstd"], s["dh"])
        return dict(dx=dx, dw=dw)

    def oracle(self, hi):
        from oracle import regions as O

        h, y, rstd = O.rmsnorm_residual_fwd(hi["x"], hi["r"], hi["w"], self.eps)
        return O.rmsnorm_residual_bwd(hi["dy"], hi["dh"], h, hi["w"], rstd)

    def reference_fn(self):
        import torch.nn.functional as F

        hidden, eps = self.hidden, self.eps

        def fn(x, r, w):
            h = x + r
            return h, F.rms_norm(h, (hidden,), w, eps)

        return fn, ("x", "r", "w"), ("dh", "dy")


class Swiglu(Workload):
    """Llama-3-8B MLP activation (configs[2] region): fwd 3 x 234 881 024 B, bwd 5 x 234 881 024 B."""

    name = "swiglu"
    rows, inter = 8192, 14336
    fwd_inputs, bwd_inputs, bwd_saved = ("a", "b"), ("a", "b", "g"), ()

    def describe(self):
        return f"swiglu_fwd_bwd bf16 [{self.rows} tokens x {self.inter}] per GPU"

    def out_nbytes(self):
        t = self.rows * self.inter * 2
        return t, 2 * t

    def make(self, g, dev):
        return dict(a=_randn(g, dev, self.rows, self.inter), b=_randn(g, dev, self.rows, self.inter), g=_randn(g, dev, self.rows, self.inter))

    def fwd(self, s):
        from b200ex import ops

        return dict(out=ops.swiglu_fwd(s["a"], s["b"]))

    def bwd(self, s, f):
        from b200ex import ops

        da, db = ops.swiglu_bwd(s["a"], s["b"], s["g"])
        return dict(da=da, db=db)

    def oracle(self, hi):
        from oracle import regions as O

        return O.swiglu_fwd(hi["a"], hi["b"]), O.swiglu_bwd(hi["a"], hi["b"], hi["g"])

    def reference_fn(self):
        import torch.nn.functional as F

        return (lambda a, b: F.silu(a) * b), ("a", "b"), ("g",)


class CrossEntropy(Workload):
    """BASELINE.json configs[3]: bf16 logits [8192, 128256]; fwd reads logits once (2.10 GB), bwd reads logits and
    writes dlogits (4.20 GB)."""

    name = "cross_entropy"
    rows, vocab = 8192, 128256
    launches_per_step = 3  # ce_fwd_kernel, ce_finalize_kernel, ce_bwd_kernel
    fwd_inputs, bwd_inputs, bwd_saved = ("logits", "target"), ("logits", "target", "g"), ("row_max", "row_logsum", "n_valid")

    def describe(self):
        return f"cross_entropy_fwd_bwd bf16 logits [{self.rows} tokens x {self.vocab} vocab] per GPU"

    def out_nbytes(self):
        return 2 + 2 * self.rows * 4 + 8, self.rows * self.vocab * 2  # (loss, row_max, row_logsum, n_valid), dlogits

    def saved_nbytes(self):
        return 2 * self.rows * 4 + 8

    def make(self, g, dev):
        t = torch.randint(0, self.vocab, (self.rows,), generator=g, device=dev)
        t[::113] = -100
        return dict(logits=_randn(g, dev, self.rows, self.vocab), target=t, g=torch.ones((), dtype=self.dtype, device=dev))

    def fwd(self, s):
        from b200ex import ops

        loss, rmax, rls, nv = ops.cross_entropy_fwd(s["logits"], s["target"], -100)
        return dict(loss=loss, row_max=rmax, row_logsum=rls, n_valid=nv)

    def bwd(self, s, f):
        from b200ex import ops

        return dict(dlogits=ops.cross_entropy_bwd(s["g"], s["logits"], s["target"], f["row_max"], f["row_logsum"], f["n_valid"], -100))

    def oracle(self, hi):
        from oracle import regions as O

        loss, logp, nv = O.cross_entropy_fwd(hi["logits"], hi["target"], -100)
        return O.cross_entropy_bwd(hi["g"], logp, hi["target"], nv, -100)

    def reference_fn(self):
        import torch.nn.functional as F

        return (lambda logits, target: F.cross_entropy(logits, target, ignore_index=-100)), ("logits",), ("g",)


class QkvRope(Workload):
    """Llama-3-8B attention slice (configs[2] region): qkv [2, 4096, 6144] -> q, k, v [2, 32, 4096, 128] (k, v
    materialised at 32 heads as the reference's expand+reshape does: 304 MB forward)."""

    name = "qkv_rope"
    B, T, n_head, groups, hs = 2, 4096, 32, 8, 128
    fwd_inputs, bwd_inputs, bwd_saved = ("qkv", "cos", "sin"), ("dq", "dk", "dv", "cos", "sin"), ()

    def describe(self):
        return f"qkv_split_rope_fwd_bwd bf16 qkv [{self.B}, {self.T}, {(self.n_head + 2 * self.groups) * self.hs}] per GPU"

    def out_nbytes(self):
        return 3 * self.B * self.n_head * self.T * self.hs * 2, self.B * self.T * (self.n_head + 2 * self.groups) * self.hs * 2

    def make(self, g, dev):
        B, T, nh, G, hs = self.B, self.T, self.n_head, self.groups, self.hs
        return dict(qkv=_randn(g, dev, B, T, (nh + 2 * G) * hs), cos=_randn(g, dev, T, hs), sin=_randn(g, dev, T, hs),
                    dq=_randn(g, dev, B, nh, T, hs), dk=_randn(g, dev, B, nh, T, hs), dv=_randn(g, dev, B, nh, T, hs))

    def fwd(self, s):
        from b200ex import ops

        q, k, v = ops.qkv_rope_fwd(s["qkv"], s["cos"], s["sin"], self.n_head, self.groups, self.hs, self.hs)
        return dict(q=q, k=k, v=v)

    def bwd(self, s, f):
        from b200ex import ops

        return dict(dqkv=ops.qkv_rope_bwd(s["dq"], s["dk"], s["dv"], s["cos"], s["sin"], self.n_head, self.groups, self.hs, self.hs))

    def oracle(self, hi):
        from oracle import regions as O

        a = (self.n_head, self.groups, self.hs, self.hs)
        return O.qkv_split_rope_fwd(hi["qkv"], hi["cos"], hi["sin"], *a), O.qkv_split_rope_bwd(hi["dq"], hi["dk"], hi["dv"], hi["cos"], hi["sin"], *a)

    def reference_fn(self):
        B, T, nh, G, hs = self.B, self.T, self.n_head, self.groups, self.hs

        def apply_rope(x, cos, sin):
            x1, x2 = x[..., : hs // 2], x[..., hs // 2 :]
            return ((x * cos) + (torch.cat((-x2, x1), dim=-1) * sin)).to(dtype=x.dtype)

        def fn(qkv, cos, sin):
            qpk = nh // G
            t = qkv.view(B, T, G, qpk + 2, hs).permute(0, 2, 3, 1, 4)
            q, k, v = t.split((qpk, 1, 1), dim=2)
            k, v = k.expand(B, G, qpk, T, hs), v.expand(B, G, qpk, T, hs)
            q, k, v = q.reshape(B, -1, T, hs), k.reshape(B, -1, T, hs), v.reshape(B, -1, T, hs)
            return apply_rope(q, cos, sin), apply_rope(k, cos, sin), v

        return fn, ("qkv",), ("dq", "dk", "dv")


class BiasGelu(Workload):
    """bias + GELU on a GPT-2-XL-like MLP activation: bf16 [8192 tokens, 6400]."""

    name = "bias_gelu"
    rows, hidden = 8192, 6400
    launches_per_step = 3
    fwd_inputs, bwd_inputs, bwd_saved = ("x", "b"), ("x", "b", "g"), ()

    def describe(self):
        return f"bias_gelu_fwd_bwd bf16 [{self.rows} tokens x {self.hidden}] per GPU"

    def out_nbytes(self):
        t = self.rows * self.hidden * 2
        return t, t + self.hidden * 2

    def make(self, g, dev):
        return dict(x=_randn(g, dev, self.rows, self.hidden), b=_randn(g, dev, self.hidden), g=_randn(g, dev, self.rows, self.hidden))

    def fwd(self, s):
        from b200ex import ops

        return dict(out=ops.bias_act_fwd(s["x"], s["b"], "gelu"))

    def bwd(self, s, f):
        from b200ex import ops

        dx, db = ops.bias_act_bwd(s["x"], s["b"], s["g"], "gelu")
        return dict(dx=dx, db=db)

    def oracle(self, hi):
        from oracle import regions as O

        return O.bias_gelu_fwd(hi["x"], hi["b"]), O.bias_gelu_bwd(hi["x"], hi["b"], hi["g"])

    def reference_fn(self):
        import torch.nn.functional as F

        return (lambda x, b: F.gelu(x + b)), ("x", "b"), ("g",)


class LayerNorm(Workload):
    """LayerNorm at the RMSNorm headline shape: fwd reads x, w, b and writes y, mean, rstd; bwd reads dy, x, w, mean,
    rstd and writes dx, dw, db."""

    name = "layernorm"
    rows, hidden, eps = 8192, 4096, 1e-5
    launches_per_step = 3  # layernorm_fwd_kernel, layernorm_bwd_tma_kernel, layernorm_finalize_kernel
    fwd_inputs, bwd_inputs, bwd_saved = ("x", "w", "b"), ("dy", "x", "w"), ("mean", "rstd")

    def describe(self):
        return f"layernorm_fwd_bwd bf16 [{self.rows} tokens x {self.hidden} hidden] per GPU"

    def out_nbytes(self):
        t = self.rows * self.hidden * 2
        return t + 2 * self.rows * 4, t + 2 * self.hidden * 2  # (y, mean, rstd), (dx, dw, db)

    def saved_nbytes(self):
        return 2 * self.rows * 4

    def make(self, g, dev):
        R, H = self.rows, self.hidden
        return dict(x=_randn(g, dev, R, H), w=_randn(g, dev, H), b=_randn(g, dev, H), dy=_randn(g, dev, R, H))

    def fwd(self, s):
        from b200ex import ops

        y, mean, rstd = ops.layernorm_fwd(s["x"], s["w"], s["b"], self.eps)
        return dict(y=y, mean=mean, rstd=rstd)

    def bwd(self, s, f):
        from b200ex import ops

        dx, dw, db = ops.layernorm_bwd(s["dy"], s["x"], s["w"], f["mean"], f["rstd"], True)
        return dict(dx=dx, dw=dw, db=db)

    def oracle(self, hi):
        from oracle import regions as O

        y, mean, rstd = O.layernorm_fwd(hi["x"], hi["w"], hi["b"], self.eps)
        return y, O.layernorm_bwd(hi["dy"], hi["x"], hi["w"], True, mean, rstd)

    def reference_fn(self):
        import torch.nn.functional as F

        hidden, eps = self.hidden, self.eps
        return (lambda x, w, b: F.layer_norm(x, (hidden,), w, b, eps)), ("x", "w", "b"), ("dy",)


class Softmax(Workload):
    """softmax over the last dimension of attention scores [2, 16, 2048, 2048] bf16: fwd reads x, writes y; bwd reads
    dy and the saved y, writes dx."""

    name = "softmax"
    rows, hidden = 65536, 2048
    launches_per_step = 2
    fwd_inputs, bwd_inputs, bwd_saved = ("x",), ("dy",), ("y",)

    def describe(self):
        return f"softmax_fwd_bwd bf16 [{self.rows} rows x {self.hidden}] per GPU"

    def out_nbytes(self):
        t = self.rows * self.hidden * 2
        return t, t

    def saved_nbytes(self):
        return self.rows * self.hidden * 2

    def make(self, g, dev):
        return dict(x=_randn(g, dev, self.rows, self.hidden, scale=2.0), dy=_randn(g, dev, self.rows, self.hidden))

    def fwd(self, s):
        from b200ex import ops

        return dict(y=ops.softmax_fwd(s["x"]))

    def bwd(self, s, f):
        from b200ex import ops

        return dict(dx=ops.softmax_bwd(s["dy"], f["y"]))

    def oracle(self, hi):
        from oracle import regions as O

        y = O.softmax_fwd(hi["x"])
        return y, O.softmax_bwd(hi["dy"], y)

    def reference_fn(self):
        import torch.nn.functional as F

        return (lambda x: F.softmax(x, -1)), ("x",), ("dy",)


WORKLOADS = {"softmax": Softmax, "layernorm": LayerNorm, "rmsnorm": RmsnormResidual, "swiglu": Swiglu, "cross_entropy": CrossEntropy, "qkv_rope": QkvRope, "bias_gelu": BiasGelu}


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


def _shrink_rows(wl, frac):
    """Bounded sample of a workload: the same region on a fraction of the rows (regions are independent per token row,
    so GB/s does not depend on the row count once the working set is far beyond the caches)."""
    if hasattr(wl, "rows"):
        wl.rows = max(256, int(wl.rows * frac) // 64 * 64)
        return f"{wl.rows} of the workload's token rows"
    wl.T = max(256, int(wl.T * frac) // 64 * 64)
    return f"seq {wl.T} of the workload's tokens"


def run_reference_arm(args, wl):
    """The reference's own CPU implementation of the path (thunder.jit with its torch executor, imported unmodified
    from baseline/_ref; the oracle port only if that is absent) on all host cores.  Each step is a bounded sample of
    the workload, sized from one timed full-size step so that the K timed steps end within about two minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    thunder = import_reference()
    kind = "reference" if thunder is not None else "port"
    make = lambda: wl.cpu_step_reference(thunder) if thunder is not None else wl.cpu_step_port()  # also sets wl.bytes_per_step
    full_desc = wl.describe()
    run = make()
    run()  # compile + first touch
    t0 = time.perf_counter()
    run()
    t_full = time.perf_counter() - t0
    sample = "full workload"
    budget_s = 120.0
    total_steps = args.steps + max(1, args.warmup)
    if t_full * total_steps > budget_s:
        sample = _shrink_rows(wl, budget_s / (t_full * total_steps)) + " per step"
        run = make()
        run()
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
        "data": "synthetic", "config": {"workload": full_desc, "device": "host CPU", "executor": "thunder.jit(executors=['torch'])" if kind == "reference" else "oracle port"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
                         "sample": f"{sample}; {args.steps} steps after {max(1, args.warmup)} warm-up, rank 0 only",
                         "note": "whole-box CPU figure: one process on ALL host cores (the cores do not multiply with --gpus); "
                                 "compare per-GPU numbers with it at N=1, the N>1 ratio is N GPUs against the same host"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# --------------------------------------------------------------------------------------------------
# Main
# --------------------------------------------------------------------------------------------------
_REAL_STDOUT = None


def _claim_stdout():
    """stdout carries exactly ONE JSON line: everything else that libraries print there (NCCL's version banner, ...)
    is sent to stderr by pointing fd 1 at fd 2 for the duration of the run."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def _emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="rmsnorm", choices=sorted(WORKLOADS) + ["llama_step"])
    # llama_step only (BASELINE.json configs[4]):
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--micro-batch", type=int, default=2)
    ap.add_argument("--seq-len", type=int, default=4096)
    ap.add_argument("--executor", default="b200", choices=["b200", "torch"])
    ap.add_argument("--fusion-type", default="dataflow", choices=["dataflow", "consecutive"])
    ap.add_argument("--spelling", default="litgpt", choices=["litgpt", "ops"],
                    help="llama_step model source: 'litgpt' = norms and the attention slice written with plain torch ops as litgpt writes them "
                         "(b200 claims the regions itself), 'ops' = F.rms_norm + b200ex.executor.qkv_split_rope (op-level kernels)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train-step", action="store_true", help="skip the Llama-3-8B-shape training step that the default run appends as `train_step`")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time the step loop through the Python launchers instead of per-step CUDA graphs")
    ap.add_argument("--profile", action="store_true", help="for ncu: W warm-up steps, then K plain steps (one launch per kernel per step), nothing else; "
                                                            "numbers printed by such a run are not bench values")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.workload == "llama_step":
        thunder = import_reference()
        if thunder is None:
            raise SystemExit("bench.py --workload llama_step needs the reference (baseline/_ref) for thunder.jit / thunder.distributed")
        if args.impl == "reference":
            args.executor = "torch"
        from baseline import train_step

        if args.steps == 2000:
            args.steps, args.warmup = 5, 3
        res = train_step.run(args, thunder, ClockSampler, emit=False)
        if res is not None:  # rank 0
            _emit(res)
        return
    wl = WORKLOADS[args.workload]()
    if os.environ.get("B200_BENCH_SHAPE") and hasattr(wl, "rows") and hasattr(wl, "hidden"):
        # A/B sweeps at other shapes (tools/); the shape shows up in config.workload through describe()
        wl.rows, wl.hidden = (int(v) for v in os.environ["B200_BENCH_SHAPE"].split(","))

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
    if args.profile:
        torch.cuda.synchronize(device)
        torch.cuda.profiler.start()  # ncu --profile-from-start off captures exactly these K steps
        for i in range(args.steps):
            wl.step(i)
        torch.cuda.synchronize(device)
        torch.cuda.profiler.stop()
        if rank == 0:
            sampler.stop()
            _emit({"profile_run": True, "workload": wl.describe(), "steps": args.steps, "warmup": args.warmup, "launches_per_step": wl.launches_per_step})
        return
    # One CUDA graph per buffer set, each holding one whole step (forward + backward launches): the timed loop
    # replays them round-robin, so a step costs one graph launch on the host instead of three Python launchers —
    # what the reference offers through CUDAGraphTransform.  --no-graph (or a failed capture) times the Python loop.
    step_graphs, launch_mode = None, "python launchers, one call per kernel"
    if not args.no_graph:
        try:
            torch.cuda.synchronize(device)
            step_graphs = []
            for i in range(NSETS):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    keep = wl.step(i)  # outputs stay alive with the graph
                step_graphs.append((g, keep))
            for g, _ in step_graphs:
                g.replay()
            torch.cuda.synchronize(device)
            launch_mode = f"one CUDA graph per step ({NSETS} graphs, one per buffer set), replayed round-robin"
        except Exception as e:
            step_graphs, launch_mode = None, f"python launchers (graph capture failed: {type(e).__name__})"
    barrier()
    sampler.mark_load(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    if step_graphs is not None:
        for i in range(args.steps):
            step_graphs[i % NSETS][0].replay()
    else:
        for i in range(args.steps):
            wl.step(i)
    ev1.record(stream)
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=device, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = ms.item()

    # ---- per-kernel timing for the roofline (same process, CUDA events on the launch stream) ------
    # Each kernel is timed twice: launched from Python one call at a time (what a trace executes), and as a CUDA
    # graph of NSETS x 4 captured launches replayed back to back, which takes the host out of the measurement — a
    # launcher costs 10-25 us of host time, so kernels shorter than that are host-bound in the first form.  The
    # roofline uses the graph number: it is the kernel's own duration.
    kernels = []
    kiters = max(200, min(args.steps, 2000))

    def timed(run, n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(device)
        a.record(stream)
        for i in range(n):
            run(i)
        b.record(stream)
        torch.cuda.synchronize(device)
        return a.elapsed_time(b) * 1e3

    for name, fn, nbytes in wl.kernels():
        for i in range(3):
            fn(i)
        eager_us = timed(fn, kiters) / kiters
        us, how = eager_us, "python launch loop"
        try:
            per_graph = NSETS * 4
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for i in range(per_graph):
                    fn(i)
            replays = max(3, kiters // per_graph)
            graph.replay()
            us, how = timed(lambda i: graph.replay(), replays) / (replays * per_graph), f"cuda graph of {per_graph} launches, {replays} replays"
            del graph
        except Exception as e:  # keep the launch-loop number
            how += f" (graph capture failed: {type(e).__name__})"
        kernels.append({"kernel": name, "us_per_launch": us, "algorithmic_bytes": nbytes, "achieved_gbps": nbytes / us / 1e3,
                        "us_per_launch_python_loop": eager_us, "timed": how})

    # ---- end to end with host buffers ---------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        # through the reference-facing plugin call: thunder.jit(fn, executors=[b200]) on host (pinned) buffers
        thunder = import_reference()
        if thunder is None:
            raise SystemExit("bench.py: baseline/_ref (the reference) is missing; run __graft_entry__.build() in the build container")
        import b200ex.executor as b200

        wl.jit_setup(thunder, b200)
        host = wl.host_inputs(99 + rank, pin=True)
        out_host = wl.e2e_out_host(use_jit=True)
        d2h_bytes = _nbytes(*out_host.values())
        e2e_steps = max(3, min(args.steps, 20))
        wl.e2e_run(host, device, out_host, 2, stream, use_jit=True)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        wl.e2e_run(host, device, out_host, e2e_steps, stream, use_jit=True)
        b.record(stream)
        barrier()
        ems = torch.tensor([a.elapsed_time(b)], device=device, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(ems, op=dist.ReduceOp.MAX)
        census = {}
        try:
            from baseline.train_step import region_census

            census = region_census(thunder, wl._jf)
        except Exception:
            pass
        e2e = {"value": world * wl.bytes_per_step * e2e_steps / (ems.item() / 1e3) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": wl.h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
               "ms_per_step": ems.item() / e2e_steps, "steps": e2e_steps, "b200_launch_kinds": census,
               "api": "thunder.jit(fn, executors=[b200]) forward + torch.autograd.grad backward on pinned host tensors (H2D of every input, "
                      "D2H of every output and gradient per step; copy-in one step ahead, copy-out behind)",
               "note": "PCIe-bound: the step moves h2d + d2h bytes over the host link for a region that takes ~0.1 ms on the device; at N > 1 all ranks "
                       "share one host's memory and PCIe root, so this column does not scale with N"}

    sampler.mark_load(False)
    clocks = sampler.stop() if rank == 0 else None

    # ---- second half of the metric: Llama-3-8B-shape training step in tokens/s (same process, same ranks) ----
    train = None
    if not args.no_train_step and args.workload == "rmsnorm":
        wl.sets = wl.saved = None
        step_graphs = None
        torch.cuda.empty_cache()
        try:
            thunder = import_reference()
            if thunder is None:
                raise RuntimeError("reference not importable (baseline/_ref missing)")
            from baseline import train_step

            targs = argparse.Namespace(layers=args.layers, micro_batch=args.micro_batch, seq_len=args.seq_len, executor="b200",
                                       fusion_type=args.fusion_type, steps=5, warmup=3, spelling=args.spelling)
            train = train_step.run(targs, thunder, ClockSampler, dist=dist, emit=False)
        except Exception as e:  # the region metric above must still be reported
            train = {"error": f"{type(e).__name__}: {e}"[:500]}
    if rank == 0:
        peak, peak_src = load_peaks()
        dom = max(kernels, key=lambda k: k["us_per_launch"])
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
                traffic = json.load(f).get(dom["kernel"])
        except Exception:
            pass
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = time_cpu(wl)
        line = {
            "metric": METRIC, "value": world * wl.bytes_per_step * args.steps / (total_ms / 1e3) / 1e9, "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": wl.describe(), "l2": f"inputs larger than L2: rotating {NSETS} buffer sets ({NSETS * wl.h2d_bytes / 1e9:.2f} GB of inputs)",
                       "sharding": "each rank runs the region on its own micro-batch; no data-path collective", "launch": launch_mode},
            "roofline": {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["achieved_gbps"], "peak": peak, "unit": UNIT,
                         "frac": dom["achieved_gbps"] / peak, "frac_of_nominal_8TBps": dom["achieved_gbps"] / 8000.0,
                         "traffic": traffic, "traffic_source": "ncu --set full capture, profiles/r2_traffic.json" if traffic else None,
                         "algorithmic_bytes": dom["algorithmic_bytes"], "peak_source": peak_src},
            "kernels": kernels,
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": wl.launches_per_step * args.steps,
            "clocks": clocks,
            "train_step": train,
        }
        _emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
