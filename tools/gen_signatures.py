"""Regenerate lightning-thunder_b200/b200ex/_signatures.py: the canonical pointwise-program signatures of the
regions that have a hand-written family kernel.

Each family's defining PyTorch expression is traced through thunder.jit with the b200 executor (CPU proxies are
accepted for claiming only; nothing is executed), the resulting fusion regions are lowered, and the signature of
the pointwise program together with the role of every operand is recorded.  Needs the reference importable
(baseline/_ref or /root/reference + baseline/shims).
"""
import os
import sys
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.dont_write_bytecode = True
sys.path[:0] = [os.path.join(ROOT, "baseline", "shims"), os.path.join(ROOT, "lightning-thunder_b200")]
for cand in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
    if os.path.isdir(os.path.join(cand, "thunder")):
        sys.path.insert(0, cand)
        break
warnings.simplefilter("ignore")

import torch
import torch.nn.functional as F

import thunder
import b200ex.executor as b200
from b200ex import pointwise


def collect(fn, args):
    """Trace fn forward+backward; return [(spec, {proxy name: role})] for every pointwise program."""
    captured = []
    orig = pointwise.PwLaunch.__init__

    def spy(self, spec, *args, **kwargs):
        captured.append(spec)
        orig(self, spec, *args, **kwargs)

    pointwise.PwLaunch.__init__ = spy
    try:
        jf = thunder.jit(fn, executors=[b200.ex])
        try:
            jf(*args)
        except RuntimeError:
            pass  # launching on CPU tensors raises by design; the traces exist
    finally:
        pointwise.PwLaunch.__init__ = orig
    fw = thunder.last_traces(jf)[-1]
    bws = thunder.last_backward_traces(jf)
    return captured, fw, (bws[-1] if bws else None)


def classify(specs, bw, table, strict=False):
    """Record every program that is exactly a SwiGLU forward (a, b -> out) or backward (a, b, g -> da, db)."""
    bw_ret = None
    if bw is not None:
        for bs in bw.bound_symbols:
            if bs.sym.name == "python_return":
                bw_ret = [p.name for p in bs.flat_proxy_args]
    for spec in specs:
        leaf_names = [n for n, _ in spec.leaves]
        out_names = [n for n, _, _ in spec.outputs]
        if len(out_names) == 1 and sorted(leaf_names) == ["a", "b"]:
            table[spec.signature] = ("swiglu_fwd", tuple(leaf_names), ("out",))
        elif len(out_names) == 2 and len(leaf_names) == 3 and {"a", "b"} <= set(leaf_names) and bw_ret and set(out_names) <= set(bw_ret):
            roles = tuple(n if n in ("a", "b") else "g" for n in leaf_names)
            out_roles = tuple("da" if n == bw_ret[0] else "db" for n in out_names)
            if set(out_roles) == {"da", "db"}:
                table[spec.signature] = ("swiglu_bwd", roles, out_roles)
        elif strict:
            raise SystemExit(f"unexpected program while calibrating swiglu: leaves {leaf_names} outs {out_names}")


def main():
    b200._testing_claim_cpu(True)
    b200._NATIVE_SIGNATURES.clear()  # capture raw programs, not already-routed ones
    table = {}
    for dt in (torch.bfloat16, torch.float16, torch.float32):
        mk = lambda grad: torch.randn(4, 64).to(dt).requires_grad_(grad)
        # (i) forward only
        specs, fw, bw = collect(lambda a, b: F.silu(a) * b, (mk(False), mk(False)))
        classify(specs, None, table, strict=True)
        # (ii) forward + backward, stand-alone (rematerialisation decides what the forward saves)
        specs, fw, bw = collect(lambda a, b: F.silu(a) * b, (mk(True), mk(True)))
        classify(specs, bw, table)
        # (iii) with other work in the same trace, as inside a transformer block
        def ctx(a, b, x, w):
            return F.silu(a) * b, F.rms_norm(x, (64,), w, 1e-5)

        specs, fw, bw = collect(ctx, (mk(True), mk(True), mk(True), torch.randn(64).to(dt).requires_grad_(True)))
        classify(specs, bw, table)
    # ---- bias + activation (gelu, relu): forward (x, b -> [pre,] out), backward (saved pre-activation a, g -> dx) ----
    for act, fn in (("gelu", lambda x, b: F.gelu(x + b)), ("relu", lambda x, b: F.relu(x + b))):
        for dt in (torch.bfloat16, torch.float16, torch.float32):
            mkx = lambda grad: torch.randn(6, 32).to(dt).requires_grad_(grad)
            mkb = lambda grad: torch.randn(32).to(dt).requires_grad_(grad)
            specs, fw, bw = collect(fn, (mkx(False), mkb(False)))
            # Outside autograd the torch executor runs gelu as ONE call (one rounding), so in 16-bit storage the region is a
            # different program from the op-by-op chain the family kernel implements: it stays on the generic path.
            if act == "gelu" and dt != torch.float32:
                specs = []
            for spec in specs:
                leaves = [n for n, _ in spec.leaves]
                if sorted(leaves) == ["b", "x"] and len(spec.outputs) == 1:
                    table[spec.signature] = (f"bias_{act}_fwd", tuple(leaves), ("out",))
            specs, fw, bw = collect(fn, (mkx(True), mkb(True)))
            fw_ret = None
            for bs in fw.bound_symbols:
                if bs.sym.name == "python_return":
                    fw_ret = bs.args[0]["flat_output"][0].name  # the function's result; the other output is the saved pre-activation
            for spec in specs:
                leaves = [n for n, _ in spec.leaves]
                outs = [n for n, _, _ in spec.outputs]
                if sorted(leaves) == ["b", "x"] and len(outs) == 2 and fw_ret in outs:
                    table[spec.signature] = (f"bias_{act}_fwd", tuple(leaves), tuple("out" if n == fw_ret else "pre" for n in outs))
                    pre_name = [n for n in outs if n != fw_ret][0]
                elif len(leaves) == 2 and len(outs) == 1 and "x" not in leaves and "b" not in leaves:
                    # backward: one leaf is the saved pre-activation (same proxy name as the forward's extra output)
                    roles = tuple("a" if n == pre_name else "g" for n in leaves)
                    if sorted(roles) == ["a", "g"]:
                        table[spec.signature] = (f"{act}_bwd", roles, ("dx",))
    path = os.path.join(ROOT, "lightning-thunder_b200", "b200ex", "_signatures.py")
    with open(path, "w") as f:
        f.write('"""Generated by tools/gen_signatures.py — do not edit."""\n\nNATIVE_SIGNATURES = {\n')
        for sig in sorted(table):
            f.write(f"    {sig!r}: {table[sig]!r},\n")
        f.write("}\n")
    print(f"wrote {len(table)} signatures to {path}")
    return table


if __name__ == "__main__":
    main()
