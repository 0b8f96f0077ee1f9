"""Generate tests/golden/regions.pt by running the UNMODIFIED reference on seeded inputs.

TEST INFRASTRUCTURE.  Run in the build container (where the reference exists):

    python oracle/gen_golden.py

The reference is imported from ``baseline/_ref`` (pip-installed copy of /root/reference) or,
failing that, from ``/root/reference``; the four pure-Python stand-ins in ``baseline/shims`` supply
its un-installable dependencies.  Every case is ``thunder.jit(fn, executors=["torch"])`` forward +
``torch.autograd.grad`` backward on CPU, i.e. the reference's torch-executor output that
BASELINE.json names as the parity target.  While generating, each case is also checked BIT-EXACTLY
against ``oracle/regions.py`` so a drift between restatement and reference fails here, loudly.
"""

from __future__ import annotations

import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.dont_write_bytecode = True
sys.path.insert(0, os.path.join(ROOT, "baseline", "shims"))
for cand in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
    if os.path.isdir(os.path.join(cand, "thunder")):
        sys.path.insert(0, cand)
        break
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import thunder  # noqa: E402

from oracle import regions as O  # noqa: E402

DT = {"bf16": torch.bfloat16, "f32": torch.float32, "f16": torch.float16}


def _rand(gen, shape, dtype, scale=1.0):
    return (torch.randn(shape, generator=gen, dtype=torch.float32) * scale).to(dtype)


def _run_ref(fn, inputs, diff_names, cotangents):
    """inputs: dict name->tensor (ordered = positional args). Returns (outs tuple, grads dict)."""
    args = []
    for k, v in inputs.items():
        t = v.clone()
        if k in diff_names:
            t.requires_grad_(True)
        args.append(t)
    jf = thunder.jit(fn, executors=["torch"])
    outs = jf(*args)
    if not isinstance(outs, (tuple, list)):
        outs = (outs,)
    diff_args = [a for a, k in zip(args, inputs) if k in diff_names]
    grads = torch.autograd.grad(list(outs), diff_args, grad_outputs=list(cotangents), allow_unused=True)
    return tuple(o.detach() for o in outs), {k: g for k, g in zip(diff_names, grads)}


def _same(a, b, what):
    if a.dtype != b.dtype or a.shape != b.shape or not torch.equal(a, b):
        diff = (a.float() - b.float()).abs().max().item() if a.shape == b.shape else "shape"
        raise AssertionError(f"oracle != reference for {what}: max abs diff {diff}")


def case_rmsnorm(gen, dtype, rows, hidden, eps=1e-5):
    x, r = _rand(gen, (rows, hidden), dtype), _rand(gen, (rows, hidden), dtype)
    w = _rand(gen, (hidden,), dtype)
    dh, dy = _rand(gen, (rows, hidden), dtype), _rand(gen, (rows, hidden), dtype)

    def fn(x, r, w):
        h = x + r
        return h, F.rms_norm(h, (hidden,), w, eps)

    (h, y), g = _run_ref(fn, {"x": x, "r": r, "w": w}, ["x", "r", "w"], [dh, dy])
    oh, oy, rstd = O.rmsnorm_residual_fwd(x, r, w, eps)
    odx, odw = O.rmsnorm_residual_bwd(dy, dh, oh, w, rstd)
    _same(oh, h, "rmsnorm h"), _same(oy, y, "rmsnorm y")
    _same(odx, g["x"], "rmsnorm dx"), _same(odx, g["r"], "rmsnorm dr"), _same(odw, g["w"], "rmsnorm dw")
    return {
        "inputs": {"x": x, "r": r, "w": w, "dh": dh, "dy": dy},
        "outputs": {"h": h, "y": y, "dx": g["x"], "dw": g["w"]},
        "meta": {"eps": eps},
    }


def case_swiglu(gen, dtype, rows, inter):
    a, b, g = (_rand(gen, (rows, inter), dtype, 2.0) for _ in range(3))

    def fn(a, b):
        return F.silu(a) * b

    (out,), gr = _run_ref(fn, {"a": a, "b": b}, ["a", "b"], [g])
    _same(O.swiglu_fwd(a, b), out, "swiglu out")
    da, db = O.swiglu_bwd(a, b, g)
    _same(da, gr["a"], "swiglu da"), _same(db, gr["b"], "swiglu db")
    return {"inputs": {"a": a, "b": b, "g": g}, "outputs": {"out": out, "da": gr["a"], "db": gr["b"]}, "meta": {}}


def case_rope(gen, dtype, B, T, n_head, groups, hs, rope_n):
    qkv = _rand(gen, (B, T, (n_head + 2 * groups) * hs), dtype)
    cos, sin = _rand(gen, (T, rope_n), dtype), _rand(gen, (T, rope_n), dtype)
    kv_heads = n_head if (groups != n_head and groups != 1) else groups
    dq = _rand(gen, (B, n_head, T, hs), dtype)
    dk, dv = (_rand(gen, (B, kv_heads, T, hs), dtype) for _ in range(2))

    def apply_rope(x, cos, sin):
        head = x.size(-1)
        x1 = x[..., : head // 2]
        x2 = x[..., head // 2 :]
        rotated = torch.cat((-x2, x1), dim=-1)
        roped = (x * cos) + (rotated * sin)
        return roped.to(dtype=x.dtype)

    def fn(qkv, cos, sin):
        q_per_kv = n_head // groups
        t = qkv.view(B, T, groups, q_per_kv + 2, hs).permute(0, 2, 3, 1, 4)
        q, k, v = t.split((q_per_kv, 1, 1), dim=2)
        if groups != n_head and groups != 1:
            k = k.expand(B, groups, q_per_kv, T, hs)
            v = v.expand(B, groups, q_per_kv, T, hs)
        q = q.reshape(B, -1, T, hs)
        k = k.reshape(B, -1, T, hs)
        v = v.reshape(B, -1, T, hs)
        q_roped = apply_rope(q[..., :rope_n], cos, sin)
        k_roped = apply_rope(k[..., :rope_n], cos, sin)
        q = torch.cat((q_roped, q[..., rope_n:]), dim=-1)
        k = torch.cat((k_roped, k[..., rope_n:]), dim=-1)
        return q, k, v

    (q, k, v), gr = _run_ref(fn, {"qkv": qkv, "cos": cos, "sin": sin}, ["qkv"], [dq, dk, dv])
    oq, ok, ov = O.qkv_split_rope_fwd(qkv, cos, sin, n_head, groups, hs, rope_n)
    _same(oq, q, "rope q"), _same(ok, k, "rope k"), _same(ov.contiguous(), v.contiguous(), "rope v")
    odqkv = O.qkv_split_rope_bwd(dq, dk, dv, cos, sin, n_head, groups, hs, rope_n)
    _same(odqkv, gr["qkv"], "rope dqkv")
    return {
        "inputs": {"qkv": qkv, "cos": cos, "sin": sin, "dq": dq, "dk": dk, "dv": dv},
        "outputs": {"q": q, "k": k, "v": v.contiguous(), "dqkv": gr["qkv"]},
        "meta": {"n_head": n_head, "n_query_groups": groups, "head_size": hs, "rope_n_elem": rope_n},
    }


def case_bias_act(gen, dtype, rows, hidden, act):
    x, g = _rand(gen, (rows, hidden), dtype, 1.5), _rand(gen, (rows, hidden), dtype)
    b = _rand(gen, (hidden,), dtype)

    def fn(x, b):
        return F.gelu(x + b) if act == "gelu" else F.relu(x + b)

    (out,), gr = _run_ref(fn, {"x": x, "b": b}, ["x", "b"], [g])
    fwd, bwd = (O.bias_gelu_fwd, O.bias_gelu_bwd) if act == "gelu" else (O.bias_relu_fwd, O.bias_relu_bwd)
    _same(fwd(x, b), out, f"bias_{act} out")
    dx, db = bwd(x, b, g)
    _same(dx, gr["x"], f"bias_{act} dx"), _same(db, gr["b"], f"bias_{act} db")
    return {"inputs": {"x": x, "b": b, "g": g}, "outputs": {"out": out, "dx": gr["x"], "db": gr["b"]}, "meta": {"act": act}}


def case_ce(gen, dtype, rows, vocab, ignore_index, n_ignored):
    logits = _rand(gen, (rows, vocab), dtype, 2.0)
    target = torch.randint(0, vocab, (rows,), generator=gen)
    for i in range(n_ignored):
        target[(3 * i + 1) % rows] = ignore_index
    g = torch.tensor(1.0 if n_ignored == 0 else 0.75, dtype=dtype)

    def fn(logits, target):
        return F.cross_entropy(logits, target, ignore_index=ignore_index)

    (loss,), gr = _run_ref(fn, {"logits": logits, "target": target}, ["logits"], [g])
    oloss, logp, nvalid = O.cross_entropy_fwd(logits, target, ignore_index)
    _same(oloss, loss, "ce loss")
    _same(O.cross_entropy_bwd(g, logp, target, nvalid, ignore_index), gr["logits"], "ce dlogits")
    return {
        "inputs": {"logits": logits, "target": target, "g": g},
        "outputs": {"loss": loss, "dlogits": gr["logits"]},
        "meta": {"ignore_index": ignore_index},
    }


def case_layernorm(gen, dtype, rows, hidden, eps=1e-5, bias=True):
    x = _rand(gen, (rows, hidden), dtype, 1.5) + 0.25
    w, b = _rand(gen, (hidden,), dtype), _rand(gen, (hidden,), dtype)
    dy = _rand(gen, (rows, hidden), dtype)

    if bias:
        def fn(x, w, b):
            return F.layer_norm(x, (hidden,), w, b, eps)
        ins, names = {"x": x, "w": w, "b": b}, ["x", "w", "b"]
    else:
        def fn(x, w):
            return F.layer_norm(x, (hidden,), w, None, eps)
        ins, names = {"x": x, "w": w}, ["x", "w"]

    (y,), g = _run_ref(fn, ins, names, [dy])
    oy, mean, rstd = O.layernorm_fwd(x, w, b if bias else None, eps)
    odx, odw, odb = O.layernorm_bwd(dy, x, w, bias, mean, rstd)
    _same(oy, y, "layernorm y"), _same(odx, g["x"], "layernorm dx"), _same(odw, g["w"], "layernorm dw")
    outs = {"y": y, "dx": g["x"], "dw": g["w"]}
    if bias:
        _same(odb, g["b"], "layernorm db")
        outs["db"] = g["b"]
    return {"inputs": {**ins, "dy": dy}, "outputs": outs, "meta": {"eps": eps, "bias": bias}}


def case_softmax(gen, dtype, rows, cols):
    x = _rand(gen, (rows, cols), dtype, 2.0)
    g = _rand(gen, (rows, cols), dtype)
    (y,), gr = _run_ref(lambda x: F.softmax(x, -1), {"x": x}, ["x"], [g])
    _same(O.softmax_fwd(x), y, "softmax y"), _same(O.softmax_bwd(g, y), gr["x"], "softmax dx")
    return {"inputs": {"x": x, "g": g}, "outputs": {"y": y, "dx": gr["x"]}, "meta": {}}


def main():
    gen = torch.Generator().manual_seed(20261106)
    cases = {}
    for name, dt in DT.items():
        cases[f"rmsnorm_{name}_16x256"] = case_rmsnorm(gen, dt, 16, 256)
    cases["rmsnorm_bf16_5x4096"] = case_rmsnorm(gen, torch.bfloat16, 5, 4096)
    cases["rmsnorm_bf16_7x100"] = case_rmsnorm(gen, torch.bfloat16, 7, 100)  # unaligned hidden -> generic kernels
    cases["rmsnorm_f32_3x1000"] = case_rmsnorm(gen, torch.float32, 3, 1000, eps=1e-6)
    for name, dt in DT.items():
        cases[f"swiglu_{name}_8x512"] = case_swiglu(gen, dt, 8, 512)
    cases["swiglu_bf16_3x77"] = case_swiglu(gen, torch.bfloat16, 3, 77)
    cases["rope_bf16_gqa"] = case_rope(gen, torch.bfloat16, 2, 16, 8, 2, 16, 16)
    cases["rope_bf16_gqa_partial"] = case_rope(gen, torch.bfloat16, 1, 12, 4, 2, 32, 16)
    cases["rope_bf16_mha"] = case_rope(gen, torch.bfloat16, 2, 8, 4, 4, 16, 16)
    cases["rope_bf16_mqa"] = case_rope(gen, torch.bfloat16, 1, 8, 4, 1, 16, 16)
    cases["rope_f32_gqa"] = case_rope(gen, torch.float32, 1, 8, 4, 2, 16, 16)
    for act in ("gelu", "relu"):
        cases[f"bias_{act}_bf16_16x128"] = case_bias_act(gen, torch.bfloat16, 16, 128, act)
        cases[f"bias_{act}_f32_9x60"] = case_bias_act(gen, torch.float32, 9, 60, act)
    cases["ce_bf16_32x517"] = case_ce(gen, torch.bfloat16, 32, 517, -100, 3)
    cases["ce_bf16_8x4096_ignm1"] = case_ce(gen, torch.bfloat16, 8, 4096, -1, 2)
    cases["ce_f32_16x300"] = case_ce(gen, torch.float32, 16, 300, -100, 0)
    cases["ce_f16_8x129"] = case_ce(gen, torch.float16, 8, 129, -100, 1)
    for name, dt in DT.items():
        cases[f"layernorm_{name}_16x256"] = case_layernorm(gen, dt, 16, 256)
    cases["layernorm_bf16_5x768"] = case_layernorm(gen, torch.bfloat16, 5, 768)
    cases["layernorm_bf16_7x100_nobias"] = case_layernorm(gen, torch.bfloat16, 7, 100, bias=False)  # unaligned hidden
    cases["layernorm_f32_3x1000"] = case_layernorm(gen, torch.float32, 3, 1000, eps=1e-6)
    for name, dt in DT.items():
        cases[f"softmax_{name}_16x256"] = case_softmax(gen, dt, 16, 256)
    cases["softmax_bf16_5x1024"] = case_softmax(gen, torch.bfloat16, 5, 1024)
    cases["softmax_bf16_7x100"] = case_softmax(gen, torch.bfloat16, 7, 100)  # unaligned width -> generic kernel

    out = os.path.join(ROOT, "tests", "golden", "regions.pt")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    torch.save({"reference": f"lightning-thunder {thunder.__version__}", "torch": str(torch.__version__), "cases": cases}, out)
    print(f"wrote {len(cases)} cases to {out} ({os.path.getsize(out)} bytes); oracle == reference bit-exact on all of them")


if __name__ == "__main__":
    main()
