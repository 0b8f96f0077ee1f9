"""GPU: every launcher is CUDA-graph capturable (no host sync, no cudaMalloc, current stream, allocator-owned scratch) —
what the reference's CUDAGraphTransform (thunder/transforms/cudagraph.py:229-387) needs from a fusion executor."""

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _step(ops, t):
    h, y, rstd = ops.rmsnorm_fwd(t["x"], t["r"], t["w"], 1e-5)
    dx, dw = ops.rmsnorm_bwd(t["dy"], h, t["w"], rstd, t["dh"])
    s = ops.swiglu_fwd(t["x"], t["r"])
    da, db = ops.swiglu_bwd(t["x"], t["r"], t["dy"])
    out, pre = ops.bias_act_fwd(t["x"], t["w"], "gelu", return_pre=True)
    gx, gb = ops.bias_act_bwd(pre, None, t["dy"], "gelu", want_dbias=True)
    loss, rmax, rls, nv = ops.cross_entropy_fwd(t["x"], t["tgt"], -100)
    dl = ops.cross_entropy_bwd(t["g"], t["x"], t["tgt"], rmax, rls, nv, -100)
    ln, mean, lrstd = ops.layernorm_fwd(t["x"], t["w"], t["w"], 1e-5)
    ldx, ldw, ldb = ops.layernorm_bwd(t["dy"], t["x"], t["w"], mean, lrstd, True)
    u = ops.uniform_philox((256, 1024), 0.0, 1.0, device=DEV, dtype=torch.bfloat16, seed=3, offset=8)
    return [h, y, rstd, dx, dw, s, da, db, out, gx, gb, loss, dl, ln, mean, lrstd, ldx, ldw, ldb, u]


def test_launchers_replay_from_a_cuda_graph():
    from b200ex import ops
    from b200ex.pointwise import PwLaunch, PwNode, compile_chain

    torch.manual_seed(0)
    mk = lambda *s: torch.randn(*s, device=DEV).to(torch.bfloat16)
    t = dict(x=mk(256, 1024), r=mk(256, 1024), w=mk(1024), dy=mk(256, 1024), dh=mk(256, 1024),
             tgt=torch.randint(0, 1024, (256,), device=DEV), g=torch.ones((), device=DEV, dtype=torch.bfloat16))
    (spec,) = compile_chain({"x": (torch.bfloat16, (256, 1024)), "r": (torch.bfloat16, (256, 1024))},
                            [PwNode("add", "o", torch.bfloat16, (256, 1024), ("x", "r"))], ["o"])
    pw = PwLaunch(spec)
    eager = _step(ops, t) + [pw.run(t)["o"]]  # also warms up lazy state (library load, plan cache, function attributes)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        _step(ops, t)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        captured = _step(ops, t) + [pw.run(t)["o"]]
    # new input values in the same buffers, replay, compare with an eager run on those values
    for k in ("x", "r", "dy", "dh"):
        t[k].copy_(mk(256, 1024))
    graph.replay()
    torch.cuda.synchronize()
    fresh = _step(ops, t) + [pw.run(t)["o"]]
    torch.cuda.synchronize()
    for i, (a, b) in enumerate(zip(captured, fresh)):
        assert torch.equal(a, b), f"output {i} differs between graph replay and eager"
    assert not all(torch.equal(a, b) for a, b in zip(eager, fresh))  # the inputs really changed


def test_reference_cudagraph_transform_over_b200_regions():
    """thunder.transforms.cudagraph.CUDAGraphTransform captures the whole forward / backward trace —
    our regions included — into CUDA graphs; replays on new input values must equal the un-graphed run."""
    from thunder_env import import_thunder

    thunder, b200 = import_thunder()
    from thunder.transforms.cudagraph import CUDAGraphTransform
    import torch.nn.functional as F

    class Block(torch.nn.Module):
        def __init__(self, d=256):
            super().__init__()
            self.ln = torch.nn.LayerNorm(d)
            self.fc, self.proj = torch.nn.Linear(d, 4 * d), torch.nn.Linear(4 * d, d)
            self.w = torch.nn.Parameter(torch.ones(d))

        def forward(self, x):
            h = x + self.proj(F.gelu(self.fc(self.ln(x))))
            return F.rms_norm(h, (h.shape[-1],), self.w, 1e-5)

    torch.manual_seed(0)
    model = Block().to(DEV).to(torch.bfloat16)
    xs = [torch.randn(64, 256, device=DEV, dtype=torch.bfloat16) for _ in range(3)]
    plain = thunder.jit(model, executors=[b200.ex])
    graphed = thunder.jit(model, executors=[b200.ex], transforms=[CUDAGraphTransform()])
    x_static = xs[0].clone().requires_grad_(True)
    for x in xs:  # same input buffer, new values: what a training loop with static buffers does
        with torch.no_grad():
            x_static.copy_(x)
        outs = []
        for jm in (plain, graphed):
            model.zero_grad(set_to_none=True)
            x_static.grad = None
            y = jm(x_static)
            y.float().square().mean().backward()
            outs.append((y.detach().clone(), x_static.grad.clone(), [p.grad.clone() for p in model.parameters()]))
        (y0, g0, p0), (y1, g1, p1) = outs
        assert torch.equal(y0, y1) and torch.equal(g0, g1)
        for a, b in zip(p0, p1):
            assert torch.equal(a, b)
