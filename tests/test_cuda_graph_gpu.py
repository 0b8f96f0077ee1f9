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
    return [h, y, rstd, dx, dw, s, da, db, out, gx, gb, loss, dl]


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
