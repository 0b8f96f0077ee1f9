"""Details for OpInfo cases: b200 vs torch reference vs the reference's torch executor.  usage: debug_opinfo.py name[:dtype] ..."""
import os, sys, warnings
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(root, "tests"), root]
warnings.simplefilter("ignore")
import torch
from thunder_env import fusions, import_thunder
thunder, b200 = import_thunder()
import thunder.core.dtypes as dtypes
from thunder.tests import opinfos
from torch.testing import assert_close
DT = {"f32": dtypes.float32, "bf16": dtypes.bfloat16, "f16": dtypes.float16}
for spec in sys.argv[1:]:
    name, _, tag = spec.partition(":")
    op = next(o for o in opinfos.opinfos if o.name == name)
    for t in ([tag] if tag else list(DT)):
        if DT[t] not in op.dtypes():
            continue
        for i, s in enumerate(op.sample_inputs("cuda", DT[t])):
            if i >= 8:
                break
            res = {}
            for label, ex in (("b200", [b200.ex]), ("torch", ["torch"])):
                try:
                    jf = thunder.jit(op.op, executors=ex, cache="no caching", disable_torch_autograd=True)
                    res[label] = jf(*s.args, **s.kwargs)
                    if label == "b200":
                        res["kinds"] = fusions(thunder.last_traces(jf)[-1])
                except Exception as e:
                    res[label] = e
            try:
                want = op.torch_reference(*s.args, **s.kwargs)
            except Exception as e:
                want = e
            def cmp(a, b):
                if isinstance(a, Exception) or isinstance(b, Exception):
                    return f"EXC {type(a).__name__ if isinstance(a, Exception) else ''} {str(a)[:80] if isinstance(a, Exception) else ''} | {str(b)[:80] if isinstance(b, Exception) else ''}"
                try:
                    (s.comp or assert_close)(a, b)
                    return "ok"
                except AssertionError as e:
                    return "MISMATCH " + " ".join(str(e).split())[:230]
            r1, r2 = cmp(res["b200"], want), cmp(res["torch"], want)
            if r1 != "ok" or r2 != "ok":
                shapes = [tuple(a.shape) if isinstance(a, torch.Tensor) else a for a in s.args]
                print(f"{name}:{t} sample {i} args={shapes} kwargs={s.kwargs} kinds={res.get('kinds')}\n   b200 : {r1}\n   torch: {r2}")
                if r1.startswith("MISMATCH") and isinstance(want, torch.Tensor) and want.numel() > 0:
                    g, w = res["b200"].flatten().float(), want.flatten().float()
                    bad = (~torch.isclose(g, w, rtol=1e-2, atol=1e-5, equal_nan=True)).nonzero().flatten()[:4].tolist()
                    flat = [a.flatten() if isinstance(a, torch.Tensor) and a.numel() == want.numel() else None for a in s.args]
                    for j in bad:
                        print("      idx", j, "in", [f[j].item() if f is not None else "-" for f in flat], "b200", g[j].item(), "ref", w[j].item(),
                              "torchex", res["torch"].flatten()[j].item() if isinstance(res["torch"], torch.Tensor) else "-")
