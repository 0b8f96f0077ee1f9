"""Details for forward+backward OpInfo cases: b200 vs the reference's torch executor.  usage: debug_opinfo_grad.py name:dtype ..."""
import os, sys, warnings, traceback
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(root, "tests"), root]
warnings.simplefilter("ignore")
import torch
import test_opinfo_gpu as T
from thunder_env import fusions
thunder = T.thunder
for spec in sys.argv[1:]:
    name, _, tag = spec.partition(":")
    op, dt = T._OPS[name], T._DTYPES[tag]
    for i, s in enumerate(op.sample_inputs("cuda", dt, requires_grad=True)):
        if i >= 3:
            break
        try:
            _, wo, wg = T._fwd_bwd(op, ["torch"], s, i)
        except Exception as e:
            print(f"{spec} sample {i}: torch arm raises {type(e).__name__}")
            continue
        shapes = [(tuple(a.shape), a.requires_grad) if isinstance(a, torch.Tensor) else a for a in s.args]
        try:
            jf, o, g = T._fwd_bwd(op, [T.b200.ex], s, i)
        except Exception as e:
            print(f"{spec} sample {i} args={shapes} kwargs={s.kwargs}: b200 raises {type(e).__name__}: {str(e)[:300]}")
            pass
            continue
        fk = fusions(thunder.last_traces(jf)[-1]); bt = thunder.last_backward_traces(jf); bk = fusions(bt[-1]) if bt else None
        def diff(a, b, what):
            if a is None or b is None:
                if (a is None) != (b is None): print(f"   {what}: None mismatch b200={a is None} torch={b is None}")
                return
            if a.dtype != b.dtype or a.shape != b.shape:
                print(f"   {what}: meta mismatch {a.dtype}{tuple(a.shape)} vs {b.dtype}{tuple(b.shape)}"); return
            try:
                torch.testing.assert_close(a, b, equal_nan=True)
            except AssertionError as e:
                print(f"   {what}: " + " ".join(str(e).split())[:260])
                fa, fb = a.flatten().float(), b.flatten().float()
                bad = (~torch.isclose(fa, fb, rtol=1e-2, atol=1e-5, equal_nan=True)).nonzero().flatten()[:3].tolist()
                for j in bad: print(f"      idx {j}: b200 {fa[j].item()} torch {fb[j].item()}")
        print(f"{spec} sample {i} args={shapes} kwargs={s.kwargs} fwd={fk} bwd={bk}")
        diff(o if isinstance(o, torch.Tensor) else None, wo if isinstance(wo, torch.Tensor) else None, "out")
        if wg is not None:
            for k, (a, b) in enumerate(zip(g, wg)): diff(a, b, f"grad[{k}]")
