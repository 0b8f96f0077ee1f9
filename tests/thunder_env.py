"""Makes the unmodified reference importable for the plugin tests: baseline/_ref (pip --target install that
travels with the repo) or /root/reference (build container only), plus the dependency shims."""

import os
import sys
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def import_thunder():
    """Returns (thunder, b200 executor module) or raises ImportError."""
    shims = os.path.join(ROOT, "baseline", "shims")
    ref = None
    for cand in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if os.path.isdir(os.path.join(cand, "thunder")):
            ref = cand
            break
    if ref is None:
        raise ImportError("reference thunder not found (baseline/_ref missing): pip --target install described in DESIGN.md")
    for p in (ref, shims, os.path.join(ROOT, "lightning-thunder_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import thunder
        import b200ex.executor as b200
    return thunder, b200


def fusions(trace):
    """[(name, launch kinds)] of the b200 fusions in a trace."""
    out = []
    for bs in trace.bound_symbols:
        if bs.sym.is_fusion and bs.sym.name.startswith("b200Fusion"):
            out.append((bs.sym.name, list(bs._call_ctx[bs.sym.name].launch_kinds)))
    return out
