"""Makes the unmodified reference importable for the plugin tests: baseline/_ref (the pip --target install that
``__graft_entry__.build()`` creates and that travels with the repo snapshot) plus the dependency shims.  Nothing here
reads /root/reference: that tree exists in the build container only."""

import os
import sys
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class ReferenceMissing(RuntimeError):
    """baseline/_ref is absent.  Deliberately not an ImportError, so no caller mistakes it for an optional import."""


def import_thunder():
    """Returns (thunder, b200 executor module); raises ReferenceMissing when baseline/_ref was never built."""
    shims = os.path.join(ROOT, "baseline", "shims")
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref, "thunder", "__init__.py")):
        raise ReferenceMissing(
            "baseline/_ref/thunder not found: run `python -c 'import __graft_entry__ as g; g.build()'` in the build "
            "container (it installs the reference there; the directory ships with the snapshot)"
        )
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
