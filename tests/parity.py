"""Comparison helpers shared by the parity tests.

Tolerances are the ones BASELINE.json's north_star states: fp32 rtol 1e-5 / atol 1e-6, bf16 rtol
1.6e-2 (atol 1e-5, torch's default for bf16), integer/index work bit-exact.  fp16 uses torch's
default (rtol 1e-3, atol 1e-5).

``ulp_scale``: some outputs are a storage-dtype rounding of an intermediate followed by more
arithmetic (``dx = dh_in + round(dh)``, ``dlogits = f(exp(round(logp)))``).  The CUDA kernels round
at the same points as the reference, but a last-bit fp32 difference upstream (reduction order,
exp/log implementation) can flip such a rounding by one ulp of the INTERMEDIATE, which is unbounded
relative to a cancelling result.  The reference's own CPU and CUDA runs differ the same way, and its
test-suite uses ``assert_closer`` for it (thunder/tests/test_ops.py:84-107).  Where that can happen the
test passes ``ulp_scale`` = magnitude of the rounded intermediate; the bound grows by one ulp of it.
"""

from __future__ import annotations

import torch

TOL = {
    torch.float32: dict(rtol=1e-5, atol=1e-6),
    torch.bfloat16: dict(rtol=1.6e-2, atol=1e-5),
    torch.float16: dict(rtol=1e-3, atol=1e-5),
}
ULP = {torch.float32: 2.0**-23, torch.bfloat16: 2.0**-7, torch.float16: 2.0**-10}


def assert_parity(actual: torch.Tensor, expected: torch.Tensor, *, what: str = "", ulp_scale: torch.Tensor | None = None, atol_extra: float = 0.0):
    assert actual.dtype == expected.dtype, f"{what}: dtype {actual.dtype} != {expected.dtype}"
    assert tuple(actual.shape) == tuple(expected.shape), f"{what}: shape {tuple(actual.shape)} != {tuple(expected.shape)}"
    a = actual.detach().cpu()
    e = expected.detach().cpu()
    if not a.dtype.is_floating_point:
        assert torch.equal(a, e), f"{what}: integer/index output is not bit-exact"
        return
    tol = TOL[a.dtype]
    a64, e64 = a.double(), e.double()
    bound = tol["atol"] + atol_extra + tol["rtol"] * e64.abs()
    if ulp_scale is not None:
        bound = bound + ULP[a.dtype] * ulp_scale.detach().cpu().double().abs()
    err = (a64 - e64).abs()
    nan_mismatch = torch.isnan(a64) != torch.isnan(e64)
    bad = (err > bound) & ~torch.isnan(e64) | nan_mismatch
    if bad.any():
        idx = bad.flatten().nonzero()[0].item()
        raise AssertionError(
            f"{what}: {int(bad.sum())} / {a.numel()} elements out of tolerance; first at flat index {idx}: "
            f"actual {a.flatten()[idx].item()} expected {e.flatten()[idx].item()} bound {bound.flatten()[idx].item():.3e}; "
            f"max abs err {err[~torch.isnan(err)].max().item():.3e}"
        )
