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


def assert_parity_or_closer(actual: torch.Tensor, expected: torch.Tensor, reference: torch.Tensor, *, what: str = ""):
    """The reference's own protocol for low-precision comparisons (thunder/tests/test_ops.py:84-107 with
    thunder/tests/framework.py:108-137): first the strict comparison of `actual` (b200) against `expected` (the
    reference's torch executor) at north_star's tolerances; where that fails, `actual` must be as close to a
    higher-precision `reference` (fp64 eager run of the same function) as `expected` is, within those tolerances —
    i.e. it may differ from the torch executor only where the torch executor itself is no nearer the truth."""
    try:
        assert_parity(actual, expected, what=what)
        return
    except AssertionError as strict:
        strict_msg = str(strict)
    a, e, r = actual.detach().cpu().double(), expected.detach().cpu().double(), reference.detach().cpu().double()
    assert a.shape == r.shape, f"{what}: reference shape {tuple(r.shape)} != {tuple(a.shape)}"
    cand, comp = (r - a).abs(), (r - e).abs()
    target = r + torch.where(r > a, -torch.minimum(cand, comp), torch.minimum(cand, comp))
    tol = TOL[actual.dtype]
    bound = tol["atol"] + tol["rtol"] * target.abs()
    err = (a - target).abs()
    bad = (err > bound) & ~torch.isnan(target) | (torch.isnan(a) != torch.isnan(r))
    if bad.any():
        idx = bad.flatten().nonzero()[0].item()
        raise AssertionError(
            f"{what}: {int(bad.sum())} / {a.numel()} elements are farther from the fp64 reference than the torch executor's; first at flat "
            f"index {idx}: b200 {a.flatten()[idx].item()} torch {e.flatten()[idx].item()} fp64 {r.flatten()[idx].item()}; strict check said: {strict_msg[:300]}"
        )


def assert_tensor_scale_parity(actual: torch.Tensor, expected: torch.Tensor, *, what: str = ""):
    """For tensors that are long sums of storage-rounded terms (parameter gradients out of GEMMs over bf16 activations
    that differ by an ulp here and there): north_star's rtol elementwise plus ONE storage ulp of the tensor's largest
    magnitude — the size of a single differently-rounded term — instead of the default atol."""
    scale = expected.detach().float().abs().max()
    assert_parity(actual, expected, what=what, ulp_scale=scale.expand_as(expected))
