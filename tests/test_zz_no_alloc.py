"""GPU: b200k_rmsnorm_bwd_noalloc — the RMSNorm backward for hosts that may not allocate device memory.

Some embedders of this library cannot call cudaMalloc (or anything that allocates) where the backward runs:
all device memory belongs to their own pool, or the call sits in a context where allocation is forbidden.  For
them the library carries its dw scratch in a fixed array of its own; this test drives that entry point straight
through the C-ABI and checks, on two consecutive calls (the array keeps the first call's contents), that it
returns what the ordinary entry point returns with caller-provided scratch.

Collected last on purpose (file name), so a run that stops at the first failure has already run everything else;
it passes both alone in a fresh process and after the rest of the suite.
"""

import pytest
import torch

from parity import assert_parity

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ptr(t):
    return None if t is None else t.data_ptr()


def _inputs(rows, hidden, dtype, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    mk = lambda *s: torch.randn(*s, generator=g, device=DEV, dtype=torch.float32).to(dtype)
    dy, h, dh = mk(rows, hidden), mk(rows, hidden), mk(rows, hidden)
    w = mk(hidden)
    rstd = torch.rsqrt(h.float().pow(2).mean(-1) + 1e-5).contiguous()
    return dy, h, w, rstd, dh


def _reference_entry(L, dy, h, w, rstd, dh, code):
    rows, hidden = h.shape
    dx, dw = torch.empty_like(h), torch.empty_like(w)
    nbytes = L.b200k_rmsnorm_bwd_scratch_bytes(rows, hidden)
    scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=DEV)
    rc = L.b200k_rmsnorm_bwd(_ptr(dy), _ptr(h), _ptr(w), _ptr(rstd), _ptr(dh), _ptr(dx), _ptr(dw), _ptr(scratch), nbytes,
                             rows, hidden, code, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.b200k_error_string(rc)
    return dx, dw


def _noalloc_entry(L, dy, h, w, rstd, dh, code):
    rows, hidden = h.shape
    dx, dw = torch.empty_like(h), torch.empty_like(w)  # outputs are the caller's; the entry point allocates nothing
    rc = L.b200k_rmsnorm_bwd_noalloc(_ptr(dy), _ptr(h), _ptr(w), _ptr(rstd), _ptr(dh), _ptr(dx), _ptr(dw), rows, hidden, code,
                                     torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.b200k_error_string(rc)
    return dx, dw


@pytest.mark.parametrize("dtype,code", [(torch.bfloat16, 1), (torch.float32, 0)])
def test_two_calls_in_a_row_match_the_allocating_entry_point(dtype, code):
    from b200ex import _lib

    L = _lib.lib()
    with torch.cuda.device(0):
        # two different problems back to back: the second call finds the first call's partial sums in the array
        for seed, (rows, hidden) in enumerate([(8192, 4096), (777, 2048)]):
            args = _inputs(rows, hidden, dtype, seed)
            dx_ref, dw_ref = _reference_entry(L, *args, code)
            dx, dw = _noalloc_entry(L, *args, code)
            torch.cuda.synchronize()
            assert_parity(dx, dx_ref, what=f"noalloc dx call {seed} {dtype}")
            assert_parity(dw, dw_ref, what=f"noalloc dw call {seed} {dtype}")
            if rows <= 1024:
                # and against the oracle itself, so that this entry point cannot inherit a fault of its sibling
                from oracle import regions as O

                dy, h, w, rstd, dh = (t.cpu() for t in args)
                edx, edw = O.rmsnorm_residual_bwd(dy, dh, h, w, rstd.view(-1, 1))
                # same derived scales as tests/test_rmsnorm_gpu.py: dx = dh_in + round(dh_norm) can flip by one ulp of
                # the rounded intermediate; dw is a reassociated fp32 sum of storage-rounded terms
                assert_parity(dx, edx, what=f"noalloc dx vs oracle, call {seed} {dtype}", ulp_scale=(edx.float() - dh.float()).abs())
                terms = ((h.float() * rstd.view(-1, 1)) * dy.float()).to(dtype).float().abs()
                scale = terms.sum(0) * 4.0 if dtype == torch.float32 else terms.max(0).values + terms.sum(0) * 2.0**-12
                assert_parity(dw, edw, what=f"noalloc dw vs oracle, call {seed} {dtype}", ulp_scale=scale)

def test_allocates_no_device_memory_and_rejects_oversized_shapes():
    from b200ex import _lib

    L = _lib.lib()
    args = _inputs(256, 4096, torch.bfloat16, 7)
    with torch.cuda.device(0):
        _noalloc_entry(L, *args, 1)  # set-up call for this device (resolves the array's address)
        torch.cuda.synchronize()
        rows, hidden = args[1].shape
        dx, dw = torch.empty_like(args[1]), torch.empty_like(args[2])
        free_before, _ = torch.cuda.mem_get_info()
        rc = L.b200k_rmsnorm_bwd_noalloc(*[_ptr(t) for t in args], _ptr(dx), _ptr(dw), rows, hidden, 1, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert rc == 0 and free_after == free_before  # the driver handed out nothing during the call
        # a hidden size whose partial rows cannot fit the fixed array is refused with the documented code
        big = 16384
        need = L.b200k_rmsnorm_bwd_scratch_bytes(8192, big)
        assert need > (16 << 20)
        t = torch.empty(8, dtype=torch.bfloat16, device=DEV)
        rc = L.b200k_rmsnorm_bwd_noalloc(_ptr(t), _ptr(t), _ptr(t), _ptr(t), None, _ptr(t), _ptr(t), 8192, big, 1, torch.cuda.current_stream().cuda_stream)
        assert rc == -5 and b"maximum" in L.b200k_error_string(rc)
