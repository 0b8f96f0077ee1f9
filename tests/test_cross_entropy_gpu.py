"""GPU parity: fused softmax cross-entropy forward/backward through the C-ABI vs the oracle and golden vectors."""

import pytest
import torch

from oracle import regions as O
from parity import assert_parity

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cuda(logits, target, g, ignore_index):
    from b200ex import ops

    loss, rmax, rls, nvalid = ops.cross_entropy_fwd(logits.to(DEV), target.to(DEV), ignore_index)
    dl = ops.cross_entropy_bwd(g.to(DEV), logits.to(DEV), target.to(DEV), rmax, rls, nvalid, ignore_index)
    torch.cuda.synchronize()
    return loss, rmax, rls, nvalid, dl


def _check(logits, target, g, ignore_index, what):
    loss, rmax, rls, nvalid, dl = _cuda(logits, target, g, ignore_index)
    eloss, logp, envalid = O.cross_entropy_fwd(logits, target, ignore_index)
    edl = O.cross_entropy_bwd(g, logp, target, envalid, ignore_index)
    assert_parity(nvalid, envalid, what=what + " n_valid (bit-exact)")
    assert_parity(rmax, logits.float().max(1).values, what=what + " row max")
    # The loss is a mean of storage-dtype values; a flipped rounding of one row's -log p moves the
    # mean by ulp(-log p) / n.
    assert_parity(loss, eloss, what=what + " loss", ulp_scale=logp.float().abs().max().reshape(()))
    # dlogits = -exp(round(logp)) * d (+ d at the target): a one-ulp flip of the rounded log p scales the
    # element by exp(ulp(logp)), i.e. an error of |dlogits| * ulp * |logp|.
    scale = edl.float().abs() * logp.float().abs() * 2
    onehot = torch.zeros_like(edl, dtype=torch.float32)
    valid = target != ignore_index
    d = (g.float() / max(int(envalid), 1)).abs()
    onehot[valid, target[valid]] = d  # at the target the result is d - exp(logp) * d: rounding of a sum of that size
    assert_parity(dl, edl, what=what + " dlogits", ulp_scale=scale + onehot)
    # ignored rows get exactly zero gradient
    if (~valid).any():
        assert torch.count_nonzero(dl.cpu()[~valid]).item() == 0


def test_golden_vectors(golden):
    for name, c in golden.items():
        if not name.startswith("ce_"):
            continue
        i, o, m = c["inputs"], c["outputs"], c["meta"]
        loss, rmax, rls, nvalid, dl = _cuda(i["logits"], i["target"], i["g"], m["ignore_index"])
        _, logp, _ = O.cross_entropy_fwd(i["logits"], i["target"], m["ignore_index"])
        assert_parity(loss, o["loss"], what=name + " loss", ulp_scale=logp.float().abs().max().reshape(()))
        t = i["target"]
        valid = t != m["ignore_index"]
        onehot = torch.zeros_like(o["dlogits"], dtype=torch.float32)
        onehot[valid, t[valid]] = (i["g"].float() / max(int(valid.sum()), 1)).abs()
        assert_parity(dl, o["dlogits"], what=name + " dlogits", ulp_scale=o["dlogits"].float().abs() * logp.float().abs() * 2 + onehot)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("rows,vocab,ignore", [(1, 8, -100), (4, 33, -100), (16, 1000, -1), (9, 4096, -100), (3, 50257, -100), (5, 32064, -10)])
def test_shapes(dtype, rows, vocab, ignore):
    g = torch.Generator().manual_seed(rows * 7 + vocab)
    logits = (torch.randn(rows, vocab, generator=g) * 3).to(dtype)
    target = torch.randint(0, vocab, (rows,), generator=g)
    if rows > 2:
        target[1] = ignore
    _check(logits, target, torch.tensor(0.5, dtype=dtype), ignore, f"{dtype} {rows}x{vocab}")


def test_all_rows_ignored_gives_nan_loss_and_zero_grad():
    logits = torch.randn(4, 64).to(torch.bfloat16)
    target = torch.full((4,), -100)
    loss, _, _, nvalid, dl = _cuda(logits, target, torch.tensor(1.0, dtype=torch.bfloat16), -100)
    assert int(nvalid) == 0 and torch.isnan(loss.cpu()).item()
    assert torch.count_nonzero(dl.cpu()).item() == 0


def test_extreme_logits():
    logits = torch.tensor([[1e4, -1e4, 0.0, 5.0] * 4, [-30000.0] * 16, [88.0, 88.0] * 8], dtype=torch.float32)
    target = torch.tensor([0, 3, 1])
    _check(logits, target, torch.tensor(1.0), -100, "extreme fp32")
    _check(logits.to(torch.bfloat16), target, torch.tensor(1.0, dtype=torch.bfloat16), -100, "extreme bf16")


def test_reference_test_shape_fp32():
    """The shape the reference's own nvFuser cross-entropy test uses (thunder/tests/test_nvfuser.py:778-822)."""
    g = torch.Generator().manual_seed(5)
    logits = torch.randn(256, 32064, generator=g)
    target = torch.randint(0, 32064, (256,), generator=g)
    target[5] = -100
    _check(logits, target, torch.tensor(1.0), -100, "256x32064 fp32")


def test_llama3_vocab_rows():
    """BASELINE.json configs[3] vocab (128256) at 512 rows against the oracle; the full 8192 rows run in
    test_full_size_properties."""
    g = torch.Generator().manual_seed(0)
    rows, vocab = 512, 128256
    logits = torch.randn(rows, vocab, generator=g).to(torch.bfloat16)
    target = torch.randint(0, vocab, (rows,), generator=g)
    target[::97] = -100
    _check(logits, target, torch.tensor(1.0, dtype=torch.bfloat16), -100, "512x128256 bf16")


def test_full_size_properties():
    """configs[3] at full size (bf16 [8192, 128256]): each valid row of dlogits sums to ~0 and equals
    (softmax - onehot) * g / n_valid; the loss matches a float64 recomputation on the GPU."""
    from b200ex import ops

    torch.manual_seed(0)
    rows, vocab = 8192, 128256
    logits = torch.randn(rows, vocab, device=DEV, dtype=torch.bfloat16)
    target = torch.randint(0, vocab, (rows,), device=DEV)
    target[::113] = -100
    loss, rmax, rls, nvalid, = ops.cross_entropy_fwd(logits, target, -100)
    dl = ops.cross_entropy_bwd(torch.ones((), device=DEV, dtype=torch.bfloat16), logits, target, rmax, rls, nvalid, -100)
    valid = target != -100
    assert int(nvalid) == int(valid.sum())
    ref = torch.nn.functional.cross_entropy(logits[:1024].double(), target[:1024], ignore_index=-100, reduction="none")
    lse = (rmax + rls)[:1024].double()
    picked = logits[:1024].double().gather(1, target[:1024].clamp(min=0).unsqueeze(1)).squeeze(1)
    torch.testing.assert_close(torch.where(valid[:1024], lse - picked, torch.zeros_like(lse)), ref, rtol=1e-5, atol=1e-5)
    full = torch.nn.functional.cross_entropy(logits.float(), target, ignore_index=-100)
    torch.testing.assert_close(loss.float(), full, rtol=1.6e-2, atol=1e-5)
    rowsum = dl.float().sum(1)
    assert rowsum.abs().max().item() < 2e-5  # |d| = 1/8119 ~ 1.2e-4; bf16 rounding noise over 128k terms
    assert torch.count_nonzero(dl[~valid]).item() == 0


def test_full_size_against_the_oracle():
    """BASELINE.json configs[3] at full size, bf16 logits [8192, 128256], against the oracle itself (not only through
    properties): loss and n_valid from one oracle forward over all rows, dlogits compared in 1024-row slabs so that the
    host-side comparison stays within a few GB."""
    from b200ex import ops

    g = torch.Generator().manual_seed(0)
    rows, vocab, ignore = 8192, 128256, -100
    logits = torch.randn(rows, vocab, generator=g, dtype=torch.float32).to(torch.bfloat16)
    target = torch.randint(0, vocab, (rows,), generator=g)
    target[::113] = ignore
    gout = torch.tensor(1.0, dtype=torch.bfloat16)
    dev_logits, dev_target = logits.to(DEV), target.to(DEV)
    loss, rmax, rls, nvalid = ops.cross_entropy_fwd(dev_logits, dev_target, ignore)
    dl = ops.cross_entropy_bwd(gout.to(DEV), dev_logits, dev_target, rmax, rls, nvalid, ignore)
    torch.cuda.synchronize()
    eloss, logp, envalid = O.cross_entropy_fwd(logits, target, ignore)
    assert_parity(nvalid, envalid, what="full size n_valid (bit-exact)")
    assert_parity(loss, eloss, what="full size loss", ulp_scale=logp.float().abs().max().reshape(()))
    d = abs(float(gout)) / max(int(envalid), 1)
    for r0 in range(0, rows, 1024):
        sl = slice(r0, r0 + 1024)
        edl = O.cross_entropy_bwd(gout, logp[sl], target[sl], envalid, ignore)
        scale = edl.float().abs() * logp[sl].float().abs() * 2
        valid = target[sl] != ignore
        onehot = torch.zeros_like(edl, dtype=torch.float32)
        onehot[valid, target[sl][valid]] = d
        got = dl[sl].cpu()
        assert_parity(got, edl, what=f"full size dlogits rows {r0}..{r0 + 1023}", ulp_scale=scale + onehot)
        assert torch.count_nonzero(got[~valid]).item() == 0


def test_out_of_range_target_traps_like_aten():
    """A class index outside [0, vocab) that is not ignore_index is an error, as in ATen's nll_loss (device assert):
    the kernel reports it and traps instead of scoring the row as zero.  Run in a child process: the trap poisons the
    CUDA context."""
    import subprocess
    import sys

    code = (
        "import sys, torch; sys.path[:0] = %r; from b200ex import ops\n"
        "lg = torch.randn(8, 64, device='cuda').bfloat16(); t = torch.randint(0, 64, (8,), device='cuda'); t[3] = 64\n"
        "try:\n"
        "    out = ops.cross_entropy_fwd(lg, t, -100); torch.cuda.synchronize(); print('NO ERROR', out[0].item())\n"
        "except Exception as e:\n"
        "    print('CUDA ERROR', type(e).__name__)\n"
    ) % ([p for p in sys.path if p.endswith("lightning-thunder_b200")],)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert "CUDA ERROR" in r.stdout and "NO ERROR" not in r.stdout, (r.stdout, r.stderr[-500:])
    assert "outside [0, 64)" in r.stdout + r.stderr


@pytest.mark.parametrize("cluster", [2, 4])
@pytest.mark.parametrize("rows,vocab", [(3, 128256), (64, 40000), (300, 128256)])
def test_forward_on_thread_block_clusters(cluster, rows, vocab):
    """Few long rows: 2 or 4 CTAs of a thread-block cluster share a row and merge their (max, sum) pairs through
    distributed shared memory.  Same results as the oracle (and as the one-CTA form) within the usual bounds."""
    import os

    g = torch.Generator().manual_seed(rows + vocab)
    logits = (torch.randn(rows, vocab, generator=g) * 3).to(torch.bfloat16)
    target = torch.randint(0, vocab, (rows,), generator=g)
    target[1] = -100
    os.environ["B200K_CE_CLUSTER"] = str(cluster)
    try:
        _check(logits, target, torch.tensor(1.0, dtype=torch.bfloat16), -100, f"cluster {cluster} {rows}x{vocab}")
        got = _cuda(logits, target, torch.tensor(1.0, dtype=torch.bfloat16), -100)
        os.environ["B200K_CE_CLUSTER"] = "1"
        one = _cuda(logits, target, torch.tensor(1.0, dtype=torch.bfloat16), -100)
    finally:
        os.environ.pop("B200K_CE_CLUSTER", None)
    assert torch.equal(got[1].cpu(), one[1].cpu())  # row max: exact either way
    torch.testing.assert_close(got[2].cpu(), one[2].cpu(), rtol=1e-6, atol=1e-6)  # log-sum: another merge order


def test_default_policy_uses_clusters_for_few_rows():
    g = torch.Generator().manual_seed(1)
    logits = (torch.randn(100, 128256, generator=g) * 2).to(torch.bfloat16)
    target = torch.randint(0, 128256, (100,), generator=g)
    _check(logits, target, torch.tensor(1.0, dtype=torch.bfloat16), -100, "100x128256 (4-CTA clusters by default)")
