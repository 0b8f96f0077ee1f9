"""CPU, world_size 2, gloo: the b200 executor in the executor list together with the reference's data-parallel
transform (``thunder.distributed.ddp``).  CPU tensors are not claimed by b200, so this checks the N > 1 host
logic that does not need a GPU: our fusion pass leaves the collective prims of the transformed trace alone,
gradients come back all-reduced and identical on both ranks, and the bench's max-over-ranks reduction works.
(``thunder.distributed.fsdp`` itself is CUDA/NCCL only in the reference — tests/distributed/test_fsdp.py:768 —
and is exercised on the GPU box by ``bench.py --workload llama_step``.)"""

import os
import tempfile

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from thunder_env import import_thunder

import_thunder()  # raises (the module errors, it does not skip) when baseline/_ref is missing


def _worker(rank, world, init_file, result_file):
    import warnings

    warnings.simplefilter("ignore")
    thunder, b200 = import_thunder()
    from thunder.distributed import ddp

    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)  # same initial weights on every rank
        model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 4))
        ref = [p.detach().clone() for p in model.parameters()]
        jm = ddp(thunder.jit(model, executors=[b200.ex]))
        g = torch.Generator().manual_seed(100 + rank)  # different data per rank
        x = torch.randn(8, 16, generator=g)
        jm(x).square().mean().backward()
        grads = [p.grad.detach().clone() for p in jm.parameters()]
        # expected: mean over ranks of the local gradients, computed without any executor or transform
        local = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 4))
        with torch.no_grad():
            for p, r in zip(local.parameters(), ref):
                p.copy_(r)
        expect = []
        for rk in range(world):
            local.zero_grad()
            xr = torch.randn(8, 16, generator=torch.Generator().manual_seed(100 + rk))
            local(xr).square().mean().backward()
            expect.append([p.grad.clone() for p in local.parameters()])
        mean = [sum(gs) / world for gs in zip(*expect)]
        ok = all(torch.allclose(a, b, rtol=1e-4, atol=1e-5) for a, b in zip(grads, mean))
        # the bench's multi-rank timing reduction: max over ranks
        t = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok = ok and t.item() == 10.0 + world - 1
        has_allreduce = "all_reduce" in str(thunder.last_backward_traces(jm)[-1])
        if rank == 0:
            with open(result_file, "w") as f:
                f.write(f"{int(ok)} {int(has_allreduce)}")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_ddp_gloo_world_size_2_with_b200_in_the_executor_list():
    with tempfile.TemporaryDirectory() as d:
        init_file, result_file = os.path.join(d, "init"), os.path.join(d, "result")
        mp.spawn(_worker, args=(2, init_file, result_file), nprocs=2, join=True)
        ok, has_allreduce = open(result_file).read().split()
    assert ok == "1", "gradients after ddp differ from the mean of the per-rank gradients"
    assert has_allreduce == "1", "the data-parallel all_reduce is missing from the backward trace"


def _fsdp_worker(rank, world, init_file, result_file):
    import warnings

    warnings.simplefilter("ignore")
    thunder, b200 = import_thunder()
    from thunder.distributed import fsdp

    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        model = torch.nn.Sequential(torch.nn.Linear(16, 32, bias=False), torch.nn.GELU(), torch.nn.Linear(32, 4, bias=False))
        ref = [p.detach().clone() for p in model.parameters()]
        jm = fsdp(thunder.jit(model, executors=[b200.ex]), device=torch.device("cpu"))
        g = torch.Generator().manual_seed(100 + rank)
        x = torch.randn(8, 16, generator=g)
        jm(x).square().mean().backward()
        shards = [p.grad.detach().clone() for p in jm.parameters()]
        local = torch.nn.Sequential(torch.nn.Linear(16, 32, bias=False), torch.nn.GELU(), torch.nn.Linear(32, 4, bias=False))
        with torch.no_grad():
            for p, r in zip(local.parameters(), ref):
                p.copy_(r)
        expect = []
        for rk in range(world):
            local.zero_grad()
            xr = torch.randn(8, 16, generator=torch.Generator().manual_seed(100 + rk))
            local(xr).square().mean().backward()
            expect.append([p.grad.clone() for p in local.parameters()])
        mean = [sum(gs) / world for gs in zip(*expect)]
        # ZeRO: every rank holds the dim-0 shard of the averaged gradient
        ok = all(torch.allclose(s, m.chunk(world, 0)[rank], rtol=1e-4, atol=1e-5) for s, m in zip(shards, mean))
        bw = str(thunder.last_backward_traces(jm)[-1])
        if rank == 0:
            with open(result_file, "w") as f:
                f.write(f"{int(ok)} {int('reduce_scatter' in bw)}")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_fsdp_gloo_world_size_2_gradient_shards_are_the_mean_over_ranks():
    """thunder.distributed.fsdp with b200 in the executor list: b200's rule for dist_prims.synchronize (the reference's
    spelling on a backend without AVG) all-gathers the parameter shards forward and reduce-scatters the gradients."""
    with tempfile.TemporaryDirectory() as d:
        init_file, result_file = os.path.join(d, "init"), os.path.join(d, "result")
        mp.spawn(_fsdp_worker, args=(2, init_file, result_file), nprocs=2, join=True)
        ok, has_rs = open(result_file).read().split()
    assert ok == "1", "gradient shards after fsdp differ from the shard of the mean of the per-rank gradients"
    assert has_rs == "1", "the reduce_scatter is missing from the backward trace"
