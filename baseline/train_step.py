"""Llama-3-8B-shape synthetic-token training step (BASELINE.json configs[4]) — benchmark infrastructure.

A LitGPT-style Llama (baseline/llama_model.py) is compiled by the UNMODIFIED reference (``thunder.jit``); the
fused elementwise/reduction regions run on the b200 executor (``--executor b200``) or on the reference's torch
executor (``--executor torch``), GEMMs and SDPA stay on the torch / sdpa executors, and for WORLD_SIZE > 1 the
reference's own ``thunder.distributed.fsdp`` (ZeRO-2: NCCL all-gather / reduce-scatter per parameter over NVLink)
shards the step.  Protocol of thunder/benchmarks/benchmark_litgpt.py:748-864: forward with the loss inside the
jitted callable, backward, optimizer step, ``loss.item()`` every iteration; tokens/s = world * B * T / step time.
"""

from __future__ import annotations

import json
import os
import time

import torch


def region_census(thunder, jm):
    """How many launches of each kind the b200 fusions of the compiled step contain."""
    out = {}
    try:
        traces = [thunder.last_traces(jm)[-1], thunder.last_backward_traces(jm)[-1]]
    except Exception:
        return out
    for tr in traces:
        for bs in tr.bound_symbols:
            if bs.sym.is_fusion and bs.sym.name.startswith("b200Fusion"):
                for k in bs._call_ctx[bs.sym.name].launch_kinds:
                    out[k] = out.get(k, 0) + 1
    return out


def _copy_stats(total_steps: int):
    """Strided operands that a family launcher had to make dense first (b200ex.ops.COPY_STATS), per step."""
    from b200ex import ops

    c = ops.COPY_STATS
    return {"copies_per_step": c["copies"] / max(total_steps, 1), "bytes_per_step": c["bytes"] / max(total_steps, 1)}


def run(args, thunder, ClockSampler, dist=None, emit=True):
    """Runs the step benchmark.  With emit=True rank 0 prints the JSON line; otherwise the dict is returned (rank 0)
    or None (other ranks), and an already initialised process group `dist` is reused and left open."""
    import warnings

    warnings.simplefilter("ignore")
    from baseline.llama_model import GPT, LlamaConfig, build_rope_cache
    from thunder.executors.sdpaex import sdpa_ex

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    own_pg = False
    if world > 1 and dist is None:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=device)
        own_pg = True
    # "litgpt": RMSNorm and the attention slice spelled with plain torch ops as litgpt does (b200 finds the regions
    # itself); "ops": F.rms_norm + b200ex.executor.qkv_split_rope (op-level route to the hand-written kernels)
    spelling = os.environ.get("B200_MODEL_SPELLING", getattr(args, "spelling", "litgpt"))
    cfg = LlamaConfig(n_layer=args.layers, spelling=spelling)
    B, T = args.micro_batch, args.seq_len
    torch.manual_seed(1234)
    t0 = time.perf_counter()
    with torch.device(device):
        model = GPT(cfg).to(torch.bfloat16)
    n_params = sum(p.numel() for p in model.parameters())
    cos, sin = build_rope_cache(T, cfg.rope_n_elem, base=cfg.rope_base, device=device, dtype=torch.bfloat16)
    # attention is a library kernel in both arms: cuDNN's fused attention unless B200_ATTENTION=flash (baseline/attention.py)
    attn = [sdpa_ex]
    attention_name = "ATen flash attention (sdpa executor)"
    if os.environ.get("B200_ATTENTION", "cudnn") == "cudnn":
        from baseline.attention import cudnn_sdpa_ex

        attn = [cudnn_sdpa_ex, sdpa_ex]
        attention_name = "cuDNN fused attention through ATen (baseline/attention.py), both arms"
    if args.executor == "b200":
        import b200ex.executor as b200

        execs = attn + [b200.ex]
    else:
        execs = attn + ["torch"]
    jm = thunder.jit(model, executors=execs, fusion_type=args.fusion_type)
    if world > 1:
        from thunder.distributed import FSDPBucketingStrategy, fsdp

        # the reference's own knob (thunder/distributed/__init__.py:337-350): NONE = one all-gather / reduce-scatter
        # per parameter (227 each per step), BLOCK = one per transformer block
        bucketing = {"none": FSDPBucketingStrategy.NONE, "layer": FSDPBucketingStrategy.LAYER, "block": FSDPBucketingStrategy.BLOCK}[
            os.environ.get("B200_FSDP_BUCKETING", getattr(args, "fsdp_bucketing", "none"))
        ]
        jm = fsdp(jm, bucketing_strategy=bucketing)
    # b200 arm: the multi-tensor AdamW kernel of this repo (b200ex.optim.AdamW, same arithmetic and state layout as
    # torch's fused AdamW, which the reference's benchmark uses); B200_TORCH_ADAMW=1 or the torch arm: torch's own.
    if args.executor == "b200" and os.environ.get("B200_TORCH_ADAMW", "0") != "1":
        from b200ex.optim import AdamW

        opt = AdamW(jm.parameters(), lr=1e-5)
        optimizer_name = "b200ex.optim.AdamW (b200k_adamw)"
    else:
        opt = torch.optim.AdamW(jm.parameters(), lr=1e-5, fused=True)
        optimizer_name = "torch.optim.AdamW(fused=True)"
    # pinned host token buffers: the step's inputs cross PCIe every iteration, as a data loader would deliver them
    g = torch.Generator().manual_seed(99 + rank)
    host = [
        (torch.randint(0, cfg.padded_vocab_size, (B, T), generator=g).pin_memory(), torch.randint(0, cfg.padded_vocab_size, (B, T), generator=g).pin_memory())
        for _ in range(4)
    ]

    def step(i):
        idx_h, tgt_h = host[i % len(host)]
        idx, tgt = idx_h.to(device, non_blocking=True), tgt_h.to(device, non_blocking=True)
        loss = jm(idx, cos, sin, tgt)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss.item()  # device -> host read of the step's result

    loss0 = step(0)  # traces + compiles
    compile_s = time.perf_counter() - t0
    for i in range(1, max(args.warmup, 3)):
        step(i)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(1.0)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(device)
    sampler.mark_load(True)
    stream = torch.cuda.current_stream(device)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    loss = loss0
    for i in range(args.steps):
        loss = step(i)
    ev1.record(stream)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(device)
    sampler.mark_load(False)
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=device, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    clocks = sampler.stop() if rank == 0 else None
    peak_mem = torch.cuda.max_memory_allocated(device) / 2**30
    if rank == 0:
        step_ms = ms.item() / args.steps
        tokens_per_s = world * B * T / (step_ms / 1e3)
        flops = 6.0 * n_params * B * T + 12.0 * cfg.n_layer * cfg.n_embd * B * T * T  # dense + causal attention estimate
        line = {
            "metric": "llama3_8b_train_tokens_per_s", "value": tokens_per_s, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"llama_step: LitGPT-style Llama-3-8B shape ({cfg.spelling} spelling), n_layer={cfg.n_layer}, micro-batch {B} x seq {T} per GPU, {optimizer_name}, attention: {attention_name}, "
                                   f"{'fsdp ZeRO-2 (thunder.distributed.fsdp, NCCL)' if world > 1 else 'single GPU'}",
                       "executors": [str(e) if isinstance(e, str) else e.name for e in execs], "fusion_type": args.fusion_type,
                       "params": n_params},
            "e2e": {"value": tokens_per_s, "unit": "tokens/s", "h2d_bytes_per_step": 2 * B * T * 8, "d2h_bytes_per_step": 2,
                    "note": "tokens and targets come from pinned host memory every step; loss.item() reads the result back"},
            "model_tflops_per_gpu": flops / (step_ms / 1e3) / 1e12,
            "compile_s": compile_s, "peak_mem_gib": peak_mem, "final_loss": loss,
            "b200_launches_per_step": region_census(thunder, jm) if args.executor == "b200" else {},
            "operand_copies": _copy_stats(args.steps + max(args.warmup, 3)) if args.executor == "b200" else None,
            "clocks": clocks,
        }
        if args.executor == "b200":
            from b200ex import jit

            line["pointwise_jit"] = dict(jit.cache_info(), enabled=jit.enabled())  # generic pointwise regions: NVRTC-specialised kernels
        line["gpu_launches"] = sum(line["b200_launches_per_step"].values()) * args.steps
        if emit:
            print(json.dumps(line), flush=True)
    else:
        line = None
    del jm, opt, model
    torch.cuda.empty_cache()
    if own_pg:
        dist.destroy_process_group()
    return line
