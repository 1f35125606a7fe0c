"""Multi-GPU measurements of the path (BASELINE configs 4 and 5), one process per GPU under torchrun:

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_multi.py [autoclip|tp|all]

autoclip: auto_clip_layer int2-g128 over all Llama-2-7B linears except q/k (160 tensors, 5.40 G weights, 512 synthetic
          calibration tokens), layers sharded over the ranks, one NCCL all-gather of the clip cache.  Wall time on the
          device = max over ranks, reported per phase.
tp:       Qwen3-8B packed int2-g128, column-parallel over the ranks: decode (batch 1) chain and prefill (M=2048) with
          the all-gather of the channel slices after every linear; parity of the gathered output against the
          single-GPU result is checked on the first layer.
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import quantizer as Q  # noqa: E402
from bitdistiller_b200 import sharded as S  # noqa: E402
from bitdistiller_b200.autoclip import auto_2clip_layer  # noqa: E402
from bitdistiller_b200.qmodule import pack_wq  # noqa: E402

LLAMA_CLIP = [("self_attn.v_proj", 4096, 4096), ("self_attn.o_proj", 4096, 4096), ("mlp.gate_proj", 11008, 4096),
              ("mlp.up_proj", 11008, 4096), ("mlp.down_proj", 4096, 11008)]
QWEN3_8B = [("q_proj", 4096, 4096), ("k_proj", 1024, 4096), ("v_proj", 1024, 4096), ("o_proj", 4096, 4096),
            ("gate_proj", 12288, 4096), ("up_proj", 12288, 4096), ("down_proj", 4096, 12288)]


def max_over_ranks(seconds, dev):
    t = torch.tensor([seconds], dtype=torch.float64, device=dev)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def bench_autoclip(dev, rank, world, n_layers=32):
    cfg = {"zero_point": True, "q_group_size": 128, "quant_type": "int"}
    # synthetic weights / activations of one layer per shape (regenerated per layer from a layer seed: the values
    # differ per layer, the memory does not grow with the model)
    def clip_layer(i):
        out = []
        for k, (name, co, ci) in enumerate(LLAMA_CLIP):
            g = torch.Generator(device=dev).manual_seed(1000 * i + k)
            w = (torch.randn(co, ci, device=dev, generator=g) * 0.02).bfloat16()
            ch = torch.exp(torch.randn(ci, device=dev, generator=g) * 0.5)
            x = (torch.randn(512, ci, device=dev, generator=g) * ch).bfloat16()
            mx, mn = auto_2clip_layer(w, x, 2, cfg)
            out.append((f"model.layers.{i}.{name}", mx, mn))
        return out

    clip_layer(0)  # warm-up (kernel attributes, allocator)
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    local = []
    for i in S.owned_layers(n_layers, rank, world):
        local += clip_layer(i)
    e[1].record()
    full = S.gather_clip_cache(local, dtype=torch.bfloat16, device=dev)
    e[2].record()
    torch.cuda.synchronize()
    t_search = max_over_ranks(e[0].elapsed_time(e[1]) * 1e-3, dev)
    t_total = max_over_ranks(e[0].elapsed_time(e[2]) * 1e-3, dev)
    assert len(full) == n_layers * len(LLAMA_CLIP)
    assert [n for n, _, _ in full] == [f"model.layers.{i}.{name}" for i in range(n_layers) for name, _, _ in LLAMA_CLIP]
    nbytes = sum(2 * mx.numel() * 2 for _, mx, _ in full)
    # parity of the gathered cache: layer 0 recomputed locally on every rank must equal what rank 0 contributed
    ref0 = clip_layer(0)
    same = all(torch.equal(a[1], b[1]) and torch.equal(a[2], b[2]) for a, b in zip(full[:len(LLAMA_CLIP)], ref0))
    return {"workload": "configs[3]: auto_clip_layer int2-g128, all Llama-2-7B linears except q/k, 512 tokens, 100 candidates",
            "n_gpus": world, "seconds_total": t_total, "seconds_search": t_search,
            "seconds_gather": t_total - t_search, "clip_cache_bytes": nbytes, "pairs": 32 * sum(co * ci // 128 for _, co, ci in LLAMA_CLIP),
            "includes": "synthetic weight/activation generation on the device (randn per tensor) inside the timed region",
            "gathered_layer0_equals_local_recompute": bool(same)}


def bench_tp(dev, rank, world, layers=36):
    from bitdistiller_b200.awq_inference_engine import gemv_forward_cuda
    from bitdistiller_b200.triton_kernels import gemm_w2
    shards, full0 = {}, {}
    for name, oc, ic in QWEN3_8B:
        if (oc, ic) in shards:
            continue
        g = torch.Generator(device=dev).manual_seed(oc * 7 + ic)
        w = (torch.randn(oc, ic, device=dev, generator=g) * ic ** -0.5).half()
        dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
        qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
        r0, r1 = S.row_shard(oc, rank, world)
        shards[(oc, ic)] = (qw[r0:r1].contiguous(), qz[r0:r1].contiguous(), qs[r0:r1].contiguous())
        full0[(oc, ic)] = (qw, qz, qs)
    # parity: gathered TP output == single-GPU output, decode and prefill, every shape
    ok, exact = True, True
    for (oc, ic), (qw, qz, qs) in full0.items():
        for M in (1, 64):
            x = torch.randn(M, ic, device=dev, generator=torch.Generator(device=dev).manual_seed(M)).half()
            sq, sz, ss = shards[(oc, ic)]
            if M == 1:
                ref, loc = gemv_forward_cuda(x, qw, qs, qz, 2, 128), gemv_forward_cuda(x, sq, ss, sz, 2, 128)
            else:
                ref, loc = gemm_w2(x, qw, qs, qz, 128), gemm_w2(x, sq, ss, sz, 128)
            got = S.all_gather_cols(loc)
            # same kernels on a row slice: bit-equal, except where the slice falls off the tcgen05 tile shape (OC/n % 256)
            ok = ok and bool(((got.float() - ref.float()).norm() / ref.float().norm()) < 2e-3)
            exact = exact and torch.equal(got, ref)
    del full0
    res = {"workload": "configs[4]: Qwen3-8B packed int2-g128, column-parallel", "n_gpus": world,
           "gathered_matches_single_gpu_rel2e-3": bool(ok), "gathered_bit_equal": bool(exact)}
    # per-layer distinct weight memory, like tools/bench_gemv.py
    ops = []
    for li in range(layers):
        for name, oc, ic in QWEN3_8B:
            sq, sz, ss = (t.clone() for t in shards[(oc, ic)])
            ops.append((sq, sz, ss, oc, ic))
    for label, M, iters in (("decode_b1", 1, 10), ("prefill_m2048", 2048, 3)):
        xs = {ic: torch.randn(M, ic, device=dev).half() for ic in (4096, 12288)}

        def token():
            for sq, sz, ss, oc, ic in ops:
                if M == 1:
                    y = gemv_forward_cuda(xs[ic], sq, ss, sz, 2, 128)
                else:
                    y = gemm_w2(xs[ic], sq, ss, sz, 128)
                S.all_gather_cols(y)
        token()
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            token()
        e1.record()
        torch.cuda.synchronize()
        t = max_over_ranks(e0.elapsed_time(e1) * 1e-3 / iters, dev)
        if M == 1:
            res[label] = {"us_per_token": t * 1e6, "tok_s": 1 / t, "launch_mode": "eager stream launches + NCCL all-gather per linear"}
        else:
            flop = 2.0 * M * sum(oc * ic for _, oc, ic in QWEN3_8B) * layers
            res[label] = {"ms": t * 1e3, "tflops_aggregate": flop / t / 1e12}
    if world > 1:
        try:
            res["fused_epilogue"] = bench_tp_fused(dev, rank, world, shards, ops, layers)
        except Exception as e:  # symmetric memory needs P2P-capable GPUs; report instead of failing the whole run
            res["fused_epilogue"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    return res


def bench_tp_fused(dev, rank, world, shards, ops, layers):
    """Same workloads with the all-gather fused into the GEMV / GEMM epilogues (NVLink peer stores into symmetric
    memory, one cross-rank barrier per launch instead of an NCCL all-gather per linear)."""
    from bitdistiller_b200.awq_inference_engine import gemv_forward_cuda
    from bitdistiller_b200.triton_kernels import gemm_w2
    out = {}
    M = 2048
    rows1 = {oc: S.SymmRows(1, oc, dev) for oc in (4096, 1024, 12288)}
    rowsM = {oc: S.SymmRows(M, oc, dev) for oc in (4096, 1024, 12288)}
    # parity against the NCCL-gathered result (bit-equal to the single-GPU result, checked above)
    ok = True
    for (oc, ic), (sq, sz, ss) in shards.items():
        x = torch.randn(1, ic, device=dev, generator=torch.Generator(device=dev).manual_seed(7)).half()
        ref = S.all_gather_cols(gemv_forward_cuda(x, sq, ss, sz, 2, 128))
        S.tp_gemv_fused(x, [(sq, ss, sz)], [rows1[oc]])
        rows1[oc].barrier()
        ok = ok and torch.equal(rows1[oc].t, ref)
        xm = torch.randn(256, ic, device=dev, generator=torch.Generator(device=dev).manual_seed(8)).half()
        if sq.shape[0] % 32 == 0:
            refm = S.all_gather_cols(gemm_w2(xm, sq, ss, sz, 128))
            S.tp_gemm_fused(xm, (sq, ss, sz), rowsM[oc])
            rowsM[oc].barrier()
            ok = ok and torch.equal(rowsM[oc].t[:256], refm)
        rows1[oc].barrier()
    out["fused_equals_nccl_gather"] = bool(ok)
    names = [n for n, _, _ in QWEN3_8B]
    x1 = {ic: torch.randn(1, ic, device=dev).half() for ic in (4096, 12288)}

    def token_decode():
        for li in range(layers):
            o = {n: ops[li * 7 + k] for k, n in enumerate(names)}
            for grp in (("q_proj", "k_proj", "v_proj"), ("o_proj",), ("gate_proj", "up_proj"), ("down_proj",)):
                sh = [(o[n][0], o[n][2], o[n][1]) for n in grp]
                # q / gate own the full-size rows; k, v and up get their own so the slices do not collide
                rws = [rows1[o[n][3]] if i == 0 else extra[(n, o[n][3])] for i, n in enumerate(grp)]
                S.tp_gemv_fused(x1[o[grp[0]][4]], sh, rws, flags=1)
                rws[0].barrier()
    extra = {("k_proj", 1024): rows1[1024], ("v_proj", 1024): S.SymmRows(1, 1024, dev), ("up_proj", 12288): S.SymmRows(1, 12288, dev)}
    token_decode()
    torch.cuda.synchronize()
    dist.barrier()
    mode = "eager"
    run = token_decode
    try:
        st = torch.cuda.Stream(dev)
        with torch.cuda.stream(st):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                token_decode()
        gr.replay()
        torch.cuda.synchronize()
        run, mode = gr.replay, "CUDA graph"
    except Exception as e:
        out["graph_capture_failed"] = f"{type(e).__name__}: {e}"[:200]
        torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    t = max_over_ranks(e0.elapsed_time(e1) * 1e-3 / 10, dev)
    out["decode_b1"] = {"us_per_token": t * 1e6, "tok_s": 1 / t, "launch_mode": mode + ", 4 fused launches + 4 symmetric-memory barriers per layer"}
    xm = {ic: torch.randn(M, ic, device=dev).half() for ic in (4096, 12288)}

    def prefill():
        for sq, sz, ss, oc, ic in ops:
            if sq.shape[0] % 32 == 0:
                S.tp_gemm_fused(xm[ic], (sq, ss, sz), rowsM[oc])
                rowsM[oc].barrier()
            else:
                S.all_gather_cols(gemm_w2(xm[ic], sq, ss, sz, 128))
    prefill()
    torch.cuda.synchronize()
    dist.barrier()
    e0.record()
    for _ in range(3):
        prefill()
    e1.record()
    torch.cuda.synchronize()
    t = max_over_ranks(e0.elapsed_time(e1) * 1e-3 / 3, dev)
    flop = 2.0 * M * sum(oc * ic for _, oc, ic in QWEN3_8B) * layers
    out["prefill_m2048"] = {"ms": t * 1e3, "tflops_aggregate": flop / t / 1e12}
    return out


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    out = {}
    if what in ("autoclip", "all"):
        out["autoclip"] = bench_autoclip(dev, rank, world)
    if what in ("tp", "all"):
        out["tp"] = bench_tp(dev, rank, world)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
