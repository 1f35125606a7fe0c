"""Quantise-before-all-gather (sharded.ShardGather), one process per GPU:

  python tools/bench_shard_gather.py                                        # 1 GPU: encode / decode kernel rooflines
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_shard_gather.py

N > 1: the seven linears of one Llama-2-7B decoder layer, bf16, dim-0 sharded over the ranks.  Three ways to get the
fake-quantised full weights on every rank, timed on the device (max over ranks), and checked bit-equal to each other:
  reference order : NCCL all-gather of the bf16 shards, then SteInt2AsymQuantizer on the full weight (2 B/weight on the wire)
  nccl payload    : encode shard -> NCCL all-gather of the 0.28 B/weight payload -> decode
  fused           : encode kernel stores its payload into every rank's symmetric-memory buffer over NVLink -> barrier -> decode
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L  # noqa: E402
from bitdistiller_b200 import sharded as S  # noqa: E402
from bitdistiller_b200.quantizer import int_fakequant  # noqa: E402

LLAMA_LAYER = [("q_proj", 4096, 4096), ("k_proj", 4096, 4096), ("v_proj", 4096, 4096), ("o_proj", 4096, 4096),
               ("gate_proj", 11008, 4096), ("up_proj", 11008, 4096), ("down_proj", 4096, 11008)]
G = 128


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6462.4


def timed(fn, iters, dev):
    torch.cuda.synchronize(dev)
    if dist.is_initialized():
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        fn(i)
    b.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([a.elapsed_time(b) / iters], dtype=torch.float64, device=dev)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def single(dev):
    """One decoder layer's flat bf16 buffer (202.4 M weights) per launch, 3 rotating buffers (>> L2)."""
    n = sum(co * ci for _, co, ci in LLAMA_LAYER)
    ws = [(torch.randn(n // 4096, 4096, device=dev) * 0.02).bfloat16() for _ in range(3)]
    gs = [S.ShardGather(n // 4096, 4096, G, torch.bfloat16, dev, symmetric=False) for _ in range(3)]
    dq = torch.empty_like(ws[0])

    def enc(i):
        gs[i % 3]._encode(ws[i % 3])

    def dec(i):
        g = gs[i % 3]
        rc = L.lib().bdq_int2_shard_decode(L.ptr(g.codes), L.ptr(g.sz), L.ptr(dq), L.i64(n), G, L.BF16, L.stream_ptr(dev))
        L.check(rc, "decode")

    def fwd(i):
        int_fakequant(ws[i % 3], 2, G, L.ROUND_HALF_AWAY, out=dq)

    for f in (enc, dec, fwd):
        timed(f, 5, dev)
    pk = peak_gbs()
    payload = n / 4 + n / G * 4
    out = {"weights": n, "payload_bytes_per_weight": payload / n}
    for name, f, nbytes in [("encode", enc, 2 * n + payload), ("decode", dec, payload + 2 * n), ("fakequant_fwd", fwd, 4 * n)]:
        ms = timed(f, 30, dev)
        out[name] = {"ms": ms, "gbs": nbytes / ms / 1e6, "frac_of_hbm_peak": nbytes / ms / 1e6 / pk, "algorithmic_bytes": nbytes}
    assert torch.equal(gs[0].forward(ws[0]).view(torch.int16), int_fakequant(ws[0], 2, G, L.ROUND_HALF_AWAY)[0].view(torch.int16))
    return out


def multi(dev, rank, world):
    gen = torch.Generator(device=dev).manual_seed(1234)     # same seed on every rank: everyone can build the reference
    full = [(torch.randn(co, ci, device=dev, generator=gen) * 0.02).bfloat16() for _, co, ci in LLAMA_LAYER]
    spans = [S.row_shard(w.shape[0], rank, world) for w in full]
    shards = [w[a:b].contiguous() for w, (a, b) in zip(full, spans)]
    fused = [S.ShardGather(co, ci, G, torch.bfloat16, dev, symmetric=True) for _, co, ci in LLAMA_LAYER]
    plain = [S.ShardGather(co, ci, G, torch.bfloat16, dev, symmetric=False) for _, co, ci in LLAMA_LAYER]
    gathered = [torch.empty_like(w) for w in full]
    even = all((b - a) * world == w.shape[0] for w, (a, b) in zip(full, spans))

    def ref_order(i, keep=None):
        for k, (w, sh) in enumerate(zip(gathered, shards)):
            if even:
                dist.all_gather_into_tensor(w, sh)
            else:
                parts = [w[a:b] for a, b in (S.row_shard(w.shape[0], r, world) for r in range(world))]
                dist.all_gather(parts, sh)
            dq = int_fakequant(w, 2, G, L.ROUND_HALF_AWAY)[0]
            if keep is not None:
                keep.append(dq)

    def nccl_payload(i, keep=None):
        for g, sh in zip(plain, shards):
            dq = g.forward(sh)
            if keep is not None:
                keep.append(dq)

    def fused_path(i, keep=None):
        # one barrier pair per layer, not per tensor: encode all, barrier, decode all
        fused[0].hdl.barrier(channel=0)
        for g, sh in zip(fused, shards):
            g._encode(sh)
        fused[0].hdl.barrier(channel=0)
        for g in fused:
            dq = g._decode(g.codes, g.sz)
            if keep is not None:
                keep.append(dq)

    grp = S.ShardGatherGroup([(co, ci) for _, co, ci in LLAMA_LAYER], G, torch.bfloat16, dev, symmetric=True)

    def fused_group(i, keep=None):
        views = grp.forward(shards)                  # 7 encodes, one barrier pair, ONE decode over the layer's flat payload
        if keep is not None:
            keep.extend(v.clone() for v in views)

    a, b, c, d = [], [], [], []
    ref_order(0, a), nccl_payload(0, b), fused_path(0, c), fused_group(0, d)
    same_grp = all(torch.equal(x.view(torch.int16), y.view(torch.int16)) for x, y in zip(a, d))
    del d
    torch.cuda.synchronize(dev)
    same = all(torch.equal(x.view(torch.int16), y.view(torch.int16)) and torch.equal(x.view(torch.int16), z.view(torch.int16))
               for x, y, z in zip(a, b, c))
    local_ref = all(torch.equal(int_fakequant(w, 2, G, L.ROUND_HALF_AWAY)[0].view(torch.int16), x.view(torch.int16))
                    for w, x in zip(full, c))
    del a, b, c
    out = {"world": world, "bit_equal_three_ways": bool(same), "bit_equal_group": bool(same_grp), "bit_equal_to_unsharded": bool(local_ref),
           "weights_per_layer": sum(w.numel() for w in full)}
    for name, f in [("reference_order_bf16_allgather_then_quant", ref_order), ("nccl_payload", nccl_payload), ("fused_peer_stores", fused_path),
                    ("fused_peer_stores_layer_group_one_decode", fused_group)]:
        timed(f, 3, dev)
        out[name + "_ms_per_layer"] = timed(f, 20, dev)
    n = out["weights_per_layer"]
    out["wire_bytes_per_rank_bf16"] = 2 * n * (world - 1) / world
    out["wire_bytes_per_rank_payload"] = (n / 4 + n / G * 4) * (world - 1) / world
    return out


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        out = multi(dev, rank, world)
    else:
        out = single(dev)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
