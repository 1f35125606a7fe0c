"""quantized_all_gather forward + backward under torchrun (data parallel, dim-0 sharded weight): the shard gradient must equal
rows [r0, r1) of the STE backward of the FULL weight applied to the sum over ranks of the gradients w.r.t. the quantised
weight -- what FSDP2 computes with gather -> SteInt2AsymQuantizer -> autograd -> reduce-scatter (quantization/qlinear.py:63-78).

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/check_shard_gather_autograd.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import sharded as S  # noqa: E402
from bitdistiller_b200.quantizer import SteInt2AsymQuantizer  # noqa: E402


def main():
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    fails = []
    for rows, cols, G in [(4096, 4096, 128), (11008, 4096, 128), (256, 1024, 64)]:
        gen = torch.Generator(device=dev).manual_seed(7)                       # same full weight on every rank
        w_full = (torch.randn(rows, cols, device=dev, generator=gen) * 0.02).bfloat16()
        x = torch.randn(64, cols, device=dev, generator=torch.Generator(device=dev).manual_seed(100 + rank)).bfloat16()   # per-rank batch
        r0, r1 = S.row_shard(rows, rank, world)
        w_shard = w_full[r0:r1].clone().requires_grad_(True)
        g = S.ShardGather(rows, cols, G, torch.bfloat16, dev, symmetric=True)
        qw = S.quantized_all_gather(w_shard, g)
        y = torch.nn.functional.linear(x, qw)
        (y.float() ** 2).sum().backward()
        # reference order on every rank: full weight through the module, gradient w.r.t. the quantised weight summed over ranks
        wf = w_full.clone().requires_grad_(True)
        q = SteInt2AsymQuantizer(G)(wf)
        q.retain_grad()
        yr = torch.nn.functional.linear(x, q)
        (yr.float() ** 2).sum().backward(inputs=[q])
        gq = q.grad.clone()
        dist.all_reduce(gq)
        wf2 = w_full.clone().requires_grad_(True)
        SteInt2AsymQuantizer(G)(wf2).backward(gq)
        same_fwd = torch.equal(qw.view(torch.int16), q.detach().view(torch.int16))
        same_bwd = torch.equal(w_shard.grad.view(torch.int16), wf2.grad[r0:r1].contiguous().view(torch.int16))
        close = float((w_shard.grad.float() - wf2.grad[r0:r1].float()).abs().max()) <= 1e-2 * float(wf2.grad.float().abs().max())
        if not (same_fwd and (same_bwd or (world > 2 and close))):
            ok = False
            fails.append(f'single {rows}x{cols} fwd={same_fwd} bwd={same_bwd} rank={rank}')           # > 2 ranks: the bf16 sum order may differ
    # parameter group (one encode launch, one decode launch), both transports, forward and backward through autograd
    shapes = [(4096, 4096), (11008, 4096), (4096, 11008), (272, 1024)]           # 272 rows: ragged shards for world > 1
    gen = torch.Generator(device=dev).manual_seed(11)
    fulls = [(torch.randn(r, c, device=dev, generator=gen) * 0.02).bfloat16() for r, c in shapes]
    gys = [torch.randn(r, c, device=dev, generator=torch.Generator(device=dev).manual_seed(50 + rank)).bfloat16() for r, c in shapes]
    refs, ref_grads = [], []
    for wfull, gy in zip(fulls, gys):
        gsum = gy.clone()
        dist.all_reduce(gsum)
        wr = wfull.clone().requires_grad_(True)
        qr = SteInt2AsymQuantizer(128)(wr)
        qr.backward(gsum)
        refs.append(qr.detach())
        ref_grads.append(wr.grad)
    for symmetric in (True, False):
        grp = S.ShardGatherGroup(shapes, 128, torch.bfloat16, dev, symmetric=symmetric)
        spans = [S.row_shard(r, rank, world) for r, _ in shapes]
        shards = [w[a:b].clone().requires_grad_(True) for w, (a, b) in zip(fulls, spans)]
        outs = S.quantized_all_gather_group(shards, grp)
        for ti, (o, r) in enumerate(zip(outs, refs)):
            if not torch.equal(o.view(torch.int16), r.view(torch.int16)):
                ok = False
                fails.append(f'fwd sym={symmetric} t={ti} rank={rank}')
        torch.autograd.backward(list(outs), gys)
        for sh, rg, (a, b) in zip(shards, ref_grads, spans):
            same = torch.equal(sh.grad.view(torch.int16), rg[a:b].contiguous().view(torch.int16))
            close = float((sh.grad.float() - rg[a:b].float()).abs().max()) <= 1e-2 * float(rg.float().abs().max())
            if not (same or (world > 2 and close)):
                ok = False
                fails.append(f'bwd sym={symmetric} shape={tuple(sh.shape)} rank={rank} maxdiff={float((sh.grad.float() - rg[a:b].float()).abs().max())}')
    if fails:
        print(rank, fails, flush=True)
    t = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"world": world, "shard_gradient_equals_full_weight_ste_backward_rows": bool(t.item()), "includes": "ShardGather + ShardGatherGroup (symmetric and collective transport, ragged shards)"}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
