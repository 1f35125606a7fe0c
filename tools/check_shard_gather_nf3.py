"""2+ ranks (torchrun): ShardGatherNF3 forward == SteN2F3Quantizer on the gathered weight on every rank (symmetric-memory and
collective transport, ragged shards), backward == rows of the full NF3 STE backward of the summed gradient; plus the kernel
rooflines of bdq_nf3_shard_encode / _decode on one decoder layer's flat buffer (rank 0)."""
import json, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L
from bitdistiller_b200 import sharded as S
from bitdistiller_b200.quantizer import SteN2F3Quantizer, nf3_fakequant

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
ok = True
if world > 1:
    for rows, cols, sym, G in [(4096, 4096, True, 128), (11008, 4096, True, 128), (48, 512, False, 128), (4096, 11008, False, 128),
                               (4096, 4096, True, 64), (1024, 2048, True, 32)]:   # 64: the train scripts' group size
        gen = torch.Generator(device=dev).manual_seed(rows + cols)
        full = (torch.randn(rows, cols, device=dev, generator=gen) * 0.02).bfloat16()
        a, b = S.row_shard(rows, rank, world)
        sg = S.ShardGatherNF3(rows, cols, G, torch.bfloat16, dev, symmetric=sym)
        for rep in range(3):
            dq = sg.forward(full[a:b].contiguous())
            ref = nf3_fakequant(full, G)[0]
            ok = ok and bool(torch.equal(dq.view(torch.int16), ref.view(torch.int16)))
        gen.manual_seed(1000 + rank)
        g = torch.randn(rows, cols, device=dev, generator=gen).bfloat16()
        gw = sg.backward(full[a:b].contiguous(), g)
        gsum = g.clone()
        dist.all_reduce(gsum)
        x = full.clone().requires_grad_(True)
        SteN2F3Quantizer(G)(x).backward(gsum)
        same = (gw.view(torch.int16) == x.grad[a:b].view(torch.int16)) | (gw.isnan() & x.grad[a:b].isnan())
        ok = ok and (bool(same.all()) if world == 2 else float((gw.float() - x.grad[a:b].float()).abs().max()) < 1e-2)
        if sym and sg.enable_fused_backward():   # gradient push over NVLink + bdq_nf3_ste_bwd_nsrc: same result, no collective
            for rep in range(2):
                gw2 = sg.backward(full[a:b].contiguous(), g)
                torch.cuda.synchronize()
                same2 = (gw2.view(torch.int16) == x.grad[a:b].view(torch.int16)) | (gw2.isnan() & x.grad[a:b].isnan())
                okf = bool(same2.all()) if world == 2 else float((gw2.float() - x.grad[a:b].float()).abs().max()) < 1e-2
                if not okf:
                    print(f"rank {rank}: fused NF3 backward mismatch on [{rows}, {cols}] rep {rep}: {int((~same2).sum())} elements", flush=True)
                ok = ok and okf
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = bool(flag.item())
res = {"world": world, "nf3_forward_bit_equal_and_backward_rows_equal_incl_fused_backward": ok}
if rank == 0:
    n = 4 * 4096 * 4096 + 3 * 11008 * 4096
    ws = [(torch.randn(n // 4096, 4096, device=dev) * 0.02).bfloat16() for _ in range(3)]
    gs = [S.ShardGatherNF3(n // 4096, 4096, 128, torch.bfloat16, dev, symmetric=False) for _ in range(3)]
    import ctypes
    dq = torch.empty_like(ws[0])
    st = L.stream_ptr(dev)

    def enc(i):
        g = gs[i % 3]
        L.lib().bdq_nf3_shard_encode(L.ptr(ws[i % 3]), L.ptr(g.buf), ctypes.c_void_p(g.buf.data_ptr() + g.ncode), L.i64(n), 128, L.BF16,
                                     L.semantics(), 0, None, None, st)

    def dec(i):
        g = gs[i % 3]
        L.lib().bdq_nf3_shard_decode(L.ptr(g.buf), ctypes.c_void_p(g.buf.data_ptr() + g.ncode), L.ptr(dq), L.i64(n), 128, L.BF16, st)

    def timed(f, it=20):
        for i in range(3):
            f(i)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(it):
            f(i)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / it

    pk = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) else 6550.7
    payload = n // 32 * 12 + n // 128 * 4
    for name, f, nb in (("encode", enc, 2 * n + payload), ("decode", dec, payload + 2 * n)):
        ms = timed(f)
        res[name] = {"ms": ms, "gbs": nb / ms / 1e6, "frac_of_hbm_peak": nb / ms / 1e6 / pk}
    res["payload_bytes_per_weight"] = payload / n
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
