"""Host-side logic of the multi-GPU path on CPU: partition plans, and the clip-cache / channel-slice all-gathers over a
world_size-2 gloo group (the same code path NCCL takes on GPUs).  No kernels are called."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bitdistiller_b200 import sharded as S


def test_row_shard_and_layer_owner_cover_everything():
    for rows, world in [(4096, 8), (11008, 8), (1024, 8), (12288, 8), (4096, 3), (48, 2), (16, 4)]:
        spans = [S.row_shard(rows, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == rows
        for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
            assert a1 == b0 and a0 % 16 == 0 and a1 % 16 == 0
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 16
    with pytest.raises(ValueError):
        S.row_shard(100, 0, 2)
    for n_layers, world in [(32, 8), (36, 8), (32, 5), (2, 4)]:
        owners = [S.layer_owner(i, n_layers, world) for i in range(n_layers)]
        assert owners == sorted(owners)
        for r in range(world):
            assert S.owned_layers(n_layers, r, world) == [i for i, o in enumerate(owners) if o == r]
    with pytest.raises(IndexError):
        S.layer_owner(32, 32, 8)


def test_gather_clip_cache_single_process_is_identity():
    e = [("a", torch.ones(4, 2, 1), -torch.ones(4, 2, 1))]
    out = S.gather_clip_cache(e, dtype=torch.float32)
    assert out[0][0] == "a" and torch.equal(out[0][1], e[0][1]) and torch.equal(out[0][2], e[0][2])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _clip_of_layer(i):
    g = torch.Generator().manual_seed(i)
    out = []
    for name, co, ng in [("v_proj", 32, 2), ("down_proj", 16, 5 + (i % 2))]:   # ragged sizes across layers
        mx = torch.rand(co, ng, 1, generator=g).bfloat16()
        out.append((f"model.layers.{i}.{name}", mx, -mx * 0.5))
    return out


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n_layers = 5
        full = S.run_clip_sharded(n_layers, _clip_of_layer, dtype=torch.bfloat16, device=torch.device("cpu"))
        ref = [e for i in range(n_layers) for e in _clip_of_layer(i)]
        ok = len(full) == len(ref) and all(a[0] == b[0] and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])
                                           for a, b in zip(full, ref))
        # channel-slice all-gather: decode row and a prefill block
        for M in (1, 3):
            whole = torch.arange(M * 64, dtype=torch.float32).view(M, 64).half()
            r0, r1 = S.row_shard(64, rank, world)
            got = S.all_gather_cols(whole[:, r0:r1].contiguous())
            ok = ok and torch.equal(got, whole)
        # quantise-then-gather payload transport: the concatenation of the rank slots is the full payload
        lay = S.payload_layout(48, 256, 128, world)
        allc = torch.arange(48 * 256 // 4, dtype=torch.int64).remainder(251).to(torch.uint8)
        alls = torch.arange(48 * 2 * 4, dtype=torch.int64).remainder(241).to(torch.uint8)
        _, _, coff, cb, soff, sb = lay[rank]
        gc_, gs_ = S.gather_payload(allc[coff:coff + cb], alls[soff:soff + sb], lay)
        ok = ok and torch.equal(gc_, allc) and torch.equal(gs_, alls)
        # gradient half of the quantise-before-all-gather path: sum over ranks, own rows back, caller's tensor untouched
        for rows in (48, 64):                                  # 48 rows -> ragged shards (16 / 32), 64 -> even
            lay2 = S.payload_layout(rows, 128, 128, world)
            mine = torch.full((rows, 128), float(rank + 1)) + torch.arange(rows, dtype=torch.float32).view(-1, 1)
            keep = mine.clone()
            got = S.reduce_scatter_rows(mine, lay2, rank)
            r0, r1 = lay2[rank][:2]
            want = (3.0 + 2 * torch.arange(rows, dtype=torch.float32).view(-1, 1)).expand(rows, 128)[r0:r1]
            ok = ok and torch.equal(got, want) and torch.equal(mine, keep)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_sharded_gathers_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_payload_layout_is_a_partition():
    for rows, cols, G, world in [(4096, 4096, 128, 8), (11008, 4096, 128, 8), (4096, 11008, 128, 8), (48, 256, 64, 2), (16, 64, 32, 3)]:
        lay = S.payload_layout(rows, cols, G, world)
        assert lay[0][0] == 0 and lay[-1][1] == rows and lay[0][2] == 0 and lay[0][4] == 0
        for a, b in zip(lay, lay[1:]):
            assert a[1] == b[0] and a[2] + a[3] == b[2] and a[4] + a[5] == b[4]
        assert lay[-1][2] + lay[-1][3] == rows * cols // 4                 # 2 bits per weight
        assert lay[-1][4] + lay[-1][5] == rows * (cols // G) * 4           # (scale, zero) 2 x 2 bytes per group
        assert all(l[2] % 8 == 0 and l[4] % 4 == 0 for l in lay)           # kernel alignment of every slot
    with pytest.raises(ValueError):
        S.payload_layout(16, 100, 128, 2)


def test_shard_gather_group_offsets_partition_the_buffer():
    """Host-side layout of ShardGatherGroup (no kernel call): tensor payloads are laid back to back, every offset keeps the
    kernels' alignment, and the flat output views tile the flat buffer in tensor order."""
    shapes = [(4096, 4096)] * 4 + [(11008, 4096)] * 2 + [(4096, 11008)]
    g = S.ShardGatherGroup(shapes, 128, torch.bfloat16, "cpu", symmetric=False)
    n = sum(r * c for r, c in shapes)
    assert g.ncode == n // 4 and g.nsz == n // 128 * 4 and g.buf.numel() == g.ncode + g.nsz
    assert g.code_base[0] == 0 and g.sz_base[0] == 0
    for t in range(1, len(shapes)):
        r, c = shapes[t - 1]
        assert g.code_base[t] - g.code_base[t - 1] == r * c // 4 and g.sz_base[t] - g.sz_base[t - 1] == r * c // 128 * 4
        assert g.code_base[t] % 8 == 0 and g.sz_base[t] % 4 == 0
    assert g.flat.numel() == n and [tuple(v.shape) for v in g.views] == shapes
    off = 0
    for v, (r, c) in zip(g.views, shapes):
        assert v.data_ptr() == g.flat.data_ptr() + 2 * off
        off += r * c


def test_grad_inbox_layout_host_logic():
    """ShardGradReducer's host-side layout: every rank's flat shard layout tiles the tensors without gaps, ragged shards
    included, and the n-1 inbox slots of a rank are addressed consistently by sender and receiver."""
    from bitdistiller_b200.sharded import grad_inbox_layout, row_shard
    shapes = [(4096, 4096), (11008, 4096), (4096, 11008), (48, 256)]
    for world in (2, 3, 8):
        spans, off, size, slot = grad_inbox_layout(shapes, world, 2)
        for t, (r, c) in enumerate(shapes):
            assert [s for s in spans[t]] == [row_shard(r, k, world) for k in range(world)]
            assert spans[t][0][0] == 0 and spans[t][-1][1] == r
            assert all(spans[t][k][1] == spans[t][k + 1][0] for k in range(world - 1))
        for k in range(world):
            o = 0
            for t, (r, c) in enumerate(shapes):
                assert off[k][t] == o and o % 32 == 0
                o += (spans[t][k][1] - spans[t][k][0]) * c * 2
            assert size[k] == o <= slot and slot % 256 == 0
        assert sum(size) == sum(r * c * 2 for r, c in shapes)
        # sender `s` writes slot index (s if s < d else s - 1) of receiver d: the n-1 senders of d get distinct slots 0..n-2
        for d in range(world):
            assert sorted((s if s < d else s - 1) for s in range(world) if s != d) == list(range(world - 1))


def test_nf3_payload_layout_host_logic():
    from bitdistiller_b200.sharded import nf3_payload_layout
    for world in (1, 2, 3, 8):
        lay = nf3_payload_layout(11008, 4096, 128, world)
        assert lay[0][0] == 0 and lay[-1][1] == 11008
        assert sum(l[3] for l in lay) == 11008 * 4096 // 32 * 12 and sum(l[5] for l in lay) == 11008 * 32 * 4
        for a, b in zip(lay, lay[1:]):
            assert a[1] == b[0] and a[2] + a[3] == b[2] and a[4] + a[5] == b[4]
