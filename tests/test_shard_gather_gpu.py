"""GPU parity tests of the quantise-before-all-gather path (bdq_int2_shard_encode / _decode, sharded.ShardGather).

Bar: the 2-bit codes, scales and zeros of the payload are bit-exact against the oracle (oracle/bdq_oracle.c, pinned by the
reference-generated golden vectors), and decode(encode(w)) is bit-equal to the fake-quantiser forward
(SteInt2AsymQuantizer, quantization/quantizer.py:153-182 / pseudo_quantize_tensor, :25-69) on the same weights."""
import numpy as np
import pytest
import torch

import oracle
from _util import DT, bits_equal, load_cases, make_weight, t

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _cpu_semantics(monkeypatch):
    monkeypatch.setenv("BDQ_SEMANTICS", "cpu")


def dev():
    return torch.device("cuda:0")


def unpack_codes(code_bytes, n):
    """Wire format of include/bdq.h: one 32-bit word per 16 elements, element j at bits [16*(j%2) + 2*(j//2), +2)."""
    words = code_bytes.cpu().numpy().view(np.uint32).astype(np.uint64)
    j = np.arange(16)
    shift = (16 * (j % 2) + 2 * (j // 2)).astype(np.uint64)
    return ((words[:, None] >> shift[None, :]) & 3).astype(np.uint8).reshape(-1)[:n]


def roundtrip(w, G, round_mode):
    from bitdistiller_b200 import sharded as S
    g = S.ShardGather(w.shape[0], w.shape[1], G, w.dtype, w.device, symmetric=False, round_mode=round_mode)
    dq = g.forward(w)
    torch.cuda.synchronize()
    sz = g.sz.view(w.dtype).view(-1, 2)
    return dq, unpack_codes(g.codes, w.numel()), sz[:, 0].clone(), sz[:, 1].clone()


@pytest.mark.parametrize("dname", ["bf16", "f16"])
@pytest.mark.parametrize("G", [32, 64, 128])
@pytest.mark.parametrize("round_mode", [0, 1])
def test_payload_matches_oracle_and_fake_quantiser(dname, G, round_mode):
    from bitdistiller_b200.quantizer import int_fakequant
    dt = DT[dname]
    for kind, seed in [("randn", 1), ("uniform", 2), ("heavy", 3)]:
        w = make_weight(kind, 48, 1024, dt, seed)
        w[3, :G] = 0.25                      # constant group: max - min clamps to 1e-5
        w[5, G:2 * G] = 0.0
        w[7, :G] = w[7, :G].abs()            # all-positive group: zero point 0
        w[9, :G] = -w[9, :G].abs()           # all-negative group: zero point 3
        wd = w.to(dev())
        dq, codes, s, z = roundtrip(wd, G, round_mode)
        o = oracle.int_fakequant(w, G, 2, round_mode=round_mode)
        assert np.array_equal(codes, o["codes"].numpy().reshape(-1)), (kind, dname, G)
        assert bits_equal(s.cpu(), o["scales"].reshape(-1)) and bits_equal(z.cpu(), o["zeros"].reshape(-1))
        assert bits_equal(dq.cpu(), o["dq"])
        ref, _, _, _ = int_fakequant(wd, 2, G, round_mode)
        assert torch.equal(dq.view(torch.int16), ref.view(torch.int16)), (kind, dname, G)


def test_golden_cases_round_trip():
    """The committed reference-generated vectors of pseudo_quantize_tensor (2-bit, 16-bit dtypes)."""
    n = 0
    for meta, c in load_cases("pq_int"):
        dname, n_bit, G = meta[0], int(meta[2]), int(meta[3])
        if n_bit != 2 or dname == "f32" or G not in (32, 64, 128):
            continue
        dt = DT[dname]
        w = t(c["w"], dt).to(dev())
        dq, _, s, z = roundtrip(w, G, 0)
        assert bits_equal(dq.cpu(), t(c["dq"], dt)), meta
        assert bits_equal(s.cpu(), t(c["s"], dt).reshape(-1)) and bits_equal(z.cpu(), t(c["z"], dt).reshape(-1)), meta
        n += 1
    assert n == 20


def test_nan_inf_and_slow_path_groups():
    from bitdistiller_b200.quantizer import int_fakequant
    w = make_weight("randn", 16, 512, torch.bfloat16, 7)
    w[0, 5] = float("nan")
    w[1, 130] = float("inf")
    w[2, :128] *= 1e30                       # huge range: quotients leave the bracketing window -> op-by-op path
    w[3, :128] *= 1e-30
    w[4, 0:128] = 3e38
    wd = w.to(dev())
    dq, codes, s, z = roundtrip(wd, 128, 1)
    ref, _, _, _ = int_fakequant(wd, 2, 128, 1)
    assert bits_equal(dq.cpu(), ref.cpu())
    fin = torch.isfinite(ref.reshape(-1, 128)).all(1).cpu()
    o = oracle.int_fakequant(w, 128, 2, round_mode=1)
    keep = fin.repeat_interleave(128).numpy()
    assert np.array_equal(codes[keep], o["codes"].numpy().reshape(-1)[keep])


def test_large_layer_bit_equal_and_linear_layout():
    """Llama-2-7B sized tensor: decode(encode(w)) == forward(w); the payload of row shards concatenates to the payload of
    the whole weight (what makes the gather a plain concatenation)."""
    from bitdistiller_b200 import sharded as S
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer
    w = (torch.randn(11008, 4096, device=dev()) * 0.02).bfloat16()
    full = S.ShardGather(11008, 4096, 128, torch.bfloat16, dev(), symmetric=False)
    dq = full.forward(w)
    assert torch.equal(dq.view(torch.int16), SteInt2AsymQuantizer(128)(w).view(torch.int16))
    lay = S.payload_layout(11008, 4096, 128, 8)
    for r in (0, 3, 7):
        r0, r1, coff, cb, soff, sb = lay[r]
        part = S.ShardGather(r1 - r0, 4096, 128, torch.bfloat16, dev(), symmetric=False)
        part.forward(w[r0:r1])
        assert torch.equal(part.codes, full.codes[coff:coff + cb]) and torch.equal(part.sz, full.sz[soff:soff + sb])


def test_autograd_matches_ste_quantizer():
    from bitdistiller_b200 import sharded as S
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer
    w = (torch.randn(64, 512, device=dev()) * 0.02).bfloat16()
    gy = torch.randn(64, 512, device=dev()).bfloat16()
    a = w.clone().requires_grad_(True)
    b = w.clone().requires_grad_(True)
    g = S.ShardGather(64, 512, 128, torch.bfloat16, dev(), symmetric=False)
    S.quantized_all_gather(a, g).backward(gy)
    SteInt2AsymQuantizer(128)(b).backward(gy)
    assert torch.equal(a.grad.view(torch.int16), b.grad.view(torch.int16))


def test_group_gather_one_decode_launch_equals_per_tensor():
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200 import sharded as S
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer
    shapes = [(64, 512), (64, 512), (176, 512), (64, 1408), (16, 128), (48, 256), (16, 4096)]
    ws = [(torch.randn(r, c, device=dev()) * 0.02).bfloat16() for r, c in shapes]
    g = S.ShardGatherGroup(shapes, 128, torch.bfloat16, dev(), symmetric=False)
    n0 = L.launch_count()
    outs = g.forward(ws)
    assert L.launch_count() - n0 == 2                        # ONE encode over all shards, ONE decode for the group
    q = SteInt2AsymQuantizer(128)
    for w, o in zip(ws, outs):
        assert o.shape == w.shape and torch.equal(o.view(torch.int16), q(w).view(torch.int16))
    many = [(32, 256)] * 11                                  # more than 8 tensors: two encode launches
    g2 = S.ShardGatherGroup(many, 64, torch.float16, dev(), symmetric=False)
    w2 = [(torch.randn(32, 256, device=dev()) * 0.02).half() for _ in many]
    n0 = L.launch_count()
    o2 = g2.forward(w2)
    assert L.launch_count() - n0 == 3
    q64 = SteInt2AsymQuantizer(64)
    assert all(torch.equal(o.view(torch.int16), q64(w).view(torch.int16)) for w, o in zip(w2, o2))


def test_group_autograd_matches_ste_quantizer():
    from bitdistiller_b200 import sharded as S
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer
    shapes = [(64, 512), (32, 1024), (16, 128)]
    ws = [(torch.randn(r, c, device=dev()) * 0.02).bfloat16() for r, c in shapes]
    gys = [torch.randn(r, c, device=dev()).bfloat16() for r, c in shapes]
    a = [w.clone().requires_grad_(True) for w in ws]
    b = [w.clone().requires_grad_(True) for w in ws]
    g = S.ShardGatherGroup(shapes, 128, torch.bfloat16, dev(), symmetric=False)
    outs = S.quantized_all_gather_group(a, g)
    torch.autograd.backward(list(outs), gys)
    q = SteInt2AsymQuantizer(128)
    for x, gy, o in zip(b, gys, outs):
        y = q(x)
        assert torch.equal(o.view(torch.int16), y.view(torch.int16))
        y.backward(gy)
    for x, y in zip(a, b):
        assert torch.equal(x.grad.view(torch.int16), y.grad.view(torch.int16))


def test_errors():
    from bitdistiller_b200 import _lib as L
    lib = L.lib()
    i64 = L.i64
    assert lib.bdq_int2_shard_encode(None, None, None, i64(100), 128, 1, 1, 0, 0, None, None, None) == -1
    assert lib.bdq_int2_shard_encode(None, None, None, i64(128), 128, 0, 1, 0, 0, None, None, None) == -2
    assert lib.bdq_int2_shard_encode(None, None, None, i64(96), 48, 1, 1, 0, 0, None, None, None) == -5
    assert lib.bdq_int2_shard_encode(None, None, None, i64(128), 128, 1, 1, 0, 8, None, None, None) == -1
    assert lib.bdq_int2_shard_decode(None, None, None, i64(0), 128, 1, None) == 0
    assert lib.bdq_int2_shard_decode(None, None, None, i64(128), 128, 1, None) == -1


@pytest.mark.parametrize("dname", ["bf16", "f16"])
@pytest.mark.parametrize("nsrc", [2, 3, 8])
def test_ste_bwd_multi_source_and_push(dname, nsrc):
    """bdq_rows_push + bdq_int_ste_bwd_nsrc on one GPU: pushing row ranges into inbox slots and summing `nsrc` gradient
    sources on the fly equals the sequential sum (one rounding per addition, source order) followed by bdq_int_ste_bwd,
    bit for bit -- including groups that take the op-by-op path (huge dynamic range) and NaN gradients."""
    import ctypes

    from bitdistiller_b200 import _lib as L
    dt = {"bf16": torch.bfloat16, "f16": torch.float16}[dname]
    dev = torch.device("cuda:0")
    rows, cols, G = 96, 1024, 128
    g = torch.Generator().manual_seed(nsrc)
    w = (torch.randn(rows, cols, generator=g) * 0.02).to(dt)
    w[3, :128] *= 1e-6          # tiny group next to ...
    w[3, 5] = 30000.0 if dname == "f16" else 3e30   # ... an extreme one: op-by-op path
    w = w.to(dev)
    srcs = [torch.randn(rows, cols, generator=g).to(dt).to(dev) for _ in range(nsrc)]
    srcs[-1][7, 9] = float("nan")
    # push every source but the first into "inbox" slots with bdq_rows_push (two segments per source: ragged split)
    inbox = torch.zeros((nsrc - 1, rows, cols), dtype=dt, device=dev)
    es = w.element_size()
    src_p, dst_p, nb = [], [], []
    for i in range(1, nsrc):
        cut = 32 * i % rows or 32
        for r0, r1 in ((0, cut), (cut, rows)):
            src_p.append(srcs[i].data_ptr() + r0 * cols * es)
            dst_p.append(inbox[i - 1].data_ptr() + r0 * cols * es)
            nb.append((r1 - r0) * cols * es)
    m = len(src_p)
    st = L.stream_ptr(dev)
    L.check(L.lib().bdq_rows_push(m, (ctypes.c_void_p * m)(*src_p), (ctypes.c_void_p * m)(*dst_p), (ctypes.c_int64 * m)(*nb), st), "push")
    assert torch.equal(inbox.view(torch.int16), torch.stack(srcs[1:]).view(torch.int16))
    gw = torch.empty_like(w)
    ptrs = [srcs[0].data_ptr()] + [inbox[i].data_ptr() for i in range(nsrc - 1)]
    L.check(L.lib().bdq_int_ste_bwd_nsrc(L.ptr(w), nsrc, (ctypes.c_void_p * nsrc)(*ptrs), L.ptr(gw), L.i64(w.numel()), G, 2,
                                         L.dtype_code(w), L.semantics(), st), "nsrc")
    acc = srcs[0].clone()
    for s in srcs[1:]:
        acc = (acc.float() + s.float()).to(dt)
    ref = torch.empty_like(w)
    L.check(L.lib().bdq_int_ste_bwd(L.ptr(w), L.ptr(acc), L.ptr(ref), L.i64(w.numel()), G, 2, L.dtype_code(w), L.semantics(), st), "bwd")
    torch.cuda.synchronize()
    same = (gw.view(torch.int16) == ref.view(torch.int16)) | (gw.isnan() & ref.isnan())
    assert bool(same.all()), int((~same).sum())
    with pytest.raises(L.BdqError):
        L.check(L.lib().bdq_int_ste_bwd_nsrc(L.ptr(w), 9, (ctypes.c_void_p * 9)(*([ptrs[0]] * 9)), L.ptr(gw), L.i64(w.numel()), G, 2,
                                             L.dtype_code(w), L.semantics(), st), "nsrc")


@pytest.mark.parametrize("dname", ["bf16", "f16"])
@pytest.mark.parametrize("nsrc,G", [(2, 128), (3, 128), (8, 128), (2, 64), (3, 32)])
def test_nf3_ste_bwd_multi_source(dname, nsrc, G):
    """bdq_nf3_ste_bwd_nsrc: summing `nsrc` gradient sources on the fly equals the sequential sum (one rounding per addition,
    source order) followed by bdq_nf3_ste_bwd, bit for bit -- including groups that take the op-by-op path (huge dynamic
    range, an all-positive group, a NaN group with max == 0) and NaN gradients."""
    import ctypes

    from bitdistiller_b200 import _lib as L
    dt = {"bf16": torch.bfloat16, "f16": torch.float16}[dname]
    dev = torch.device("cuda:0")
    rows, cols = 96, 1024
    g = torch.Generator().manual_seed(100 + nsrc)
    w = (torch.randn(rows, cols, generator=g) * 0.02).to(dt)
    w[3, :128] *= 1e-6
    w[3, 5] = 30000.0 if dname == "f16" else 3e30      # op-by-op path
    w[4, 128:256] = w[4, 128:256].abs() + 0.001        # all-positive group
    w[5, :128] = torch.minimum(w[5, :128], torch.zeros(128, dtype=dt))
    w[5, 0] = 0.0                                      # max == 0: NaN group
    w = w.to(dev)
    srcs = [torch.randn(rows, cols, generator=g).to(dt).to(dev) for _ in range(nsrc)]
    srcs[-1][7, 9] = float("nan")
    st = L.stream_ptr(dev)
    gw = torch.empty_like(w)
    ptrs = [t_.data_ptr() for t_ in srcs]
    L.check(L.lib().bdq_nf3_ste_bwd_nsrc(L.ptr(w), nsrc, (ctypes.c_void_p * nsrc)(*ptrs), L.ptr(gw), L.i64(w.numel()), G,
                                         L.dtype_code(w), L.semantics(), st), "nsrc")
    acc = srcs[0].clone()
    for s_ in srcs[1:]:
        acc = (acc.float() + s_.float()).to(dt)
    ref = torch.empty_like(w)
    L.check(L.lib().bdq_nf3_ste_bwd(L.ptr(w), L.ptr(acc), L.ptr(ref), L.i64(w.numel()), G, L.dtype_code(w), L.semantics(), st), "bwd")
    torch.cuda.synchronize()
    same = (gw.view(torch.int16) == ref.view(torch.int16)) | (gw.isnan() & ref.isnan())
    assert bool(same.all()), int((~same).sum())
    with pytest.raises(L.BdqError):
        L.check(L.lib().bdq_nf3_ste_bwd_nsrc(L.ptr(w), 9, (ctypes.c_void_p * 9)(*([ptrs[0]] * 9)), L.ptr(gw), L.i64(w.numel()), G,
                                             L.dtype_code(w), L.semantics(), st), "nsrc")
    w32 = w.float()
    with pytest.raises(L.BdqError):     # fp32 has no multi-source kernel
        L.check(L.lib().bdq_nf3_ste_bwd_nsrc(L.ptr(w32), 2, (ctypes.c_void_p * 2)(*ptrs[:2]), L.ptr(gw), L.i64(w.numel()), G,
                                             L.dtype_code(w32), L.semantics(), st), "nsrc")


@pytest.mark.parametrize("dname", ["bf16", "f16"])
@pytest.mark.parametrize("G", [128, 64, 32])
def test_nf3_shard_payload_roundtrip(dname, G):
    """bdq_nf3_shard_encode / _decode (the NF3 payload of quantise-before-all-gather): level indices against the oracle's NF3
    codes, scales against its (|max|, |min|), and decode(encode(w)) bit-equal to the NF3 fake-quantiser forward -- incl. the
    groups that take the op-by-op path, the NaN groups (max or min exactly 0) and clipped (tie-heavy) weights."""
    import ctypes

    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.sharded import ShardGatherNF3
    dt = DT[dname]
    rows, cols = 64, 1024
    w = make_weight("randn", rows, cols, dt, 5)
    lim = w.float().reshape(rows, -1, G).abs().amax(-1, keepdim=True) * 0.8
    w[:16] = torch.clamp(w[:16].float().reshape(16, -1, G), -lim[:16], lim[:16]).reshape(16, cols).to(dt)   # ties at the extremes
    w[20, :G] = w[20, :G].abs()
    w[20, 3] = 0.0                      # min exactly 0 -> NaN group in the reference
    w[21, :G] = 0.0                     # all-zero group
    w[22, :G] *= 1e-6                   # tiny scales: op-by-op path
    w[23, 5] = 3e4 if dname == "f16" else 3e30
    wd = w.to(dev())
    sg = ShardGatherNF3(rows, cols, G, dt, dev(), symmetric=False)
    dq = sg.forward(wd)
    ref_dq, sp, sn, codes = Q.nf3_fakequant(wd, G, want_codes=True, want_scales=True)
    assert bits_equal(dq.cpu(), ref_dq.cpu())
    orc = oracle.nf3_fakequant(w, G, sem=0)
    assert bits_equal(dq.cpu(), orc["dq"])
    # wire format: level index e per element from the three words of every 32-element chunk
    words = sg.codes.cpu().numpy().view(np.uint32).reshape(-1, 3).astype(np.uint64)
    j = np.arange(32)
    lo = np.where(j[None, :] < 16, words[:, 0:1], words[:, 1:2])
    e = ((lo >> (2 * (j % 16)).astype(np.uint64)[None, :]) & 3) | (((words[:, 2:3] >> j.astype(np.uint64)[None, :]) & 1) << 2)
    e = e.reshape(rows, cols)
    oc = orc["codes"].numpy().astype(np.int64)          # 0 = zero, 1..4 = positive levels, 5..7 = negative levels, 255 = NaN group
    stair = np.where(oc == 0, 3, np.where(oc <= 4, oc + 3, 7 - oc))
    ok = oc != 255
    assert np.array_equal(e[ok], stair[ok])
    sc = sg.sc.view(torch.int16).cpu().view(dt).float().reshape(rows, cols // G, 2)
    nan_g = torch.from_numpy(~ok).reshape(rows, cols // G, G).any(-1)
    assert bool(sc[nan_g].isnan().all())
    assert torch.equal(sc[..., 0][~nan_g], sp.float().cpu().reshape(rows, -1)[~nan_g])
    assert torch.equal(sc[..., 1][~nan_g], sn.float().cpu().reshape(rows, -1)[~nan_g])
    # larger, through the flat layer launch shape
    big = make_weight("heavy", 512, 4096, dt, 9).to(dev())
    sgb = ShardGatherNF3(512, 4096, G, dt, dev(), symmetric=False)
    assert bits_equal(sgb.forward(big).cpu(), Q.nf3_fakequant(big, G)[0].cpu())
    with pytest.raises(L.BdqError):
        L.check(L.lib().bdq_nf3_shard_decode(L.ptr(sgb.codes), L.ptr(sgb.sc), L.ptr(big), L.i64(100), G, L.dtype_code(big), L.stream_ptr(dev())), "x")
