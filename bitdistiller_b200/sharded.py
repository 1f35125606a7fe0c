"""Multi-GPU side of the hot path: one process per GPU, torch.distributed for the plumbing.

* fake-quant forward / STE backward / packing shard by rows with NO collective (groups lie along the last dimension, so
  any dim-0 split keeps whole groups): `row_shard`.
* autoclip (quantization/autoclip.py:86-107, 190-199) shards by decoder layer: every rank searches the clip ratios of
  the layers it owns (`layer_owner`) and ONE all-gather per run distributes the clip cache (`gather_clip_cache`): the
  reference's `{"clip": [(name, max_val, min_val), ...]}` list comes back complete and in the reference's order on
  every rank.
* packed GEMV / GEMM are column-parallel over output channels, matching how WQLinear rows are laid out
  (quantization/qmodule.py:62-66): rank r holds qweight[r*OC/n:(r+1)*OC/n], the matching scales / zero rows, and the
  partial outputs [M, OC/n] are all-gathered (`TPWQLinear`).

The collectives go through torch.distributed (NCCL on GPUs; the same code runs on gloo/CPU tensors, which is how the
host-side logic is tested without a GPU).  Nothing here touches the kernels' arithmetic.
"""
import os

import torch
import torch.distributed as dist
import torch.nn as nn


# ------------------------------------------------------------------------------------------------- partitioning
def row_shard(rows, rank, world, multiple=16):
    """Rows [r0, r1) of a [rows, cols] tensor owned by `rank`: contiguous, sizes differ by at most `multiple`, every
    boundary a multiple of `multiple` (16 = the GEMV / pack row-block; FSDP2-style dim-0 shards)."""
    if rows % multiple != 0:
        raise ValueError(f"rows={rows} must be a multiple of {multiple}")
    blocks = rows // multiple
    b0 = (blocks * rank) // world
    b1 = (blocks * (rank + 1)) // world
    return b0 * multiple, b1 * multiple


def layer_owner(layer, n_layers, world):
    """Rank that owns decoder layer `layer` when n_layers are split into contiguous, near-equal runs."""
    if not 0 <= layer < n_layers:
        raise IndexError(layer)
    for r in range(world):
        if (n_layers * r) // world <= layer < (n_layers * (r + 1)) // world:
            return r
    raise AssertionError("unreachable")


def owned_layers(n_layers, rank, world):
    return list(range((n_layers * rank) // world, (n_layers * (rank + 1)) // world))


# ------------------------------------------------------------------------------------------------- clip cache gather
def _flatten_clip(entries, dtype, device):
    """[(name, max_val[co,ng,1], min_val[co,ng,1])] -> (metadata list, one flat tensor: max then min per entry)."""
    meta, parts = [], []
    for name, mx, mn in entries:
        assert mx.shape == mn.shape and mx.dim() == 3 and mx.shape[2] == 1
        meta.append((name, int(mx.shape[0]), int(mx.shape[1])))
        parts += [mx.reshape(-1).to(dtype), mn.reshape(-1).to(dtype)]
    flat = torch.cat(parts) if parts else torch.empty(0, dtype=dtype, device=device)
    return meta, flat


def gather_clip_cache(local_entries, dtype=torch.bfloat16, device=None, group=None):
    """All-gather of the per-rank clip lists.  `local_entries`: the (name, max_val, min_val) tuples this rank computed,
    in the reference's order (autoclip.py:199 `clip_list += ...`); names must be globally unique (the reference prefixes
    them with the layer, autoclip.py:204).  Returns the concatenation over ranks 0..n-1 on every rank -- with
    layer-contiguous ownership that IS the reference's single-process order.  One object all-gather for the few hundred
    bytes of metadata, one tensor all-gather for the values (169 MB for Llama-2-7B int2-g128, 21 MB per rank at n=8)."""
    if device is None:
        device = local_entries[0][1].device if local_entries else torch.device("cpu")
    meta, flat = _flatten_clip(local_entries, dtype, device)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return [(n, mx.to(dtype), mn.to(dtype)) for n, mx, mn in local_entries]
    world = dist.get_world_size(group)
    metas = [None] * world
    dist.all_gather_object(metas, meta, group=group)
    sizes = [sum(2 * co * ng for _, co, ng in m) for m in metas]
    pad = max(sizes)
    send = torch.zeros(pad, dtype=dtype, device=device)
    send[: flat.numel()] = flat
    recv = torch.empty(world * pad, dtype=dtype, device=device)
    dist.all_gather_into_tensor(recv, send, group=group)
    out = []
    for r, m in enumerate(metas):
        off = r * pad
        for name, co, ng in m:
            n = co * ng
            out.append((name, recv[off:off + n].view(co, ng, 1), recv[off + n:off + 2 * n].view(co, ng, 1)))
            off += 2 * n
    return out


def run_clip_sharded(n_layers, clip_layer_fn, dtype=torch.bfloat16, device=None, group=None):
    """Layer-sharded equivalent of the reference's run_clip layer loop (autoclip.py:165-206): `clip_layer_fn(i)` returns
    the clip list of decoder layer i with globally unique names (what auto_clip_block + the name prefixing of
    autoclip.py:199-204 produce); it is called only for the layers this rank owns.  Returns the full list on every
    rank."""
    rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    local = []
    for i in owned_layers(n_layers, rank, world):
        local += clip_layer_fn(i)
    return gather_clip_cache(local, dtype=dtype, device=device, group=group)


# ------------------------------------------------------------------------------------------------- tensor parallel
def all_gather_cols(local_out, group=None):
    """[M, OC/n] per rank -> [M, OC] (rank-major along the channel dimension) on every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local_out
    world = dist.get_world_size(group)
    M, ocl = local_out.shape
    buf = torch.empty((world, M, ocl), dtype=local_out.dtype, device=local_out.device)
    dist.all_gather_into_tensor(buf.view(-1), local_out.contiguous().view(-1), group=group)
    if M == 1:
        return buf.view(1, world * ocl)          # decode: the gathered buffer already is the row
    return buf.permute(1, 0, 2).reshape(M, world * ocl)


class TPWQLinear(nn.Module):
    """Column-parallel shard of a WQLinear (quantization/qmodule.py:41-173): rows [r0, r1) of qweight / qzeros / scales
    / bias; forward = local packed GEMV (one row) or tcgen05 GEMM (several rows) + all-gather of the channel slices.
    `from_full` slices an existing WQLinear (or any module with the same buffers)."""

    def __init__(self, qweight, qzeros, scales, bias, group_size, out_features, group=None):
        super().__init__()
        self.register_buffer("qweight", qweight)
        self.register_buffer("qzeros", qzeros)
        self.register_buffer("scales", scales)
        self.register_buffer("bias", bias) if bias is not None else setattr(self, "bias", None)
        self.group_size, self.out_features, self.group = group_size, out_features, group
        self.in_features = qweight.shape[1] * 16

    @classmethod
    def from_full(cls, full, rank=None, world=None, group=None):
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
            world = dist.get_world_size(group) if dist.is_initialized() else 1
        oc = full.qweight.shape[0]
        if oc % (16 * world) != 0:
            raise ValueError(f"out_features={oc} must be a multiple of 16*world_size for equal channel slices")
        r0, r1 = row_shard(oc, rank, world)
        bias = full.bias[r0:r1].clone() if getattr(full, "bias", None) is not None else None
        return cls(full.qweight[r0:r1].clone(), full.qzeros[r0:r1].clone(), full.scales[r0:r1].clone(), bias,
                   full.group_size, oc, group)

    @torch.no_grad()
    def forward(self, x):
        from .awq_inference_engine import gemv_forward_cuda
        from .triton_kernels import gemm_w2
        x2 = x.reshape(-1, x.shape[-1])
        if x2.shape[0] == 1:
            y = gemv_forward_cuda(x2, self.qweight, self.scales, self.qzeros, 2, self.group_size)
        else:
            y = gemm_w2(x2, self.qweight, self.scales, self.qzeros, self.group_size)
        if self.bias is not None:
            y = y + self.bias
        y = all_gather_cols(y, self.group)
        return y.reshape(x.shape[:-1] + (self.out_features,))


# ------------------------------------------------------------------------------------------------- fused all-gather
class SymmRows:
    """[M, OC] fp16 output rows allocated in torch symmetric memory: the same buffer exists on every rank of the group
    and every rank holds peer-mapped pointers to all of them (NVLink).  The tensor-parallel GEMV / GEMM epilogues store
    their channel slice straight into all of these buffers (bdq_gemv_w2_tp / bdq_gemm_w2_tp), so no collective runs --
    only `barrier()` between the producers and the consumers of the gathered rows."""

    def __init__(self, M, OC, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.t = symm_mem.empty((M, OC), dtype=torch.float16, device=device)
        self.hdl = symm_mem.rendezvous(self.t, self.group)
        self.rank, self.world = self.hdl.rank, self.hdl.world_size
        self.M, self.OC = M, OC
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]

    def slice_ptrs(self, col0):
        """(own pointer, [peer pointers]) of column col0 of row 0."""
        off = col0 * 2
        return self.ptrs[self.rank] + off, [self.ptrs[p] + off for p in range(self.world) if p != self.rank]

    def barrier(self):
        self.hdl.barrier(channel=0)


def tp_gemv_fused(x, shards, rows, flags=0):
    """Decode: up to 4 column-parallel shards [(qweight, scales, qzeros)] sharing x [1, IC] in ONE launch whose epilogue
    writes every slice into all ranks' `rows[i]` (SymmRows [1, OC_i]).  Call rows[0].barrier() before reading them."""
    import ctypes

    from . import _lib as L
    n = len(shards)
    ic = x.shape[-1]
    qw = (ctypes.c_void_p * n)(*[s[0].data_ptr() for s in shards])
    sc = (ctypes.c_void_p * n)(*[s[1].data_ptr() for s in shards])
    qz = (ctypes.c_void_p * n)(*[s[2].data_ptr() for s in shards])
    ocl = (ctypes.c_int64 * n)(*[s[0].shape[0] for s in shards])
    own, peers = [], []
    for s, r in zip(shards, rows):
        o, p = r.slice_ptrs(r.rank * s[0].shape[0])
        own.append(o)
        peers += p
    npeer = rows[0].world - 1
    outp = (ctypes.c_void_p * n)(*own)
    peerp = (ctypes.c_void_p * max(len(peers), 1))(*peers)
    with L.on(x.device):
        rc = L.lib().bdq_gemv_w2_tp(L.ptr(x), n, qw, sc, qz, outp, ocl, L.i64(ic), 128, int(flags), npeer, peerp,
                                    L.stream_ptr(x.device))
    L.check(rc, "bdq_gemv_w2_tp")


def tp_gemm_fused(x, shard, rows, group_size=128):
    """Prefill: x [M, IC] times one column-parallel shard; the tcgen05 epilogue writes the [M, OC/n] block at the
    full-width row pitch into all ranks' `rows` (SymmRows [>=M, OC])."""
    import ctypes

    from . import _lib as L
    qweight, scales, qzeros = shard
    M, ic = x.shape
    ocl = qweight.shape[0]
    own, peers = rows.slice_ptrs(rows.rank * ocl)
    peerp = (ctypes.c_void_p * max(len(peers), 1))(*peers)
    with L.on(x.device):
        rc = L.lib().bdq_gemm_w2_tp(L.ptr(x), L.ptr(qweight), L.ptr(scales), L.ptr(qzeros), ctypes.c_void_p(own), L.i64(M),
                                    L.i64(ic), L.i64(ocl), int(group_size), L.i64(rows.OC), len(peers), peerp,
                                    L.stream_ptr(x.device))
    L.check(rc, "bdq_gemm_w2_tp")


# ------------------------------------------------------------------------------------------------- quantise, then gather
def payload_layout(rows, cols, group_size, world, multiple=16):
    """Byte layout of the gathered int2 payload of a [rows, cols] weight split on dim 0 by `row_shard`: per rank
    (row0, row1, codes byte offset, codes bytes, scale/zero byte offset, scale/zero bytes).  Codes take 2 bits per weight,
    every group one (scale, zero) pair of 2-byte values; shards are whole rows, hence whole groups and whole 32-element
    code words, so the concatenation of the rank payloads IS the payload of the full weight."""
    if cols % group_size != 0 or group_size % 32 != 0:
        raise ValueError(f"cols={cols} must be a multiple of group_size={group_size}, itself a multiple of 32")
    out = []
    for r in range(world):
        r0, r1 = row_shard(rows, r, world, multiple if world > 1 else 1)
        out.append((r0, r1, r0 * cols // 4, (r1 - r0) * cols // 4, r0 * (cols // group_size) * 4,
                    (r1 - r0) * (cols // group_size) * 4))
    return out


def gather_payload(local_codes, local_sz, layout, group=None):
    """Collective transport of the payload (the path without peer-mapped memory; also what the gloo tests run): uint8
    tensors of this rank's slot -> the full (codes, sz) byte tensors on every rank."""
    world = len(layout)
    if world == 1 or not (dist.is_available() and dist.is_initialized()):
        return local_codes, local_sz
    dev = local_codes.device
    cmax, smax = max(l[3] for l in layout), max(l[5] for l in layout)
    send = torch.zeros(cmax + smax, dtype=torch.uint8, device=dev)       # ragged shards: pad every slot to the largest
    send[: local_codes.numel()] = local_codes
    send[cmax: cmax + local_sz.numel()] = local_sz
    recv = torch.empty(world * (cmax + smax), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(recv, send, group=group)
    recv = recv.view(world, cmax + smax)
    return (torch.cat([recv[r, : l[3]] for r, l in enumerate(layout)]),
            torch.cat([recv[r, cmax: cmax + l[5]] for r, l in enumerate(layout)]))


def reduce_scatter_rows(g_full, layout, rank, group=None, average=False):
    """Sum over ranks of a [rows, cols] gradient, rows [r0, r1) of `layout[rank]` returned (the gradient half of the
    quantise-before-all-gather path).  Even shards: one reduce-scatter; ragged shards: all-reduce of a COPY (the caller's
    tensor is never written) and a slice.  Works on gloo/CPU tensors too."""
    world = len(layout)
    r0, r1 = layout[rank][:2]
    if world == 1 or not (dist.is_available() and dist.is_initialized()):
        return g_full
    sizes = [l[1] - l[0] for l in layout]
    if len(set(sizes)) == 1 and dist.get_backend(group) != "gloo":
        g = torch.empty((r1 - r0, g_full.shape[1]), dtype=g_full.dtype, device=g_full.device)
        dist.reduce_scatter_tensor(g, g_full.contiguous(), op=dist.ReduceOp.SUM, group=group)
    else:
        tmp = g_full.clone()
        dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=group)
        g = tmp[r0:r1].contiguous()
    if average:
        g /= world
    return g


class PeerBarrier:
    """Cross-GPU barrier on flags in symmetric memory (bdq_peer_barrier: one 32-thread kernel, a few microseconds) instead of
    torch symmetric memory's barrier (~40 us on 2 x B200).  BDQ_TORCH_BARRIER=1 keeps the torch barrier (A/B comparisons)."""

    def __init__(self, device, group=None):
        import os

        import torch.distributed._symmetric_memory as symm_mem
        self.device = torch.device(device)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.buf = symm_mem.empty(64, dtype=torch.uint8, device=self.device)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.epoch = 0
        self.use_torch = os.environ.get("BDQ_TORCH_BARRIER", "0") == "1"
        torch.cuda.synchronize(self.device)
        self.hdl.barrier(channel=0)          # every rank's flags are zero before the first epoch goes out
        torch.cuda.synchronize(self.device)

    def __call__(self):
        import ctypes

        from . import _lib as L
        if self.use_torch:
            self.hdl.barrier(channel=0)
            return
        self.epoch += 1
        with L.on(self.device):
            rc = L.lib().bdq_peer_barrier((ctypes.c_void_p * self.world)(*self.ptrs), self.rank, self.world,
                                          ctypes.c_uint32(self.epoch & 0xffffffff), L.stream_ptr(self.device))
        L.check(rc, "bdq_peer_barrier")


class ShardGather:
    """Quantise-before-all-gather of ONE int2 QLinear weight [rows, cols] held as dim-0 shards (FSDP2 layout):
    `forward(w_shard)` returns the fake-quantised FULL weight on every rank, bit-equal to
    SteInt2AsymQuantizer(group)(all_gather(w_shard)) (quantization/quantizer.py:153-182), moving 0.28 B/weight over NVLink
    instead of 2.  With `symmetric=True` the encode kernel stores its payload straight into every rank's gather buffer
    (torch symmetric memory, bdq_int2_shard_encode's peer destinations) and only a barrier separates it from the decode;
    otherwise the payload goes through torch.distributed all_gather.  `backward(w_shard, g_full)`: reduce-scatter of the
    gradient w.r.t. the quantised weight, then the fused STE backward on the shard (the STE backward is linear in the
    incoming gradient and group-local, so it commutes with the sum over ranks)."""

    def __init__(self, rows, cols, group_size, dtype, device, group=None, symmetric=True, round_mode=None):
        from . import _lib as L
        self.rows, self.cols, self.group_size, self.dtype, self.device = rows, cols, group_size, dtype, torch.device(device)
        self.pg = group
        on = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if on else 0
        self.world = dist.get_world_size(group) if on else 1
        self.layout = payload_layout(rows, cols, group_size, self.world)
        self.round_mode = L.ROUND_HALF_AWAY if round_mode is None else round_mode
        ncode = rows * cols // 4
        nsz = rows * (cols // group_size) * 4
        self.symmetric = bool(symmetric) and self.world > 1
        if self.symmetric:
            import torch.distributed._symmetric_memory as symm_mem
            self.buf = symm_mem.empty(ncode + nsz, dtype=torch.uint8, device=self.device)
            self.hdl = symm_mem.rendezvous(self.buf, self.pg if self.pg is not None else dist.group.WORLD)
            self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        else:
            self.buf = torch.empty(ncode + nsz, dtype=torch.uint8, device=self.device)
            self.ptrs = [self.buf.data_ptr()]
        self.codes = self.buf[:ncode]
        self.sz = self.buf[ncode:]
        self._sz_off = ncode

    def _encode(self, w_shard):
        import ctypes

        from . import _lib as L
        r0, r1, coff, cbytes, soff, sbytes = self.layout[self.rank]
        assert w_shard.shape == (r1 - r0, self.cols) and w_shard.dtype == self.dtype
        L.require_cuda(w_shard)
        w = w_shard.detach().contiguous()
        base = self.buf.data_ptr()
        peers = [p for i, p in enumerate(self.ptrs) if i != self.rank] if self.symmetric else []
        pc = (ctypes.c_void_p * max(len(peers), 1))(*[p + coff for p in peers])
        ps = (ctypes.c_void_p * max(len(peers), 1))(*[p + self._sz_off + soff for p in peers])
        with L.on(self.device):
            rc = L.lib().bdq_int2_shard_encode(L.ptr(w), ctypes.c_void_p(base + coff), ctypes.c_void_p(base + self._sz_off + soff),
                                               L.i64(w.numel()), int(self.group_size), L.dtype_code(w), int(self.round_mode),
                                               L.semantics(), len(peers), pc, ps, L.stream_ptr(self.device))
        L.check(rc, "bdq_int2_shard_encode")

    def _decode(self, codes, sz):
        from . import _lib as L
        dq = torch.empty((self.rows, self.cols), dtype=self.dtype, device=self.device)
        with L.on(self.device):
            rc = L.lib().bdq_int2_shard_decode(L.ptr(codes), L.ptr(sz), L.ptr(dq), L.i64(dq.numel()), int(self.group_size),
                                               L.dtype_code(dq), L.stream_ptr(self.device))
        L.check(rc, "bdq_int2_shard_decode")
        return dq

    @torch.no_grad()
    def forward(self, w_shard):
        if self.symmetric:
            self.hdl.barrier(channel=0)          # every rank is done reading the previous payload
            self._encode(w_shard)
            self.hdl.barrier(channel=0)          # every slot of this rank's buffer has been written
            return self._decode(self.codes, self.sz)
        self._encode(w_shard)
        if self.world == 1:
            return self._decode(self.codes, self.sz)
        _, _, coff, cbytes, soff, sbytes = self.layout[self.rank]
        codes, sz = gather_payload(self.codes[coff:coff + cbytes], self.sz[soff:soff + sbytes], self.layout, self.pg)
        return self._decode(codes, sz)

    @torch.no_grad()
    def backward(self, w_shard, g_full, average=False):
        from . import _lib as L
        g = reduce_scatter_rows(g_full.to(self.dtype).contiguous(), self.layout, self.rank, self.pg, average)
        w = w_shard.detach().contiguous()
        gw = torch.empty_like(w)
        with L.on(w.device):
            rc = L.lib().bdq_int_ste_bwd(L.ptr(w), L.ptr(g), L.ptr(gw), L.i64(w.numel()), int(self.group_size), 2,
                                         L.dtype_code(w), L.semantics(), L.stream_ptr(w.device))
        L.check(rc, "bdq_int_ste_bwd")
        return gw


def nf3_payload_layout(rows, cols, group_size, world, multiple=16):
    """payload_layout for the NF3 wire format (include/bdq.h: 12 bytes of level indices per 32 weights, one (|max|, |min|) pair
    of 2-byte values per group): per rank (row0, row1, codes byte offset, codes bytes, scales byte offset, scales bytes)."""
    if cols % group_size != 0 or group_size % 32 != 0:
        raise ValueError(f"cols={cols} must be a multiple of group_size={group_size}, itself a multiple of 32")
    out = []
    for r in range(world):
        r0, r1 = row_shard(rows, r, world, multiple if world > 1 else 1)
        out.append((r0, r1, r0 * cols // 32 * 12, (r1 - r0) * cols // 32 * 12, r0 * (cols // group_size) * 4,
                    (r1 - r0) * (cols // group_size) * 4))
    return out


class ShardGatherNF3:
    """ShardGather for the NF3 quantiser -- QLinear's class default `quant_type="ste-n2f3"` (quantization/qlinear.py:51) and
    the reference's recommended 3-bit recipe: `forward(w_shard)` returns SteN2F3Quantizer(group)(all_gather(w_shard))
    (quantization/quantizer.py:184-242) bit for bit on every rank, moving 0.41 B/weight instead of 2 (bdq_nf3_shard_encode
    stores 3-bit level indices + the two scales of every group into every rank's symmetric-memory buffer, one barrier,
    bdq_nf3_shard_decode).  `backward`: reduce-scatter of the gradient w.r.t. the quantised weight, then bdq_nf3_ste_bwd on the
    shard (the NF3 STE backward is linear in the incoming gradient and group-local, like the integer one); after
    enable_fused_backward() both are one pass over NVLink peer memory (ShardGradReducer)."""

    def __init__(self, rows, cols, group_size, dtype, device, group=None, symmetric=True):
        self.rows, self.cols, self.group_size, self.dtype, self.device = rows, cols, group_size, dtype, torch.device(device)
        self.pg = group
        on = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if on else 0
        self.world = dist.get_world_size(group) if on else 1
        self.layout = nf3_payload_layout(rows, cols, group_size, self.world)
        self.ncode = rows * cols // 32 * 12
        self.nsc = rows * (cols // group_size) * 4
        self.symmetric = bool(symmetric) and self.world > 1
        if self.symmetric:
            import torch.distributed._symmetric_memory as symm_mem
            self.half = (self.ncode + self.nsc + 511) // 512 * 512
            self.buf2 = symm_mem.empty(2 * self.half, dtype=torch.uint8, device=self.device)     # double-buffered: one barrier per call
            self.hdl = symm_mem.rendezvous(self.buf2, self.pg if self.pg is not None else dist.group.WORLD)
            self.ptrs2 = [int(p) for p in self.hdl.buffer_ptrs]
            self.flip = 0
            self.barrier = PeerBarrier(self.device, self.pg)
        else:
            self.buf = torch.empty(self.ncode + self.nsc, dtype=torch.uint8, device=self.device)

    @torch.no_grad()
    def forward(self, w_shard):
        import ctypes

        from . import _lib as L
        r0, r1, coff, cbytes, soff, sbytes = self.layout[self.rank]
        assert w_shard.shape == (r1 - r0, self.cols) and w_shard.dtype == self.dtype
        L.require_cuda(w_shard)
        w = w_shard.detach().contiguous()
        if self.symmetric:
            off = self.flip * self.half
            self.flip ^= 1
            buf = self.buf2[off:off + self.ncode + self.nsc]
            peers = [p + off for i, p in enumerate(self.ptrs2) if i != self.rank]
        else:
            buf, peers = self.buf, []
        base = buf.data_ptr()
        pc = (ctypes.c_void_p * max(len(peers), 1))(*[p + coff for p in peers])
        ps = (ctypes.c_void_p * max(len(peers), 1))(*[p + self.ncode + soff for p in peers])
        st = L.stream_ptr(self.device)
        with L.on(self.device):
            rc = L.lib().bdq_nf3_shard_encode(L.ptr(w), ctypes.c_void_p(base + coff), ctypes.c_void_p(base + self.ncode + soff),
                                              L.i64(w.numel()), int(self.group_size), L.dtype_code(w), L.semantics(), len(peers), pc, ps, st)
            L.check(rc, "bdq_nf3_shard_encode")
            codes, sc = buf[:self.ncode], buf[self.ncode:]
            if self.symmetric:
                self.barrier()
            elif self.world > 1:
                codes, sc = gather_payload(codes[coff:coff + cbytes], sc[soff:soff + sbytes], self.layout, self.pg)
            dq = torch.empty((self.rows, self.cols), dtype=self.dtype, device=self.device)
            rc = L.lib().bdq_nf3_shard_decode(L.ptr(codes), L.ptr(sc), L.ptr(dq), L.i64(dq.numel()), int(self.group_size),
                                              L.dtype_code(dq), st)
            L.check(rc, "bdq_nf3_shard_decode")
        self.codes, self.sc = codes, sc
        return dq

    def enable_fused_backward(self):
        """Use ShardGradReducer(quant="nf3") for `backward`: gradient rows pushed over NVLink peer memory + the multi-source NF3
        STE backward (bdq_nf3_ste_bwd_nsrc), no collective, the reduced gradient never materialised."""
        if self.symmetric and getattr(self, "reducer", None) is None:
            self.reducer = ShardGradReducer([(self.rows, self.cols)], self.group_size, self.dtype, self.device, self.pg, quant="nf3")
        return getattr(self, "reducer", None) is not None

    @torch.no_grad()
    def backward(self, w_shard, g_full, average=False):
        from . import _lib as L
        if getattr(self, "reducer", None) is not None and not average:
            return self.reducer.backward([w_shard], [g_full.to(self.dtype)])[0]
        g = reduce_scatter_rows(g_full.to(self.dtype).contiguous(), self.layout, self.rank, self.pg, average)
        w = w_shard.detach().contiguous()
        gw = torch.empty_like(w)
        with L.on(w.device):
            rc = L.lib().bdq_nf3_ste_bwd(L.ptr(w), L.ptr(g), L.ptr(gw), L.i64(w.numel()), int(self.group_size), L.dtype_code(w),
                                         L.semantics(), L.stream_ptr(w.device))
        L.check(rc, "bdq_nf3_ste_bwd")
        return gw


class ShardGatherGroup:
    """ShardGather for a GROUP of weights that are gathered together (FSDP2 gathers the parameters of one decoder layer in
    one all-gather): one gather buffer for the group, laid out as the concatenation of the tensors' payloads -- which is
    the payload of the concatenated (flat) weight, so ONE decode launch rebuilds every tensor of the group into one flat
    buffer; the encode kernel runs once per tensor shard and stores into that tensor's slot of every rank's buffer.
    `forward(list of w_shard)` -> list of fake-quantised full weights (views of the flat buffer, overwritten by the next
    call).  The symmetric gather buffer is double-buffered, so ONE barrier per call suffices: passing the barrier of call i
    (all encodes of call i are done) implies every rank has finished the decode of call i-1, whose buffer call i+1 reuses."""

    def __init__(self, shapes, group_size, dtype, device, group=None, symmetric=True, round_mode=None):
        from . import _lib as L
        self.shapes, self.group_size, self.dtype, self.device, self.pg = list(shapes), group_size, dtype, torch.device(device), group
        on = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if on else 0
        self.world = dist.get_world_size(group) if on else 1
        self.round_mode = L.ROUND_HALF_AWAY if round_mode is None else round_mode
        self.layouts = [payload_layout(r, c, group_size, self.world) for r, c in self.shapes]
        self.elems = [r * c for r, c in self.shapes]
        self.code_base, self.sz_base, co, so = [], [], 0, 0
        for (r, c) in self.shapes:
            self.code_base.append(co)
            self.sz_base.append(so)
            co += r * c // 4
            so += r * (c // group_size) * 4
        self.ncode, self.nsz = co, so
        self.symmetric = bool(symmetric) and self.world > 1
        if self.symmetric:
            import torch.distributed._symmetric_memory as symm_mem
            self.half = (co + so + 511) // 512 * 512
            self.buf2 = symm_mem.empty(2 * self.half, dtype=torch.uint8, device=self.device)
            self.hdl = symm_mem.rendezvous(self.buf2, self.pg if self.pg is not None else dist.group.WORLD)
            self.ptrs2 = [int(p) for p in self.hdl.buffer_ptrs]
            self.flip = 0
            self._select(0)
            self.barrier = PeerBarrier(self.device, self.pg)
        else:
            self.buf = torch.empty(co + so, dtype=torch.uint8, device=self.device)
            self.ptrs = [self.buf.data_ptr()]
            self.codes, self.sz = self.buf[:co], self.buf[co:]
        self.flat = torch.empty(sum(self.elems), dtype=dtype, device=self.device)
        self.views, off = [], 0
        for (r, c), n in zip(self.shapes, self.elems):
            self.views.append(self.flat[off:off + n].view(r, c))
            off += n

    def _select(self, which):
        off = which * self.half
        self.buf = self.buf2[off:off + self.ncode + self.nsz]
        self.ptrs = [p + off for p in self.ptrs2]
        self.codes, self.sz = self.buf[:self.ncode], self.buf[self.ncode:]

    def _encode_all(self, shards):
        """All shards of the group in ONE launch (bdq_int2_shard_encode_multi, up to 8 segments per launch)."""
        import ctypes

        from . import _lib as L
        base = self.buf.data_ptr()
        peers = [p for i, p in enumerate(self.ptrs) if i != self.rank] if self.symmetric else []
        st = L.stream_ptr(self.device)
        ws, coffs, soffs = [], [], []
        for t, w_shard in enumerate(shards):
            r0, r1, coff, cbytes, soff, sbytes = self.layouts[t][self.rank]
            assert w_shard.shape == (r1 - r0, self.shapes[t][1]) and w_shard.dtype == self.dtype
            L.require_cuda(w_shard)
            ws.append(w_shard.detach().contiguous())
            coffs.append(self.code_base[t] + coff)
            soffs.append(self.sz_base[t] + soff)
        pc = (ctypes.c_void_p * max(len(peers), 1))(*peers)
        ps = (ctypes.c_void_p * max(len(peers), 1))(*[p + self.ncode for p in peers])
        with L.on(self.device):
            for i in range(0, len(ws), 8):
                n = len(ws[i:i + 8])
                rc = L.lib().bdq_int2_shard_encode_multi(
                    n, (ctypes.c_void_p * n)(*[w.data_ptr() for w in ws[i:i + 8]]), (ctypes.c_int64 * n)(*[w.numel() for w in ws[i:i + 8]]),
                    (ctypes.c_int64 * n)(*coffs[i:i + 8]), (ctypes.c_int64 * n)(*soffs[i:i + 8]), ctypes.c_void_p(base),
                    ctypes.c_void_p(base + self.ncode), int(self.group_size), L.dtype_code(ws[0]), int(self.round_mode), L.semantics(),
                    len(peers), pc, ps, st)
                L.check(rc, "bdq_int2_shard_encode_multi")

    @torch.no_grad()
    def forward(self, shards, record=None):
        """`record`: optional list; a (start, stop) CUDA-event pair around the decode launch is appended (bench.py)."""
        from . import _lib as L
        if self.symmetric:
            self._select(self.flip)
            self.flip ^= 1
            self._encode_all(shards)
            self.barrier()
        else:
            self._encode_all(shards)
            if self.world > 1:                       # collective transport: per tensor, ragged slots padded
                for t in range(len(self.shapes)):
                    _, _, coff, cb, soff, sb = self.layouts[t][self.rank]
                    c0, s0 = self.code_base[t], self.sz_base[t]
                    codes, sz = gather_payload(self.codes[c0 + coff:c0 + coff + cb], self.sz[s0 + soff:s0 + soff + sb],
                                               self.layouts[t], self.pg)
                    self.codes[c0:c0 + codes.numel()] = codes
                    self.sz[s0:s0 + sz.numel()] = sz
        with L.on(self.device):
            if record is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            rc = L.lib().bdq_int2_shard_decode(L.ptr(self.codes), L.ptr(self.sz), L.ptr(self.flat), L.i64(self.flat.numel()),
                                               int(self.group_size), L.dtype_code(self.flat), L.stream_ptr(self.device))
            if record is not None:
                e1.record()
                record.append((e0, e1))
        L.check(rc, "bdq_int2_shard_decode")
        return self.views


    def enable_fused_backward(self):
        """Use ShardGradReducer (gradient push over NVLink peer memory + multi-source STE backward, no collective, the reduced
        gradient never materialised) for `backward` instead of NCCL reduce-scatter + bdq_int_ste_bwd."""
        if self.symmetric and getattr(self, "reducer", None) is None:
            self.reducer = ShardGradReducer(self.shapes, self.group_size, self.dtype, self.device, self.pg)
        return getattr(self, "reducer", None) is not None

    @torch.no_grad()
    def backward(self, shards, g_fulls, average=False):
        """Per tensor: reduce-scatter of the gradient w.r.t. the quantised weight, then the fused STE backward on the shard
        (see ShardGather.backward); with enable_fused_backward() the two are one pass (ShardGradReducer)."""
        from . import _lib as L
        if getattr(self, "reducer", None) is not None and not average and all(g is not None for g in g_fulls):
            return self.reducer.backward(list(shards), [g.to(self.dtype) for g in g_fulls])
        out = []
        for t, (w_shard, g_full) in enumerate(zip(shards, g_fulls)):
            if g_full is None:
                out.append(None)
                continue
            g = reduce_scatter_rows(g_full.to(self.dtype).contiguous(), self.layouts[t], self.rank, self.pg, average)
            w = w_shard.detach().contiguous()
            gw = torch.empty_like(w)
            with L.on(w.device):
                rc = L.lib().bdq_int_ste_bwd(L.ptr(w), L.ptr(g), L.ptr(gw), L.i64(w.numel()), int(self.group_size), 2,
                                             L.dtype_code(w), L.semantics(), L.stream_ptr(w.device))
            L.check(rc, "bdq_int_ste_bwd")
            out.append(gw)
        return out


class _ShardGatherGroupFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gatherer, average, *shards):
        ctx.gatherer, ctx.average = gatherer, average
        ctx.save_for_backward(*shards)
        # the views alias the gatherer's flat buffer, which the next forward overwrites: hand out copies that autograd may keep
        return tuple(v.clone() for v in gatherer.forward(list(shards)))

    @staticmethod
    def backward(ctx, *g_fulls):
        return (None, None) + tuple(ctx.gatherer.backward(list(ctx.saved_tensors), list(g_fulls), ctx.average))


def quantized_all_gather_group(shards, gatherer, average=False):
    """Autograd entry for a parameter group: list of shard parameters -> tuple of fake-quantised full weights
    (ShardGatherGroup: one encode launch, one barrier, one decode launch per call)."""
    return _ShardGatherGroupFn.apply(gatherer, average, *shards)


class _ShardGatherFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, w_shard, gatherer, average):
        ctx.gatherer, ctx.average = gatherer, average
        ctx.save_for_backward(w_shard)
        return gatherer.forward(w_shard)

    @staticmethod
    def backward(ctx, g_full):
        (w_shard,) = ctx.saved_tensors
        return ctx.gatherer.backward(w_shard, g_full, ctx.average), None, None


def quantized_all_gather(w_shard, gatherer, average=False):
    """Autograd entry: shard parameter -> fake-quantised full weight (see ShardGather)."""
    return _ShardGatherFn.apply(w_shard, gatherer, average)


# ------------------------------------------------------------------------------------------------- gradient push + fused STE
def grad_inbox_layout(shapes, world, elem_size, multiple=16):
    """Host-side layout of ShardGradReducer: spans[t][k] = rows of tensor t owned by rank k, off[k][t] = byte offset of tensor
    t inside rank k's flat shard layout, size[k] = bytes of that layout, slot = bytes of one inbox slot (256-aligned max)."""
    spans = [[row_shard(r, k, world, multiple) for k in range(world)] for r, _ in shapes]
    off = [[0] * len(shapes) for _ in range(world)]
    size = [0] * world
    for k in range(world):
        o = 0
        for t, (_, c) in enumerate(shapes):
            off[k][t] = o
            o += (spans[t][k][1] - spans[t][k][0]) * c * elem_size
        size[k] = o
    return spans, off, size, (max(size) + 255) // 256 * 256


class ShardGradReducer:
    """The backward half of the sharded QAT step as ONE pass over NVLink peer memory (the reference leaves it to ZeRO-3 / FSDP:
    full-weight STE backward on every rank, then a reduce-scatter; quantization/qlinear.py:63-78):
      1. `bdq_rows_push`: every rank stores the rows of its gradient (w.r.t. the fake-quantised weight) that belong to rank
         r's dim-0 shard into r's symmetric-memory inbox -- the kernel's stores are the NVLink transfer, one launch per
         parameter group;
      2. one symmetric-memory barrier (the inbox is double-buffered across calls, like ShardGatherGroup's payload);
      3. `bdq_int_ste_bwd_nsrc` (`bdq_nf3_ste_bwd_nsrc` with quant="nf3") per tensor: the STE backward of the shard with the incoming gradient summed on the fly from the
         rank's own rows and the n-1 inbox slots, in rank order, one rounding to the dtype per addition (for two ranks: the
         rounded exact sum, bit-equal to an NCCL reduce-scatter; in general what a ring reduction in that dtype computes).
    The reduced gradient is never materialised.  `shapes`: [(rows, cols)] of the group's tensors."""

    def __init__(self, shapes, group_size, dtype, device, group=None, quant="int2"):
        import torch.distributed._symmetric_memory as symm_mem
        if quant not in ("int2", "nf3"):
            raise ValueError("ShardGradReducer: quant must be 'int2' (SteInt2AsymQuantizer) or 'nf3' (SteN2F3Quantizer)")
        self.quant = quant
        self.shapes, self.group_size, self.dtype, self.device, self.pg = list(shapes), group_size, dtype, torch.device(device), group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if not 2 <= self.world <= 8:
            raise ValueError("ShardGradReducer needs 2..8 ranks")
        es = torch.empty(0, dtype=dtype).element_size()
        self.spans, self.off, self.size, self.slot = grad_inbox_layout(self.shapes, self.world, es)
        self.half = (self.world - 1) * self.slot
        self.inbox = symm_mem.empty(2 * self.half, dtype=torch.uint8, device=self.device)
        self.hdl = symm_mem.rendezvous(self.inbox, self.pg if self.pg is not None else dist.group.WORLD)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.flip = 0
        self.es = es
        self.barrier = PeerBarrier(self.device, self.pg)

    def _slot_of(self, src, dst):
        """Index of source rank `src` among the n-1 inbox slots of rank `dst`."""
        return src if src < dst else src - 1

    @torch.no_grad()
    def push(self, g_fulls):
        """Stage 1 + 2 on the reducer's own stream (so that the NVLink transfer of one parameter group overlaps the HBM-bound
        STE backward of the previous one): push this rank's rows to their owners, then the barrier.  Returns a ticket for
        `finish`.  Calls must alternate push / finish per group, at most one push ahead."""
        import ctypes

        from . import _lib as L
        if not hasattr(self, "comm"):
            # high priority: the push kernel's CTAs are placed before the HBM-bound STE launches they overlap with, so the NVLink
            # transfer (the longer of the two) is never the one that waits for SM slots (BDQ_PUSH_PRIORITY=0: default priority)
            prio = -1 if os.environ.get("BDQ_PUSH_PRIORITY", "-1") != "0" else 0
            self.comm = torch.cuda.Stream(self.device, priority=prio)
            self._done = None                 # event: the STE launches of the previous `finish` have completed
        base = self.flip * self.half
        self.flip ^= 1
        me, n, es = self.rank, self.world, self.es
        src, dst, nb = [], [], []
        gs = [g.contiguous() for g in g_fulls]
        for t, ((_, c), g) in enumerate(zip(self.shapes, gs)):
            L.require_cuda(g)
            assert g.dtype == self.dtype and tuple(g.shape) == tuple(self.shapes[t])
            for k in range(n):
                if k == me:
                    continue
                r0, r1 = self.spans[t][k]
                if r1 > r0:
                    src.append(g.data_ptr() + r0 * c * es)
                    dst.append(self.ptrs[k] + base + self._slot_of(me, k) * self.slot + self.off[k][t])
                    nb.append((r1 - r0) * c * es)
        m = len(src)
        cur = torch.cuda.current_stream(self.device)
        self.comm.wait_stream(cur)               # the gradients were produced on the caller's stream
        with L.on(self.device), torch.cuda.stream(self.comm):
            rc = L.lib().bdq_rows_push(m, (ctypes.c_void_p * max(m, 1))(*src), (ctypes.c_void_p * max(m, 1))(*dst),
                                       (ctypes.c_int64 * max(m, 1))(*nb), L.stream_ptr(self.device))
            L.check(rc, "bdq_rows_push")
            # passing this barrier tells every peer that this rank is done READING the inbox half the peers write next:
            # wait for the previous group's STE launches (they read the other half, which the next push overwrites)
            if self._done is not None:
                self.comm.wait_event(self._done)
            self.barrier()                       # every rank's rows have landed in this rank's inbox
            landed = torch.cuda.Event()
            landed.record(self.comm)
        return (base, gs, landed)

    @torch.no_grad()
    def finish(self, ticket, shards, outs=None):
        """Stage 3 on the caller's stream: the multi-source STE backward of every tensor of the group."""
        import ctypes

        from . import _lib as L
        base, gs, landed = ticket
        me, n, es = self.rank, self.world, self.es
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(landed)
        st = L.stream_ptr(self.device)
        res = []
        with L.on(self.device):
            for t, ((_, c), w, g) in enumerate(zip(self.shapes, shards, gs)):
                r0, r1 = self.spans[t][me]
                w = w.detach().contiguous()
                gw = torch.empty_like(w) if outs is None else outs[t]
                srcs = []
                for k in range(n):               # rank order: every shard is reduced in the same order as an all-reduce would
                    if k == me:
                        srcs.append(g.data_ptr() + r0 * c * es)
                    else:
                        srcs.append(self.inbox.data_ptr() + base + self._slot_of(k, me) * self.slot + self.off[me][t])
                if self.quant == "nf3":
                    rc = L.lib().bdq_nf3_ste_bwd_nsrc(L.ptr(w), n, (ctypes.c_void_p * n)(*srcs), L.ptr(gw), L.i64(w.numel()),
                                                      int(self.group_size), L.dtype_code(w), L.semantics(), st)
                    L.check(rc, "bdq_nf3_ste_bwd_nsrc")
                else:
                    rc = L.lib().bdq_int_ste_bwd_nsrc(L.ptr(w), n, (ctypes.c_void_p * n)(*srcs), L.ptr(gw), L.i64(w.numel()),
                                                      int(self.group_size), 2, L.dtype_code(w), L.semantics(), st)
                    L.check(rc, "bdq_int_ste_bwd_nsrc")
                res.append(gw)
        self._done = torch.cuda.Event()
        self._done.record(cur)
        for g in gs:                             # pushed on the reducer's stream: keep the allocator from reusing them early
            g.record_stream(self.comm)
        return res

    def backward(self, shards, g_fulls, outs=None):
        """shards[t]: this rank's rows of weight t; g_fulls[t]: this rank's [rows, cols] gradient w.r.t. the fake-quantised
        weight t.  Returns the gradients of the shards (written into `outs` if given)."""
        return self.finish(self.push(g_fulls), shards, outs)
