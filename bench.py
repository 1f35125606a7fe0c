#!/usr/bin/env python
"""bench.py -- headline benchmark of the low-bit weight-quantisation hot path (BASELINE.json).

Workload: BASELINE.json configs[1] -- one QAT step's worth of fake-quant forward (SteInt2AsymQuantizer, int2, group 128)
+ straight-through backward over ALL converted linears of a random-init Llama-2-7B (224 bf16 tensors, 6.476 G weights).

N = 1: the whole model on one GPU, one forward + one backward launch per decoder layer over the layer's flat buffer.
N > 1: the model is SHARDED the way the reference trains it under ZeRO-3 / FSDP2 (train/config/zero3.json, dim-0 shards
  of every linear, `sharded.row_shard`) -> "scaling": "strong" (one model, N GPUs).  Per decoder layer and step:
    forward : every rank quantises its shard ONCE, the encode kernel stores the 0.28 B/weight int2 payload into every
              rank's symmetric-memory buffer over NVLink (no collective), one barrier, one decode launch rebuilds the
              fake-quantised FULL weights on every rank (what QLinear.forward needs, qlinear.py:75) -- ShardGatherGroup;
    backward: every rank pushes the rows of its gradient (every rank has its own batch's gradient) into the owning rank's
              symmetric-memory inbox over NVLink (bdq_rows_push, no collective), one barrier, then the STE backward of the
              shard sums the N sources on the fly (bdq_int_ste_bwd_nsrc: the STE backward is linear in the incoming gradient
              and group-local, so it commutes with the sum over ranks; the reduced gradient is never materialised).
  `sharded.reference_order` in the line = the reference's order on the same shards (NCCL all-gather of the bf16 shards,
  full fake-quant and full STE backward on every rank, then the same reduce-scatter), timed the same way.

  value   = algorithmic bytes of the model step (10 B/weight: fwd 2R+2W, bwd 2R+2R+2W) / step time, inputs resident in HBM
  e2e     = same metric through the Python drop-in API with pinned HOST buffers: H2D of weights and upstream grads,
            kernels, D2H of the fake-quantised weights and the weight gradients, all inside the timed region
  roofline= the dominant kernel (int_ste_bwd_kernel, 6 B/weight) timed per launch with CUDA events on its stream
  cpu_baseline / --impl reference = the CPU oracle port (oracle/, all host threads) on a bounded sample
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fake-quant+STE weight GB/s (Llama-2-7B int2-g128 fwd+bwd per QAT step)"
UNIT = "GB/s"
GROUP = 128
BITS = 2
LAYERS = 32
# (name, out_features, in_features) of the converted linears of one Llama-2-7B decoder layer
LAYER_SHAPES = [("q_proj", 4096, 4096), ("k_proj", 4096, 4096), ("v_proj", 4096, 4096), ("o_proj", 4096, 4096),
                ("gate_proj", 11008, 4096), ("up_proj", 11008, 4096), ("down_proj", 4096, 11008)]
WEIGHTS_PER_LAYER = sum(o * i for _, o, i in LAYER_SHAPES)
FWD_BYTES, BWD_BYTES = 4, 6  # per bf16 weight


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_port_sample(threads):
    """Time the CPU oracle port on ONE decoder layer (7 tensors, 202.4 M weights): int2-g128 forward (half-away
    rounding, quantizer.py:153-182) + closed-form STE backward, rows split over `threads` host threads."""
    import numpy as np
    import ctypes
    import oracle
    lib = oracle.lib()
    rng = np.random.default_rng(0)
    total_w, t_total = 0, 0.0
    for _, o, i in LAYER_SHAPES:
        w = (rng.standard_normal((o, i), dtype=np.float32) * 0.02)
        # bf16-representable values (the oracle takes fp32 arrays holding T values)
        w = (w.view(np.uint32) & np.uint32(0xffff0000)).view(np.float32)
        gy = rng.standard_normal((o, i), dtype=np.float32)
        gy = (gy.view(np.uint32) & np.uint32(0xffff0000)).view(np.float32)
        dq = np.empty_like(w)
        gw = np.empty((o, i), np.float64)
        bounds = np.linspace(0, o, threads + 1).astype(int)

        def work(k):
            r0, r1 = int(bounds[k]), int(bounds[k + 1])
            if r1 <= r0:
                return
            n = ctypes.c_int64((r1 - r0) * i)
            p = lambda a: a[r0:r1].ctypes.data_as(ctypes.c_void_p)
            lib.orc_int_fakequant(p(w), n, GROUP, BITS, 1, 1, 0, None, None, p(dq), None, None, None, None)
            lib.orc_int_ste_bwd(p(w), p(gy), n, GROUP, BITS, 1, 0, p(gw))

        ths = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        t_total += time.perf_counter() - t0
        total_w += o * i
    return total_w, t_total


def run_reference(args):
    """--impl reference: the reference's CPU path for this workload = the oracle port (the reference itself is
    Python/PyTorch and cannot travel to the GPU box; the port is pinned to it by tests/golden)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.build()
    threads = os.cpu_count() or 1
    times = []
    for s in range(args.warmup + args.steps):
        w, t = cpu_port_sample(threads)
        if s >= args.warmup:
            times.append(t)
    t = sum(times) / len(times)
    gbs = w * (FWD_BYTES + BWD_BYTES) / t / 1e9
    sample = f"1 of {LAYERS} decoder layers per step ({w/1e6:.1f} M weights, fwd+bwd), rate extrapolates linearly"
    line = {"impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(n):
    cfg = {"workload": "configs[1]: Llama-2-7B random-init, all 224 converted linears, int2-g128 "
                       "SteInt2AsymQuantizer forward + STE backward per QAT step (bf16)",
           "weights": WEIGHTS_PER_LAYER * LAYERS, "group": GROUP, "bits": BITS}
    if n == 1:
        cfg.update({"parallelism": "1 GPU, whole model",
                    "launches": "one forward + one backward launch per decoder layer over the layer's flat parameter "
                                "buffer (7 tensors, 202.4 M weights)",
                    "l2": "inputs are 25.9 GB per step, >> 126 MB L2 (no flush needed)"})
    else:
        cfg.update({"parallelism": f"fsdp{n}: dim-0 shards of every linear (ZeRO-3 / FSDP2 layout), quantise-before-"
                                   "all-gather forward over NVLink peer stores, reduce-scatter + shard STE backward",
                    "launches": "per decoder layer: 1 encode (7 shards) + 1 barrier + 1 decode forward; 1 gradient push + "
                                "1 barrier + 7 multi-source STE backward launches",
                    "l2": "every rank reads its 12.95 GB upstream gradient and writes 12.95 GB of fake-quantised "
                          "weights per step, >> 126 MB L2 (no flush needed)"})
    return cfg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="bdq", choices=["bdq", "reference"])
    ap.add_argument("--layers", type=int, default=LAYERS, help=argparse.SUPPRESS)  # dev only; default = full model
    ap.add_argument("--no-cpu-baseline", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import torch.distributed as dist
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nlayers = args.layers
    lib = L.lib()
    st = L.stream_ptr(dev)
    if world > 1:
        try:
            run_sharded(args, torch, dist, L, dev, rank, world, local)
        finally:
            dist.destroy_process_group()
        return

    # ---- synthetic model: weights and upstream gradients resident in HBM.  The 7 converted linears of a decoder layer
    # live in ONE flat bf16 buffer (the layout a flat-parameter wrapper -- DeepSpeed ZeRO / FSDP per-layer wrapping, what
    # the reference trains under -- gives them); quantisation groups are 128 contiguous elements and every tensor size
    # is a multiple of 128, so one launch over the flat buffer fake-quantises all 7 tensors exactly as 7 launches would.
    gen = torch.Generator(device=dev)
    Ws, Gs = [], []
    for li in range(nlayers):
        wf = torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16)
        gf = torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16)
        off = 0
        for ti, (_, o, i) in enumerate(LAYER_SHAPES):
            gen.manual_seed(1000 * li + ti)
            wf[off:off + o * i] = (torch.randn(o * i, device=dev, generator=gen, dtype=torch.float32) * 0.02).to(torch.bfloat16)
            gf[off:off + o * i] = torch.randn(o * i, device=dev, generator=gen, dtype=torch.float32).to(torch.bfloat16)
            off += o * i
        Ws.append(wf)
        Gs.append(gf)
    nbuf = 2   # output buffers of one layer, rotated (2 x 405 MB each, far beyond L2)
    out_dq = [torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    out_gw = [torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    weights = sum(w.numel() for w in Ws)
    bwd_events = []

    def step(record=False):
        for k, w in enumerate(Ws):
            rc = lib.bdq_int_fakequant_fwd(L.ptr(w), L.ptr(out_dq[k % nbuf]), None, None, None, None, None, L.i64(w.numel()),
                                           GROUP, BITS, L.BF16, L.ROUND_HALF_AWAY, L.semantics(), st)
            L.check(rc, "bdq_int_fakequant_fwd")
        for k, (w, g) in enumerate(zip(Ws, Gs)):
            if record:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            rc = lib.bdq_int_ste_bwd(L.ptr(w), L.ptr(g), L.ptr(out_gw[k % nbuf]), L.i64(w.numel()), GROUP, BITS, L.BF16,
                                     L.semantics(), st)
            L.check(rc, "bdq_int_ste_bwd")
            if record:
                e1.record()
                bwd_events.append((e0, e1, w.numel()))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    n0 = L.launch_count()
    e_start, e_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e_start.record()
    for _ in range(args.steps):
        step(record=True)
    e_stop.record()
    barrier()
    launches = L.launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None
    elapsed = torch.tensor([e_start.elapsed_time(e_stop) * 1e-3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
    elapsed = float(elapsed.item())
    step_s = elapsed / args.steps
    value = weights * (FWD_BYTES + BWD_BYTES) * world / step_s / 1e9

    # roofline of the dominant kernel from the per-launch events
    bwd_t = sum(a.elapsed_time(b) for a, b, _ in bwd_events) * 1e-3
    bwd_b = sum(n for _, _, n in bwd_events) * BWD_BYTES
    peak, peak_src = peaks()
    achieved = bwd_b / bwd_t / 1e9
    traffic = traffic_src = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            tj = json.load(f)
        traffic = tj.get("int_ste_bwd_kernel_bytes_per_launch")
        # NOT a property of this run: dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this
        # kernel at this launch shape, recorded in profiles/ together with the capture it came from
        traffic_src = tj.get("source", "profiles/roofline_traffic.json (ncu --set full capture of an earlier build)")
    roofline = {"bound": "hbm", "kernel": "int_thr_bwd_kernel<bf16,128> (bdq_int_ste_bwd)", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src, "algorithmic_bytes_per_weight": BWD_BYTES,
                "algorithmic_bytes_per_launch": BWD_BYTES * WEIGHTS_PER_LAYER, "bwd_share_of_step": bwd_t / elapsed,
                "launches_timed": len(bwd_events)}

    # ---- what the timed launches wrote, checked outside the timed region: the last layer's outputs are still in the
    # rotating buffers
    lw = nlayers - 1
    verified, verified_how = verify_slabs(torch, L, Ws[lw], Gs[lw], out_dq[lw % nbuf], out_gw[lw % nbuf], L.semantics())

    # ---- e2e: Python drop-in API, pinned host buffers in and out, copies inside the timed region
    e2e = run_e2e(torch, dev, world, nlayers, SteInt2AsymQuantizer, min(args.steps, 3))

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(world), "roofline": roofline,
            "e2e": e2e, "gpu_launches": int(launches) * world, "clocks": clocks,
            "frac_of_hbm_peak_per_gpu": value / world / peak, "verified": verified, "verified_how": verified_how,
            "semantics": "cuda" if L.semantics() else "cpu"}
    if nlayers != LAYERS:
        line["config"]["DEV_ONLY_reduced_layers"] = nlayers
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            import oracle
            oracle.build()
            threads = os.cpu_count() or 1
            w, t = cpu_port_sample(threads)
            line["cpu_baseline"] = {"value": w * (FWD_BYTES + BWD_BYTES) / t / 1e9, "unit": UNIT, "cores": threads,
                                    "kind": "port", "sample": f"1 of {LAYERS} decoder layers ({w/1e6:.1f} M weights, "
                                    f"fwd+bwd) in {t:.1f} s on {threads} threads"}
            if nlayers == LAYERS:
                # the other two numbers BASELINE.json's metric names, measured live on this GPU (not part of `value`)
                del Ws, Gs, out_dq, out_gw
                torch.cuda.empty_cache()
                try:
                    line["secondary"] = secondary_metrics(torch, dev, peak)
                except Exception as e:  # never lose the headline line to a secondary measurement
                    line["secondary"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def verify_slabs(torch, L, w_flat, gy_flat, dq_flat, gw_flat, sem, nslab=6, slab=32 * 1024):
    """Outside the timed region: slabs of what the timed launches wrote (first / last / random, group-aligned) against the
    CPU oracle -- forward bit-exact, backward within 2 ulp + 1e-5 * sum|gy| of the group (the bound of tests/
    test_fakequant_gpu.py).  The oracle is the checker here, never the thing measured."""
    import numpy as np
    import oracle
    oracle.build()
    n = w_flat.numel()
    rs = np.random.RandomState(1)
    starts = [0, n - slab] + [int(x) * GROUP for x in rs.randint(0, (n - slab) // GROUP, size=max(nslab - 2, 0))]
    checked = 0
    for s0 in starts:
        w = w_flat[s0:s0 + slab].cpu()
        if dq_flat is not None:
            ref = oracle.int_fakequant(w, GROUP, BITS, round_mode=1, sem=sem)["dq"].reshape(-1)
            if not torch.equal(dq_flat[s0:s0 + slab].cpu().float(), ref.float()):
                return False, f"forward mismatch in slab at {s0}"
        if gw_flat is not None:
            gy = gy_flat[s0:s0 + slab].cpu()
            ref = oracle.int_ste_bwd(w, gy, GROUP, BITS, sem=sem).reshape(-1)
            got = gw_flat[s0:s0 + slab].cpu().double()
            ulp = torch.pow(torch.tensor(2.0, dtype=torch.float64), torch.floor(torch.log2(ref.abs().clamp(min=2.0 ** -126))) - 7)
            gsum = gy.double().abs().reshape(-1, GROUP).sum(1, keepdim=True).expand(-1, GROUP).reshape(-1)
            if bool(((got - ref).abs() > 2.01 * ulp + 1e-5 * gsum + 1e-30).any()):
                return False, f"backward mismatch in slab at {s0}"
        checked += slab
    return True, f"{len(starts)} slabs x {slab} weights of the last timed layer vs oracle (sem={'cuda' if sem else 'cpu'})"


def run_sharded(args, torch, dist, L, dev, rank, world, local):
    """N > 1: the sharded QAT step described in the module docstring.  One process per GPU."""
    from bitdistiller_b200 import sharded as S
    lib, st, nlayers = L.lib(), L.stream_ptr(dev), args.layers
    sem = L.semantics()
    shapes = [(o, i) for _, o, i in LAYER_SHAPES]
    spans = [S.row_shard(o, rank, world) for o, _ in shapes]
    sh_elems = [(b - a) * i for (a, b), (_, i) in zip(spans, shapes)]
    nshard = sum(sh_elems)
    gen = torch.Generator(device=dev)
    Wsh, Gfull = [], []          # per layer: flat shard weights (tensor-major), flat FULL upstream gradient of this rank's batch
    for li in range(nlayers):
        wf = torch.empty(nshard, device=dev, dtype=torch.bfloat16)
        gf = torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16)
        so = fo = 0
        for ti, ((o, i), (a, b)) in enumerate(zip(shapes, spans)):
            gen.manual_seed(1000 * li + ti)       # the same model as N = 1: rows [a, b) of the same seeded tensors
            full = (torch.randn(o * i, device=dev, generator=gen, dtype=torch.float32) * 0.02).to(torch.bfloat16)
            wf[so:so + (b - a) * i] = full[a * i:b * i]
            gen.manual_seed(7_000_000 + 100_000 * rank + 1000 * li + ti)   # every rank has its own batch -> its own gradient
            gf[fo:fo + o * i] = torch.randn(o * i, device=dev, generator=gen, dtype=torch.float32).to(torch.bfloat16)
            so += (b - a) * i
            fo += o * i
            del full
        Wsh.append(wf)
        Gfull.append(gf)

    def views(flat, full):
        out, off = [], 0
        for (o, i), (a, b) in zip(shapes, spans):
            r = o if full else (b - a)
            out.append(flat[off:off + r * i].view(r, i))
            off += r * i
        return out

    Wsh_v = [views(w, False) for w in Wsh]
    Gfull_v = [views(g, True) for g in Gfull]
    grp = S.ShardGatherGroup(shapes, GROUP, torch.bfloat16, dev, symmetric=True)
    # backward: gradient push over NVLink peer memory + multi-source STE backward (sharded.ShardGradReducer); the NCCL
    # reduce-scatter + bdq_int_ste_bwd order is timed next to it after the main measurement
    red = S.ShardGradReducer(shapes, GROUP, torch.bfloat16, dev) if os.environ.get("BDQ_BENCH_NCCL_RS", "0") != "1" else None
    nbuf = 2
    g_sh = [torch.empty(nshard, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]     # reduce-scattered gradient shards
    g_sh_v = [views(g, False) for g in g_sh]
    out_gw = [torch.empty(nshard, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    out_gw_v = [views(g, False) for g in out_gw]
    even = all((b - a) * world == o for (a, b), (o, _) in zip(spans, shapes))
    if not even:
        raise SystemExit("bench.py: the Llama-2-7B shapes shard evenly for N in {2, 4, 8}")
    dec_events, phase = [], []
    nccl_bwd = [False]

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def step(record=False):
        t0 = ev() if record else None
        for li in range(nlayers):
            grp.forward(Wsh_v[li], record=dec_events if record else None)
        t1 = ev() if record else None
        rs_ev = []
        fused = red is not None and not nccl_bwd[0]
        ticket = red.push(Gfull_v[0]) if fused else None                   # NVLink push of layer 0 (the reducer's own stream)
        for li in range(nlayers):
            b = li % nbuf
            r0 = ev() if record else None
            if fused:
                nxt = red.push(Gfull_v[li + 1]) if li + 1 < nlayers else None   # overlaps this layer's HBM-bound STE launches
                red.finish(ticket, Wsh_v[li], outs=out_gw_v[b])             # 7 multi-source STE launches
                ticket = nxt
                r1 = r0
            else:
                for t in range(len(shapes)):
                    dist.reduce_scatter_tensor(g_sh_v[b][t], Gfull_v[li][t], op=dist.ReduceOp.SUM)
                r1 = ev() if record else None
                rc = lib.bdq_int_ste_bwd(L.ptr(Wsh[li]), L.ptr(g_sh[b]), L.ptr(out_gw[b]), L.i64(nshard), GROUP, BITS, L.BF16, sem, st)
                L.check(rc, "bdq_int_ste_bwd")
            if record:
                rs_ev.append((r0, r1, ev()))
        if record:
            phase.append((t0, t1, ev(), rs_ev))

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    n0 = L.launch_count()
    e_start, e_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e_start.record()
    for _ in range(args.steps):
        step(record=True)
    e_stop.record()
    barrier()
    launches = L.launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None

    def maxr(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    elapsed = maxr(e_start.elapsed_time(e_stop) * 1e-3)
    step_s = elapsed / args.steps
    weights = WEIGHTS_PER_LAYER * nlayers
    value = weights * (FWD_BYTES + BWD_BYTES) / step_s / 1e9
    fwd_ms = maxr(sum(a.elapsed_time(b) for a, b, _, _ in phase) / len(phase))
    bwd_ms = maxr(sum(b.elapsed_time(c) for _, b, c, _ in phase) / len(phase))
    rs_ms = maxr(sum(r0.elapsed_time(r1) for p in phase for r0, r1, _ in p[3]) / len(phase))
    ste_ms = maxr(sum(r1.elapsed_time(r2) for p in phase for _, r1, r2 in p[3]) / len(phase))
    dec_ms = maxr(sum(a.elapsed_time(b) for a, b in dec_events) / len(phase))

    # roofline of the dominant kernel of OURS at N > 1: the decode (every rank rebuilds the full fake-quantised weights)
    payload_pw = 0.25 + 4.0 / GROUP
    dec_bytes = (payload_pw + 2.0) * WEIGHTS_PER_LAYER
    dec_t = sum(a.elapsed_time(b) for a, b in dec_events) * 1e-3 / len(dec_events)
    peak, peak_src = peaks()
    achieved = dec_bytes / dec_t / 1e9
    roofline = {"bound": "hbm", "kernel": "int2_decode_kernel<bf16,128>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "traffic_source": None, "peak_source": peak_src,
                "algorithmic_bytes_per_weight": payload_pw + 2.0, "share_of_step": dec_ms / (step_s * 1e3),
                "launches_timed": len(dec_events)}

    # ---- correctness of what the timed region produced (outside it): the decode output of the LAST layer must be bit-equal
    # to fake-quantising the all-gathered bf16 weight, and this rank's gradient shard must match the oracle on slabs.
    li = nlayers - 1
    full_w = torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16)
    for t, fv in enumerate(views(full_w, True)):
        dist.all_gather_into_tensor(fv, Wsh_v[li][t])
    dq_ref = torch.empty_like(full_w)
    L.check(lib.bdq_int_fakequant_fwd(L.ptr(full_w), L.ptr(dq_ref), None, None, None, None, None, L.i64(full_w.numel()), GROUP,
                                      BITS, L.BF16, L.ROUND_HALF_AWAY, sem, st), "verify fwd")
    ok_gather = bool(torch.equal(dq_ref.view(torch.int16), grp.flat.view(torch.int16)))
    ok_f, msg_f = verify_slabs(torch, L, full_w, None, grp.flat, None, sem)
    b = li % nbuf
    # reference reduced gradient of this rank's shard: every rank's rows summed in rank order, one bf16 rounding per addition
    # (what the fused backward computes; for 2 ranks also what NCCL computes)
    r_me = []
    for t, ((o, i), (a, bb_)) in enumerate(zip(shapes, spans)):
        parts = [torch.empty_like(Gfull_v[li][t]) for _ in range(world)]
        dist.all_gather(parts, Gfull_v[li][t])
        acc = parts[0][a:bb_].clone()
        for k in range(1, world):
            acc = (acc.float() + parts[k][a:bb_].float()).to(torch.bfloat16)
        r_me.append(acc.reshape(-1))
        del parts
    g_ref = torch.cat(r_me)
    gw_ref = torch.empty_like(g_ref)
    L.check(lib.bdq_int_ste_bwd(L.ptr(Wsh[li]), L.ptr(g_ref), L.ptr(gw_ref), L.i64(nshard), GROUP, BITS, L.BF16, sem, st), "verify bwd")
    ok_red = bool(torch.equal(gw_ref.view(torch.int16), out_gw[b].view(torch.int16))) if red is not None else True
    ok_b, msg_b = verify_slabs(torch, L, Wsh[li], g_ref if red is not None else g_sh[b], None, out_gw[b], sem)
    ok_b = ok_b and ok_red
    flag = torch.tensor([int(ok_gather and ok_f and ok_b)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    verified = bool(flag.item())
    del full_w, dq_ref

    # ---- the reference's order on the same shards (secondary, 2 timed steps): all-gather bf16 -> full fake-quant; full STE
    # backward -> reduce-scatter.  Generous to it: the gathered weights are kept for the backward (ZeRO-3 / FSDP2 re-gather).
    ref_order = None
    try:
        Wfull = [torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16) for _ in range(nlayers)]
        Wfull_v = [views(w, True) for w in Wfull]
        dq_full = [torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
        gw_full = [torch.empty(WEIGHTS_PER_LAYER, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
        gw_full_v = [views(g, True) for g in gw_full]

        def ref_step():
            for li in range(nlayers):
                for t in range(len(shapes)):
                    dist.all_gather_into_tensor(Wfull_v[li][t], Wsh_v[li][t])
                L.check(lib.bdq_int_fakequant_fwd(L.ptr(Wfull[li]), L.ptr(dq_full[li % nbuf]), None, None, None, None, None,
                                                  L.i64(WEIGHTS_PER_LAYER), GROUP, BITS, L.BF16, L.ROUND_HALF_AWAY, sem, st), "ref fwd")
            for li in range(nlayers):
                b = li % nbuf
                L.check(lib.bdq_int_ste_bwd(L.ptr(Wfull[li]), L.ptr(Gfull[li]), L.ptr(gw_full[b]), L.i64(WEIGHTS_PER_LAYER), GROUP,
                                            BITS, L.BF16, sem, st), "ref bwd")
                for t in range(len(shapes)):
                    dist.reduce_scatter_tensor(g_sh_v[b][t], gw_full_v[b][t], op=dist.ReduceOp.SUM)

        ref_step()
        barrier()
        a, bb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(2):
            ref_step()
        bb.record()
        barrier()
        ref_s = maxr(a.elapsed_time(bb) * 1e-3) / 2
        ref_order = {"ms_per_step": ref_s * 1e3, "value": weights * (FWD_BYTES + BWD_BYTES) / ref_s / 1e9, "unit": UNIT,
                     "speedup_of_ours": ref_s / step_s,
                     "what": "NCCL all-gather of the bf16 shards + bdq_int_fakequant_fwd on the full weight + bdq_int_ste_bwd on "
                             "the full weight + the same reduce-scatter (qlinear.py:63-78 under ZeRO-3 / FSDP2; gathered weights "
                             "kept for the backward, i.e. without the re-gather those frameworks do by default)"}
        del Wfull, dq_full, gw_full
        torch.cuda.empty_cache()
    except Exception as e:  # the headline line must survive
        ref_order = {"error": f"{type(e).__name__}: {e}"[:300]}

    # ---- our forward with the NCCL reduce-scatter + bdq_int_ste_bwd backward (the round-1 order), 2 timed steps
    nccl_variant = None
    if red is not None:
        try:
            nccl_bwd[0] = True
            step()
            barrier()
            a, bb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(2):
                step()
            bb.record()
            barrier()
            nccl_variant = {"ms_per_step": maxr(a.elapsed_time(bb) * 1e-3) / 2 * 1e3,
                            "what": "same forward, backward = 7 NCCL reduce_scatter_tensor calls + one bdq_int_ste_bwd per layer"}
        except Exception as e:
            nccl_variant = {"error": f"{type(e).__name__}: {e}"[:300]}
        finally:
            nccl_bwd[0] = False

    # ---- e2e through the public autograd entry with pinned host buffers (shards in, shard results out)
    try:
        e2e = run_e2e_sharded(torch, dist, S, dev, rank, world, nlayers, shapes, spans, min(args.steps, 2))
    except Exception as e:
        e2e = {"error": f"{type(e).__name__}: {e}"[:300]}

    bytes_w = weights * 2
    sharded = {
        "fwd_ms": fwd_ms, "bwd_ms": bwd_ms, "decode_ms": dec_ms, "encode_plus_barrier_ms": fwd_ms - dec_ms,
        "backward": ("fused: bdq_rows_push (NVLink peer stores into the owners' symmetric-memory inboxes) + 1 barrier + "
                     "bdq_int_ste_bwd_nsrc per tensor (gradient summed on the fly, never materialised)") if red is not None else
                    "NCCL reduce_scatter_tensor per tensor + bdq_int_ste_bwd per layer",
        "reduce_scatter_ms": rs_ms if red is None else None, "ste_bwd_shard_ms": ste_ms if red is None else None,
        "nccl_reduce_scatter_variant": nccl_variant,
        "nvlink_bytes_per_rank_per_step": {
            "payload_peer_stores_out": payload_pw * weights / world * (world - 1),
            "reduce_scatter_out_min": bytes_w * (world - 1) / world,
            "reference_order_all_gather_in": bytes_w * (world - 1) / world},
        "limiter": "NVLink: every rank must ship (N-1)/N of its 12.95 GB bf16 weight gradient per step (inherent to data-parallel "
                   "training and identical in the reference's order); of the forward, the full-weight decode every rank must run "
                   "(2.28 B/weight, HBM-bound) -- see fwd_ms / bwd_ms",
        "reference_order": ref_order}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(world), "roofline": roofline,
            "e2e": e2e, "gpu_launches": int(launches) * world, "clocks": clocks, "verified": verified,
            "verified_how": f"decode(all ranks' payload) bit-equal to fake-quant(all-gather(bf16 shards)) on the last layer: {ok_gather}; "
                            f"fwd {msg_f}; shard bwd {msg_b}",
            "sharded": sharded, "semantics": "cuda" if sem else "cpu"}
    if nlayers != LAYERS:
        line["config"]["DEV_ONLY_reduced_layers"] = nlayers
    if rank == 0:
        print(json.dumps(line), flush=True)


def run_e2e_sharded(torch, dist, S, dev, rank, world, nlayers, shapes, spans, steps):
    """N > 1 end to end through the public autograd entry `sharded.quantized_all_gather_group` with pinned HOST buffers of
    one decoder layer reused for all layers: H2D(weight shards, this rank's full upstream gradients) -> forward (full
    fake-quantised weights on every rank) -> backward (reduce-scatter + shard STE) -> D2H(this rank's rows of dq, its
    gradient shards).  All copies inside the timed region; ranks move in lockstep (the path has collectives)."""
    hW = [torch.empty(b - a, i, dtype=torch.bfloat16).normal_(0, 0.02).pin_memory() for (a, b), (_, i) in zip(spans, shapes)]
    hG = [torch.empty(o, i, dtype=torch.bfloat16).normal_(0, 1).pin_memory() for o, i in shapes]
    hDQ = [torch.empty_like(t).pin_memory() for t in hW]
    hGW = [torch.empty_like(t).pin_memory() for t in hW]
    grp = S.ShardGatherGroup(shapes, GROUP, torch.bfloat16, dev, symmetric=True)
    grp.enable_fused_backward()
    dW = [torch.empty_like(t, device=dev) for t in hW]
    dG = [torch.empty_like(t, device=dev) for t in hG]
    t_copy = [0.0]

    def one_step():
        for li in range(nlayers):
            c0 = time.perf_counter()
            for k in range(len(hW)):
                dW[k].copy_(hW[k], non_blocking=True)
                dG[k].copy_(hG[k], non_blocking=True)
            shards = [w.requires_grad_(True) for w in dW]
            outs = S.quantized_all_gather_group(shards, grp)
            torch.autograd.backward(outs, dG)
            for k, ((a, b), w) in enumerate(zip(spans, shards)):
                hDQ[k].copy_(outs[k].detach()[a:b], non_blocking=True)
                hGW[k].copy_(w.grad, non_blocking=True)
                w.grad = None
                w.requires_grad_(False)
            t_copy[0] += time.perf_counter() - c0

    one_step()
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        one_step()
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_s = float(t.item()) / steps
    h2d = (sum(t.numel() for t in hW) + sum(t.numel() for t in hG)) * 2 * nlayers
    d2h = 2 * sum(t.numel() for t in hW) * 2 * nlayers
    return {"value": WEIGHTS_PER_LAYER * nlayers * (FWD_BYTES + BWD_BYTES) / step_s / 1e9, "unit": UNIT,
            "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world, "ms_per_step": step_s * 1e3, "steps": steps,
            "host_copy_gbs_per_rank": (h2d + d2h) / step_s / 1e9,
            "api": "bitdistiller_b200.sharded.quantized_all_gather_group (autograd) over ShardGatherGroup",
            "note": "every rank copies its weight shards AND its own batch's full upstream gradient (12.95 GB per step) from "
                    "pinned host memory; all ranks share one host's memory / PCIe complex, which bounds this number"}


def secondary_metrics(torch, dev, peak):
    """BASELINE configs[3] and configs[4] at N=1: seconds for the int2-g128 clip search over all Llama-2-7B linears
    (tools/bench_multi.py) and tok/s of the Qwen3-8B packed int2 decode GEMV chain (tools/bench_gemv.py: CUDA graph +
    PDL, 4 fused launches per layer, dependent inputs, 1.859 GB of packed weights per token = L2-cold)."""
    from tools import bench_gemv, bench_multi, bench_nf3, gemv_big
    out = {}
    nf = {r["kernel"]: r for r in bench_nf3.run(dev, peak)}
    out["nf3_g128_llama2_7b_layer"] = {
        "what": "configs[2]: one decoder layer's flat bf16 buffer (202.4 M weights) per launch, 3 rotating buffers",
        "fwd_with_clip_cache_gbs": nf["nf3 fwd + clip (config 3)"]["gbs"],
        "fwd_with_clip_cache_frac_of_hbm_peak": nf["nf3 fwd + clip (config 3)"]["frac_of_hbm_peak"],
        "fwd_gbs": nf["nf3 fwd"]["gbs"], "fwd_frac_of_hbm_peak": nf["nf3 fwd"]["frac_of_hbm_peak"],
        "ste_bwd_gbs": nf["nf3 ste bwd"]["gbs"], "ste_bwd_frac_of_hbm_peak": nf["nf3 ste bwd"]["frac_of_hbm_peak"]}
    try:
        n64 = {r["kernel"]: r for r in bench_nf3.run(dev, peak, 64)}
        out["group64_llama2_7b_layer"] = {
            "what": "the same launch shape at q_group_size 64 (the group size of the reference's train scripts, scripts/train/train.sh:52)",
            **{k.replace(" ", "_").replace("+", "and").replace("(", "").replace(")", "") + "_frac_of_hbm_peak": v["frac_of_hbm_peak"]
               for k, v in n64.items()}}
    except Exception as e:
        out["group64_llama2_7b_layer"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    ac = bench_multi.bench_autoclip(dev, 0, 1)
    out["autoclip_7b_int2_g128"] = {"seconds": ac["seconds_total"], "n_gpus": 1, "pairs": ac["pairs"], "candidates": 100,
                                    "tokens": 512}
    ops = bench_gemv.build(dev, batch=1, chain=True)
    nbytes = bench_gemv.bytes_per_token(ops)
    t_graph = bench_gemv.graph_time(ops, 1, True, dev, True)
    t_chain = bench_gemv.chain_time(ops, 0)
    t = min(t_graph, t_chain)
    out["gemv_int2_decode_qwen3_8b"] = {"tok_s": 1.0 / t, "us_per_token": t * 1e6, "bytes_per_token": nbytes,
                                        "frac_of_hbm_peak": nbytes / t / 1e9 / peak,
                                        "persistent_chain_us_per_token": t_chain * 1e6,
                                        "graph_pdl_us_per_token": t_graph * 1e6,
                                        "mode": "best of: CUDA graph + PDL with 144 dependent launches | ONE persistent launch per token "
                                                "(bdq_gemv_w2_chain: 144 dependent phases separated by grid barriers on an L2 counter)"}
    big = gemv_big.run(dev, peak, 1, shapes=((151936, 4096),))[0]
    out["gemv_int2_kernel_streaming"] = {"shape": [big["OC"], big["IC"]], "mb": big["MB"], "us_per_launch": big["us_per_launch"],
                                         "gbs": big["gbs"], "frac_of_hbm_peak": big["frac_of_hbm_peak"],
                                         "what": "the same batch-1 kernel on one lm_head-sized launch (cold, distinct copies)"}
    try:
        from tools import microbench
        out["config0_fp32_4096x4096_int2_g128"] = microbench.config0(dev, peak)
    except Exception as e:
        out["config0_fp32_4096x4096_int2_g128"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    try:
        from tools import bench_nf3_gemv
        r = bench_nf3_gemv.run(dev, peak, shapes=[(151936, 4096)])[0]
        out["gemv_nf3_kernel_streaming"] = {"shape": r["shape"], "mb": r["mb"], "us_per_launch": r["us"], "gbs": r["gbs"],
                                            "frac_of_hbm_peak": r["frac_of_hbm_peak"], "bytes_per_weight": 0.40625,
                                            "what": "packed NF3 (3 bits + two fp16 scales per group of 128) batch-1 GEMV on one lm_head-sized "
                                                    "launch, L2-cold distinct copies, CUDA graph: PRMT level lookup + INT8 mma.sync "
                                                    "(gemv_nf3_stream.cu); the reference has no packed NF3 kernel"}
    except Exception as e:
        out["gemv_nf3_kernel_streaming"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    try:
        # the reference's own kernels on this GPU (baseline/_ref, shipped by baseline/build_ref.py): the perf bars of SURVEY 8d 5a/5b
        from tools import bench_vs_reference
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        vr = bench_vs_reference.run(dev, peak, float(pk.get("bf16_tflops_sustained", 1416.1)))
        out["gemm_prefill_qwen3_8b"] = vr["gemm_prefill_qwen3_8b"]
        out["gemv_int2_decode_qwen3_8b"]["vs_ref_kernel"] = vr["gemv_decode_qwen3_8b"]["vs_ref_kernel"]
        out["gemv_int2_decode_qwen3_8b"]["reference_kernel"] = vr["gemv_decode_qwen3_8b"]
    except Exception as e:
        out["vs_reference_error"] = f"{type(e).__name__}: {e}"[:300]
    try:
        from tools import bench_shard_gather
        sg = bench_shard_gather.single(dev)
        out["int2_shard_gather_llama2_7b_layer"] = {
            "what": "SURVEY 8f N4, quantise-before-all-gather kernels on one decoder layer's flat bf16 buffer (202.4 M weights) "
                    "per launch: encode = 2 B read + 0.28 B payload written per weight, decode = the reverse; decode(encode(w)) "
                    "is checked bit-equal to the fake-quantiser in this run",
            "encode_gbs": sg["encode"]["gbs"], "encode_frac_of_hbm_peak": sg["encode"]["gbs"] / peak,
            "decode_gbs": sg["decode"]["gbs"], "decode_frac_of_hbm_peak": sg["decode"]["gbs"] / peak,
            "payload_bytes_per_weight": sg["payload_bytes_per_weight"]}
    except Exception as e:
        out["int2_shard_gather_llama2_7b_layer"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    return out


def run_e2e(torch, dev, world, nlayers, Quantizer, steps):
    """One decoder layer's worth of pinned host buffers is reused for all layers (the copies are real every time):
    H2D(W, gY) -> SteInt2AsymQuantizer forward -> autograd backward -> D2H(dq, gW), pipelined over 3 streams."""
    import torch.distributed as dist
    quant = Quantizer(GROUP)
    hW = [torch.empty(o, i, dtype=torch.bfloat16).normal_(0, 0.02).pin_memory() for _, o, i in LAYER_SHAPES]
    hG = [torch.empty(o, i, dtype=torch.bfloat16).normal_(0, 1).pin_memory() for _, o, i in LAYER_SHAPES]
    hDQ = [torch.empty(o, i, dtype=torch.bfloat16).pin_memory() for _, o, i in LAYER_SHAPES]
    hGW = [torch.empty(o, i, dtype=torch.bfloat16).pin_memory() for _, o, i in LAYER_SHAPES]
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    main = torch.cuda.current_stream(dev)
    nbuf = 3
    dW = [[torch.empty_like(t, device=dev) for t in hW] for _ in range(nbuf)]
    dG = [[torch.empty_like(t, device=dev) for t in hG] for _ in range(nbuf)]
    free_ev = [[None] * len(hW) for _ in range(nbuf)]

    def one_step():
        it = 0
        for li in range(nlayers):
            b = li % nbuf
            for k in range(len(hW)):
                with torch.cuda.stream(s_in):
                    if free_ev[b][k] is not None:
                        s_in.wait_event(free_ev[b][k])
                    dW[b][k].copy_(hW[k], non_blocking=True)
                    dG[b][k].copy_(hG[k], non_blocking=True)
                    ready = torch.cuda.Event()
                    ready.record(s_in)
                main.wait_event(ready)
                x = dW[b][k].requires_grad_(True)
                y = quant(x)
                y.backward(dG[b][k])
                done = torch.cuda.Event()
                done.record(main)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(done)
                    hDQ[k].copy_(y.detach(), non_blocking=True)
                    hGW[k].copy_(x.grad, non_blocking=True)
                    fe = torch.cuda.Event()
                    fe.record(s_out)
                y.record_stream(s_out)
                x.grad.record_stream(s_out)
                free_ev[b][k] = fe
                x.grad = None
                dW[b][k].requires_grad_(False)
                it += 1
        return it

    one_step()  # warm-up
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        one_step()
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_s = float(t.item()) / steps
    nbytes = WEIGHTS_PER_LAYER * nlayers * 2
    return {"value": WEIGHTS_PER_LAYER * nlayers * (FWD_BYTES + BWD_BYTES) * world / step_s / 1e9, "unit": UNIT,
            "h2d_bytes_per_step": 2 * nbytes, "d2h_bytes_per_step": 2 * nbytes, "ms_per_step": step_s * 1e3,
            "steps": steps, "api": "bitdistiller_b200.quantizer.SteInt2AsymQuantizer(128) forward + autograd backward",
            "note": "pinned host buffers of one decoder layer reused for all layers; bound by PCIe, not by the kernels"}


if __name__ == "__main__":
    main()
