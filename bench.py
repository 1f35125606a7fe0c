#!/usr/bin/env python
"""bench.py -- headline benchmark of the low-bit weight-quantisation hot path (BASELINE.json).

Workload (N = 1 and every N): BASELINE.json configs[1] -- one QAT step's worth of fake-quant forward
(SteInt2AsymQuantizer, int2, group 128) + straight-through backward over ALL converted linears of a random-init
Llama-2-7B (224 bf16 tensors, 6.476 G weights).  Every rank holds and processes the full set, exactly like the
reference under DeepSpeed ZeRO-2 (each rank fake-quantises the full weights every step) -> "scaling": "weak".

  value   = algorithmic bytes (10 B/weight: fwd 2R+2W, bwd 2R+2R+2W) * weights * n_gpus / step time, inputs resident in HBM
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
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(n):
    return {"workload": "configs[1]: Llama-2-7B random-init, all 224 converted linears, int2-g128 "
                        "SteInt2AsymQuantizer forward + STE backward per QAT step (bf16)",
            "weights_per_rank": WEIGHTS_PER_LAYER * LAYERS, "group": GROUP, "bits": BITS,
            "parallelism": f"replicated x{n} (ZeRO-2 style: every rank fake-quantises the full weights)",
            "launches": "one forward + one backward launch per decoder layer over the layer's flat parameter buffer "
                        "(7 tensors, 202.4 M weights)",
            "l2": "inputs are 25.9 GB per step per rank, >> 126 MB L2 (no flush needed)"}


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
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            traffic = json.load(f).get("int_ste_bwd_kernel_bytes_per_launch")
    roofline = {"bound": "hbm", "kernel": "int_ste_bwd_kernel<bf16,128>", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_weight": BWD_BYTES, "bwd_share_of_step": bwd_t / elapsed,
                "launches_timed": len(bwd_events)}

    # ---- e2e: Python drop-in API, pinned host buffers in and out, copies inside the timed region
    e2e = run_e2e(torch, dev, world, nlayers, SteInt2AsymQuantizer, min(args.steps, 3))

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(world), "roofline": roofline,
            "e2e": e2e, "gpu_launches": int(launches) * world, "clocks": clocks,
            "frac_of_hbm_peak_per_gpu": value / world / peak}
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
    ac = bench_multi.bench_autoclip(dev, 0, 1)
    out["autoclip_7b_int2_g128"] = {"seconds": ac["seconds_total"], "n_gpus": 1, "pairs": ac["pairs"], "candidates": 100,
                                    "tokens": 512}
    ops = bench_gemv.build(dev, batch=1, chain=True)
    nbytes = bench_gemv.bytes_per_token(ops)
    t = bench_gemv.graph_time(ops, 1, True, dev, True)
    out["gemv_int2_decode_qwen3_8b"] = {"tok_s": 1.0 / t, "us_per_token": t * 1e6, "bytes_per_token": nbytes,
                                        "frac_of_hbm_peak": nbytes / t / 1e9 / peak,
                                        "mode": "CUDA graph + PDL, fused q/k/v and gate/up launches, L2 prefetch hint; "
                                                "144 dependent launches per token (latency-bound)"}
    big = gemv_big.run(dev, peak, 1, shapes=((151936, 4096),))[0]
    out["gemv_int2_kernel_streaming"] = {"shape": [big["OC"], big["IC"]], "mb": big["MB"], "us_per_launch": big["us_per_launch"],
                                         "gbs": big["gbs"], "frac_of_hbm_peak": big["frac_of_hbm_peak"],
                                         "what": "the same batch-1 kernel on one lm_head-sized launch (cold, distinct copies)"}
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
