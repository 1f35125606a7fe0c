"""The reference's OWN kernels timed on this GPU next to ours ("the kernel to beat"):

  decode : gemv_kernel_g128_2bit (inference/kernels/csrc/quantization/gemv_cuda.cu:202-272, compiled for sm_100 into
           baseline/_ref/awq_inference_engine.so) on all 36 x 7 Qwen3-8B linears, one launch per linear as
           WQLinear.forward issues them (qmodule.py:166), eager and captured in a CUDA graph -- vs bdq_gemv_w2 (CUDA graph + PDL,
           fused q/k/v and gate/up launches), same packed tensors, distinct weight memory per layer (1.86 GB/token, L2-cold)
  prefill: Triton quant_matmul_v2 (quantization/triton_kernels.py:14-134, the reference's only 2-bit GEMM) at M = 2048 on the
           Qwen3-8B shapes: the kernel alone, and as WQLinear.forward calls it with its three .T.contiguous() copies per call
           (qmodule.py:152-165) -- vs bdq_gemm_w2 (tcgen05)

  python tools/bench_vs_reference.py            # prints a table + one JSON line
Same CUDA-event method for both sides (median of per-iteration events after warm-up)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools import bench_gemv  # noqa: E402


def _time(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2] * 1e-3


def _time_graph(fn, reps=10, iters=10):
    """GPU time of one call: `reps` calls captured in a CUDA graph (no Python / launch gaps between them: prefill GEMMs of
    20-50 us are shorter than an eager launch from Python), median over `iters` replays."""
    for _ in range(3):
        fn()            # warm-up; Triton autotunes here
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    return _time(g.replay, iters=iters, warm=2) / reps


def decode(dev, ref, peak):
    ops = bench_gemv.build(dev, batch=1, chain=True)
    nbytes = bench_gemv.bytes_per_token(ops)
    eng = ref["engine"]

    def ref_token():
        for x, qw, qs, qz, out, b, ic, oc in ops:
            eng.gemv_forward_cuda(x, qw, qs, qz, 2, 128)     # allocates its output inside, like the reference flow

    # the reference launches on the legacy default stream (gemv_cuda.cu:318-341 `<<<num_blocks, num_threads>>>`), so it cannot
    # be captured in a CUDA graph; its kernels run 10-25 us each, longer than a launch, so back-to-back eager launches are
    # GPU-bound and the event time is the kernels' time
    t_ref = _time(ref_token, iters=10)
    t_ours = bench_gemv.graph_time(ops, 1, True, dev, True)
    # parity of the two on the first layer's tensors (static inputs)
    x, qw, qs, qz, out, b, ic, oc = ops[3]
    xs = torch.randn(1, ic, device=dev).half()
    from bitdistiller_b200 import awq_inference_engine as E
    a, r = E.gemv_forward_cuda(xs, qw, qs, qz, 2, 128).float(), eng.gemv_forward_cuda(xs, qw, qs, qz, 2, 128).float()
    rel = float((a - r).norm() / r.norm())
    return {"bytes_per_token": nbytes, "ref_kernel_us_per_token": t_ref * 1e6,
            "ref_kernel_frac_of_hbm_peak": nbytes / t_ref / 1e9 / peak,
            "ours_us_per_token": t_ours * 1e6, "ours_frac_of_hbm_peak": nbytes / t_ours / 1e9 / peak,
            "vs_ref_kernel": t_ref / t_ours, "rel_frobenius_diff_o_proj": rel,
            "what": "252 back-to-back launches of the reference's gemv_kernel_g128_2bit per token (eager: it launches on the "
                    "legacy default stream and cannot be graph-captured; GPU-bound at ~18 us per launch) vs bdq_gemv_w2 (CUDA "
                    "graph + PDL, 144 launches), same packed tensors, L2-cold"}


def prefill(dev, ref, bf16_peak, M=2048):
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.qmodule import pack_wq
    from bitdistiller_b200.triton_kernels import gemm_w2
    tk = ref["triton_kernels"]
    rows, tot = {}, {"ours": 0.0, "triton": 0.0, "triton_module": 0.0, "flop": 0.0}
    for name, oc, ic, mult in [("q/o_proj", 4096, 4096, 2), ("k/v_proj", 1024, 4096, 2), ("gate/up_proj", 12288, 4096, 2),
                               ("down_proj", 4096, 12288, 1)]:
        w = (torch.randn(oc, ic, device=dev) * 0.02).half()
        dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
        qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
        x = torch.randn(M, ic, device=dev).half()
        kw = dict(M=M, N=oc, K=ic, pack_num=16, group_size=128, w_bit=2, offset=3)
        qwT, qzT, qsT = qw.T.contiguous(), qz.T.contiguous(), qs.T.contiguous()
        t_tr = _time_graph(lambda: tk.quant_matmul_v2(x, qwT, qzT, qsT, **kw), reps=4, iters=5)
        t_mod = _time_graph(lambda: tk.quant_matmul_v2(x, qw.T.contiguous(), qz.T.contiguous(), qs.T.contiguous(), **kw), reps=4, iters=5)
        t_us = _time_graph(lambda: gemm_w2(x, qw, qs, qz, 128))
        t_eager = _time(lambda: gemm_w2(x, qw, qs, qz, 128), iters=10)
        r, o = tk.quant_matmul_v2(x, qwT, qzT, qsT, **kw).float(), gemm_w2(x, qw, qs, qz, 128).float()
        f = 2.0 * M * oc * ic
        rows[name] = {"shape": [oc, ic], "ours_eager_launch_us": t_eager * 1e6, "ours_us": t_us * 1e6, "ours_tflops": f / t_us / 1e12, "triton_ref_us": t_tr * 1e6,
                      "triton_ref_tflops": f / t_tr / 1e12, "triton_ref_as_called_us": t_mod * 1e6, "vs_triton_ref": t_tr / t_us,
                      "rel_frobenius_diff": float((o - r).norm() / r.norm())}
        tot["ours"] += t_us * mult
        tot["triton"] += t_tr * mult
        tot["triton_module"] += t_mod * mult
        tot["flop"] += f * mult
    # q/k/v as ONE fused WQLinear (packed buffers concatenated along out_features, inference/modules/fused_attn.py:312-337):
    # the configuration the reference's own fused attention module runs
    oc, ic = 6144, 4096
    w = (torch.randn(oc, ic, device=dev) * 0.02).half()
    dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
    qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
    x = torch.randn(M, ic, device=dev).half()
    t_qkv = _time_graph(lambda: gemm_w2(x, qw, qs, qz, 128))
    f_qkv = 2.0 * M * oc * ic
    t_fused = tot["ours"] - rows["q/o_proj"]["ours_us"] * 1e-6 - 2 * rows["k/v_proj"]["ours_us"] * 1e-6 + t_qkv
    fused = {"qkv_us": t_qkv * 1e6, "qkv_tflops": f_qkv / t_qkv / 1e12, "qkv_frac": f_qkv / t_qkv / 1e12 / bf16_peak,
             "ms": t_fused * 36 * 1e3, "tflops": tot["flop"] / t_fused / 1e12, "frac": tot["flop"] / t_fused / 1e12 / bf16_peak,
             "what": "the same model with q/k/v as one fused [6144, 4096] WQLinear (fused_attn.py:312-337): 4 GEMMs per layer"}
    L = 36
    return {"M": M, "fused_qkv": fused, "ms": tot["ours"] * L * 1e3, "tflops": tot["flop"] / tot["ours"] / 1e12,
            "frac": tot["flop"] / tot["ours"] / 1e12 / bf16_peak, "triton_ref_ms": tot["triton"] * L * 1e3,
            "triton_ref_as_called_ms": tot["triton_module"] * L * 1e3, "vs_triton_ref": tot["triton"] / tot["ours"],
            "vs_triton_ref_as_called": tot["triton_module"] / tot["ours"], "per_shape": rows,
            "timing": "every kernel (ours and the reference's Triton kernel) timed as 4-10 back-to-back calls captured in one CUDA graph, "
                      "median of the replays: GPU time without Python launch gaps; ours_eager_launch_us = the same call issued eagerly",
            "what": "all 36 x 7 Qwen3-8B linears at M = 2048, q/k/v separate; `frac` = of MEASURED_PEAKS bf16_tflops_sustained; "
                    "triton_ref = quant_matmul_v2 kernel alone on pre-transposed operands, as_called = with WQLinear.forward's "
                    "three .T.contiguous() copies per call"}


def run(dev, hbm_peak, bf16_peak):
    from baseline.ref_loader import load
    ref = load()
    return {"gemv_decode_qwen3_8b": decode(dev, ref, hbm_peak), "gemm_prefill_qwen3_8b": prefill(dev, ref, bf16_peak)}


if __name__ == "__main__":
    pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    res = run(torch.device("cuda:0"), pk.get("hbm_gbs", 6462.4), pk.get("bf16_tflops_sustained", 1416.1))
    d, p = res["gemv_decode_qwen3_8b"], res["gemm_prefill_qwen3_8b"]
    print(f"decode  : reference kernel {d['ref_kernel_us_per_token']:.0f} us/token ({d['ref_kernel_frac_of_hbm_peak']*100:.1f}% of HBM peak)  "
          f"ours {d['ours_us_per_token']:.0f} us/token ({d['ours_frac_of_hbm_peak']*100:.1f}%)  -> {d['vs_ref_kernel']:.2f}x")
    for k, r in p["per_shape"].items():
        print(f"prefill {k:13s} ours {r['ours_us']:8.1f} us {r['ours_tflops']:7.1f} TF/s | triton ref {r['triton_ref_us']:9.1f} us "
              f"{r['triton_ref_tflops']:6.1f} TF/s (as called {r['triton_ref_as_called_us']:9.1f} us) -> {r['vs_triton_ref']:.1f}x  rel diff {r['rel_frobenius_diff']:.2e}")
    fq = p["fused_qkv"]
    print(f"prefill fused q/k/v [6144 x 4096]: {fq['qkv_us']:.1f} us {fq['qkv_tflops']:.0f} TF/s ({fq['qkv_frac']*100:.1f}%); model with fused q/k/v: "
          f"{fq['ms']:.2f} ms {fq['tflops']:.0f} TF/s ({fq['frac']*100:.1f}%)")
    print(f"prefill model: ours {p['ms']:.2f} ms ({p['tflops']:.0f} TF/s, {p['frac']*100:.1f}%)  triton ref {p['triton_ref_ms']:.1f} ms -> {p['vs_triton_ref']:.1f}x")
    print(json.dumps(res))
