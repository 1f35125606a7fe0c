import torch, sys
sys.path.insert(0, '.')
from bitdistiller_b200.nf3_packed import gemm_f16, WQLinearNF3
dev = torch.device("cuda:0")
def t(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)/it*1e-3
M=2048
for oc,ic in [(4096,4096),(12288,4096),(4096,12288),(6144,4096)]:
    w=(torch.randn(oc,ic,device=dev)*0.02).half(); x=torch.randn(M,ic,device=dev).half()
    f=2.0*M*oc*ic
    t1=t(lambda: gemm_f16(x,w)); t2=t(lambda: torch.matmul(x,w.T))
    lin=torch.nn.Linear(ic,oc,bias=False,dtype=torch.float16,device=dev); lin.weight.data=w
    m=WQLinearNF3.from_linear(lin,128); t3=t(lambda: m(x))
    print(f"[{oc}x{ic}] M={M}: bdq_gemm_f16 {t1*1e6:7.1f} us {f/t1/1e12:7.1f} TF/s | cuBLAS {t2*1e6:7.1f} us {f/t2/1e12:7.1f} TF/s | WQLinearNF3.forward (dequant + gemm_f16) {t3*1e6:7.1f} us {f/t3/1e12:7.1f} TF/s")
