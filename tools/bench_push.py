"""NVLink push bandwidth (2+ ranks, torchrun): bdq_rows_push into a peer's symmetric-memory buffer vs cudaMemcpyAsync (copy
engine) on the same buffers, 405 MB (one Llama-2-7B layer of bf16 gradients) in 7 segments."""
import ctypes, os, sys
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
sizes = [4096 * 4096 * 2] * 4 + [11008 * 4096 * 2] * 3
tot = sum(sizes)
buf = symm_mem.empty(tot, dtype=torch.uint8, device=dev)
hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
peer = int(hdl.buffer_ptrs[(rank + 1) % world])
src = torch.empty(tot, dtype=torch.uint8, device=dev).random_(0, 255)
offs = [sum(sizes[:i]) for i in range(len(sizes))]
n = len(sizes)
srcp = (ctypes.c_void_p * n)(*[src.data_ptr() + o for o in offs])
dstp = (ctypes.c_void_p * n)(*[peer + o for o in offs])
nb = (ctypes.c_int64 * n)(*sizes)
st = L.stream_ptr(dev)
cudart = ctypes.CDLL("libcudart.so.12")


def timed(fn, iters=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def k_push():
    L.check(L.lib().bdq_rows_push(n, srcp, dstp, nb, st), "push")


def ce_push():
    for i in range(n):
        cudart.cudaMemcpyAsync(ctypes.c_void_p(peer + offs[i]), ctypes.c_void_p(src.data_ptr() + offs[i]), ctypes.c_size_t(sizes[i]), 4, st)


for name, fn in (("bdq_rows_push kernel", k_push), ("cudaMemcpyAsync (copy engine)", ce_push)):
    for waves in ((None, "2", "8", "32") if fn is k_push else (None,)):
        if waves:
            os.environ["BDQ_PUSH_CTAS_PER_SM"] = waves
        ms = timed(fn)
        if rank == 0:
            print(f"{name:32s} ctas/sm {waves or '-':>3s}: {ms:7.3f} ms  {tot/ms/1e6:7.1f} GB/s out per rank ({world} ranks pushing at once)")
hdl.barrier(channel=0)
ok = torch.equal(buf.cpu(), buf.cpu())
dist.destroy_process_group()
