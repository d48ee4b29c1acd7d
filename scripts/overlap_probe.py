"""Does a small elementwise kernel get placed beside the persistent tcgen05 GEMM CTAs? (probe)"""
import ctypes as C, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from evi_rag_b200 import lib
L = lib.load()
dev = torch.device("cuda", 0)
M = 400_000
x = torch.randn(M, 1024, device=dev).half()
w = torch.randn(1024, 1024, device=dev).half()
bias = torch.zeros(1024, device=dev)
out = torch.empty(M, 1024, device=dev, dtype=torch.half)
src = torch.randn(4 * 1024 * 1024 * 1024 // 4, device=dev)
dst = torch.empty(src.numel(), device=dev, dtype=torch.half)
acts = (C.c_int * 8)(*([0] * 8))
sms = torch.cuda.get_device_properties(dev).multi_processor_count
sa, sb = torch.cuda.Stream(priority=-1), torch.cuda.Stream(priority=0)

def gemm(stream):
    lib.check(L.evi_project(x.data_ptr(), M, w.data_ptr(), 1024, 1024, bias.data_ptr(), acts, out.data_ptr(), None, 0, sms, stream.cuda_stream), "proj")

def conv(stream):
    lib.check(L.evi_convert_f32(src.data_ptr(), dst.data_ptr(), src.numel(), 0, stream.cuda_stream), "conv")

def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)

cur = torch.cuda.current_stream()
for _ in range(2):
    gemm(cur); conv(cur)
print("gemm alone ms", timed(lambda: gemm(cur)))
print("conv alone ms", timed(lambda: conv(cur)))
def both():
    ev = torch.cuda.Event()
    ev.record(cur)
    sa.wait_event(ev); sb.wait_event(ev)
    gemm(sa)
    conv(sb)
    e1, e2 = torch.cuda.Event(), torch.cuda.Event()
    e1.record(sa); e2.record(sb)
    cur.wait_event(e1); cur.wait_event(e2)
print("both on two streams ms", timed(both))
print("both on two streams ms", timed(both))
def both_rev():
    ev = torch.cuda.Event()
    ev.record(cur)
    sa.wait_event(ev); sb.wait_event(ev)
    conv(sb)
    gemm(sa)
    e1, e2 = torch.cuda.Event(), torch.cuda.Event()
    e1.record(sa); e2.record(sb)
    cur.wait_event(e1); cur.wait_event(e2)
print("conv first, two streams ms", timed(both_rev))
print("conv first, two streams ms", timed(both_rev))
