"""The scoring GEMM shape against cuBLAS on the same shape, same process, interleaved (same power / clock state).

Shape: the `W_sc sc` launch of one chunk, [M, 1024] x [1024, 1024]^T, 16-bit operands, fp32 accumulate.
  ours      evi_project -> gemm2_kernel<EpiF32Tma> (fp32 rows out, what the tail kernel consumes)
  cublas16  torch.mm fp16 -> fp16 output (half the output bytes of ours)
  cublas32  torch.mm(..., out_dtype=torch.float32) when this torch has it (same bytes as ours)
and the large square bf16 GEMM MEASURED_PEAKS.json's sustained figure is taken from, for the clock state.
Usage (GPU box): python scripts/gemm_vs_cublas.py [--rows 316000] [--iters 30]
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from evi_rag_b200 import lib as _lib  # noqa: E402


def timed(fn, iters):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=316000)
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--rounds", type=int, default=3)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    M, D = args.rows, 1024
    x = (torch.randn(M, D, device=dev) * 0.5).half()
    w = (torch.randn(D, D, device=dev) * 0.03).half()
    out32 = torch.empty(M, D, dtype=torch.float32, device=dev)
    L = _lib.load()
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    acts = (C.c_int * 8)(*([0] * 8))

    def ours():
        st = L.evi_project(x.data_ptr(), M, None, w.data_ptr(), D, D, D, None, acts, None, out32.data_ptr(),
                           _lib.FMT_F16, sms, _lib.stream_handle())
        _lib.check(st, "evi_project")

    wt = w.t()
    out16 = torch.empty(M, D, dtype=torch.float16, device=dev)

    def cublas16():
        torch.mm(x, wt, out=out16)

    have32 = True
    try:
        torch.mm(x[:256], wt, out_dtype=torch.float32)
    except TypeError:
        have32 = False

    def cublas32():
        torch.mm(x, wt, out_dtype=torch.float32)

    n = 8192
    a_sq = torch.randn(n, n, device=dev, dtype=torch.bfloat16)
    b_sq = torch.randn(n, n, device=dev, dtype=torch.bfloat16)

    def square():
        torch.mm(a_sq, b_sq)

    ours()
    ref = x.float() @ w.float().t()
    err = (out32 - ref).norm() / ref.norm()
    flops = 2.0 * M * D * D
    legs = [("ours_f32out", ours, flops), ("cublas_f16out", cublas16, flops)]
    if have32:
        legs.append(("cublas_f32out", cublas32, flops))
    legs.append(("cublas_square_bf16_8192", square, 2.0 * n ** 3))
    for _, fn, _ in legs:  # warm-up
        for _ in range(5):
            fn()
    torch.cuda.synchronize()
    res = {name: [] for name, _, _ in legs}
    for _ in range(args.rounds):
        for name, fn, fl in legs:
            ms = timed(fn, args.iters)
            res[name].append(fl / (ms * 1e-3) / 1e12)
    print(json.dumps({"rows": M, "rel_err_vs_fp32": float(err),
                      "tflops": {k: [round(v, 1) for v in vs] for k, vs in res.items()}}))


if __name__ == "__main__":
    main()
