"""Times the tcgen05 GEMM on the L-model shapes next to cuBLAS (torch.matmul, bf16). TSB200_GEMM_PAIR=0/1 selects one CTA per
tile or CTA pairs; TSB_FORCE_BN overrides the tile width. python scratch/gemm_pair.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchsystem_b200 import _lib
SHAPES = [('cube (the shape of MEASURED_PEAKS.json bf16_tflops)', 0, 0, 8192, 8192, 8192), ('fwd', 0, 0, 8192, 4096, 4096), ('dX', 0, 1, 8192, 4096, 4096), ('dW', 1, 1, 4096, 4096, 8192), ('fwd1', 0, 0, 8192, 4096, 784)]
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for name, am, bm, M, N, K in SHAPES:
    A = torch.randn((M, K) if am == 0 else (K, M), device='cuda').bfloat16()
    B = torch.randn((N, K) if bm == 0 else (K, N), device='cuda').bfloat16()
    out = torch.empty(M, N, dtype=torch.float32, device='cuda')
    def mine():
        rc = _lib.lib.ts_test_gemm_bf16(A.data_ptr(), am, A.stride(0), B.data_ptr(), bm, B.stride(0), M, N, K, out.data_ptr(), N, 1, M * N, None, st)
        assert rc == 0
    Al = A if am == 0 else A.t()
    Bl = B.t() if bm == 0 else B
    def blas(): return Al @ Bl
    res = {}
    for label, fn in (('mine', mine), ('cublas', blas)):
        for _ in range(3): fn()
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort(); res[label] = ts[len(ts) // 2]
    ref = (Al.float() @ Bl.float())
    err = ((out - ref).abs().max() / ref.abs().max()).item()
    fl = 2.0 * M * N * K
    print(f'{name}: {M}x{N}x{K} mine {res["mine"]*1e3:.1f} us = {fl/res["mine"]/1e9:.0f} TF/s (fp32 out), cuBLAS bf16-out {res["cublas"]*1e3:.1f} us = {fl/res["cublas"]/1e9:.0f} TF/s, '
          f'ratio {res["cublas"]/res["mine"]:.3f}, rel err {err:.2e}, pair={os.environ.get("TSB200_GEMM_PAIR","1")}', flush=True)
