import sys, torch
sys.path.insert(0, '/root/repo')
from torchsystem_b200 import _lib
def run(a_major, b_major, M, N, K, splits=1, reps=3, flag=0):
    g = torch.Generator(device='cuda').manual_seed(1)
    A_log = torch.randn(M, K, generator=g, device='cuda').bfloat16(); B_log = torch.randn(K, N, generator=g, device='cuda').bfloat16()
    A = A_log.contiguous() if a_major == 0 else A_log.t().contiguous()
    B = B_log.t().contiguous() if b_major == 0 else B_log.contiguous()
    out = torch.zeros(splits, M, N, device='cuda')
    st = torch.cuda.current_stream().cuda_stream
    for r in range(reps):
        dbg = torch.zeros(64, dtype=torch.int64, device='cuda'); dbg[60] = 2**62; dbg[50] = flag
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(_lib.lib.ts_test_gemm_bf16(A.data_ptr(), a_major, A.stride(0), B.data_ptr(), b_major, B.stride(0), M, N, K, out.data_ptr(), N, splits, M*N, dbg.data_ptr(), st))
        e1.record(); torch.cuda.synchronize()
    d = dbg.cpu().tolist()
    # back-to-back launches: steady-state per-kernel time
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(20):
        _lib.lib.ts_test_gemm_bf16(A.data_ptr(), a_major, A.stride(0), B.data_ptr(), b_major, B.stride(0), M, N, K, out.data_ptr(), N, splits, M*N, None, st)
    e1.record(); torch.cuda.synchronize()
    print(f'flag {flag} --- majors {a_major}{b_major} M{M} N{N} K{K} s{splits}: single event {e0.elapsed_time(e1)*0:.0f} grid span {(d[61]-d[60])/1e3:.2f} us, start skew {(d[62]-d[60])/1e3:.2f} us, cta0 cycles {d[7]-d[0]}, slowest CTA {d[63]/1e3:.1f} us, mean CTA {d[55]/1e3/(((M+127)//128)*max(1,(N+255)//256)*splits):.1f} us (if BN=256), back-to-back {e0.elapsed_time(e1)/20*1e3:.2f} us/launch')
import os
for bn in ('256','192','128'):
    os.environ['TSB_FORCE_BN'] = bn
    print('BN', bn)
    run(0, 0, 8192, 4096, 784)
    run(0, 0, 32768, 256, 784)
    run(0, 0, 4096, 4096, 4096)
