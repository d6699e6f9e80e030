import sys, torch, os
sys.path.insert(0, '/root/repo')
from torchsystem_b200 import _lib
def run(a_major, b_major, M, N, K, splits=1, reps=2):
    g = torch.Generator(device='cuda').manual_seed(1)
    A_log = torch.randn(M, K, generator=g, device='cuda').bfloat16(); B_log = torch.randn(K, N, generator=g, device='cuda').bfloat16()
    A = A_log.contiguous() if a_major == 0 else A_log.t().contiguous()
    B = B_log.t().contiguous() if b_major == 0 else B_log.contiguous()
    out = torch.zeros(splits, M, N, device='cuda')
    st = torch.cuda.current_stream().cuda_stream
    for r in range(reps):
        dbg = torch.zeros(64, dtype=torch.int64, device='cuda'); dbg[60] = 2**62
        _lib.check(_lib.lib.ts_test_gemm_bf16(A.data_ptr(), a_major, A.stride(0), B.data_ptr(), b_major, B.stride(0), M, N, K, out.data_ptr(), N, splits, M*N, dbg.data_ptr(), st))
        torch.cuda.synchronize()
    d = dbg.cpu().tolist(); t0 = d[0]
    print(f'--- majors {a_major}{b_major} M{M} N{N} K{K} BN={os.environ.get("TSB_FORCE_BN")}: span {(d[61]-d[60])/1e3:.1f} us slowest CTA {d[63]/1e3:.1f} us')
    print('   setup', d[1]-t0, 'tmem_full', d[6]-t0, 'epi_done', d[7]-t0)
    print('   producerA issue:', [x - t0 for x in d[40:48] if x])
    print('   mma pass full  :', [x - t0 for x in d[8:24] if x])
    print('   mma after issue:', [x - t0 for x in d[24:40] if x])
run(0, 0, 4096, 256, 784)
run(1, 1, 256, 784, 4096, 8)
run(0, 1, 4096, 256, 256)
run(0, 0, 4096, 4096, 4096)
run(0, 0, 8192, 4096, 4096)
run(1, 1, 4096, 4096, 8192)
run(0, 1, 8192, 4096, 4096)
