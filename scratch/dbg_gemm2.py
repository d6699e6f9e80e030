import sys, torch
sys.path.insert(0, '/root/repo')
from torchsystem_b200 import _lib
def run(a_major, b_major, M, N, K, splits=1, reps=2):
    g = torch.Generator(device='cuda').manual_seed(1)
    A_log = torch.randn(M, K, generator=g, device='cuda').bfloat16(); B_log = torch.randn(K, N, generator=g, device='cuda').bfloat16()
    A = A_log.contiguous() if a_major == 0 else A_log.t().contiguous()
    B = B_log.t().contiguous() if b_major == 0 else B_log.contiguous()
    out = torch.zeros(splits, M, N, device='cuda'); dbg = torch.zeros(64, dtype=torch.int64, device='cuda')
    st = torch.cuda.current_stream().cuda_stream
    for r in range(reps):
        _lib.check(_lib.lib.ts_test_gemm_bf16(A.data_ptr(), a_major, A.stride(0), B.data_ptr(), b_major, B.stride(0), M, N, K, out.data_ptr(), N, splits, M*N, dbg.data_ptr(), st))
        torch.cuda.synchronize()
    d = dbg.cpu().tolist(); t0 = d[0]
    print(f'--- majors {a_major}{b_major} M{M} N{N} K{K}')
    print('   kblock pass:', [x - t0 for x in d[8:8+8] if x])
    for kb in range(0):
        print('   kb', kb, 'mma issue clocks + before/after commit:', [x - t0 for x in d[32+8*kb:32+8*kb+6]])
for n in (16, 64, 256):
    run(0, 0, 128, n, 1024)
