import sys, torch
sys.path.insert(0, '/root/repo')
from torchsystem_b200 import _lib
def run(a_major, b_major, M, N, K, splits=1, reps=3):
    g = torch.Generator(device='cuda').manual_seed(1)
    A_log = torch.randn(M, K, generator=g, device='cuda').bfloat16(); B_log = torch.randn(K, N, generator=g, device='cuda').bfloat16()
    A = A_log.contiguous() if a_major == 0 else A_log.t().contiguous()
    B = B_log.t().contiguous() if b_major == 0 else B_log.contiguous()
    out = torch.zeros(splits, M, N, device='cuda'); dbg = torch.zeros(32, dtype=torch.int64, device='cuda')
    st = torch.cuda.current_stream().cuda_stream
    for r in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(_lib.lib.ts_test_gemm_bf16(A.data_ptr(), a_major, A.stride(0), B.data_ptr(), b_major, B.stride(0), M, N, K, out.data_ptr(), N, splits, M*N, dbg.data_ptr(), st))
        e1.record(); torch.cuda.synchronize()
    d = dbg.cpu().tolist(); t0 = d[0]
    print(f'--- majors {a_major}{b_major} M{M} N{N} K{K} splits {splits}: event {e0.elapsed_time(e1)*1e3:.1f} us')
    print('   setup', d[1]-t0, 'first_tma_issued', d[2]-t0, 'all_tma_issued', d[3]-t0, 'mma_done_issue', d[5]-t0, 'tmem_full', d[6]-t0, 'epi_done', d[7]-t0)
    print('   full-barrier pass times per k-block:', [x - t0 for x in d[8:8+16] if x])
run(0, 0, 4096, 256, 784)
run(1, 1, 256, 784, 4096, 10)
pass
pass
print("==== single CTA sweeps")
for n in (16, 32, 64, 128, 256):
    run(0, 0, 128, n, 2048, reps=2)
run(1, 1, 128, 64, 2048, reps=2)
run(0, 1, 128, 64, 2048, reps=2)
print("==== multi CTA same tile shape N=64")
run(0, 0, 128*8, 64, 2048, reps=2)
run(0, 0, 128*64, 64, 2048, reps=2)
run(0, 0, 128*128, 64, 2048, reps=2)
