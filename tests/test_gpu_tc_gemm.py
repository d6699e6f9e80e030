"""The tcgen05/TMA GEMM on its own (C ABI test hook ts_test_gemm_bf16) against a plain torch fp32 matmul of the
same bf16 operands: all operand-major combinations a Linear layer needs, ragged M/N/K tails, split-K.
Tolerance: fp32 accumulation of exactly representable bf16 products in a different order -> 1e-4 relative to
the row/column scale (this is a floating-point kernel; the fp32 torch reference is the oracle for it)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def run_gemm(A, a_major, B, b_major, M, N, K, splits=1):
    from torchsystem_b200 import _lib
    dev = A.device
    out = torch.full((splits, M, N), float('nan'), dtype=torch.float32, device=dev)
    rc = _lib.lib.ts_test_gemm_bf16(A.data_ptr(), a_major, A.stride(0), B.data_ptr(), b_major, B.stride(0), M, N, K,
                                   out.data_ptr(), N, splits, M * N, None, torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(rc, 'ts_test_gemm_bf16')
    torch.cuda.synchronize()
    return out.sum(0)


CASES = [
    # (a_major, b_major, M, N, K, splits)
    (0, 0, 128, 64, 64, 1),            # one tile, one k-block
    (0, 0, 4096, 256, 784, 1),         # S-model forward: K tail of 16
    (0, 0, 1000, 256, 784, 1),         # ragged M
    (0, 0, 512, 4096, 4096, 1),        # L-model forward shape class
    (0, 1, 256, 784, 256, 1),          # dX = dY W   (B stored [K][N])
    (0, 1, 4096, 128, 64, 1),
    (1, 1, 256, 784, 4096, 10),        # S-model dW1 = dH^T X, batch split across CTAs
    (1, 1, 256, 784, 1000, 3),         # ragged contraction
    (1, 1, 64, 48, 200, 1),            # small hidden sizes (golden 'deep' case class)
    (1, 0, 128, 128, 128, 1),          # remaining combination
    (0, 0, 130, 24, 40, 1),            # tiny N (multiple of 8), K < 64
    (0, 0, 1000, 16, 512, 1),          # tensor-core head: logits against the [16, K] padded last-layer weights
    (0, 1, 1000, 384, 16, 1),          # tensor-core head: dA = dZ W_last, contraction over the 16 padded classes
    (1, 1, 384, 16, 1000, 3),          # tensor-core head: dW_last^T = A^T dZ, batch split
    # 256-wide tiles with M >= 256: the CTA-pair (cta_group::2) kernel, every operand-major combination, ragged tails
    (0, 0, 2048, 4096, 1024, 1),       # L-model forward
    (0, 1, 2048, 4096, 1024, 1),       # L-model dX
    (1, 1, 4096, 4096, 2048, 1),       # L-model dW
    (1, 0, 2048, 4096, 512, 1),
    (0, 0, 2000, 4000, 1000, 1),       # ragged M (the peer CTA's rows run out), N and K
    (1, 1, 2040, 4000, 1000, 1),
    (0, 0, 300, 37888, 64, 1),         # second CTA of the last pair mostly out of range; many column tiles per pair
    (1, 1, 4096, 4096, 2048, 2),       # pairs with split-K
    (1, 1, 4096, 784, 2048, 1),        # L-model dW1: 4 x 256 columns in one wave of pairs, the last peer half entirely out of range
    (0, 0, 65536, 256, 784, 1),        # tall and one tile wide
]


@pytest.mark.parametrize('a_major,b_major,M,N,K,splits', CASES)
def test_tc_gemm_matches_fp32_matmul(a_major, b_major, M, N, K, splits):
    g = torch.Generator(device='cuda').manual_seed(M * 31 + N * 7 + K)
    A_log = torch.randn(M, K, generator=g, device='cuda').bfloat16()       # logical A[M,K]
    B_log = torch.randn(K, N, generator=g, device='cuda').bfloat16()       # logical B[K,N]
    A = A_log.contiguous() if a_major == 0 else A_log.t().contiguous()       # stored [M][K] or [K][M]
    B = B_log.t().contiguous() if b_major == 0 else B_log.contiguous()       # stored [N][K] or [K][N]
    got = run_gemm(A, a_major, B, b_major, M, N, K, splits)
    want = A_log.float() @ B_log.float()
    err = (got - want).abs().max().item()
    scale = want.abs().max().item()
    assert torch.isfinite(got).all(), 'unwritten or non-finite outputs'
    assert err <= 1e-4 * scale + 1e-3, f'max abs err {err:.4g} vs scale {scale:.4g}'


def test_tc_gemm_identity_picks_out_layout_errors():
    """A = I: the output must equal B exactly, column for column (catches swizzle / descriptor mistakes that a
    random test could only report as 'wrong')."""
    M = N = K = 128
    A = torch.eye(M, device='cuda').bfloat16()
    B_log = (torch.arange(K * N, device='cuda').reshape(K, N) % 251).float().bfloat16()
    for a_major in (0, 1):
        for b_major in (0, 1):
            B = B_log.t().contiguous() if b_major == 0 else B_log.contiguous()
            got = run_gemm(A, a_major, B, b_major, M, N, K)
            assert torch.equal(got, B_log.float()), (a_major, b_major)
