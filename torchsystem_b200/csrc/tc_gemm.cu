// tcgen05 / TMEM / TMA GEMM for the bf16 precision mode (sm_100a).
//
//   D[M,N] (fp32 accumulate in tensor memory) = A[M,K] * B[K,N],  bf16 operands
//
// One CTA computes one 128 x BN output tile (BN a multiple of 16, <= 256) over a K range (split-K across
// blockIdx.z). Warp roles: warp 0 = TMA producer (one lane), warp 1 = TMEM allocator + tcgen05.mma issuer (one
// lane), warps 2..5 = epilogue (TMEM -> registers -> fused epilogue -> global). Operands are staged by
// cp.async.bulk.tensor (TMA) into 128-byte-swizzled shared-memory tiles behind a ring of mbarriers; the same
// tile layout serves K-major and MN-major operands, which is what lets the three contractions of a Linear layer
// run without any transposed copy:
//   forward  Y  = X  W^T : A = X  [M][K] K-major,   B = W  [N][K] K-major
//   dX          = dY W   : A = dY [M][K] K-major,   B = W  [K][N] MN-major
//   dW          = dY^T X : A = dY [K][M] MN-major,  B = X  [K][N] MN-major   (K = batch, split across CTAs)
#include <cuda.h>

#include <mutex>
#include <vector>

#include "kernels.h"

namespace tsb {

namespace {

constexpr int BM = 128;        // UMMA M (cta_group::1)
constexpr int BK = 64;         // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int kThreads = 192;
constexpr int kMaxStages = 8;
constexpr uint32_t kABytes = BM * BK * 2;   // 16 KiB per stage

struct TcParams {
    int M, N, K, BN;
    int num_kblocks, kblocks_per_split;
    int a_major, b_major;
    uint32_t idesc;
    uint32_t b_bytes;
    int stages;
    uint32_t tmem_cols;
    void *out;
    int64_t ldo, split_stride;
    const float *bias;
    const __nv_bfloat16 *A_prev;
    DropoutSpec drop;
    int training;
};

// ------------------------------------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug traps (kernel error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (uint32_t spin = 0;; ++spin) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) return;
        if (spin > (1u << 24)) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap *map, uint64_t *bar, void *dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols));
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrives on the mbarrier once all tcgen05.mma issued so far by this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor, 128-byte swizzle (cute::UMMA::SmemDescriptor bit layout):
//   [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout=2 (SWIZZLE_128B)
//   K-major : rows of 128 B, 8-row groups SBO = 1024 B apart (LBO unused, 1)
//   MN-major: 64-element MN atoms LBO = BK*128 B apart, 8-row K groups SBO = 1024 B apart
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, int major) {
    const uint64_t lbo = (major == TC_MAJOR_K) ? 1u : (uint32_t)(BK * 128) >> 4;
    const uint64_t sbo = 1024u >> 4;
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (2ull << 61);
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages];
    __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ uint32_t tmem_base_slot;

    uint8_t *tiles = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t stage_bytes = kABytes + p.b_bytes;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * p.BN;
    const int kb_begin = blockIdx.z * p.kblocks_per_split;
    const int kb_end = min(p.num_kblocks, kb_begin + p.kblocks_per_split);

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&tmem_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_b)) : "memory");
    }
    if (warp == 1) tmem_alloc(&tmem_base_slot, p.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0 && lane == 0) {
        // ===================== TMA producer =====================
        int s = 0;
        uint32_t phase = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            mbar_wait(&empty_bar[s], phase ^ 1);
            uint8_t *sa = tiles + (size_t)s * stage_bytes, *sb = sa + kABytes;
            mbar_expect_tx(&full_bar[s], stage_bytes);
            const int k0 = kb * BK;
            if (p.a_major == TC_MAJOR_K) {
                tma_load_2d(&tma_a, &full_bar[s], sa, k0, m0);                      // box 64(K) x 128(M)
            } else {
                tma_load_2d(&tma_a, &full_bar[s], sa, m0, k0);                      // two boxes 64(M) x 64(K)
                tma_load_2d(&tma_a, &full_bar[s], sa + BK * 128, m0 + 64, k0);
            }
            if (p.b_major == TC_MAJOR_K) {
                tma_load_2d(&tma_b, &full_bar[s], sb, k0, n0);                      // box 64(K) x BN(N)
            } else {
                for (int j = 0; j * 64 < p.BN; ++j)                                 // boxes 64(N) x 64(K)
                    tma_load_2d(&tma_b, &full_bar[s], sb + (size_t)j * BK * 128, n0 + j * 64, k0);
            }
            if (++s == p.stages) { s = 0; phase ^= 1; }
        }
    } else if (warp == 1 && lane == 0) {
        // ===================== MMA issuer =====================
        int s = 0;
        uint32_t phase = 0;
        const uint32_t a_step = (p.a_major == TC_MAJOR_K) ? (UMMA_K * 2) >> 4 : (UMMA_K * 128) >> 4;
        const uint32_t b_step = (p.b_major == TC_MAJOR_K) ? (UMMA_K * 2) >> 4 : (UMMA_K * 128) >> 4;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            mbar_wait(&full_bar[s], phase);
            tc_fence_after();
            const uint32_t sa = smem_u32(tiles + (size_t)s * stage_bytes);
            const uint64_t adesc = make_desc(sa, p.a_major), bdesc = make_desc(sa + kABytes, p.b_major);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
                umma_bf16(tmem_base, adesc + (uint64_t)(k * a_step), bdesc + (uint64_t)(k * b_step), p.idesc,
                          (kb > kb_begin || k > 0) ? 1u : 0u);
            umma_commit(&empty_bar[s]);          // frees the smem stage when these MMAs retire
            if (++s == p.stages) { s = 0; phase ^= 1; }
        }
        if (kb_begin < kb_end) umma_commit(&tmem_full_bar);   // accumulator complete
    } else if (warp >= 2) {
        // ===================== epilogue: TMEM -> registers -> global =====================
        const int q = warp & 3;                  // TMEM lane quarter this warp may read
        const int row = m0 + q * 32 + lane;
        const bool has_k = kb_begin < kb_end;    // an empty split range contributes zeros
        if (has_k) { mbar_wait(&tmem_full_bar, 0); tc_fence_after(); }
        const bool row_ok = row < p.M;
        for (int c = 0; c < p.BN; c += 16) {
            float v[16];
            if (has_k) {
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);   // warp-collective: no early exit
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0.0f;
            }
            const int n = n0 + c;
            if (!row_ok || n >= p.N) continue;
            const bool full = (n + 16 <= p.N);
            if (EPI == TC_EPI_STORE_F32) {
                float *o = reinterpret_cast<float *>(p.out) + (int64_t)blockIdx.z * p.split_stride + (int64_t)row * p.ldo + n;
                if (full && (p.ldo & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4 *>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                } else {
                    for (int j = 0; j < 16 && n + j < p.N; ++j) o[j] = v[j];
                }
            } else {
                __nv_bfloat16 *o = reinterpret_cast<__nv_bfloat16 *>(p.out) + (int64_t)row * p.ldo + n;
                if (EPI == TC_EPI_FWD) {
                    // Linear bias, Dropout, ReLU (mlp.py:14-16), all in fp32 on the accumulator
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        if (n + j < p.N) {
                            float u = v[j] + p.bias[n + j];
                            if (p.training) u *= p.drop.factor(row, n + j, p.N);
                            v[j] = fmaxf(u, 0.0f);
                        }
                    }
                } else {
                    // dH = dA * relu'(A_prev) / (1-p): A_prev > 0 implies the unit was kept
                    const __nv_bfloat16 *ap = p.A_prev + (int64_t)row * p.ldo + n;
                    const float sc = p.training ? p.drop.scale : 1.0f;
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (n + j < p.N) v[j] = (__bfloat162float(ap[j]) > 0.0f) ? v[j] * sc : 0.0f;
                }
                if (full && (p.ldo & 7) == 0) {
                    __nv_bfloat162 h[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) h[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                    *reinterpret_cast<uint4 *>(o) = *reinterpret_cast<uint4 *>(&h[0]);
                    *reinterpret_cast<uint4 *>(o + 8) = *reinterpret_cast<uint4 *>(&h[4]);
                } else {
                    for (int j = 0; j < 16 && n + j < p.N; ++j) o[j] = __float2bfloat16_rn(v[j]);
                }
            }
        }
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, p.tmem_cols);
    }
}

// ------------------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
    });
    return fn;
}

struct MapKey {
    const void *ptr;
    uint64_t inner, outer, pitch;
    uint32_t box_inner, box_outer;
    bool operator==(const MapKey &o) const {
        return ptr == o.ptr && inner == o.inner && outer == o.outer && pitch == o.pitch && box_inner == o.box_inner &&
               box_outer == o.box_outer;
    }
};
std::vector<std::pair<MapKey, CUtensorMap>> g_maps;   // small per-process cache (buffers live in fixed workspaces)
std::mutex g_maps_mu;

int get_map(const void *ptr, uint64_t inner, uint64_t outer, uint64_t pitch_elems, uint32_t box_inner, uint32_t box_outer,
            CUtensorMap *out) {
    const MapKey key{ptr, inner, outer, pitch_elems, box_inner, box_outer};
    std::lock_guard<std::mutex> lock(g_maps_mu);
    for (auto &e : g_maps)
        if (e.first == key) { *out = e.second; return TS_OK; }
    EncodeTiledFn fn = encode_fn();
    TS_REQUIRE(fn != nullptr, TS_E_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    TS_REQUIRE(((uintptr_t)ptr & 15) == 0 && (pitch_elems * 2) % 16 == 0, TS_E_INVALID,
               "TMA operand needs a 16-byte aligned base and row pitch (ptr %p, pitch %llu elements)", ptr,
               (unsigned long long)pitch_elems);
    const cuuint64_t dims[2] = {inner, outer};
    const cuuint64_t strides[1] = {pitch_elems * 2};
    const cuuint32_t box[2] = {box_inner, box_outer};
    const cuuint32_t estr[2] = {1, 1};
    CUtensorMap m;
    const CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    TS_REQUIRE(r == CUDA_SUCCESS, TS_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (dims %llu x %llu, box %u x %u)",
               (int)r, (unsigned long long)inner, (unsigned long long)outer, box_inner, box_outer);
    if (g_maps.size() > 256) g_maps.clear();
    g_maps.emplace_back(key, m);
    *out = m;
    return TS_OK;
}

// Tile width: among widths that waste <= 12.5 % of the columns, the one with the best wave efficiency
// (CTAs / (waves * SMs)), ties to the wider tile.
int pick_bn(int M, int N, int splits, int sm_count) {
    static const int cand[] = {256, 224, 192, 160, 128, 112, 96, 80, 64, 48, 32, 16};
    const int tiles_m = (M + BM - 1) / BM;
    int best = 16;
    double best_eff = -1.0;
    for (int bn : cand) {
        const int tiles_n = (N + bn - 1) / bn;
        const int waste = tiles_n * bn - N;
        if (waste * 8 > N && bn > 16) continue;
        const long ctas = (long)tiles_m * tiles_n * splits;
        const long waves = (ctas + sm_count - 1) / sm_count;
        const double eff = (double)ctas / (double)(waves * sm_count) * (double)N / (double)(tiles_n * bn);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = bn; }
    }
    return best;
}

}  // namespace

int launch_tc_gemm(const TcGemmArgs &a, int sm_count, cudaStream_t st) {
    TS_REQUIRE(a.M >= 1 && a.N >= 1 && a.K >= 1, TS_E_INVALID, "tc gemm: bad shape %d x %d x %d", a.M, a.N, a.K);
    TS_REQUIRE(a.splits >= 1 && (a.splits == 1 || a.epi == TC_EPI_STORE_F32), TS_E_INVALID, "split-K needs the fp32 store epilogue");
    TcParams p{};
    p.M = a.M; p.N = a.N; p.K = a.K;
    p.a_major = a.a_major; p.b_major = a.b_major;
    p.num_kblocks = (a.K + BK - 1) / BK;
    TS_REQUIRE(a.k_chunk % BK == 0, TS_E_INVALID, "tc gemm: k_chunk %d is not a multiple of %d", a.k_chunk, BK);
    p.kblocks_per_split = a.k_chunk > 0 ? a.k_chunk / BK : (p.num_kblocks + a.splits - 1) / a.splits;
    const int splits = a.splits;
    TS_REQUIRE((long)splits * p.kblocks_per_split >= p.num_kblocks, TS_E_INVALID, "tc gemm: %d splits of %d k-blocks do not cover K=%d",
               splits, p.kblocks_per_split, a.K);
    p.BN = pick_bn(a.M, a.N, splits, sm_count);
    p.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a.a_major << 15) | ((uint32_t)a.b_major << 16) |
              ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    p.b_bytes = (a.b_major == TC_MAJOR_K) ? (uint32_t)p.BN * 128u : (uint32_t)((p.BN + 63) / 64) * (BK * 128);
    p.b_bytes = (p.b_bytes + 1023u) & ~1023u;
    const uint32_t stage = kABytes + p.b_bytes;
    p.stages = (int)((200u * 1024u) / stage);
    if (p.stages > kMaxStages) p.stages = kMaxStages;
    if (p.stages > p.kblocks_per_split) p.stages = p.kblocks_per_split < 2 ? 2 : p.kblocks_per_split;
    p.tmem_cols = 32;
    while ((int)p.tmem_cols < p.BN) p.tmem_cols <<= 1;
    p.out = a.out; p.ldo = a.ldo; p.split_stride = a.split_stride;
    p.bias = a.bias; p.A_prev = a.A_prev; p.drop = a.drop; p.training = a.training;

    CUtensorMap ma, mb;
    int rc;
    if (a.a_major == TC_MAJOR_K) rc = get_map(a.A, (uint64_t)a.K, (uint64_t)a.M, (uint64_t)a.lda, BK, BM, &ma);
    else rc = get_map(a.A, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)a.lda, 64, BK, &ma);
    if (rc) return rc;
    if (a.b_major == TC_MAJOR_K) rc = get_map(a.B, (uint64_t)a.K, (uint64_t)a.N, (uint64_t)a.ldb, BK, (uint32_t)p.BN, &mb);
    else rc = get_map(a.B, (uint64_t)a.N, (uint64_t)a.K, (uint64_t)a.ldb, 64, BK, &mb);
    if (rc) return rc;

    const size_t smem = (size_t)p.stages * stage + 1024;
    dim3 grid((a.N + p.BN - 1) / p.BN, (a.M + BM - 1) / BM, splits);
#define TS_TC_LAUNCH(E)                                                                                             \
    do {                                                                                                            \
        static bool attr_set[64] = {false};                                                                         \
        int dev_ = 0;                                                                                               \
        TS_CUDA_CHECK(cudaGetDevice(&dev_));                                                                        \
        if (!attr_set[dev_ & 63]) {                                                                                 \
            TS_CUDA_CHECK(cudaFuncSetAttribute(tc_gemm_kernel<E>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024)); \
            attr_set[dev_ & 63] = true;                                                                             \
        }                                                                                                           \
        tc_gemm_kernel<E><<<grid, kThreads, smem, st>>>(ma, mb, p);                                                 \
    } while (0)
    if (a.epi == TC_EPI_STORE_F32) TS_TC_LAUNCH(TC_EPI_STORE_F32);
    else if (a.epi == TC_EPI_FWD) TS_TC_LAUNCH(TC_EPI_FWD);
    else TS_TC_LAUNCH(TC_EPI_ACTGRAD);
#undef TS_TC_LAUNCH
    TS_LAUNCH_CHECK();
    return TS_OK;
}

}  // namespace tsb
