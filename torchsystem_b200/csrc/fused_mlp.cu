// Fused forward + head (+ backward to dH) for the one-hidden-layer aggregate -- the shape the reference ships
// (examples/tinysys/main.py:60: MLP(784, 256, 10)). A cluster of 1, 2 or 4 CTAs owns 128 batch rows (each CTA a slice
// of the hidden width; partial logits all-gathered through distributed shared memory), so a whole row of the network
// lives on chip from the input tile to the loss:
//
//   H  = X W1^T                     tcgen05.mma, X and W1 tiles by TMA, accumulator in TMEM (cols 0..H)
//   A  = relu(dropout(H + b1))      epilogue threads (one row each), written as bf16 into a swizzled smem tile
//   Z  = A W2^T                     tcgen05.mma again: that smem tile IS a valid K-major operand; Z -> TMEM cols 256..271
//   loss, argmax, dZ                per-row softmax in registers; loss sum / correct count reduced warp->CTA->grid
//   dA = dZ W2                      third tcgen05.mma (K = 16 padded classes), overwrites TMEM cols 0..H
//   dH = dA * relu'(A) / (1-p)      epilogue threads, bf16, coalesced store
//   dW2^T, db1 partials             A^T dZ and dH^T 1 as MN-major tcgen05.mma over the 128 rows (TMEM cols 272..)
// Inputs: bf16 by TMA (every CTA of a cluster fetches the slab's X tile itself; the repeats are L2 hits), or fp32
// converted on chip when a slab is one CTA.
//
// classifier.py:22-27 (forward, loss), :31-35 (fit), :37-39 (evaluate); mlp.py:14-17; metrics.py:14-15.
// Restrictions (else the engine uses the layer-by-layer path): hidden width 64, 128 or 256,
// at most 16 classes, ReLU, bf16 mode.
#include <cstdlib>

#include "tc_common.cuh"

namespace tsb {

using namespace tc;

namespace {

constexpr int kThreads = 352;   // warp 0: X producer, 1: MMA issuer + TMEM owner, 2: W1/W2 producer, 3..10: epilogue
constexpr int kEpiThreads = 256; // two warps per TMEM lane quarter, each owning half of the hidden columns
constexpr int kMaxStages = 6;
constexpr int kXStages = 4;      // most fp32 staging slots (32 KiB per 128 x 64 tile)
constexpr uint32_t kXBytes = 128 * 64 * 4;
constexpr int CP = 16;          // classes padded to one UMMA N / K step
constexpr int kZCol = 256;      // TMEM column of the logits accumulator
constexpr int kGCol = 272;      // TMEM columns of the gradient partials: dW2^T at +0/+16, db1 at +32/+48

struct FusedParams {
    int B, D, H, C;
    int ns, BN;                // CTAs per cluster splitting the hidden width, hidden columns per CTA (H / ns)
    int num_kblocks, stages, xstages;
    int x_f32;                 // inputs are fp32: TMA them into a staging ring and convert to bf16 on chip
    __nv_bfloat16 *x_bf16_out; // fp32 inputs + training: the bf16 copy the dW1 GEMM reads (written by cluster rank 0), or nullptr
    uint32_t b_bytes;
    uint32_t idesc_main, idesc_z, idesc_da;
    const float *b1, *b2;
    const int64_t *y;
    int64_t *pred;
    float *dz;                 // [B, C] fp32 (training) or nullptr
    __nv_bfloat16 *A_out;      // [B, H] (training) or nullptr
    __nv_bfloat16 *dH_out;     // [B, H] (training) or nullptr
    // per-CTA gradient partials computed on the tensor cores (training, H >= 128, C <= 15), else nullptr
    float *w2_partial, *b2_partial, *b1_partial;
    int64_t w2_stride, b2_stride, b1_stride;
    uint32_t idesc_g;
    DropoutSpec drop;
    int training;
    HeadPartial *partials;
    unsigned int *ticket;
    ts_batch_record *record;
    uint32_t *bad_target;
};

// clock64 trace of CTA 0 of the most recent launch (read back by the ts_test_fused_trace hook)
__device__ unsigned long long g_fused_trace[32];
#define TRACE(i) do { if (blockIdx.x < 2) g_fused_trace[16 * blockIdx.x + (i)] = clock64(); } while (0)

// thread-block-cluster helpers (the hidden width of one 128-row slab is split over `ns` CTAs)
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire;" ::: "memory"); }
__device__ __forceinline__ uint32_t map_to_rank(uint32_t local_saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

template <bool TRAIN>
__global__ void __launch_bounds__(kThreads, 1)
fused_mlp_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_w1,
                 const __grid_constant__ CUtensorMap tma_w2, const FusedParams p) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages];
    __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t xf_full_bar[kXStages], xf_empty_bar[kXStages];
    __shared__ __align__(8) uint64_t h_full_bar, w2_full_bar, a_ready_bar, z_full_bar, dz_ready_bar, da_full_bar, dh_ready_bar, g_full_bar;
    __shared__ float s_db2[4][CP];
    __shared__ uint32_t tmem_base_slot;
    __shared__ float s_loss[4];
    __shared__ int s_corr[4];
    __shared__ bool s_last;

    uint8_t *tiles = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t a_bytes = BM * 128;                       // one 64-wide K atom of X: 128 rows x 128 B
    const uint32_t stage_bytes = a_bytes + p.b_bytes;
    const int atoms = p.BN / KATOM;                          // 64-column atoms of this CTA's share of the hidden width
    uint8_t *a_stage = tiles;                                // after the main loop: A (then dH) tile, atoms x 16 KiB
    uint8_t *dh_stage = tiles + (size_t)atoms * a_bytes;     // dH tile, same layout (kept apart from A: both feed the gradient MMAs)
    uint8_t *dz_tile = tiles + (size_t)2 * atoms * a_bytes;  // 128 rows x 128 B, first 32 B of each row used
    uint8_t *w2_tile = tiles + (size_t)p.stages * stage_bytes;           // atoms x [16 rows x 128 B]
    float *b1s = reinterpret_cast<float *>(w2_tile + (size_t)atoms * 2048);
    float *zbuf = b1s + p.BN;                                // [ns][128 rows][16] partial logits of every CTA of the cluster (ns > 1)
    // fp32 inputs: staging ring behind everything else, 1024-byte aligned for the TMA swizzle
    uint8_t *xstage = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(zbuf + (p.ns > 1 ? p.ns * BM * CP : 0)) + 1023) & ~(uintptr_t)1023);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int ns = p.ns;
    const uint32_t rank = ns > 1 ? cluster_rank() : 0u;       // which slice of the hidden width
    const int slab = blockIdx.x / ns;                         // which 128 batch rows
    const int m0 = slab * BM;
    const int n0 = (int)rank * p.BN;
    if (threadIdx.x == 0) TRACE(0);

    if (threadIdx.x == 0) {
        // a stage is full when the W1 tile has landed and the X tile has landed (bf16 inputs: TMA) or been converted
        // (fp32 inputs: one arrival per converter warp -- 256 per-thread arrivals would serialise on the barrier word)
        // Every CTA of a cluster loads the whole X tile for itself, so all of these barriers are CTA-local.
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], p.x_f32 ? 1 + kEpiThreads / 32 : 2); mbar_init(&empty_bar[s], 1); }
        for (int s = 0; s < p.xstages; ++s) { mbar_init(&xf_full_bar[s], 1); mbar_init(&xf_empty_bar[s], kEpiThreads / 32); }
        mbar_init(&h_full_bar, 1);
        mbar_init(&w2_full_bar, 1);
        mbar_init(&a_ready_bar, kEpiThreads / 32);      // one arrival per warp: 256 per-thread arrivals serialise on the barrier word
        mbar_init(&z_full_bar, 1);
        mbar_init(&dz_ready_bar, 4);
        mbar_init(&da_full_bar, 1);
        mbar_init(&dh_ready_bar, kEpiThreads / 32);
        mbar_init(&g_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_x)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_w1)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_w2)) : "memory");
    }
    if (warp == 1) tmem_alloc(&tmem_base_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    if (ns > 1) { cluster_arrive(); cluster_wait(); }          // every CTA of the cluster is running: its shared memory may be written
    // Dependents are released only AFTER the wait: the early-running cast of the next batch (engine.cu) relies on
    // "a kernel behind this one has started" implying "everything in front of this one has completed".
    pdl_wait();          // the prologue above overlapped the predecessor's tail; global memory is touched from here on
    pdl_launch_dependents();
    if (threadIdx.x == 0) TRACE(1);

    if (ns > 1 && warp < 3) cluster_arrive();                  // phase 1 (logit exchange): these warps only have to have arrived
    if (warp == 0) {
        // ===================== TMA producer: X tiles (A operand) =====================
        // In a cluster every CTA needs the same X tile and fetches it itself (the source is the caller's pointer).
        int s = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < p.num_kblocks; ++kb) {
            if (elect_one()) {
                if (!p.x_f32) {
                    mbar_wait(&empty_bar[s], phase ^ 1);
                    mbar_expect_tx(&full_bar[s], a_bytes);
                    tma_load_2d(&tma_x, &full_bar[s], tiles + (size_t)s * stage_bytes, kb * KATOM, m0);
                } else {
                    // fp32 tile = two boxes of 32 columns (128 B rows) into the staging ring; the epilogue warps convert
                    const int xs = kb % p.xstages;
                    mbar_wait(&xf_empty_bar[xs], ((kb / p.xstages) & 1) ^ 1);
                    mbar_expect_tx(&xf_full_bar[xs], kXBytes);
                    uint8_t *dst = xstage + (size_t)xs * kXBytes;
                    tma_load_2d(&tma_x, &xf_full_bar[xs], dst, kb * KATOM, m0);
                    tma_load_2d(&tma_x, &xf_full_bar[xs], dst + kXBytes / 2, kb * KATOM + 32, m0);
                }
            }
            __syncwarp();
            if (++s == p.stages) { s = 0; phase ^= 1; }
        }
    } else if (warp == 2) {
        // ===================== TMA producer: W2 tile once, then W1 tiles (B operand) =====================
        if (elect_one()) {
            mbar_expect_tx(&w2_full_bar, (uint32_t)atoms * 2048u);
            for (int a = 0; a < atoms; ++a) tma_load_2d(&tma_w2, &w2_full_bar, w2_tile + (size_t)a * 2048, n0 + a * KATOM, 0);
        }
        __syncwarp();
        int s = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < p.num_kblocks; ++kb) {
            if (elect_one()) {
                mbar_wait(&empty_bar[s], phase ^ 1);
                mbar_expect_tx(&full_bar[s], p.b_bytes);
                tma_load_2d(&tma_w1, &full_bar[s], tiles + (size_t)s * stage_bytes + a_bytes, kb * KATOM, n0);
            }
            __syncwarp();
            if (++s == p.stages) { s = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        int s = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < p.num_kblocks; ++kb) {
            if (elect_one()) {
                mbar_wait(&full_bar[s], phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(tiles + (size_t)s * stage_bytes);
                const uint64_t adesc = make_desc(sa, TC_MAJOR_K, KATOM), bdesc = make_desc(sa + a_bytes, TC_MAJOR_K, KATOM);
#pragma unroll
                for (int k = 0; k < KATOM / UMMA_K; ++k)
                    umma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), p.idesc_main, (kb > 0 || k > 0) ? 1u : 0u);
                umma_commit(&empty_bar[s]);
                if (kb == p.num_kblocks - 1) umma_commit(&h_full_bar);
                if (kb == 0) TRACE(2);
                if (kb == p.num_kblocks - 1) TRACE(3);
            }
            __syncwarp();
            if (++s == p.stages) { s = 0; phase ^= 1; }
        }
        if (elect_one()) {
            // ---- Z = A W2^T : A tile written by the epilogue threads (K-major, `atoms` atoms), W2 tile K-major (N = 16)
            mbar_wait(&w2_full_bar, 0);
            mbar_wait(&a_ready_bar, 0);
            tc_fence_after();
            const uint32_t sa = smem_u32(a_stage), sw = smem_u32(w2_tile);
            for (int a = 0; a < atoms; ++a) {
                const uint64_t adesc = make_desc(sa + a * a_bytes, TC_MAJOR_K, KATOM);
                const uint64_t bdesc = make_desc(sw + a * 2048, TC_MAJOR_K, KATOM);
#pragma unroll
                for (int k = 0; k < KATOM / UMMA_K; ++k)
                    umma_bf16(tmem_base + kZCol, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), p.idesc_z, (a > 0 || k > 0) ? 1u : 0u);
            }
            umma_commit(&z_full_bar);
            TRACE(6);
            if (TRAIN) {
                // ---- dA = dZ W2 : dZ tile K-major with K = 16; the same W2 tile read as an MN-major operand
                //      (K = 16 class rows, N atoms of 64 columns 2 KiB apart)
                mbar_wait(&dz_ready_bar, 0);
                tc_fence_after();
                umma_bf16(tmem_base, make_desc(smem_u32(dz_tile), TC_MAJOR_K, KATOM), make_desc(sw, TC_MAJOR_MN, CP), p.idesc_da, 0u);
                umma_commit(&da_full_bar);
                TRACE(10);
                if (p.w2_partial != nullptr) {
                    // ---- per-CTA gradient partials, contraction over this CTA's 128 batch rows (K = rows, 8 steps):
                    //        dW2^T[h, c] = sum_r A[r, h] dZ[r, c]           TMEM cols 272.. (one 16-column block per 128 hidden units)
                    //        db1[h]      = sum_r dH[r, h] * 1              = column 15 of dH^T dZ', dZ' column 15 being 1 -> cols 304..
                    //      The A / dH tiles are read as MN-major operands (M = hidden), the dZ tile as an MN-major B (N = 16).
                    mbar_wait(&dh_ready_bar, 0);
                    tc_fence_after();
                    const uint32_t sd = smem_u32(dz_tile), sh = smem_u32(dh_stage);
                    for (int mh = 0; mh < (p.BN + 127) / 128; ++mh) {   // a 64-wide slice still issues M = 128: lanes 64.. are ignored
#pragma unroll
                        for (int k = 0; k < BM / UMMA_K; ++k) {
                            const uint64_t bdesc = make_desc(sd + k * 2048, TC_MAJOR_MN, BM);
                            umma_bf16(tmem_base + kGCol + 16 * mh, make_desc(sa + (2 * mh) * a_bytes + k * 2048, TC_MAJOR_MN, BM), bdesc, p.idesc_g, k > 0 ? 1u : 0u);
                            umma_bf16(tmem_base + kGCol + 32 + 16 * mh, make_desc(sh + (2 * mh) * a_bytes + k * 2048, TC_MAJOR_MN, BM), bdesc, p.idesc_g, k > 0 ? 1u : 0u);
                        }
                    }
                    umma_commit(&g_full_bar);
                }
            }
        }
        __syncwarp();
    } else {
        // ===================== epilogue warps: thread = (batch row, half of the hidden columns) =====================
        const int q = warp & 3;                          // TMEM lane quarter
        const int half = (warp - 3) >> 2;                // which half of the hidden columns
        const int r_local = q * 32 + lane;               // row within the 128-row tile == TMEM lane
        const int64_t row = (int64_t)m0 + r_local;
        const bool row_ok = row < p.B;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const int et = threadIdx.x - 96;                 // 0..255
        const int cbeg = half * (p.BN >> 1), cend = cbeg + (p.BN >> 1);     // local columns
        for (int i = et; i < p.BN; i += kEpiThreads) b1s[i] = p.b1[n0 + i];
        named_bar_sync(1, kEpiThreads);

        if (p.x_f32) {
            // ---- during the main loop these warps are the fp32 -> bf16 converters: thread = (row, 32-column half)
            // tile row 0..127, which fp32 box. 8 consecutive threads = 8 consecutive rows, so their 16-byte accesses fall into 8
            // different bank groups of the swizzled tiles (row pairs sharing a chunk index were 2-way bank conflicts)
            const int cr = et & 127, ch = et >> 7;
            const uint32_t xsrc0 = smem_u32(xstage) + (uint32_t)ch * (kXBytes / 2) + (uint32_t)cr * 128;
            const uint32_t dst0 = smem_u32(tiles) + (uint32_t)cr * 128;
            const uint32_t sw = (uint32_t)(cr & 7);
            int s = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < p.num_kblocks; ++kb) {
                const int xs = kb % p.xstages;
                // every lane observes the TMA barrier itself: bytes written through the async proxy are only guaranteed
                // visible to threads that waited on the mbarrier
                mbar_wait(&xf_full_bar[xs], (kb / p.xstages) & 1);          // fp32 tile landed
                const uint32_t src = xsrc0 + (uint32_t)xs * kXBytes;
                float4 f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = lds128(src + (((uint32_t)j ^ sw) << 4));
                uint4 pk[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    pk[t].x = pack_bf16(f[2 * t].x, f[2 * t].y);
                    pk[t].y = pack_bf16(f[2 * t].z, f[2 * t].w);
                    pk[t].z = pack_bf16(f[2 * t + 1].x, f[2 * t + 1].y);
                    pk[t].w = pack_bf16(f[2 * t + 1].z, f[2 * t + 1].w);
                }
                mbar_wait(&empty_bar[s], phase ^ 1);                        // bf16 stage free (its MMAs retired)
                const uint32_t dst = dst0 + (uint32_t)s * (uint32_t)stage_bytes;
#pragma unroll
                for (int t = 0; t < 4; ++t) sts128(dst + ((((uint32_t)ch * 4 + t) ^ sw) << 4), pk[t]);
                // the staging slot is released only after its values were consumed by the stores above: an arrive does
                // not wait for loads that are still in flight, so releasing right behind the loads lets the next TMA
                // fill overtake them
                __syncwarp();
                if (lane == 0) mbar_arrive(&xf_empty_bar[xs]);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&full_bar[s]);
                if (p.x_bf16_out != nullptr && rank == 0 && (int64_t)m0 + cr < p.B) {
                    // the dW1 = dH^T X GEMM of this step reads the bf16 copy: 64 contiguous bytes per thread
                    const int col = kb * KATOM + ch * 32;
                    __nv_bfloat16 *g = p.x_bf16_out + ((int64_t)m0 + cr) * p.D + col;
#pragma unroll
                    for (int t = 0; t < 4; ++t)
                        if (col + t * 8 < p.D) *reinterpret_cast<uint4 *>(g + t * 8) = pk[t];
                }
                if (++s == p.stages) { s = 0; phase ^= 1; }
            }
        }

        // ---- dropout keep decisions of this thread's row x columns, one bit per hidden unit. They depend on indices
        //      only (counter-based hash) or on the caller's mask, so they are drawn here, while the main loop is still
        //      running on the tensor cores. Rolled loops on purpose: this code runs once per thread, and straight-line
        //      code that is executed once costs about a cycle per instruction in instruction fetch alone.
        uint32_t keepb[4] = {~0u, ~0u, ~0u, ~0u};
        const bool dropout_on = TRAIN && p.training && p.drop.p > 0.0f;
        if (dropout_on && row_ok) {
#pragma unroll 1
            for (int ci = 0; ci < 4; ++ci) {
                const int c0 = cbeg + 32 * ci;
                if (c0 >= cend) break;
                uint32_t kb = 0;
                if (p.drop.mask == nullptr) {
#pragma unroll 2
                    for (int j = 0; j < 32; j += 2) {
                        const uint32_t h = dropout_bits(p.drop.key, row, (n0 + c0 + j) >> 1);
                        kb |= ((h & 0xFFFFu) >= p.drop.thresh ? 1u : 0u) << j;
                        kb |= ((h >> 16) >= p.drop.thresh ? 1u : 0u) << (j + 1);
                    }
                } else {
                    const uint8_t *mrow = p.drop.mask + row * (int64_t)p.H + n0 + c0;
#pragma unroll 4
                    for (int j = 0; j < 32; ++j) kb |= (mrow[j] != 0 ? 1u : 0u) << j;
                }
                if (ci == 0) keepb[0] = kb; else if (ci == 1) keepb[1] = kb; else if (ci == 2) keepb[2] = kb; else keepb[3] = kb;
            }
        }
        // keep the draws on this side of the wait below (the compiler would otherwise sink them next to their use)
        asm volatile("" ::"r"(keepb[0]), "r"(keepb[1]), "r"(keepb[2]), "r"(keepb[3]) : "memory");

        // ---- A = relu(dropout(H + b1)) -> bf16 -> swizzled K-major tile
        uint32_t sign[4] = {0, 0, 0, 0};
        if (elect_one()) mbar_wait(&h_full_bar, 0);
        __syncwarp();
        tc_fence_after();
        if (et == 0) TRACE(4);
#pragma unroll 1
        for (int c0 = cbeg; c0 < cend; c0 += 32) {
            float v[32];
            tmem_ld16(t_lane + (uint32_t)c0, *reinterpret_cast<float(*)[16]>(&v[0]));
            tmem_ld16(t_lane + (uint32_t)(c0 + 16), *reinterpret_cast<float(*)[16]>(&v[16]));
            uint32_t bits = 0;
            const int ci_k = (c0 - cbeg) >> 5;
            const uint32_t kb = ci_k == 0 ? keepb[0] : ci_k == 1 ? keepb[1] : ci_k == 2 ? keepb[2] : keepb[3];
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                float keep[4] = {1.0f, 1.0f, 1.0f, 1.0f};
                if (dropout_on) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) keep[i] = ((kb >> (j + i)) & 1u) ? p.drop.scale : 0.0f;
                }
                const float4 bb = *reinterpret_cast<const float4 *>(&b1s[c0 + j]);
                const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float a = round_bf16(fmaxf((v[j + i] + bv[i]) * keep[i], 0.0f));
                    v[j + i] = a;
                    bits |= (a > 0.0f ? 1u : 0u) << (j + i);
                }
            }
            const int ci = (c0 - cbeg) >> 5;
            if (ci == 0) sign[0] = bits; else if (ci == 1) sign[1] = bits; else if (ci == 2) sign[2] = bits; else sign[3] = bits;
            uint8_t *rowp = a_stage + (size_t)(c0 >> 6) * a_bytes + (size_t)r_local * 128;
            const int cbase = (c0 & 63) >> 3;            // first 16-byte chunk of this 32-column group: 0 or 4
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                uint4 pk;
                pk.x = pack_bf16(v[8 * t], v[8 * t + 1]);
                pk.y = pack_bf16(v[8 * t + 2], v[8 * t + 3]);
                pk.z = pack_bf16(v[8 * t + 4], v[8 * t + 5]);
                pk.w = pack_bf16(v[8 * t + 6], v[8 * t + 7]);
                *reinterpret_cast<uint4 *>(rowp + (((cbase + t) ^ (r_local & 7)) << 4)) = pk;
            }
        }
        fence_proxy_async_smem();          // the tensor core reads this tile through the async proxy
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_ready_bar);
        if (et == 0) TRACE(5);

        // ---- coalesced copy of this warp's rows x columns of a bf16 tile held in a_stage to a global [B, H] matrix
        const int npieces = p.BN >> 4;                   // 16-byte pieces per row in this warp's column half (4, 8 or 16)
        const int np_shift = 31 - __clz(npieces);
        auto copy_out = [&](__nv_bfloat16 *dst, const uint8_t *tile) {
            for (int idx = lane; idx < 32 * npieces; idx += 32) {
                const int rr = q * 32 + (idx >> np_shift);
                const int pc = half * npieces + (idx & (npieces - 1));   // piece index within the full row
                const uint4 t = *reinterpret_cast<const uint4 *>(tile + (size_t)(pc >> 3) * a_bytes + (size_t)rr * 128 + (((pc & 7) ^ (rr & 7)) << 4));
                if ((int64_t)m0 + rr < p.B) *reinterpret_cast<uint4 *>(dst + ((int64_t)m0 + rr) * p.H + n0 + pc * 8) = t;
            }
        };
        if (TRAIN && p.A_out != nullptr) {
            __syncwarp();
            copy_out(p.A_out, a_stage);                  // dW2 = dZ^T A reads it when the gradient partials are not fused
        }
        if (et == 0) TRACE(7);

        float loss_acc = 0.0f;
        int corr_acc = 0;
        if (half == 0) {
            // ---- logits, cross-entropy, argmax, dZ (classifier.py:27,35): one row per thread, all in registers.
            //      bf16 mode tolerates the fast exp/log (ex2/lg2.approx, ~2 ulp); argmax is taken on the raw logits.
            const int yy = row_ok ? (int)p.y[row] : 0;            // issued before the wait: latency hidden behind the logits MMA
            float b2v[CP];
#pragma unroll
            for (int c = 0; c < CP; ++c) b2v[c] = (c < p.C) ? p.b2[c] : 0.0f;
            if (elect_one()) mbar_wait(&z_full_bar, 0);
            __syncwarp();
            tc_fence_after();
            if (et == 0) TRACE(8);
            float z[CP];
            tmem_ld16(t_lane + (uint32_t)kZCol, z);
            if (ns > 1) {
                // every CTA holds the logits of its slice of the hidden units only: all-gather the partial logits
                // through distributed shared memory, then every CTA adds them up in the same (rank) order
                // (this warp sends to the lower half of the ranks, its idle twin of the other column half to the upper)
                const uint32_t mine = smem_u32(zbuf + ((size_t)rank * BM + r_local) * CP);
                for (int t = 0; t < (ns >> 1); ++t) {
                    const uint32_t dst = map_to_rank(mine, (uint32_t)t);
#pragma unroll
                    for (int c = 0; c < CP; c += 4) st_cluster_v4(dst + c * 4, z[c], z[c + 1], z[c + 2], z[c + 3]);
                }
                if (et == 0) TRACE(14);
                cluster_arrive();
                cluster_wait();
                if (et == 0) TRACE(15);
#pragma unroll
                for (int c = 0; c < CP; ++c) z[c] = 0.0f;
                for (int t = 0; t < ns; ++t) {
                    const float4 *src = reinterpret_cast<const float4 *>(zbuf + ((size_t)t * BM + r_local) * CP);
#pragma unroll
                    for (int c = 0; c < CP; c += 4) {
                        const float4 v4 = src[c >> 2];
                        z[c] += v4.x; z[c + 1] += v4.y; z[c + 2] += v4.z; z[c + 3] += v4.w;
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < CP; ++c) z[c] = (c < p.C) ? z[c] + b2v[c] : 0.0f;
            float mx = z[0];
            int arg = 0;
#pragma unroll
            for (int c = 1; c < CP; ++c)
                if (c < p.C && z[c] > mx) { mx = z[c]; arg = c; }      // strict '>': first maximum wins (torch.argmax)
            float e[CP];
            float ssum = 0.0f;
#pragma unroll
            for (int c = 0; c < CP; ++c) { e[c] = (c < p.C) ? __expf(z[c] - mx) : 0.0f; ssum += e[c]; }
            const float lse = mx + __logf(ssum);
            const float inv_sum = 1.0f / ssum;
            float zy = 0.0f;
#pragma unroll
            for (int c = 0; c < CP; ++c)
                if (c == yy) zy = z[c];
            float row_loss = lse - zy;
            if (yy < 0 || yy >= p.C) {                       // target outside [0, C): poison the loss, tell the host
                row_loss = __int_as_float(0x7fc00000);
                if (row_ok && rank == 0 && p.bad_target != nullptr) *reinterpret_cast<volatile uint32_t *>(p.bad_target) = (uint32_t)row + 1u;
            }
            loss_acc = row_ok ? row_loss : 0.0f;
            corr_acc = (row_ok && arg == yy) ? 1 : 0;
            if (row_ok && rank == 0) p.pred[row] = arg;

            if (TRAIN) {
                const float inv_B = 1.0f / (float)p.B;
                float dzv[CP];
#pragma unroll
                for (int c = 0; c < CP; ++c)
                    dzv[c] = (c < p.C && row_ok) ? (e[c] * inv_sum - (c == yy ? 1.0f : 0.0f)) * inv_B : 0.0f;
                if (p.dz != nullptr && row_ok && rank == 0) {
#pragma unroll
                    for (int c = 0; c < CP; ++c)
                        if (c < p.C) p.dz[row * p.C + c] = dzv[c];
                }
                float db2v[CP];
#pragma unroll
                for (int c = 0; c < CP; ++c) db2v[c] = dzv[c];
                if (p.w2_partial != nullptr) dzv[CP - 1] = row_ok ? 1.0f : 0.0f;   // the "ones" column that turns dH^T dZ' into column sums of dH
                // dZ tile: K-major, 16 bf16 (two 16-byte chunks) per row
                uint8_t *drow = dz_tile + (size_t)r_local * 128;
                uint4 c0v, c1v;
                c0v.x = pack_bf16(dzv[0], dzv[1]);   c0v.y = pack_bf16(dzv[2], dzv[3]);
                c0v.z = pack_bf16(dzv[4], dzv[5]);   c0v.w = pack_bf16(dzv[6], dzv[7]);
                c1v.x = pack_bf16(dzv[8], dzv[9]);   c1v.y = pack_bf16(dzv[10], dzv[11]);
                c1v.z = pack_bf16(dzv[12], dzv[13]); c1v.w = pack_bf16(dzv[14], dzv[15]);
                *reinterpret_cast<uint4 *>(drow + ((0 ^ (r_local & 7)) << 4)) = c0v;
                *reinterpret_cast<uint4 *>(drow + ((1 ^ (r_local & 7)) << 4)) = c1v;
                fence_proxy_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&dz_ready_bar);
                if (et == 0) TRACE(9);
                if (p.w2_partial != nullptr && rank == 0) {
                    // db2[c] = sum_r dz[r, c] (fp32), off the critical path: the dA MMA is already running
#pragma unroll
                    for (int c = 0; c < CP; ++c) {
                        const float t = warp_sum(db2v[c]);
                        if (lane == 0) s_db2[q][c] = t;
                    }
                }
            }
        } else if (ns > 1) {
            // the other column half's warps have nothing to compute here: they carry the same rows' partial logits to
            // the upper half of the cluster (the same TMEM lanes are readable by both warps of a lane quarter)
            if (elect_one()) mbar_wait(&z_full_bar, 0);
            __syncwarp();
            tc_fence_after();
            float z[CP];
            tmem_ld16(t_lane + (uint32_t)kZCol, z);
            const uint32_t mine = smem_u32(zbuf + ((size_t)rank * BM + r_local) * CP);
            for (int t = ns >> 1; t < ns; ++t) {
                const uint32_t dst = map_to_rank(mine, (uint32_t)t);
#pragma unroll
                for (int c = 0; c < CP; c += 4) st_cluster_v4(dst + c * 4, z[c], z[c + 1], z[c + 2], z[c + 3]);
            }
            cluster_arrive();                            // phase 1: these warps only have to have arrived
        }

        if (TRAIN) {
            // ---- dH = dA * relu'(A) / (1-p) -> bf16 -> staged in this warp's part of the (now free) A tile
            if (elect_one()) mbar_wait(&da_full_bar, 0);
            __syncwarp();
            tc_fence_after();
            if (et == 0) TRACE(11);
            const float sc = p.training ? p.drop.scale : 1.0f;
#pragma unroll 1
            for (int c0 = cbeg; c0 < cend; c0 += 32) {
                float v[32];
                tmem_ld16(t_lane + (uint32_t)c0, *reinterpret_cast<float(*)[16]>(&v[0]));
                tmem_ld16(t_lane + (uint32_t)(c0 + 16), *reinterpret_cast<float(*)[16]>(&v[16]));
                const int ci = (c0 - cbeg) >> 5;
                const uint32_t bits = ci == 0 ? sign[0] : ci == 1 ? sign[1] : ci == 2 ? sign[2] : sign[3];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = ((bits >> j) & 1u) ? v[j] * sc : 0.0f;
                uint8_t *rowp = dh_stage + (size_t)(c0 >> 6) * a_bytes + (size_t)r_local * 128;
                const int cbase = (c0 & 63) >> 3;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    uint4 pk;
                    pk.x = pack_bf16(v[8 * t], v[8 * t + 1]);
                    pk.y = pack_bf16(v[8 * t + 2], v[8 * t + 3]);
                    pk.z = pack_bf16(v[8 * t + 4], v[8 * t + 5]);
                    pk.w = pack_bf16(v[8 * t + 6], v[8 * t + 7]);
                    *reinterpret_cast<uint4 *>(rowp + (((cbase + t) ^ (r_local & 7)) << 4)) = pk;
                }
            }
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&dh_ready_bar);
            if (p.dH_out != nullptr) copy_out(p.dH_out, dh_stage);

            if (p.w2_partial != nullptr) {
                // ---- gradient partials: TMEM lane = hidden unit; this warp group owns hidden units [128*half, 128*half+128)
                if (elect_one()) mbar_wait(&g_full_bar, 0);
                __syncwarp();
                tc_fence_after();
                if (half * 128 + r_local < p.BN) {
                    const int h = n0 + half * 128 + r_local;
                    float gw[CP], gb[CP];
                    tmem_ld16(t_lane + (uint32_t)(kGCol + 16 * half), gw);
                    tmem_ld16(t_lane + (uint32_t)(kGCol + 32 + 16 * half), gb);
                    float *wp = p.w2_partial + (int64_t)slab * p.w2_stride + h;
#pragma unroll
                    for (int c = 0; c < CP; ++c)
                        if (c < p.C) wp[(int64_t)c * p.H] = gw[c];
                    p.b1_partial[(int64_t)slab * p.b1_stride + h] = gb[CP - 1];
                }
            }
        }
        if (et == 0) TRACE(12);

        // ---- loss sum / correct count: warp -> CTA -> grid (last CTA finalises in a fixed order)
        if (half == 0 && rank == 0) {
            loss_acc = warp_sum(loss_acc);
            corr_acc = warp_sum(corr_acc);
            if (lane == 0) { s_loss[q] = loss_acc; s_corr[q] = corr_acc; }
            named_bar_sync(2, 128);
            if (TRAIN && p.w2_partial != nullptr && q == 0 && lane < p.C)      // s_db2 was filled before this barrier
                p.b2_partial[(int64_t)slab * p.b2_stride + lane] = (s_db2[0][lane] + s_db2[1][lane]) + (s_db2[2][lane] + s_db2[3][lane]);
            if (q == 3 && lane == 0) {
                p.partials[slab].loss_sum = s_loss[0] + s_loss[1] + s_loss[2] + s_loss[3];
                p.partials[slab].correct = s_corr[0] + s_corr[1] + s_corr[2] + s_corr[3];
                __threadfence();
                s_last = (atomicAdd(p.ticket, 1u) == gridDim.x / ns - 1);
            }
            named_bar_sync(2, 128);
            if (s_last && q == 3) {
                __threadfence();
                float l = 0.0f;
                int c = 0;
                for (int b = lane; b < (int)gridDim.x / ns; b += 32) {
                    const volatile HeadPartial *hp = p.partials + b;
                    l += hp->loss_sum;
                    c += hp->correct;
                }
                l = warp_sum(l);
                c = warp_sum(c);
                if (lane == 0) {
                    p.record->loss = l / (float)p.B;
                    p.record->n_correct = c;
                    p.record->n_rows = (int32_t)p.B;
                    p.record->reserved = 0;
                    *p.ticket = 0u;
                }
            }
        }
    }
    __syncwarp();
    if (ns > 1 && (warp < 3 || ((warp - 3) >> 2) == 1)) cluster_wait();   // complete phase 1 for the warps that only arrived
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
    if (threadIdx.x == 0) TRACE(13);
}

}  // namespace

int read_fused_trace(unsigned long long *out_host) {
    TS_CUDA_CHECK(cudaMemcpyFromSymbol(out_host, g_fused_trace, sizeof(unsigned long long) * 32));
    return TS_OK;
}

bool fused_mlp_supported(int D, int H, int C) { return (H == 64 || H == 128 || H == 256) && C >= 1 && C <= CP && D % 8 == 0; }

// Split the hidden width of each 128-row slab over a cluster of ns CTAs while the whole grid still fits in one wave:
// small batches would otherwise leave most SMs idle (batch 4096 = 32 slabs -> 4 x 32 = 128 CTAs).
int fused_cluster_size(int64_t B, int H, int sm_count) {
    const int64_t slabs = (B + BM - 1) / BM;
    for (int ns = 4; ns >= 2; ns >>= 1)
        if (H / ns >= KATOM && slabs * ns <= sm_count) return ns;
    return 1;
}

int launch_fused_mlp(const FusedMlpArgs &a, int sm_count, cudaStream_t st) {
    TS_REQUIRE(fused_mlp_supported(a.D, a.H, a.C), TS_E_UNSUPPORTED, "fused MLP kernel: unsupported shape D=%d H=%d C=%d", a.D, a.H, a.C);
    TS_REQUIRE(a.B >= 1, TS_E_INVALID, "fused MLP kernel: batch %lld", (long long)a.B);
    FusedParams p{};
    p.B = (int)a.B; p.D = a.D; p.H = a.H; p.C = a.C;
    const int slabs = (int)((a.B + BM - 1) / BM);
    TS_REQUIRE(slabs <= kMaxHeadBlocks, TS_E_TOO_LARGE, "fused MLP kernel: batch %lld needs %d row slabs, more than the %d partial slots",
               (long long)a.B, slabs, kMaxHeadBlocks);
    p.ns = fused_cluster_size(a.B, a.H, sm_count);
    TS_REQUIRE(p.ns == 1 || p.ns == 2 || p.ns == 4, TS_E_INVALID, "cluster size %d", p.ns);
    TS_REQUIRE(a.H / p.ns >= KATOM, TS_E_INVALID, "cluster size %d leaves fewer than 64 hidden columns per CTA", p.ns);
    p.BN = a.H / p.ns;
    p.num_kblocks = (a.D + KATOM - 1) / KATOM;
    p.b_bytes = (uint32_t)p.BN * 128u;
    const uint32_t stage = BM * 128 + p.b_bytes;
    const int atoms = p.BN / KATOM;
    const uint32_t alias_need = 2u * (uint32_t)atoms * BM * 128 + BM * 128;   // A tile + dH tile + dZ tile reuse the stage memory
    p.x_f32 = a.X_f32 != nullptr ? 1 : 0;
    TS_REQUIRE((a.X != nullptr) != (a.X_f32 != nullptr), TS_E_INVALID, "fused MLP kernel: exactly one of X (bf16) and X_f32");
    p.x_bf16_out = p.x_f32 ? a.x_bf16_out : nullptr;
    // shared memory: pipeline stages + W2 tile + b1 + (cluster) partial logits + (fp32 inputs) staging ring
    p.xstages = 2;
    if (const char *f = getenv("TSB_XSTAGES")) p.xstages = atoi(f);
    TS_REQUIRE(p.xstages >= 1 && p.xstages <= kXStages, TS_E_UNSUPPORTED, "fused MLP kernel: bad staging depth %d", p.xstages);
    const size_t fixed = (size_t)atoms * 2048 + (size_t)p.BN * 4 + (p.ns > 1 ? (size_t)p.ns * BM * CP * 4 : 0) +
                         (p.x_f32 ? (size_t)p.xstages * kXBytes + 1024 : 0) + 1024 + 64;
    p.stages = (int)((226u * 1024u - fixed) / stage);
    if (p.stages > kMaxStages) p.stages = kMaxStages;
    TS_REQUIRE((uint32_t)p.stages * stage >= alias_need && p.stages >= 2, TS_E_UNSUPPORTED,
               "fused MLP kernel: shared memory budget does not fit H=%d ns=%d (stages %d)", a.H, p.ns, p.stages);
    p.idesc_main = make_idesc(BM, p.BN, TC_MAJOR_K, TC_MAJOR_K);
    p.idesc_z = make_idesc(BM, CP, TC_MAJOR_K, TC_MAJOR_K);
    p.idesc_da = make_idesc(BM, p.BN, TC_MAJOR_K, TC_MAJOR_MN);
    p.idesc_g = make_idesc(BM, CP, TC_MAJOR_MN, TC_MAJOR_MN);
    const bool fuse_grads = a.w2_partial != nullptr;
    TS_REQUIRE(!fuse_grads || (a.H >= 128 && a.C < CP && a.b1_partial && a.b2_partial), TS_E_INVALID,
               "fused gradient partials need H >= 128, C <= 15 and all three partial buffers");
    p.w2_partial = a.w2_partial; p.b2_partial = a.b2_partial; p.b1_partial = a.b1_partial;
    p.w2_stride = a.w2_stride; p.b2_stride = a.b2_stride; p.b1_stride = a.b1_stride;
    p.b1 = a.b1; p.b2 = a.b2; p.y = a.y; p.pred = a.pred; p.dz = a.dz; p.A_out = a.A_out; p.dH_out = a.dH_out;
    p.drop = a.drop; p.training = a.training;
    p.partials = a.partials; p.ticket = a.ticket; p.record = a.record; p.bad_target = a.bad_target;

    CUtensorMap mx, mw1, mw2;
    // X box: all BM rows of the slab -- every CTA of a cluster loads the whole tile for itself
    int rc = p.x_f32 ? get_map(a.X_f32, (uint64_t)a.D, (uint64_t)a.B, (uint64_t)a.D, 32, (uint32_t)BM, &mx, 4)
                     : get_map(a.X, (uint64_t)a.D, (uint64_t)a.B, (uint64_t)a.D, KATOM, (uint32_t)BM, &mx);
    if (rc) return rc;
    rc = get_map(a.W1, (uint64_t)a.D, (uint64_t)a.H, (uint64_t)a.D, KATOM, (uint32_t)p.BN, &mw1);
    if (rc) return rc;
    rc = get_map(a.W2p, (uint64_t)a.H, (uint64_t)CP, (uint64_t)a.H, KATOM, CP, &mw2);
    if (rc) return rc;

    const size_t smem = (size_t)p.stages * stage + fixed;
    const bool train = a.dH_out != nullptr;
    static bool attr_set[2][64] = {{false}};
    int dev = 0;
    TS_CUDA_CHECK(cudaGetDevice(&dev));
    if (!attr_set[train][dev & 63]) {
        if (train) TS_CUDA_CHECK(cudaFuncSetAttribute(fused_mlp_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        else TS_CUDA_CHECK(cudaFuncSetAttribute(fused_mlp_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        attr_set[train][dev & 63] = true;
    }
    const dim3 grid((unsigned)(slabs * p.ns));
    if (train) TS_CUDA_CHECK(launch_pdl(fused_mlp_kernel<true>, grid, dim3(kThreads), smem, st, (unsigned)p.ns, mx, mw1, mw2, p));
    else TS_CUDA_CHECK(launch_pdl(fused_mlp_kernel<false>, grid, dim3(kThreads), smem, st, (unsigned)p.ns, mx, mw1, mw2, p));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

}  // namespace tsb
