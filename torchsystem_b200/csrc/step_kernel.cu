// One kernel for a whole fit step on a small batch (at most 64 rows): the shape the reference ships and runs
// (examples/tinysys/main.py:60-65: MLP(784, 256, 10), Adam, batch 32). The per-batch loop it serves is unchanged
// (examples/tinysys/tinysys/services/training.py:32-35). At this size the four-kernel sequence of fused_mlp.cu ->
// tc_gemm.cu -> optimizer is a chain of launch gaps; here one cluster of H/64 CTAs does everything, and nothing but the
// inputs, the parameters and the optimizer state ever touches global memory:
//
//   X            fp32 rows converted on chip (or bf16 rows by TMA) into a RESIDENT swizzled tile set, 64 columns per block
//   H  = X W1^T  tcgen05.mma, W1 slice (64 hidden units per CTA) streamed by TMA; accumulator in TMEM
//   A, Z, loss, argmax, dZ, dA, dH, dW2^T, db1   exactly as fused_mlp.cu (partial logits all-gathered over DSMEM)
//   dW1^T slice  = X^T dH: the resident X blocks re-read as MN-major A operands (M = input feature, K = batch row), the dH
//                tile as the MN-major B operand; 7 M-tiles x 64 columns in TMEM
//   gradients    epilogue threads store them from TMEM into `grads` (TMEM lane = input feature d, so a warp writes 32
//                consecutive elements of one W1 row), then the cluster raises a flag
//   update       the optimizer state is 30 bytes per parameter of memory traffic, far more than four SMs can move in a few
//                microseconds (measured: 52 us when the compute cluster did it alone). So the SAME launch carries ~128 more
//                CTAs, one per SM, that do nothing but the update: they fetch their share of the parameters and moments into
//                registers while the compute cluster is still working, wait for the flag, read their share of the gradient
//                from L2, apply SGD / Adam (opt_math.cuh: the arithmetic of the flat optimizer kernel) and refresh the bf16
//                shadows. One launch, one dependency edge inside it, no second kernel.
//
// classifier.py:29-35; mlp.py:14-17; torch/optim/adam.py, sgd.py.
#include <cstdlib>

#include "opt_math.cuh"
#include "tc_common.cuh"

namespace tsb {

using namespace tc;

namespace {

constexpr int kThreads = 352;    // warp 0: X producer (bf16 inputs), 1: MMA issuer + TMEM owner, 2: W2 / W1 producer, 3..10: epilogue
constexpr int kEpiThreads = 256;
constexpr int kMaxStages = 13;   // the whole W1 slice in flight when shared memory allows
constexpr int CP = 16;           // classes padded to one UMMA N / K step
constexpr int BN = 64;           // hidden units per CTA
constexpr int kXBlocks = 14;     // 64-column blocks of X kept resident (13 hold data for D <= 832, one more is over-read only)
constexpr int kMTiles = 7;       // 128-feature tiles of dW1^T
constexpr int kZCol = 448;       // TMEM: dW1^T tiles at columns 0..447 (tile 0 reuses the H / dA accumulator), logits,
constexpr int kGwCol = 464;      //       dW2^T slice,
constexpr int kGbCol = 480;      //       db1 slice (column 15 of dH^T dZ')

struct StepParams {
    int B, D, H, C;
    int ns;                    // CTAs in the cluster = H / 64
    int rbox;                  // batch rows padded to 32 or 64: rows of every resident X block
    int num_kblocks, stages;
    const float *x_f32;        // fp32 inputs [B, D], or nullptr when the bf16 tensor map is used
    uint32_t idesc_main, idesc_z, idesc_da, idesc_g, idesc_w;
    const int64_t *y;
    int64_t *pred;
    DropoutSpec drop;
    int training;              // dropout active
    ts_batch_record *record;
    int64_t off_w1, off_b1, off_w2, off_b2;       // element offsets in the flat buffers
    int64_t sh_w1, sh_w2;                         // element offsets in the bf16 shadow buffer
    OptArgs opt;                                  // flat buffers, tensor table (gradients final in `grads`) + hyper-parameters
    uint32_t *bad_target;
    unsigned int *flag;                           // monotonic counter: += ns per step once the gradients are in `grads`
    unsigned int flag_target;
    int n_chunks;                                 // 1024-element chunks of the flat buffers
    int n_updaters;                               // CTAs behind the compute cluster
};

__device__ unsigned long long g_step_trace[32];
#define STRACE(i) do { if (blockIdx.x == 0) g_step_trace[(i)] = clock64(); } while (0)

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
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

constexpr int kUpdQuads = 4;                      // float4 quads an updater thread can own

// ---- updater CTAs: parameters and moments into registers early, then gradient -> update -> store once the flag is up
__device__ __forceinline__ void updater_role(const StepParams &p, int u) {
    const OptArgs &o = p.opt;
    const bool adam = o.kind != TS_OPT_SGD;
    const bool use_m = adam || o.momentum != 0.0f;
    const bool use_x = adam && o.amsgrad;
    const int64_t nq = (int64_t)p.n_chunks * (kChunk / 4);
    const int64_t stride = (int64_t)p.n_updaters * blockDim.x;
    float4 w[kUpdQuads], m[kUpdQuads], v[kUpdQuads], x[kUpdQuads];
    int tix[kUpdQuads];
    pdl_wait();                 // the previous step's updaters have written the state this one reads
    pdl_launch_dependents();
#pragma unroll
    for (int r = 0; r < kUpdQuads; ++r) {
        const int64_t q = (int64_t)u * blockDim.x + threadIdx.x + r * stride;
        tix[r] = -1;
        w[r] = m[r] = v[r] = x[r] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (q < nq) {
            const int t = o.chunk_tensor[q / (kChunk / 4)];
            if (t >= 0 && q * 4 - o.t[t].offset < o.t[t].count) {
                tix[r] = t;
                w[r] = *reinterpret_cast<const float4 *>(o.params + q * 4);
                if (use_m) m[r] = *reinterpret_cast<const float4 *>(o.exp_avg + q * 4);
                if (adam) v[r] = *reinterpret_cast<const float4 *>(o.exp_avg_sq + q * 4);
                if (use_x) x[r] = *reinterpret_cast<const float4 *>(o.max_exp_avg_sq + q * 4);
            }
        }
    }
    if (threadIdx.x == 0) {
        unsigned long long t0 = 0;
        for (uint32_t spins = 0;; ++spins) {
            unsigned int seen;
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.flag) : "memory");
            if ((int)(seen - p.flag_target) >= 0) break;
            if ((spins & 4095u) == 4095u) {          // only a protocol bug gets here: fail loudly instead of hanging
                unsigned long long now;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now; else if (now - t0 > 10000000000ull) __trap();
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kUpdQuads; ++r) {
        if (tix[r] < 0) continue;
        const int64_t i0 = ((int64_t)u * blockDim.x + threadIdx.x + r * stride) * 4;
        const TensorDesc d = o.t[tix[r]];
        const int64_t j0 = i0 - d.offset;
        const int n = (int)min((int64_t)4, d.count - j0);
        const float4 g4 = __ldcg(reinterpret_cast<const float4 *>(o.grads + i0));     // written by another SM: read it from L2
        const float g[4] = {g4.x, g4.y, g4.z, g4.w};
        float ww[4] = {w[r].x, w[r].y, w[r].z, w[r].w}, mm[4] = {m[r].x, m[r].y, m[r].z, m[r].w};
        float vv[4] = {v[r].x, v[r].y, v[r].z, v[r].w}, xx[4] = {x[r].x, x[r].y, x[r].z, x[r].w};
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (e < n) opt_update_element(o, g[e], ww[e], mm[e], vv[e], xx[e]);
        *reinterpret_cast<float4 *>(o.params + i0) = make_float4(ww[0], ww[1], ww[2], ww[3]);
        if (use_m) *reinterpret_cast<float4 *>(o.exp_avg + i0) = make_float4(mm[0], mm[1], mm[2], mm[3]);
        if (adam) *reinterpret_cast<float4 *>(o.exp_avg_sq + i0) = make_float4(vv[0], vv[1], vv[2], vv[3]);
        if (use_x) *reinterpret_cast<float4 *>(o.max_exp_avg_sq + i0) = make_float4(xx[0], xx[1], xx[2], xx[3]);
        if (d.shadow_off >= 0) {
            __nv_bfloat16 *sh = o.shadow + d.shadow_off + j0;
            if (n == 4 && ((reinterpret_cast<uintptr_t>(sh) & 7) == 0)) {
                uint2 pk;
                pk.x = pack_bf16(ww[0], ww[1]);
                pk.y = pack_bf16(ww[2], ww[3]);
                *reinterpret_cast<uint2 *>(sh) = pk;
            } else {
                for (int e = 0; e < n; ++e) sh[e] = __float2bfloat16_rn(ww[e]);
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads, 1)
fit_step_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_w1,
                const __grid_constant__ CUtensorMap tma_w2, const StepParams p) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages], empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t x_ready_bar, h_full_bar, w2_full_bar, a_ready_bar, z_full_bar, dz_ready_bar, da_full_bar, dh_ready_bar,
        g_full_bar;
    __shared__ float s_db2[4][CP];
    __shared__ uint32_t tmem_base_slot;
    __shared__ float s_loss[4];
    __shared__ int s_corr[4];

    // shared memory: [A tile | dH tile | dZ tile] (16 KiB each) | W2 tile 2 KiB | b1 slice | partial logits [ns][rbox][16] |
    //                resident X: kXBlocks x (rbox x 128 B) | W1 ring: stages x 8 KiB. The forward MMAs address every X block
    //                as 128 rows; what they read past the block's rbox rows (the next blocks, the head of the ring) only
    //                reaches TMEM lanes >= rbox, which nobody reads.
    uint8_t *tiles = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t tile_bytes = BM * 128;
    uint8_t *a_stage = tiles, *dh_stage = tiles + tile_bytes, *dz_tile = tiles + 2 * tile_bytes;
    uint8_t *w2_tile = tiles + 3 * tile_bytes;                      // [16 rows x 128 B]
    float *b1s = reinterpret_cast<float *>(w2_tile + 2048);
    float *zbuf = b1s + BN;
    const uint32_t xkb_bytes = (uint32_t)p.rbox * 128u;
    uint8_t *xres = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(zbuf + (p.ns > 1 ? p.ns * p.rbox * CP : 0)) + 1023) & ~(uintptr_t)1023);
    uint8_t *ring = xres + (size_t)kXBlocks * xkb_bytes;            // 1024-aligned: xkb_bytes is a multiple of 4096
    const uint32_t b_bytes = BN * 128;
    const bool w1_resident = p.stages >= p.num_kblocks;

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int ns = p.ns;
    if ((int)blockIdx.x >= ns) {           // every cluster behind the first one only updates
        updater_role(p, (int)blockIdx.x - ns);
        return;
    }
    const uint32_t rank = ns > 1 ? cluster_rank() : 0u;
    const int n0 = (int)rank * BN;
    if (threadIdx.x == 0) STRACE(0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&x_ready_bar, p.x_f32 ? kEpiThreads / 32 : 1);
        mbar_init(&h_full_bar, 1);
        mbar_init(&w2_full_bar, 1);
        mbar_init(&a_ready_bar, kEpiThreads / 32);
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
    if (ns > 1) { cluster_arrive(); cluster_wait(); }
    pdl_wait();                // the predecessor (the previous step: it wrote the weights this one reads) has completed
    pdl_launch_dependents();
    if (threadIdx.x == 0) STRACE(1);

    if (ns > 1 && warp < 3) cluster_arrive();                  // phase 1 (logit exchange): these warps only have to have arrived
    if (warp == 0) {
        // ===================== bf16 inputs: all X blocks by TMA (rows >= B and columns >= D arrive as zeros) =====================
        if (!p.x_f32 && elect_one()) {
            mbar_expect_tx(&x_ready_bar, (uint32_t)p.num_kblocks * xkb_bytes);
            for (int kb = 0; kb < p.num_kblocks; ++kb) tma_load_2d(&tma_x, &x_ready_bar, xres + (size_t)kb * xkb_bytes, kb * KATOM, 0);
        }
        __syncwarp();
    } else if (warp == 2) {
        // ===================== TMA producer: W2 tile once, then the W1 slice through the ring =====================
        if (elect_one()) {
            mbar_expect_tx(&w2_full_bar, 2048u);
            tma_load_2d(&tma_w2, &w2_full_bar, w2_tile, n0, 0);
        }
        __syncwarp();
        if (w1_resident) {
            // the whole W1 slice fits: every block goes out at once against one barrier, no stage is ever reused
            if (elect_one()) {
                mbar_expect_tx(&full_bar[0], (uint32_t)p.num_kblocks * b_bytes);
                for (int kb = 0; kb < p.num_kblocks; ++kb) tma_load_2d(&tma_w1, &full_bar[0], ring + (size_t)kb * b_bytes, kb * KATOM, n0);
            }
            __syncwarp();
        } else {
            int s = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < p.num_kblocks; ++kb) {
                if (elect_one()) {
                    mbar_wait(&empty_bar[s], phase ^ 1);
                    mbar_expect_tx(&full_bar[s], b_bytes);
                    tma_load_2d(&tma_w1, &full_bar[s], ring + (size_t)s * b_bytes, kb * KATOM, n0);
                }
                __syncwarp();
                if (++s == p.stages) { s = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            mbar_wait(&x_ready_bar, 0);
            tc_fence_after();
        }
        __syncwarp();
        int s = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < p.num_kblocks; ++kb) {
            if (elect_one()) {
                if (!w1_resident || kb == 0) {
                    mbar_wait(&full_bar[s], phase);
                    tc_fence_after();
                }
                const uint64_t adesc = make_desc(smem_u32(xres + (size_t)kb * xkb_bytes), TC_MAJOR_K, KATOM);
                const uint64_t bdesc = make_desc(smem_u32(ring + (size_t)(w1_resident ? kb : s) * b_bytes), TC_MAJOR_K, KATOM);
#pragma unroll
                for (int k = 0; k < KATOM / UMMA_K; ++k)
                    umma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), p.idesc_main, (kb > 0 || k > 0) ? 1u : 0u);
                if (!w1_resident) umma_commit(&empty_bar[s]);
                if (kb == p.num_kblocks - 1) umma_commit(&h_full_bar);
                if (kb == 0) STRACE(2);
            }
            __syncwarp();
            if (++s == p.stages) { s = 0; phase ^= 1; }
        }
        if (elect_one()) {
            STRACE(3);
            const int ksteps = p.rbox / UMMA_K;            // contractions over the batch rows stop at the padded batch
            // ---- Z = A W2^T
            mbar_wait(&w2_full_bar, 0);
            mbar_wait(&a_ready_bar, 0);
            tc_fence_after();
            const uint32_t sa = smem_u32(a_stage), sw = smem_u32(w2_tile);
            {
                const uint64_t adesc = make_desc(sa, TC_MAJOR_K, KATOM), bdesc = make_desc(sw, TC_MAJOR_K, KATOM);
#pragma unroll
                for (int k = 0; k < KATOM / UMMA_K; ++k)
                    umma_bf16(tmem_base + kZCol, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), p.idesc_z, k > 0 ? 1u : 0u);
            }
            umma_commit(&z_full_bar);
            // ---- dA = dZ W2 (K = 16 padded classes; the W2 tile re-read as an MN-major operand)
            mbar_wait(&dz_ready_bar, 0);
            tc_fence_after();
            umma_bf16(tmem_base, make_desc(smem_u32(dz_tile), TC_MAJOR_K, KATOM), make_desc(sw, TC_MAJOR_MN, CP), p.idesc_da, 0u);
            umma_commit(&da_full_bar);
            // ---- dW2^T slice = A^T dZ, db1 slice = dH^T 1 (column 15 of dZ' is 1), contraction over the batch rows
            mbar_wait(&dh_ready_bar, 0);
            tc_fence_after();
            const uint32_t sd = smem_u32(dz_tile), sh = smem_u32(dh_stage);
            for (int k = 0; k < ksteps; ++k) {
                const uint64_t bdesc = make_desc(sd + k * 2048, TC_MAJOR_MN, BM);
                umma_bf16(tmem_base + kGwCol, make_desc(sa + k * 2048, TC_MAJOR_MN, BM), bdesc, p.idesc_g, k > 0 ? 1u : 0u);
                umma_bf16(tmem_base + kGbCol, make_desc(sh + k * 2048, TC_MAJOR_MN, BM), bdesc, p.idesc_g, k > 0 ? 1u : 0u);
            }
            // ---- dW1^T slice [D x 64] = X^T dH: X blocks 2t, 2t+1 are the two 64-feature atoms of M-tile t (one block apart)
            const uint32_t sx = smem_u32(xres);
            for (int t = 0; t < kMTiles; ++t)
                for (int k = 0; k < ksteps; ++k)
                    umma_bf16(tmem_base + (uint32_t)(t * BN), make_desc(sx + (uint32_t)(2 * t) * xkb_bytes + k * 2048, TC_MAJOR_MN, p.rbox),
                              make_desc(sh + k * 2048, TC_MAJOR_MN, BM), p.idesc_w, k > 0 ? 1u : 0u);
            umma_commit(&g_full_bar);
            STRACE(10);
        }
        __syncwarp();
    } else {
        // ===================== epilogue warps =====================
        const int q = warp & 3;                          // TMEM lane quarter
        const int half = (warp - 3) >> 2;                // which 32 of this CTA's 64 hidden columns
        const int r_local = q * 32 + lane;               // batch row == TMEM lane (forward / backward phases)
        const int64_t row = r_local;
        const bool row_ok = r_local < p.B;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const int et = threadIdx.x - 96;                 // 0..255
        const int cbeg = half * 32;
        for (int i = et; i < BN; i += kEpiThreads) b1s[i] = p.opt.params[p.off_b1 + n0 + i];

        if (p.x_f32) {
            // ---- fp32 -> bf16 into the resident blocks: 8 features (one 16-byte chunk) per thread and iteration; rows >= B
            //      and features >= D become zeros (they meet zero-filled W1 columns / zero dH rows, which must not see NaNs)
            const int chunks = p.num_kblocks * 8;
            const int total = p.rbox * chunks;
#pragma unroll 4
            for (int idx = et; idx < total; idx += kEpiThreads) {
                const int r = idx / chunks, cc = idx - r * chunks;
                const int d = cc * 8;
                uint4 pk = make_uint4(0u, 0u, 0u, 0u);
                if (r < p.B && d < p.D) {
                    const float4 f0 = __ldg(reinterpret_cast<const float4 *>(p.x_f32 + (int64_t)r * p.D + d));
                    const float4 f1 = __ldg(reinterpret_cast<const float4 *>(p.x_f32 + (int64_t)r * p.D + d + 4));
                    pk.x = pack_bf16(f0.x, f0.y); pk.y = pack_bf16(f0.z, f0.w);
                    pk.z = pack_bf16(f1.x, f1.y); pk.w = pack_bf16(f1.z, f1.w);
                }
                sts128(smem_u32(xres) + (uint32_t)(cc >> 3) * xkb_bytes + (uint32_t)r * 128u + ((uint32_t)((cc & 7) ^ (r & 7)) << 4), pk);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&x_ready_bar);
        }
        named_bar_sync(1, kEpiThreads);                  // b1s visible to all epilogue threads
        if (et == 0) STRACE(4);

        // ---- dropout keep bits of this thread's row x 32 columns (drawn while the main loop runs)
        uint32_t keepb = ~0u;
        const bool dropout_on = p.training && p.drop.p > 0.0f;
        if (dropout_on && row_ok) {
            uint32_t kb = 0;
            if (p.drop.mask == nullptr) {
#pragma unroll 2
                for (int j = 0; j < 32; j += 2) {
                    const uint32_t h = dropout_bits(p.drop.key, row, (n0 + cbeg + j) >> 1);
                    kb |= ((h & 0xFFFFu) >= p.drop.thresh ? 1u : 0u) << j;
                    kb |= ((h >> 16) >= p.drop.thresh ? 1u : 0u) << (j + 1);
                }
            } else {
                const uint8_t *mrow = p.drop.mask + row * (int64_t)p.H + n0 + cbeg;
#pragma unroll 4
                for (int j = 0; j < 32; ++j) kb |= (mrow[j] != 0 ? 1u : 0u) << j;
            }
            keepb = kb;
        }
        asm volatile("" ::"r"(keepb) : "memory");

        // ---- A = relu(dropout(H + b1)) -> bf16 -> swizzled K-major tile
        uint32_t sign = 0;
        if (elect_one()) mbar_wait(&h_full_bar, 0);
        __syncwarp();
        tc_fence_after();
        if (et == 0) STRACE(5);
        {
            float v[32];
            tmem_ld16(t_lane + (uint32_t)cbeg, *reinterpret_cast<float(*)[16]>(&v[0]));
            tmem_ld16(t_lane + (uint32_t)(cbeg + 16), *reinterpret_cast<float(*)[16]>(&v[16]));
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const float keep = dropout_on ? (((keepb >> j) & 1u) ? p.drop.scale : 0.0f) : 1.0f;
                const float a = row_ok ? round_bf16(fmaxf((v[j] + b1s[cbeg + j]) * keep, 0.0f)) : 0.0f;   // rows >= B: zeros (they enter dW2^T)
                v[j] = a;
                sign |= (a > 0.0f ? 1u : 0u) << j;
            }
            uint8_t *rowp = a_stage + (size_t)r_local * 128;
            const int cbase = cbeg >> 3;
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
        if (lane == 0) mbar_arrive(&a_ready_bar);

        float loss_acc = 0.0f;
        int corr_acc = 0;
        if (half == 0) {
            // ---- logits, cross-entropy, argmax, dZ: one row per thread
            const int yy = row_ok ? (int)p.y[row] : 0;
            float b2v[CP];
#pragma unroll
            for (int c = 0; c < CP; ++c) b2v[c] = (c < p.C) ? p.opt.params[p.off_b2 + c] : 0.0f;
            if (elect_one()) mbar_wait(&z_full_bar, 0);
            __syncwarp();
            tc_fence_after();
            float z[CP];
            tmem_ld16(t_lane + (uint32_t)kZCol, z);
            if (ns > 1) {
                const bool in_box = r_local < p.rbox;             // rows past the padded batch hold nothing anybody reads
                const uint32_t mine = smem_u32(zbuf + ((size_t)rank * p.rbox + (in_box ? r_local : 0)) * CP);
                for (int t = 0; t < (ns >> 1) && in_box; ++t) {
                    const uint32_t dst = map_to_rank(mine, (uint32_t)t);
#pragma unroll
                    for (int c = 0; c < CP; c += 4) st_cluster_v4(dst + c * 4, z[c], z[c + 1], z[c + 2], z[c + 3]);
                }
                cluster_arrive();
                cluster_wait();
#pragma unroll
                for (int c = 0; c < CP; ++c) z[c] = 0.0f;
                for (int t = 0; t < ns && in_box; ++t) {
                    const float4 *src = reinterpret_cast<const float4 *>(zbuf + ((size_t)t * p.rbox + r_local) * CP);
#pragma unroll
                    for (int c = 0; c < CP; c += 4) {
                        const float4 v4 = src[c >> 2];
                        z[c] += v4.x; z[c + 1] += v4.y; z[c + 2] += v4.z; z[c + 3] += v4.w;
                    }
                }
            }
            if (et == 0) STRACE(6);
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

            const float inv_B = 1.0f / (float)p.B;
            float dzv[CP];
#pragma unroll
            for (int c = 0; c < CP; ++c)
                dzv[c] = (c < p.C && row_ok) ? (e[c] * inv_sum - (c == yy ? 1.0f : 0.0f)) * inv_B : 0.0f;
            float db2v[CP];
#pragma unroll
            for (int c = 0; c < CP; ++c) db2v[c] = dzv[c];
            dzv[CP - 1] = row_ok ? 1.0f : 0.0f;              // the "ones" column: dH^T dZ' column 15 = column sums of dH
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
            if (rank == 0) {
                // db2[c] = sum_r dz[r, c] (fp32)
#pragma unroll
                for (int c = 0; c < CP; ++c) {
                    const float t = warp_sum(db2v[c]);
                    if (lane == 0) s_db2[q][c] = t;
                }
            }
        } else if (ns > 1) {
            // the other column half's warps carry the same rows' partial logits to the upper half of the cluster
            if (elect_one()) mbar_wait(&z_full_bar, 0);
            __syncwarp();
            tc_fence_after();
            float z[CP];
            tmem_ld16(t_lane + (uint32_t)kZCol, z);
            const bool in_box = r_local < p.rbox;
            const uint32_t mine = smem_u32(zbuf + ((size_t)rank * p.rbox + (in_box ? r_local : 0)) * CP);
            for (int t = ns >> 1; t < ns && in_box; ++t) {
                const uint32_t dst = map_to_rank(mine, (uint32_t)t);
#pragma unroll
                for (int c = 0; c < CP; c += 4) st_cluster_v4(dst + c * 4, z[c], z[c + 1], z[c + 2], z[c + 3]);
            }
            cluster_arrive();
        }

        // ---- dH = dA * relu'(A) / (1-p) -> bf16 tile (rows >= B come out as zeros: their dZ is zero)
        if (elect_one()) mbar_wait(&da_full_bar, 0);
        __syncwarp();
        tc_fence_after();
        if (et == 0) STRACE(7);
        {
            const float sc = p.training ? p.drop.scale : 1.0f;
            float v[32];
            tmem_ld16(t_lane + (uint32_t)cbeg, *reinterpret_cast<float(*)[16]>(&v[0]));
            tmem_ld16(t_lane + (uint32_t)(cbeg + 16), *reinterpret_cast<float(*)[16]>(&v[16]));
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = (row_ok && ((sign >> j) & 1u)) ? v[j] * sc : 0.0f;
            uint8_t *rowp = dh_stage + (size_t)r_local * 128;
            const int cbase = cbeg >> 3;
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

        // ---- loss / correct count of the batch (single slab: cluster rank 0's four half-0 warps hold all rows)
        if (half == 0 && rank == 0) {
            loss_acc = warp_sum(loss_acc);
            corr_acc = warp_sum(corr_acc);
            if (lane == 0) { s_loss[q] = loss_acc; s_corr[q] = corr_acc; }
            named_bar_sync(2, 128);
            if (q == 3 && lane == 0) {
                p.record->loss = ((s_loss[0] + s_loss[1]) + (s_loss[2] + s_loss[3])) / (float)p.B;
                p.record->n_correct = s_corr[0] + s_corr[1] + s_corr[2] + s_corr[3];
                p.record->n_rows = (int32_t)p.B;
                p.record->reserved = 0;
            }
        }

        // ===================== gradients: TMEM -> `grads`, then the flag the updater CTAs wait for =====================
        if (elect_one()) mbar_wait(&g_full_bar, 0);
        __syncwarp();
        tc_fence_after();
        if (et == 0) STRACE(11);
        float *grads = p.opt.grads;
        // W2 slice and b1 slice: TMEM lane = hidden unit (lanes 0..63), held by the half-0 warps of quarters 0 and 1
        if (half == 0 && r_local < BN) {
            const int h = n0 + r_local;
            float gw[CP], gb[CP];
            tmem_ld16(t_lane + (uint32_t)kGwCol, gw);
            tmem_ld16(t_lane + (uint32_t)kGbCol, gb);
#pragma unroll
            for (int c = 0; c < CP; ++c)
                if (c < p.C) grads[p.off_w2 + (int64_t)c * p.H + h] = gw[c];
            grads[p.off_b1 + h] = gb[CP - 1];
        }
        if (rank == 0 && half == 0 && q == 0 && lane < p.C)           // s_db2 was filled before named barrier 2 above
            grads[p.off_b2 + lane] = (s_db2[0][lane] + s_db2[1][lane]) + (s_db2[2][lane] + s_db2[3][lane]);
        // W1 slice: M-tile t, TMEM lane = input feature d, this warp's 32 hidden columns: 32 lanes = 128 contiguous bytes
#pragma unroll 1
        for (int t = 0; t < kMTiles; ++t) {
            const int d = t * BM + r_local;
            if (t * BM + q * 32 >= p.D) break;               // warp-uniform: this quarter's features are all padding
            float g[32];
            tmem_ld16(t_lane + (uint32_t)(t * BN + cbeg), *reinterpret_cast<float(*)[16]>(&g[0]));
            tmem_ld16(t_lane + (uint32_t)(t * BN + cbeg + 16), *reinterpret_cast<float(*)[16]>(&g[16]));
            if (d < p.D) {
                float *dst = grads + p.off_w1 + (int64_t)(n0 + cbeg) * p.D + d;
#pragma unroll
                for (int j = 0; j < 32; ++j) dst[(int64_t)j * p.D] = g[j];
            }
        }
        __threadfence();                                     // this thread's gradient stores before the flag
        named_bar_sync(3, kEpiThreads);
        if (et == 0) atomicAdd(p.flag, 1u);
        if (et == 0) STRACE(12);
    }
    __syncwarp();
    if (ns > 1 && (warp < 3 || ((warp - 3) >> 2) == 1)) cluster_wait();   // complete phase 1 for the warps that only arrived
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
    if (threadIdx.x == 0) STRACE(13);
}

}  // namespace

int read_step_trace(unsigned long long *out_host) {
    TS_CUDA_CHECK(cudaMemcpyFromSymbol(out_host, g_step_trace, sizeof(unsigned long long) * 32));
    return TS_OK;
}

bool fit_step_supported(int D, int H, int C, int64_t B) {
    return (H == 64 || H == 128 || H == 256) && C >= 1 && C < CP && D % 8 == 0 && D <= (kXBlocks - 1) * KATOM && B >= 1 && B <= 64;
}

int launch_fit_step(const FitStepArgs &a, cudaStream_t st) {
    TS_REQUIRE(fit_step_supported(a.D, a.H, a.C, a.B), TS_E_UNSUPPORTED, "single-kernel step: unsupported shape D=%d H=%d C=%d B=%lld", a.D,
               a.H, a.C, (long long)a.B);
    TS_REQUIRE((a.X != nullptr) != (a.X_f32 != nullptr), TS_E_INVALID, "single-kernel step: exactly one of X (bf16) and X_f32");
    StepParams p{};
    p.B = (int)a.B; p.D = a.D; p.H = a.H; p.C = a.C;
    p.ns = a.H / BN;
    p.rbox = a.B <= 32 ? 32 : 64;
    p.num_kblocks = (a.D + KATOM - 1) / KATOM;
    p.x_f32 = a.X_f32;
    p.idesc_main = make_idesc(BM, BN, TC_MAJOR_K, TC_MAJOR_K);
    p.idesc_z = make_idesc(BM, CP, TC_MAJOR_K, TC_MAJOR_K);
    p.idesc_da = make_idesc(BM, BN, TC_MAJOR_K, TC_MAJOR_MN);
    p.idesc_g = make_idesc(BM, CP, TC_MAJOR_MN, TC_MAJOR_MN);
    p.idesc_w = make_idesc(BM, BN, TC_MAJOR_MN, TC_MAJOR_MN);
    p.y = a.y; p.pred = a.pred; p.drop = a.drop; p.training = a.training; p.record = a.record;
    p.off_w1 = a.off_w1; p.off_b1 = a.off_b1; p.off_w2 = a.off_w2; p.off_b2 = a.off_b2; p.sh_w1 = a.sh_w1; p.sh_w2 = a.sh_w2;
    p.opt = a.opt;
    TS_REQUIRE(p.opt.params && p.opt.grads && p.opt.shadow && p.opt.chunk_tensor && a.flag, TS_E_INVALID, "single-kernel step: buffers missing");
    p.flag = a.flag; p.flag_target = a.flag_target; p.n_chunks = a.n_chunks; p.bad_target = a.bad_target;
    // updater CTAs: one per SM next to the compute cluster (every CTA of the launch reserves the same ~200 KB of shared
    // memory, so they land on distinct SMs and all of them are resident: the compute cluster never waits for an SM)
    int upd_clusters = (a.sm_count - 16) / p.ns - 1;
    if (upd_clusters * p.ns > 96) upd_clusters = 96 / p.ns;      // leaves whole GPCs' worth of SMs free: the compute cluster always finds its four
    const int64_t quads = (int64_t)a.n_chunks * (kChunk / 4);
    const int need = (int)((quads + (int64_t)kThreads * kUpdQuads - 1) / ((int64_t)kThreads * kUpdQuads));
    TS_REQUIRE(upd_clusters >= 1 && upd_clusters * p.ns >= need, TS_E_UNSUPPORTED, "single-kernel step: %d updater CTAs cannot cover %lld parameters",
               upd_clusters * p.ns, (long long)quads * 4);
    p.n_updaters = upd_clusters * p.ns;

    const uint32_t xkb_bytes = (uint32_t)p.rbox * 128u;
    const size_t fixed = 3 * (size_t)BM * 128 + 2048 + (size_t)BN * 4 + (p.ns > 1 ? (size_t)p.ns * p.rbox * CP * 4 : 0) + 1024 +
                         (size_t)kXBlocks * xkb_bytes + 1024 + 64;
    p.stages = (int)((226u * 1024u - fixed) / (BN * 128));
    if (p.stages > kMaxStages) p.stages = kMaxStages;
    // the forward MMAs over-read up to 16 KiB - rbox * 128 B past the last X block: that must still be inside the ring
    TS_REQUIRE(p.stages >= 2 && (size_t)p.stages * BN * 128 >= (size_t)BM * 128, TS_E_UNSUPPORTED,
               "single-kernel step: shared memory budget does not fit H=%d B=%lld (stages %d)", a.H, (long long)a.B, p.stages);

    CUtensorMap mx, mw1, mw2;
    int rc;
    if (a.X != nullptr) rc = get_map(a.X, (uint64_t)a.D, (uint64_t)a.B, (uint64_t)a.D, KATOM, (uint32_t)p.rbox, &mx);
    else rc = get_map(a.W1, (uint64_t)a.D, (uint64_t)a.H, (uint64_t)a.D, KATOM, (uint32_t)BN, &mx);   // placeholder, never used
    if (rc) return rc;
    rc = get_map(a.W1, (uint64_t)a.D, (uint64_t)a.H, (uint64_t)a.D, KATOM, (uint32_t)BN, &mw1);
    if (rc) return rc;
    rc = get_map(a.W2p, (uint64_t)a.H, (uint64_t)CP, (uint64_t)a.H, KATOM, CP, &mw2);
    if (rc) return rc;

    const size_t smem = fixed + (size_t)p.stages * BN * 128;
    static bool attr_set[64] = {false};
    int dev = 0;
    TS_CUDA_CHECK(cudaGetDevice(&dev));
    if (!attr_set[dev & 63]) {
        TS_CUDA_CHECK(cudaFuncSetAttribute(fit_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        attr_set[dev & 63] = true;
    }
    TS_CUDA_CHECK(launch_pdl(fit_step_kernel, dim3((unsigned)(p.ns + p.n_updaters)), dim3(kThreads), smem, st, (unsigned)p.ns, mx, mw1, mw2, p));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

}  // namespace tsb
