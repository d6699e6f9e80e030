// tcgen05 / TMEM / TMA GEMM for the bf16 precision mode (sm_100a).
//
//   D[M,N] (fp32 accumulate in tensor memory) = A[M,K] * B[K,N],  bf16 operands
//
// Persistent: one CTA per SM walks the 128 x BN output tiles (BN a multiple of 16, <= 256; split-K ranges are further
// tiles). Warp roles: warp 0 = TMA producer of A, warp 2 = TMA producer of B (one elected lane each), warp 1 = TMEM
// allocator + tcgen05.mma issuer, warps 3..10 = epilogue (TMEM -> registers -> fused epilogue -> global). The
// accumulator alternates between two TMEM buffers, so a tile's epilogue overlaps the next tile's main loop.
// Operands are staged by
// cp.async.bulk.tensor (TMA) into 128-byte-swizzled shared-memory tiles behind a ring of mbarriers; the same
// tile layout serves K-major and MN-major operands, which is what lets the three contractions of a Linear layer
// run without any transposed copy:
//   forward  Y  = X  W^T : A = X  [M][K] K-major,   B = W  [N][K] K-major
//   dX          = dY W   : A = dY [M][K] K-major,   B = W  [K][N] MN-major
//   dW          = dY^T X : A = dY [K][M] MN-major,  B = X  [K][N] MN-major   (K = batch, split across CTAs)
#include <cstdlib>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "tc_common.cuh"

namespace tsb {

using namespace tc;

namespace {

constexpr int kThreads = 352;  // warp 0: A producer, 1: MMA issuer + TMEM owner, 2: B producer, 3..10: epilogue
constexpr int kEpiWarps = 8;   // two per TMEM lane quarter, taking alternate 32-column chunks
constexpr int kMaxStages = 8;

struct TcParams {
    int M, N, K, BN;
    int BNh;                   // columns of B this CTA holds: BN, or BN / 2 in a CTA pair
    int BK;                    // K elements per pipeline stage: 64 or 128
    int num_kblocks, kblocks_per_split;
    int tiles_m, tiles_n, splits;
    int a_major, b_major;
    uint32_t idesc;
    uint32_t a_bytes, b_bytes; // per stage
    int stages;
    uint32_t tmem_cols, acc_stride;   // two accumulators, acc_stride columns apart
    void *out;
    int64_t ldo, split_stride;
    const float *bias;
    const __nv_bfloat16 *A_prev;
    __nv_bfloat16 *U_out;
    int act;
    float *colsum_out;
    int64_t colsum_stride;
    DropoutSpec drop;
    int training;
    unsigned long long *dbg;   // optional trace buffer (test hook), or nullptr
    int wide_epi;              // bf16 epilogues: 64-column chunks (16-byte stores, whole 128-byte lines per row); staging 8 KiB per warp
    uint32_t epi_stage_bytes;  // staging per epilogue warp: 4096, or 8192 with wide_epi
};

// Persistent: one CTA per SM walks the tiles t = blockIdx.x, + gridDim.x, ...; the accumulator alternates between two
// TMEM buffers, so the epilogue of tile i (TMEM -> registers -> bias/dropout/ReLU or ReLU' -> global) runs while the
// producers and the MMA issuer are already in the main loop of tile i+1. With one tile per CTA the epilogue was fully
// exposed: 333 us instead of 241 us for the forward layer of the L model, 411 us for its dX layer.
// At most 128 registers per thread (352 x 128 = 45 K of the SM's 64 K): a small second kernel -- the per-layer NVLS exchange of
// data-parallel training, 256 threads x 64 registers -- can then be resident on the same SM next to a GEMM CTA.
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 b = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&b);
}

// TWO: a CTA pair (cluster of 2) works on a 256 x BN tile with tcgen05.mma.cta_group::2 -- each CTA stages its own 128 rows
// of A and HALF of B (BN / 2 columns), so the shared-memory traffic per FLOP of the B operand halves (one CTA alone moves
// 96 KB through shared memory per 64-wide K block of a 128 x 256 tile, ~750 cycles at 128 B/cycle against 512 cycles of tensor
// time; a pair moves 64 KB per CTA). The leader (cluster rank 0) owns the full barriers and issues every MMA; completion is
// committed to the barriers of both CTAs; each CTA drains its own rows of the accumulator from its own tensor memory.
template <int EPI, bool TWO>
__global__ void __maxnreg__(128)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages];
    __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t tmem_full_bar[2], tmem_empty_bar[2];
    __shared__ uint32_t tmem_base_slot;

    uint8_t *tiles = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t stage_bytes = p.a_bytes + p.b_bytes;
    uint8_t *staging = tiles + (size_t)p.stages * stage_bytes;          // 4 KiB per epilogue warp, apart from the live pipeline
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform role id
    const int katoms = p.BK / KATOM;
    const int tiles_mn = p.tiles_m * p.tiles_n, total_tiles = tiles_mn * p.splits;
    const uint32_t cta_rank = TWO ? cluster_ctarank() : 0u;
    const int unit = TWO ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;          // which CTA (pair) of the persistent grid
    const int n_units = TWO ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    constexpr int BMT = TWO ? 2 * BM : BM;                                    // rows of a tile
    unsigned long long *dbg = (p.dbg && blockIdx.x == 0) ? p.dbg : nullptr;     // CTA 0, first tile only
    unsigned long long cta_t0 = 0;
    if (dbg && threadIdx.x == 0) dbg[0] = clock64();
    if (p.dbg && threadIdx.x == 0) {   // all CTAs: earliest / latest start on the global timer (ns)
        cta_t0 = global_ns();
        atomicMin(&p.dbg[60], cta_t0);
        atomicMax(&p.dbg[62], cta_t0);
    }

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 2); mbar_init(&empty_bar[s], 1); }   // A + B producers arrive
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full_bar[b], 1); mbar_init(&tmem_empty_bar[b], TWO ? 2 * kEpiWarps : kEpiWarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_b)) : "memory");
    }
    if (TWO) cluster_sync_all();      // the peer's barriers exist before anything can signal them
    if (warp == 1) { if (TWO) tmem_alloc2(&tmem_base_slot, p.tmem_cols); else tmem_alloc(&tmem_base_slot, p.tmem_cols); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    pdl_launch_dependents();
    pdl_wait();          // prologue above overlapped the predecessor's tail; operands and outputs are touched from here on
    if (dbg && threadIdx.x == 0) dbg[1] = clock64();

    // tile t -> (split z, row block tm, column block tn); consecutive tiles share the A row block
#define TS_TILE_COORDS(t)                                                       \
    const int z_ = (t) / tiles_mn, r_ = (t) - z_ * tiles_mn;                    \
    const int tm_ = r_ / p.tiles_n, tn_ = r_ - tm_ * p.tiles_n;                 \
    const int m0 = tm_ * BMT + (int)cta_rank * BM, n0 = tn_ * p.BN;             \
    const int kb_begin = z_ * p.kblocks_per_split;                              \
    const int kb_end = min(p.num_kblocks, kb_begin + p.kblocks_per_split);      \
    (void)m0; (void)n0;

    if (warp == 0) {
        // ===================== TMA producer, operand A (whole warp loops; one elected lane issues) =====================
        int s = 0;
        uint32_t phase = 0;
        for (int t = unit, it = 0; t < total_tiles; t += n_units, ++it) {
            TS_TILE_COORDS(t)
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                if (elect_one()) {
                    mbar_wait(&empty_bar[s], phase ^ 1);
                    uint8_t *sa = tiles + (size_t)s * stage_bytes;
                    // pair: the leader arms its barrier for both CTAs' bytes; the peer's loads are counted there too
                    if (!TWO) mbar_expect_tx(&full_bar[s], p.a_bytes);
                    else if (cta_rank == 0) mbar_expect_tx(&full_bar[s], 2u * p.a_bytes);
                    const int k0 = kb * p.BK;
                    auto load = [&](void *dst, int c0, int c1) {
                        if (TWO) tma_load_2d_pair(&tma_a, &full_bar[s], dst, c0, c1); else tma_load_2d(&tma_a, &full_bar[s], dst, c0, c1);
                    };
                    if (p.a_major == TC_MAJOR_K) {
                        for (int a = 0; a < katoms; ++a)                                   // boxes 64(K) x 128(M), one per K atom
                            load(sa + (size_t)a * BM * 128, k0 + a * KATOM, m0);
                    } else {
                        load(sa, m0, k0);                                                  // two boxes 64(M) x BK(K)
                        load(sa + (size_t)p.BK * 128, m0 + 64, k0);
                    }
                    if (dbg && it == 0 && kb - kb_begin < 8) dbg[40 + (kb - kb_begin)] = clock64();
                }
                __syncwarp();
                if (++s == p.stages) { s = 0; phase ^= 1; }
            }
        }
    } else if (warp == 2) {
        // ===================== TMA producer, operand B =====================
        int s = 0;
        uint32_t phase = 0;
        for (int t = unit; t < total_tiles; t += n_units) {
            TS_TILE_COORDS(t)
            const int nb = n0 + (int)cta_rank * p.BNh;                                     // pair: this CTA's half of the tile's columns
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                if (elect_one()) {
                    mbar_wait(&empty_bar[s], phase ^ 1);
                    uint8_t *sb = tiles + (size_t)s * stage_bytes + p.a_bytes;
                    if (!TWO) mbar_expect_tx(&full_bar[s], p.b_bytes);
                    else if (cta_rank == 0) mbar_expect_tx(&full_bar[s], 2u * p.b_bytes);
                    const int k0 = kb * p.BK;
                    auto load = [&](void *dst, int c0, int c1) {
                        if (TWO) tma_load_2d_pair(&tma_b, &full_bar[s], dst, c0, c1); else tma_load_2d(&tma_b, &full_bar[s], dst, c0, c1);
                    };
                    if (p.b_major == TC_MAJOR_K) {
                        for (int a = 0; a < katoms; ++a)                                   // boxes 64(K) x BNh(N), one per K atom
                            load(sb + (size_t)a * p.BNh * 128, k0 + a * KATOM, nb);
                    } else {
                        for (int j = 0; j * 64 < p.BNh; ++j)                               // boxes 64(N) x BK(K)
                            load(sb + (size_t)j * p.BK * 128, nb + j * 64, k0);
                    }
                }
                __syncwarp();
                if (++s == p.stages) { s = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1 && cta_rank == 0) {
        // ===================== MMA issuer (pair: the leader CTA only) =====================
        int s = 0;
        uint32_t phase = 0;
        for (int t = unit, it = 0; t < total_tiles; t += n_units, ++it) {
            TS_TILE_COORDS(t)
            const int buf = it & 1;
            const uint32_t tmem_d = tmem_base + (uint32_t)buf * p.acc_stride;
            if (elect_one()) {
                mbar_wait(&tmem_empty_bar[buf], ((uint32_t)(it >> 1) & 1u) ^ 1u);   // the epilogue has drained this accumulator
                tc_fence_after();
                if (kb_begin >= kb_end) { if (TWO) umma2_commit(&tmem_full_bar[buf]); else umma_commit(&tmem_full_bar[buf]); }   // empty split range
            }
            __syncwarp();
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                if (elect_one()) {
                    mbar_wait(&full_bar[s], phase);
                    tc_fence_after();
                    if (dbg && it == 0 && kb - kb_begin < 16) dbg[8 + (kb - kb_begin)] = clock64();
                    const uint32_t sa = smem_u32(tiles + (size_t)s * stage_bytes), sb = sa + p.a_bytes;
                    for (int a = 0; a < katoms; ++a) {
                        // K-major: K atom `a` is its own [rows x 128 B] block; MN-major: 64 more K rows further down
                        const uint32_t a_at = (p.a_major == TC_MAJOR_K) ? sa + a * (BM * 128) : sa + a * (KATOM * 128);
                        const uint32_t b_at = (p.b_major == TC_MAJOR_K) ? sb + a * (p.BNh * 128) : sb + a * (KATOM * 128);
                        const uint64_t adesc = make_desc(a_at, p.a_major, p.BK), bdesc = make_desc(b_at, p.b_major, p.BK);
                        const uint32_t a_step = (p.a_major == TC_MAJOR_K) ? (UMMA_K * 2) >> 4 : (UMMA_K * 128) >> 4;
                        const uint32_t b_step = (p.b_major == TC_MAJOR_K) ? (UMMA_K * 2) >> 4 : (UMMA_K * 128) >> 4;
#pragma unroll
                        for (int k = 0; k < KATOM / UMMA_K; ++k) {
                            if (TWO) umma2_bf16(tmem_d, adesc + (uint64_t)(k * a_step), bdesc + (uint64_t)(k * b_step), p.idesc,
                                                (kb > kb_begin || a > 0 || k > 0) ? 1u : 0u);
                            else umma_bf16(tmem_d, adesc + (uint64_t)(k * a_step), bdesc + (uint64_t)(k * b_step), p.idesc,
                                           (kb > kb_begin || a > 0 || k > 0) ? 1u : 0u);
                        }
                    }
                    if (TWO) {      // frees the stage / publishes the accumulator in BOTH CTAs when these MMAs retire
                        umma2_commit(&empty_bar[s]);
                        if (kb == kb_end - 1) umma2_commit(&tmem_full_bar[buf]);
                    } else {
                        umma_commit(&empty_bar[s]);                               // frees the smem stage when these MMAs retire
                        if (kb == kb_end - 1) umma_commit(&tmem_full_bar[buf]);   // accumulator complete
                    }
                    if (dbg && it == 0 && kb - kb_begin < 16) dbg[24 + (kb - kb_begin)] = clock64();
                }
                __syncwarp();
                if (++s == p.stages) { s = 0; phase ^= 1; }
            }
        }
    } else if (warp >= 3) {
        // ===================== epilogue: TMEM -> registers -> smem (transpose) -> coalesced global =====================
        const int q = warp & 3;                  // TMEM lane quarter this warp may read
        const int half = (warp - 3) >> 2;        // which of the two warps of that quarter: chunks half, half + 2, ...
        float4 *stage = reinterpret_cast<float4 *>(staging + (size_t)(warp - 3) * p.epi_stage_bytes);
        const int sub_row = lane >> 3, piece = lane & 7;     // phase-2 mapping: 4 rows x 8 float4 per instruction
        auto release_acc = [&](int buf) {      // this warp's reads of the accumulator are done; the pair's barrier lives in the leader
            if (TWO && cta_rank != 0) mbar_arrive_remote(&tmem_empty_bar[buf], 0u); else mbar_arrive(&tmem_empty_bar[buf]);
        };
        // ---- bf16 outputs (FWD, ACTGRAD): 64-column chunks. A 32-column chunk is 64 bytes of a bf16 row: every store and every
        //      A_prev load touched half a 128-byte line, and the kernel runs at the L2's throughput -- the epilogue's
        //      half-line requests were paid at the bottleneck. Here 8 lanes cover one whole line of a row (16 bytes each); the
        //      accumulator chunk goes through an 8 KiB staging tile [32 rows][16 float4] in two TMEM rounds of 32 columns. The
        //      two halves of a row are XOR-swizzled differently so that the 32-byte reads of phase 2 stay bank-conflict free.
        if (EPI != TC_EPI_STORE_F32 && p.wide_epi) {
            for (int t = unit, it = 0; t < total_tiles; t += n_units, ++it) {
                TS_TILE_COORDS(t)
                const int buf = it & 1;
                const uint32_t tmem_acc = tmem_base + (uint32_t)buf * p.acc_stride + ((uint32_t)(q * 32) << 16);
                const bool has_k = kb_begin < kb_end;
                mbar_wait(&tmem_full_bar[buf], (uint32_t)(it >> 1) & 1u);
                tc_fence_after();
                if (half * 64 >= p.BN) {             // narrow tile: nothing for the second warp of the quarter
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) release_acc(buf);
                }
                for (int c = half * 64; c < p.BN; c += 128) {
                    const int n = n0 + c + piece * 8;                      // this lane's 8 columns in phase 2
                    const bool col_in = c + piece * 8 < p.BN && n < p.N;   // BN and the piece are multiples of 8
                    const int nv = col_in ? min(8, p.N - n) : 0;
                    const bool vec = nv == 8 && (p.ldo & 7) == 0;
                    uint4 prev[8];
                    if (EPI == TC_EPI_ACTGRAD) {      // the layer's own output (ReLU') / stored pre-activation (GELU'): issued before the TMEM rounds
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int row = m0 + q * 32 + i * 4 + sub_row;
                            prev[i] = make_uint4(0u, 0u, 0u, 0u);
                            if (vec && row < p.M) prev[i] = *reinterpret_cast<const uint4 *>(p.A_prev + (int64_t)row * p.ldo + n);
                        }
                    }
                    float bias8[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) bias8[j] = (EPI == TC_EPI_FWD && j < nv) ? p.bias[n + j] : 0.0f;
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int cc = c + 32 * hh;
                        float v[32];
                        if (has_k && cc < p.BN) {
                            tmem_ld16(tmem_acc + (uint32_t)cc, *reinterpret_cast<float(*)[16]>(&v[0]));
                            if (cc + 16 < p.BN) tmem_ld16(tmem_acc + (uint32_t)(cc + 16), *reinterpret_cast<float(*)[16]>(&v[16]));
                            else {
#pragma unroll
                                for (int j = 16; j < 32; ++j) v[j] = 0.0f;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = 0.0f;
                        }
                        // phase 1: lane = row; 16-byte pieces XOR-swizzled by row (and by the half)
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            stage[lane * 16 + 8 * hh + ((j ^ (lane & 7)) ^ hh)] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    if (c + 128 >= p.BN) {       // this warp's last TMEM read of the tile is in shared memory: hand the accumulator back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) release_acc(buf);
                    }
                    __syncwarp();
                    // phase 2: 8 consecutive lanes cover 128 contiguous bytes (64 bf16 columns) of one output row
                    float cs[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
                    const int hh2 = piece >> 2, j0 = (2 * piece) & 7;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int rr = i * 4 + sub_row;
                        const int row = m0 + q * 32 + rr;
                        const float4 t0 = stage[rr * 16 + 8 * hh2 + ((j0 ^ (rr & 7)) ^ hh2)];
                        const float4 t1 = stage[rr * 16 + 8 * hh2 + (((j0 + 1) ^ (rr & 7)) ^ hh2)];
                        if (row >= p.M || nv == 0) continue;
                        float w[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
                        __nv_bfloat16 *o = reinterpret_cast<__nv_bfloat16 *>(p.out) + (int64_t)row * p.ldo + n;
                        float pf[8];            // ACTGRAD: A_prev (or the stored pre-activation) of these 8 columns
                        if (EPI == TC_EPI_ACTGRAD) {
                            if (vec) {
                                const uint32_t pw[4] = {prev[i].x, prev[i].y, prev[i].z, prev[i].w};
#pragma unroll
                                for (int e = 0; e < 4; ++e) {      // bf16 -> fp32 is a shift
                                    pf[2 * e] = __uint_as_float(pw[e] << 16); pf[2 * e + 1] = __uint_as_float(pw[e] & 0xffff0000u);
                                }
                            } else {
                                const __nv_bfloat16 *ap = p.A_prev + (int64_t)row * p.ldo + n;
#pragma unroll
                                for (int j = 0; j < 8; ++j) pf[j] = j < nv ? __bfloat162float(ap[j]) : 0.0f;
                            }
                        }
                        if (EPI == TC_EPI_FWD) {
                            // Linear bias, Dropout, ReLU (mlp.py:14-16), all in fp32 on the accumulator
                            float k0[4] = {1.0f, 1.0f, 1.0f, 1.0f}, k1[4] = {1.0f, 1.0f, 1.0f, 1.0f};
                            if (p.training) {
                                p.drop.factor4((int64_t)row, n, p.N, k0);
                                p.drop.factor4((int64_t)row, n + 4, p.N, k1);
                            }
                            const float keep[8] = {k0[0], k0[1], k0[2], k0[3], k1[0], k1[1], k1[2], k1[3]};
                            if (p.act == TS_ACT_RELU) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) w[j] = fmaxf((w[j] + bias8[j]) * keep[j], 0.0f);
                            } else {
                                // GELU: the post-dropout pre-activation is kept (bf16) for the backward pass
                                float u[8];
#pragma unroll
                                for (int j = 0; j < 8; ++j) { u[j] = (w[j] + bias8[j]) * keep[j]; w[j] = gelu_f(u[j]); }
                                __nv_bfloat16 *uo = p.U_out + (int64_t)row * p.ldo + n;
                                if (vec) {
                                    *reinterpret_cast<uint4 *>(uo) = make_uint4(pack_bf16x2(u[0], u[1]), pack_bf16x2(u[2], u[3]),
                                                                                pack_bf16x2(u[4], u[5]), pack_bf16x2(u[6], u[7]));
                                } else {
                                    #pragma unroll
                                    for (int j = 0; j < 8; ++j)
                                        if (j < nv) uo[j] = __float2bfloat16_rn(u[j]);
                                }
                            }
                        } else if (p.act != TS_ACT_RELU) {
                            // dH = dA * gelu'(u) * keep/(1-p), u = the stored pre-activation (a dropped unit has u = 0 and
                            // gelu'(0) = 1/2, so the keep factor is applied explicitly)
                            float k0[4] = {1.0f, 1.0f, 1.0f, 1.0f}, k1[4] = {1.0f, 1.0f, 1.0f, 1.0f};
                            if (p.training) {
                                p.drop.factor4((int64_t)row, n, p.N, k0);
                                p.drop.factor4((int64_t)row, n + 4, p.N, k1);
                            }
                            const float keep[8] = {k0[0], k0[1], k0[2], k0[3], k1[0], k1[1], k1[2], k1[3]};
#pragma unroll
                            for (int j = 0; j < 8; ++j) w[j] = w[j] * gelu_grad_f(pf[j]) * keep[j];
                        } else {
                            // dH = dA * relu'(A_prev) / (1-p): A_prev > 0 implies the unit was kept
                            const float sc = p.training ? p.drop.scale : 1.0f;
#pragma unroll
                            for (int j = 0; j < 8; ++j) w[j] = (pf[j] > 0.0f) ? w[j] * sc : 0.0f;
                        }
                        uint32_t pk[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            pk[e] = pack_bf16x2(w[2 * e], w[2 * e + 1]);
                            if (EPI == TC_EPI_ACTGRAD) {   // sums of the ROUNDED values: what a column sum of the stored dH gives
                                if (2 * e < nv) cs[2 * e] += __uint_as_float(pk[e] << 16);
                                if (2 * e + 1 < nv) cs[2 * e + 1] += __uint_as_float(pk[e] & 0xffff0000u);
                            }
                        }
                        if (vec) {
                            *reinterpret_cast<uint4 *>(o) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        } else {
                            #pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (j < nv) o[j] = __float2bfloat16_rn(w[j]);
                        }
                    }
                    if (EPI == TC_EPI_ACTGRAD && p.colsum_out != nullptr) {
                        // the 4 lanes that share a column group (sub_row 0..3) fold their 8 rows each: 32 rows per warp
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 8);
                            cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 16);
                        }
                        if (sub_row == 0 && nv > 0) {
                            float *o = p.colsum_out + (int64_t)((m0 / BM) * 4 + q) * p.colsum_stride + n;
                            #pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (j < nv) o[j] = cs[j];
                        }
                    }
                    __syncwarp();
                }
            }
        } else
        for (int t = unit, it = 0; t < total_tiles; t += n_units, ++it) {
            TS_TILE_COORDS(t)
            const int buf = it & 1;
            const uint32_t tmem_acc = tmem_base + (uint32_t)buf * p.acc_stride + ((uint32_t)(q * 32) << 16);
            const bool has_k = kb_begin < kb_end;    // an empty split range contributes zeros
            mbar_wait(&tmem_full_bar[buf], (uint32_t)(it >> 1) & 1u);
            tc_fence_after();
            if (dbg && it == 0 && warp == 3 && lane == 0) dbg[6] = clock64();
            // Global operands of the epilogue (the layer's own output for ReLU') are fetched one 32-column chunk ahead
            // and for all 8 row groups at once: issue is in order, so a load used right away costs a memory round trip.
            uint2 prev_next[8];
            auto fetch_prev = [&](int c) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int row = m0 + q * 32 + i * 4 + sub_row;
                    const int n = n0 + c + piece * 4;
                    prev_next[i] = make_uint2(0u, 0u);
                    if (c < p.BN && row < p.M && n + 4 <= p.N && c + piece * 4 < p.BN && (p.ldo & 3) == 0)
                        prev_next[i] = *reinterpret_cast<const uint2 *>(p.A_prev + (int64_t)row * p.ldo + n);
                }
            };
            if (half * 32 >= p.BN) {             // narrow tile: nothing for the second warp of the quarter
                tc_fence_before();
                __syncwarp();
                if (lane == 0) release_acc(buf);
            }
            if (EPI == TC_EPI_ACTGRAD) fetch_prev(half * 32);
            for (int c = half * 32; c < p.BN; c += 64) {
                uint2 prev[8];
                if (EPI == TC_EPI_ACTGRAD) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) prev[i] = prev_next[i];
                    fetch_prev(c + 64);
                }
                float v[32];
                if (has_k) {
                    tmem_ld16(tmem_acc + (uint32_t)c, *reinterpret_cast<float(*)[16]>(&v[0]));
                    if (c + 16 < p.BN) tmem_ld16(tmem_acc + (uint32_t)(c + 16), *reinterpret_cast<float(*)[16]>(&v[16]));
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = 0.0f;
                }
                if (c + 64 >= p.BN) {
                    // this warp's last TMEM read of the tile is in registers: hand the accumulator back to the MMA issuer
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) release_acc(buf);
                }
                // phase 1: lane = row; 16-byte pieces XOR-swizzled by row so both phases are bank-conflict free
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    stage[lane * 8 + (j ^ (lane & 7))] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                __syncwarp();
                // phase 2: 8 consecutive lanes cover 128 contiguous bytes of one output row.
                float cs[4] = {0.0f, 0.0f, 0.0f, 0.0f};      // ACTGRAD: column sums of what this lane stores (bias gradient)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int rr = i * 4 + sub_row;
                    const int row = m0 + q * 32 + rr;
                    const int n = n0 + c + piece * 4;
                    float4 tv = stage[rr * 8 + (piece ^ (rr & 7))];
                    if (row >= p.M || n >= p.N || c + piece * 4 >= p.BN) continue;
                    float w[4] = {tv.x, tv.y, tv.z, tv.w};
                    const int nv = min(4, p.N - n);
                    if (EPI == TC_EPI_STORE_F32) {
                        float *o = reinterpret_cast<float *>(p.out) + (int64_t)z_ * p.split_stride + (int64_t)row * p.ldo + n;
                        if (nv == 4 && (p.ldo & 3) == 0) *reinterpret_cast<float4 *>(o) = tv;
                        else for (int j = 0; j < nv; ++j) o[j] = w[j];
                    } else {
                        __nv_bfloat16 *o = reinterpret_cast<__nv_bfloat16 *>(p.out) + (int64_t)row * p.ldo + n;
                        if (EPI == TC_EPI_FWD) {
                            // Linear bias, Dropout, ReLU (mlp.py:14-16), all in fp32 on the accumulator
                            float keep[4] = {1.0f, 1.0f, 1.0f, 1.0f};
                            if (p.training) p.drop.factor4((int64_t)row, n, p.N, keep);
                            if (p.act == TS_ACT_RELU) {
#pragma unroll
                                for (int j = 0; j < 4; ++j)
                                    if (j < nv) w[j] = fmaxf((w[j] + p.bias[n + j]) * keep[j], 0.0f);
                            } else {
                                // GELU: the post-dropout pre-activation is kept (bf16) for the backward pass
                                float u[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                                for (int j = 0; j < 4; ++j)
                                    if (j < nv) { u[j] = (w[j] + p.bias[n + j]) * keep[j]; w[j] = gelu_f(u[j]); }
                                __nv_bfloat16 *uo = p.U_out + (int64_t)row * p.ldo + n;
                                if (nv == 4 && (p.ldo & 3) == 0) {
                                    const __nv_bfloat162 ulo = __floats2bfloat162_rn(u[0], u[1]), uhi = __floats2bfloat162_rn(u[2], u[3]);
                                    uint2 upk;
                                    upk.x = *reinterpret_cast<const uint32_t *>(&ulo);
                                    upk.y = *reinterpret_cast<const uint32_t *>(&uhi);
                                    *reinterpret_cast<uint2 *>(uo) = upk;
                                } else {
                                    for (int j = 0; j < nv; ++j) uo[j] = __float2bfloat16_rn(u[j]);
                                }
                            }
                        } else if (p.act != TS_ACT_RELU) {
                            // dH = dA * gelu'(u) * keep/(1-p), u = the stored pre-activation (a dropped unit has u = 0 and
                            // gelu'(0) = 1/2, so the keep factor is applied explicitly)
                            float keep[4] = {1.0f, 1.0f, 1.0f, 1.0f};
                            if (p.training) p.drop.factor4((int64_t)row, n, p.N, keep);
                            float u[4];
                            if (nv == 4 && (p.ldo & 3) == 0) {
                                const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162 *>(&prev[i].x);
                                const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162 *>(&prev[i].y);
                                u[0] = __low2float(lo); u[1] = __high2float(lo); u[2] = __low2float(hi); u[3] = __high2float(hi);
                            } else {
                                const __nv_bfloat16 *ap = p.A_prev + (int64_t)row * p.ldo + n;
                                for (int j = 0; j < 4; ++j) u[j] = j < nv ? __bfloat162float(ap[j]) : 0.0f;
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j) w[j] = w[j] * gelu_grad_f(u[j]) * keep[j];
                        } else {
                            // dH = dA * relu'(A_prev) / (1-p): A_prev > 0 implies the unit was kept
                            const float sc = p.training ? p.drop.scale : 1.0f;
                            if (nv == 4 && (p.ldo & 3) == 0) {
                                const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162 *>(&prev[i].x);
                                const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162 *>(&prev[i].y);
                                w[0] = (__low2float(lo) > 0.0f) ? w[0] * sc : 0.0f;
                                w[1] = (__high2float(lo) > 0.0f) ? w[1] * sc : 0.0f;
                                w[2] = (__low2float(hi) > 0.0f) ? w[2] * sc : 0.0f;
                                w[3] = (__high2float(hi) > 0.0f) ? w[3] * sc : 0.0f;
                            } else {
                                const __nv_bfloat16 *ap = p.A_prev + (int64_t)row * p.ldo + n;
                                for (int j = 0; j < nv; ++j) w[j] = (__bfloat162float(ap[j]) > 0.0f) ? w[j] * sc : 0.0f;
                            }
                        }
                        __nv_bfloat162 lo = __floats2bfloat162_rn(w[0], w[1]), hi = __floats2bfloat162_rn(w[2], w[3]);
                        if (EPI == TC_EPI_ACTGRAD) {      // sums of the ROUNDED values: what a column sum of the stored dH gives
                            cs[0] += __low2float(lo); cs[1] += __high2float(lo); cs[2] += __low2float(hi); cs[3] += __high2float(hi);
                        }
                        if (nv == 4 && (p.ldo & 3) == 0) {
                            uint2 pk;
                            pk.x = *reinterpret_cast<uint32_t *>(&lo);
                            pk.y = *reinterpret_cast<uint32_t *>(&hi);
                            *reinterpret_cast<uint2 *>(o) = pk;
                        } else {
                            for (int j = 0; j < nv; ++j) o[j] = __float2bfloat16_rn(w[j]);
                        }
                    }
                }
                if (EPI == TC_EPI_ACTGRAD && p.colsum_out != nullptr) {
                    // the 4 lanes that share a column quad (sub_row 0..3) fold their 8 rows each: 32 rows per warp
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 8);
                        cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 16);
                    }
                    const int n = n0 + c + piece * 4;
                    if (sub_row == 0 && c + piece * 4 < p.BN) {
                        float *o = p.colsum_out + (int64_t)((m0 / BM) * 4 + q) * p.colsum_stride + n;
                        for (int j = 0; j < 4; ++j)
                            if (n + j < p.N) o[j] = cs[j];
                    }
                }
                __syncwarp();
            }
            if (dbg && it == 0 && warp == 3 && lane == 0) dbg[7] = clock64();
        }
    }
#undef TS_TILE_COORDS
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (TWO) cluster_sync_all();      // neither CTA leaves (or frees tensor memory) while the other may still use its shared memory
    if (warp == 1) {
        tc_fence_after();
        if (TWO) tmem_dealloc2(tmem_base, p.tmem_cols); else tmem_dealloc(tmem_base, p.tmem_cols);
    }
    if (p.dbg && threadIdx.x == 0) {
        const unsigned long long t = global_ns();
        atomicMax(&p.dbg[61], t);
        atomicMax(&p.dbg[63], t - cta_t0);                 // slowest CTA (ns)
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
    uint32_t box_inner, box_outer, elem_bytes;
    bool operator==(const MapKey &o) const {
        return ptr == o.ptr && inner == o.inner && outer == o.outer && pitch == o.pitch && box_inner == o.box_inner &&
               box_outer == o.box_outer && elem_bytes == o.elem_bytes;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey &k) const {
        uint64_t h = (uint64_t)(uintptr_t)k.ptr * 0x9E3779B97F4A7C15ull;
        h ^= (k.inner * 0xC2B2AE3D27D4EB4Full) ^ (k.outer << 17) ^ (k.pitch << 29) ^ ((uint64_t)k.box_outer << 41) ^ k.elem_bytes;
        return (size_t)(h ^ (h >> 31));
    }
};
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;   // per-process cache: loaders hand out a bounded set of batch views
std::mutex g_maps_mu;

}  // namespace

int get_map(const void *ptr, uint64_t inner, uint64_t outer, uint64_t pitch_elems, uint32_t box_inner, uint32_t box_outer,
            CUtensorMap *out, uint32_t elem_bytes) {
    const MapKey key{ptr, inner, outer, pitch_elems, box_inner, box_outer, elem_bytes};
    std::lock_guard<std::mutex> lock(g_maps_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) { *out = it->second; return TS_OK; }
    EncodeTiledFn fn = encode_fn();
    TS_REQUIRE(fn != nullptr, TS_E_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    TS_REQUIRE(elem_bytes == 2 || elem_bytes == 4, TS_E_INVALID, "TMA element size %u", elem_bytes);
    TS_REQUIRE(((uintptr_t)ptr & 15) == 0 && (pitch_elems * elem_bytes) % 16 == 0, TS_E_INVALID,
               "TMA operand needs a 16-byte aligned base and row pitch (ptr %p, pitch %llu elements)", ptr,
               (unsigned long long)pitch_elems);
    const cuuint64_t dims[2] = {inner, outer};
    const cuuint64_t strides[1] = {pitch_elems * elem_bytes};
    const cuuint32_t box[2] = {box_inner, box_outer};
    const cuuint32_t estr[2] = {1, 1};
    CUtensorMap m;
    const CUresult r = fn(&m, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                          const_cast<void *>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    TS_REQUIRE(r == CUDA_SUCCESS, TS_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (dims %llu x %llu, box %u x %u)",
               (int)r, (unsigned long long)inner, (unsigned long long)outer, box_inner, box_outer);
    if (g_maps.size() > 8192) g_maps.clear();
    g_maps.emplace(key, m);
    *out = m;
    return TS_OK;
}

namespace {

// Tile width from a small cost model fitted to in-kernel clock traces on B200 (cycles per CTA):
//   ~2000 prologue + per 64 K-elements max(2*BN tensor time, issue cadence: ~440 with 128-wide stages, ~684 with
//   64-wide stages) + ~12 per output column of epilogue; total = waves * per-CTA. Widths that would pad the
//   problem by more than 12.5 % are not considered.
// A CTA pair on a 256 x 256 tile runs close to the tensor time (~560 cycles per 64 K-elements measured against 684 for one
// CTA on 128 x 256) and is costed next to the single-CTA widths -- without the padding rule: a padded tile that saves a wave
// is still a win (dW1 of the 784 -> 4096 layer: N = 784 as 4 x 256 in ONE wave of pairs instead of two waves of 112-wide tiles).
int pick_bn(int M, int N, int K, int splits, int sm_count, bool pairs_allowed, bool *two) {
    static const int cand[] = {256, 224, 192, 160, 128, 112, 96, 80, 64, 48, 32, 16};
    *two = false;
    if (const char *f = getenv("TSB_FORCE_BN")) {   // experiments; TSB_FORCE_PAIR=1: CTA pairs at that width
        const char *fp = getenv("TSB_FORCE_PAIR");
        *two = pairs_allowed && M >= 2 * BM && (atoi(f) == 256 || (fp && atoi(fp) != 0));
        return atoi(f);
    }
    const int tiles_m = (M + BM - 1) / BM;
    const long k64 = ((K + KATOM - 1) / KATOM + splits - 1) / splits;
    int best = 16;
    double best_cost = 1e30;
    for (int bn : cand) {
        const int tiles_n = (N + bn - 1) / bn;
        const int waste = tiles_n * bn - N;
        if (waste * 8 > N && bn > 16) continue;
        const long ctas = (long)tiles_m * tiles_n * splits;
        const long waves = (ctas + sm_count - 1) / sm_count;
        const double cadence = bn <= 128 ? 440.0 : 684.0;
        const double per_cta = 2000.0 + (double)k64 * (2.0 * bn > cadence ? 2.0 * bn : cadence) + 12.0 * bn;
        const double cost = (double)waves * per_cta;
        if (cost < best_cost - 1e-9) { best_cost = cost; best = bn; }
    }
    if (pairs_allowed && M >= 2 * BM && N > 128 && sm_count >= 2) {
        const long tiles = (long)((M + 2 * BM - 1) / (2 * BM)) * ((N + 255) / 256) * splits;
        const long units = sm_count / 2;
        const long waves = (tiles + units - 1) / units;
        const double cost = (double)waves * (2000.0 + (double)k64 * 560.0 + 12.0 * 256);
        if (cost < best_cost) { best_cost = cost; best = 256; *two = true; }
    }
    return best;
}

}  // namespace

int launch_tc_gemm(const TcGemmArgs &a, int sm_count, cudaStream_t st) {
    TS_REQUIRE(a.M >= 1 && a.N >= 1 && a.K >= 1, TS_E_INVALID, "tc gemm: bad shape %d x %d x %d", a.M, a.N, a.K);
    TS_REQUIRE(a.splits >= 1 && (a.splits == 1 || a.epi == TC_EPI_STORE_F32), TS_E_INVALID, "split-K needs the fp32 store epilogue");
    TS_REQUIRE(a.k_chunk % KATOM == 0, TS_E_INVALID, "tc gemm: k_chunk %d is not a multiple of %d", a.k_chunk, KATOM);
    TS_REQUIRE(a.epi != TC_EPI_FWD || a.act == TS_ACT_RELU || a.U_out != nullptr, TS_E_INVALID, "tc gemm: the GELU forward epilogue needs U_out");
    TcParams p{};
    p.M = a.M; p.N = a.N; p.K = a.K;
    p.a_major = a.a_major; p.b_major = a.b_major;
    const int splits = a.splits;
    // CTA pairs for the widest tile: there one CTA alone is bound by shared-memory bandwidth, not by the tensor pipe
    static const int pair_mode = [] { const char *e = getenv("TSB200_GEMM_PAIR"); return e ? atoi(e) : 1; }();
    bool two = false;
    p.BN = pick_bn(a.M, a.N, a.K, splits, sm_count, pair_mode != 0, &two);
    p.BNh = two ? p.BN / 2 : p.BN;
    const int BMT = two ? 2 * BM : BM;
    // K range per split in elements (a multiple of 64), then K per pipeline stage: 128 for narrow tiles, where the
    // per-stage issue overhead (~300 cycles of barrier + TMA issue) dominates, 64 for wide ones (MMA-bound)
    const int k_atoms = (a.K + KATOM - 1) / KATOM;
    const int atoms_per_split = a.k_chunk > 0 ? a.k_chunk / KATOM : (k_atoms + splits - 1) / splits;
    TS_REQUIRE((long)splits * atoms_per_split >= k_atoms, TS_E_INVALID, "tc gemm: %d splits of %d K-atoms do not cover K=%d", splits,
               atoms_per_split, a.K);
    p.BK = (p.BN <= 128 && k_atoms >= 2 && (splits == 1 || atoms_per_split % 2 == 0)) ? 128 : 64;
    p.num_kblocks = (a.K + p.BK - 1) / p.BK;
    p.kblocks_per_split = (atoms_per_split * KATOM + p.BK - 1) / p.BK;   // split boundaries stay stage-aligned
    p.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a.a_major << 15) | ((uint32_t)a.b_major << 16) |
              ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(BMT >> 4) << 24);
    p.a_bytes = (uint32_t)BM * (uint32_t)p.BK * 2u;                       // per CTA
    p.b_bytes = (a.b_major == TC_MAJOR_K) ? (uint32_t)p.BNh * (uint32_t)p.BK * 2u
                                          : (uint32_t)((p.BNh + 63) / 64) * (uint32_t)(p.BK * 128);
    const uint32_t stage = p.a_bytes + p.b_bytes;     // always a multiple of 1024
    p.tiles_m = (a.M + BMT - 1) / BMT; p.tiles_n = (a.N + p.BN - 1) / p.BN; p.splits = splits;
    const long total_tiles = (long)p.tiles_m * p.tiles_n * splits;
    const int units = two ? sm_count / 2 : sm_count;                      // CTAs, or CTA pairs
    const int grid_units = (int)(total_tiles < units ? total_tiles : units);
    const int grid_x = two ? 2 * grid_units : grid_units;
    const long tiles_per_cta = (total_tiles + grid_units - 1) / grid_units;
    // pipeline stages + 32 (64: bf16 epilogues with 64-column chunks) KiB of epilogue staging (the pipeline stays live while a
    // tile's epilogue runs) + alignment
    static const int wide_mode = [] { const char *e = getenv("TSB200_GEMM_WIDE_EPI"); return e ? atoi(e) : 1; }();
    p.wide_epi = (wide_mode != 0 && a.epi != TC_EPI_STORE_F32 && two) ? 1 : 0;   // pairs: 32 KiB stages, five of them still fit
    p.epi_stage_bytes = p.wide_epi ? 8192u : 4096u;
    p.stages = (int)((226u * 1024u - 1024u - (uint32_t)kEpiWarps * p.epi_stage_bytes) / stage);
    if (p.stages > kMaxStages) p.stages = kMaxStages;
    const long kblocks_per_cta = (long)p.kblocks_per_split * tiles_per_cta;
    if (p.stages > kblocks_per_cta) p.stages = kblocks_per_cta < 2 ? 2 : (int)kblocks_per_cta;
    p.acc_stride = 32;
    while ((int)p.acc_stride < p.BN) p.acc_stride <<= 1;
    p.tmem_cols = 2 * p.acc_stride;
    p.out = a.out; p.ldo = a.ldo; p.split_stride = a.split_stride;
    p.bias = a.bias; p.A_prev = a.A_prev; p.U_out = a.U_out; p.act = a.act; p.colsum_out = a.colsum_out; p.colsum_stride = a.colsum_stride;
    p.drop = a.drop; p.training = a.training;
    p.dbg = a.dbg;

    CUtensorMap ma, mb;
    int rc;
    if (a.a_major == TC_MAJOR_K) rc = get_map(a.A, (uint64_t)a.K, (uint64_t)a.M, (uint64_t)a.lda, KATOM, BM, &ma);
    else rc = get_map(a.A, (uint64_t)a.M, (uint64_t)a.K, (uint64_t)a.lda, 64, (uint32_t)p.BK, &ma);
    if (rc) return rc;
    if (a.b_major == TC_MAJOR_K) rc = get_map(a.B, (uint64_t)a.K, (uint64_t)a.N, (uint64_t)a.ldb, KATOM, (uint32_t)p.BNh, &mb);
    else rc = get_map(a.B, (uint64_t)a.N, (uint64_t)a.K, (uint64_t)a.ldb, 64, (uint32_t)p.BK, &mb);
    if (rc) return rc;

    const size_t smem = (size_t)p.stages * stage + (size_t)kEpiWarps * p.epi_stage_bytes + 1024;
    dim3 grid((unsigned)grid_x);
#define TS_TC_LAUNCH_K(KERNEL, CLUSTER)                                                                              \
    do {                                                                                                            \
        static bool attr_set[64] = {false};                                                                         \
        int dev_ = 0;                                                                                               \
        TS_CUDA_CHECK(cudaGetDevice(&dev_));                                                                        \
        if (!attr_set[dev_ & 63]) {                                                                                 \
            TS_CUDA_CHECK(cudaFuncSetAttribute(KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));   \
            attr_set[dev_ & 63] = true;                                                                             \
        }                                                                                                           \
        TS_CUDA_CHECK(launch_pdl(KERNEL, grid, dim3(kThreads), smem, st, CLUSTER, ma, mb, p));                      \
    } while (0)
#define TS_TC_LAUNCH(E)                                                                                             \
    do {                                                                                                            \
        if (two) TS_TC_LAUNCH_K((tc_gemm_kernel<E, true>), 2); else TS_TC_LAUNCH_K((tc_gemm_kernel<E, false>), 1);  \
    } while (0)
    if (a.epi == TC_EPI_STORE_F32) TS_TC_LAUNCH(TC_EPI_STORE_F32);
    else if (a.epi == TC_EPI_FWD) TS_TC_LAUNCH(TC_EPI_FWD);
    else TS_TC_LAUNCH(TC_EPI_ACTGRAD);
#undef TS_TC_LAUNCH
#undef TS_TC_LAUNCH_K
    TS_LAUNCH_CHECK();
    return TS_OK;
}

}  // namespace tsb
