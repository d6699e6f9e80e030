// dW1 = dH^T X and the optimizer update of W1 in one kernel, for the one-hidden-layer fused path at training batches
// (classifier.py:31-34: backward + optimizer.step). The batch is the contraction, so it is split over the CTAs of a
// thread-block cluster; what used to be split-K partials in global memory, folded by a separate optimizer launch, is now
//
//   main loop   every CTA of the cluster (rank z) contracts its batch rows [z * k_chunk, ...) of one 128 x 128 tile of dW1:
//               tcgen05.mma, dH and X tiles by TMA as MN-major operands (no transposed copy), accumulator in TMEM
//   reduce      the accumulators go to shared memory, one cluster barrier, then rank z sums rows [z * 128/S, ...) of the tile over
//               all S CTAs through distributed shared memory, in rank order (bit-reproducible)
//   update      SGD / Adam (opt_math.cuh) on those elements: parameters and moments were fetched into registers while the
//               main loop ran; the new master weights, moments, `grads` and the bf16 shadow go straight out
//
// One more cluster of the same launch does what is left of the optimizer kernel: it folds the per-slab partials of W2, b1 and
// b2 that fused_mlp.cu produced and updates them. No gradient of W1 ever reaches global memory except as `param.grad`.
//
// Data parallel (DP = true, dp_mode 'p2p'): the cross-rank gradient sum happens in the same launch. After the cluster reduce
// every CTA stores its share of this rank's gradient into the rank's peer-visible buffer (NVLink / NVSwitch mappings of every
// rank's buffer are in DpArgs); the last CTA of the grid to do so publishes "step s complete" into every rank's flag array
// (st.release.sys), everybody waits until the local flag array shows step s for every rank (ld.acquire.sys, bounded by
// wall-clock time), then re-reads ITS OWN quads from all ranks' buffers, sums them in rank order (bit-identical replicas),
// scales by 1/world and updates. What the separate dp_optimizer_kernel launch cost (a kernel boundary, a fold of eight split
// partials, a second pass over the optimizer state) is gone; what remains is the handshake and the remote reads.
#include <cstdlib>

#include "opt_math.cuh"
#include "tc_common.cuh"

namespace tsb {

using namespace tc;

namespace {

constexpr int kThreads = 352;     // warp 0: dH producer, 1: MMA issuer + TMEM owner, 2: X producer, 3..10: epilogue / reduce / update
constexpr int kEpiThreads = 256;
constexpr int kStages = 5;        // 32 KiB each
constexpr int TN = 128;           // features per tile
constexpr int BK = 64;            // batch rows per pipeline stage

struct WgParams {
    int H, D;
    int64_t B;
    int S;                        // cluster size = number of batch splits
    int tiles_n, n_tiles;         // feature tiles per hidden-row tile; tiles in all
    int kblocks, kblocks_per_split;
    uint32_t idesc;
    int64_t off_w1, sh_w1;
    int small_chunk_begin, small_chunks;   // 1024-element chunks of the other tensors (b1, W2, b2), folded by the last cluster
    OptArgs opt;
    DpArgs dp;                    // DP = true: peer mappings of every rank's gradient buffer (this step's parity) and flag array
};

// Data parallel handshake of one 256-thread group (tid 0..255, named barrier `bar`): every thread has stored its share of
// this rank's gradient. The last CTA of the grid to get here publishes the step into every rank's flag array; then the
// group waits until its own rank's flag array shows the step for every rank. Returns false if a peer never arrived within
// dp.timeout_ns (the error word is set; the caller skips its update).
__device__ __forceinline__ bool dp_publish_and_wait(const DpArgs &dp, int tid, int bar, bool *s_last, bool *s_timed_out) {
    uint32_t *own = dp.flags[dp.rank];
    __threadfence();                                               // this thread's gradient stores, device-wide, before the ticket
    named_bar_sync(bar, 256);
    if (tid == 0) {
        *s_timed_out = false;
        const uint32_t arrived = atomicAdd(own + 8, 1u);
        *s_last = arrived == gridDim.x - 1;
        if (*s_last) own[8] = 0u;                                  // the next step's kernel starts from zero
    }
    named_bar_sync(bar, 256);
    if (*s_last && tid < dp.world)    // one thread per destination rank: the system-scope releases run side by side
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dp.flags[tid] + dp.rank), "r"(dp.step) : "memory");
    if (tid < dp.world && !wait_for_flag(own + tid, dp.step, dp)) *s_timed_out = true;
    named_bar_sync(bar, 256);
    return !*s_timed_out;
}

// ---- low-latency exchange (ts_dp_bind_ll). The pull exchange above crosses NVLink three times on the critical path (publish,
// remote load request, reply) behind a grid-wide ticket. Here the data carries its own validity, as in a collective library's
// low-latency protocol: a 16-byte packet {three floats, step number} written with ONE store; the receiver polls its own memory
// until the fourth word shows this step (a 16-byte aligned store or load is a single transaction, so a packet is seen whole
// or not at all: a hardware property -- the PTX memory model only promises it per scalar -- of the kind the 128-byte
// low-latency protocols of collective libraries rely on; the replicas are compared bit for bit by tests and bench). No fence, no separate flag, no ticket, and a late CTA delays only its own share. Two hops keep the wire
// volume flat in the world size: CTA c's share of the gradient is owned by rank c % world; the other ranks send their share
// to the owner's inbox, the owner adds them in rank order (its own term in its place) and sends the sum to everybody -- all
// replicas apply bit-identical gradients. (A world of two skips the second hop: each rank sends its share to the other and both
// add the same two terms in the same order.) Slots alternate with the step parity: a rank can be one step ahead of a peer's
// reads, never two (to finish a step it needs every peer's packets of that step, sent after they had read the step before).
__device__ __forceinline__ void ll_store(float4 *dst, float a, float b, float c, uint32_t step) {
    asm volatile("st.volatile.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(__uint_as_float(step)) : "memory");
}
__device__ __forceinline__ float4 ll_load(const float4 *src) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(src) : "memory");
    return v;
}
// packets k = 0..NQ-1 of up to NB sources (nullptr: skip), all loads in flight together; polls until every packet shows `step`
template <int NB, int NQ>
__device__ __forceinline__ bool ll_recv(const float4 *const (&src)[NB], float4 (&v)[NB][NQ], uint32_t step, const DpArgs &dp) {
    uint32_t pending = 0;
#pragma unroll
    for (int b = 0; b < NB; ++b)
#pragma unroll
        for (int k = 0; k < NQ; ++k) {
            v[b][k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (src[b] != nullptr) pending |= 1u << (b * NQ + k);
        }
    unsigned long long t0 = 0;
    for (uint32_t spins = 0; pending != 0u; ++spins) {
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int k = 0; k < NQ; ++k)
                if ((pending >> (b * NQ + k)) & 1u) v[b][k] = ll_load(src[b] + k * 256);
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int k = 0; k < NQ; ++k)
                if (((pending >> (b * NQ + k)) & 1u) && __float_as_uint(v[b][k].w) == step) pending &= ~(1u << (b * NQ + k));
        if (pending != 0u && spins > 64u) __nanosleep(32);
        if ((spins & 1023u) == 1023u) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > dp.timeout_ns) {
                if (dp.err != nullptr) *reinterpret_cast<volatile uint32_t *>(dp.err) = dp.step ? dp.step : 1u;
                return false;
            }
        }
    }
    return true;
}
// One 256-thread group (tid, named barrier `bar`) exchanges NF floats per thread; tot = sum over the ranks, identical everywhere.
// Inactive threads (nothing to send: out-of-range quads) only take part in the barriers. False after a timeout, for the whole group.
template <int NF>
__device__ __forceinline__ bool ll_exchange(const DpArgs &dp, int cta, int tid, bool active, const float (&mine)[NF], float (&tot)[NF],
                                            int bar, bool *s_timed_out) {
    constexpr int NQ = (NF + 2) / 3;
    static_assert(NQ <= kLlQuads, "packets per thread");
    const int W = dp.world, R = dp.rank, owner = cta % W;
    if (tid == 0) *s_timed_out = false;
    named_bar_sync(bar, 256);
    bool ok = true;
    auto send = [&](float4 *dst, const float (&val)[NF]) {
#pragma unroll
        for (int k = 0; k < NQ; ++k)
            ll_store(dst + k * 256, val[3 * k], 3 * k + 1 < NF ? val[3 * k + 1] : 0.0f, 3 * k + 2 < NF ? val[3 * k + 2] : 0.0f, dp.step);
    };
    auto unpack = [&](const float4 (&v)[NQ], float (&val)[NF]) {
#pragma unroll
        for (int k = 0; k < NQ; ++k) {
            val[3 * k] = v[k].x;
            if (3 * k + 1 < NF) val[3 * k + 1] = v[k].y;
            if (3 * k + 2 < NF) val[3 * k + 2] = v[k].z;
        }
    };
    if (active) {
        if (dp.ll_direct) {          // world of two
            send(dp.ll_peer[R ^ 1] + ll_index(R, cta, tid), mine);
            const float4 *src[1] = {dp.ll_own + ll_index(R ^ 1, cta, tid)};
            float4 v[1][NQ];
            ok = ll_recv<1, NQ>(src, v, dp.step, dp);
            float val[NF];
            unpack(v[0], val);
#pragma unroll
            for (int e = 0; e < NF; ++e) tot[e] = R == 0 ? (0.0f + mine[e]) + val[e] : (0.0f + val[e]) + mine[e];
        } else if (owner != R) {
            send(dp.ll_peer[owner] + ll_index(R, cta, tid), mine);
            const float4 *src[1] = {dp.ll_own + ll_index(kMaxRanks, cta, tid)};
            float4 v[1][NQ];
            ok = ll_recv<1, NQ>(src, v, dp.step, dp);
            unpack(v[0], tot);
        } else {
#pragma unroll
            for (int e = 0; e < NF; ++e) tot[e] = 0.0f;
#pragma unroll
            for (int rb = 0; rb < kMaxRanks; rb += 4) {       // four sources' packets in flight together; the sum stays in rank order
                if (rb >= W) break;
                const float4 *src[4];
#pragma unroll
                for (int b = 0; b < 4; ++b) src[b] = (rb + b < W && rb + b != R) ? dp.ll_own + ll_index(rb + b, cta, tid) : nullptr;
                float4 v[4][NQ];
                ok = ll_recv<4, NQ>(src, v, dp.step, dp) && ok;
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    if (rb + b >= W) continue;
                    float val[NF];
                    unpack(v[b], val);
#pragma unroll
                    for (int e = 0; e < NF; ++e) tot[e] += (rb + b == R) ? mine[e] : val[e];
                }
            }
            if (ok) {
#pragma unroll
                for (int r = 0; r < kMaxRanks; ++r)
                    if (r < W && r != R) send(dp.ll_peer[r] + ll_index(kMaxRanks, cta, tid), tot);
            }
        }
    }
    if (!ok) *s_timed_out = true;
    named_bar_sync(bar, 256);
    return !*s_timed_out;
}

// clock64 trace of CTA 0 of the most recent launch (read back by the ts_test_wgrad_trace hook)
__device__ unsigned long long g_wgrad_trace[16];
#define WTRACE(i) do { if (blockIdx.x == 0) g_wgrad_trace[(i)] = clock64(); } while (0)

__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire;" ::: "memory"); }
__device__ __forceinline__ float4 ld_cluster_v4(uint32_t local_saddr, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_saddr), "r"(rank));
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(remote) : "memory");
    return v;
}

// staging of one accumulator tile: [128 rows][32 chunks of 16 B], chunk index XOR-swizzled by the row so that both the
// row-per-lane writes and the row-contiguous reads are free of bank conflicts
__device__ __forceinline__ uint32_t acc_off(int row, int chunk) { return (uint32_t)row * 512u + ((uint32_t)(chunk ^ (row & 7)) << 4); }

template <int S, bool DP>      // cluster size = number of batch splits: 4 or 8; DP: cross-rank gradient sum over peer mappings
__global__ void __launch_bounds__(kThreads, 1)
wgrad_update_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const WgParams p) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kStages], empty_bar[kStages];
    __shared__ __align__(8) uint64_t acc_full_bar;
    __shared__ uint32_t tmem_base_slot;
    __shared__ bool s_last, s_timed_out;

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int cluster_id = (int)blockIdx.x / S;
    const uint32_t rank = cluster_rank();
    const OptArgs &o = p.opt;
    const bool ll = DP && p.dp.ll_own != nullptr;     // low-latency packets instead of publish + remote reads

    if (cluster_id >= p.n_tiles) {
        // ===================== the other tensors: per-slab partials from the fused kernel -> sum -> update =====================
        float4 *slots = reinterpret_cast<float4 *>((reinterpret_cast<uintptr_t>(smem_raw) + 15) & ~(uintptr_t)15);
        pdl_launch_dependents();
        pdl_wait();
        if (threadIdx.x >= 256) return;
        if (DP && ll) {
            // low-latency exchange: at most one chunk per CTA (the launcher checks small_chunks <= S); every thread joins the barriers
            const int c = (int)rank;
            const int chunk = p.small_chunk_begin + c;
            const int tix = c < p.small_chunks ? o.chunk_tensor[chunk] : -1;
            TensorDesc d{};
            if (tix >= 0) d = o.t[tix];
            const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
            const int64_t j0 = i0 - d.offset;
            const bool active = tix >= 0 && j0 < d.count;
            const int n = active ? (int)min((int64_t)4, d.count - j0) : 0;
            float g[4] = {0.0f, 0.0f, 0.0f, 0.0f}, tot[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            if (active) {
                if (d.nsplit > 0) {
                    sum_partials(o.partials + d.partial_off + j0, d.partial_stride, d.nsplit, g, slots, o.partial_depth);
                    if (n < 2) g[1] = 0.0f;
                    if (n < 3) g[2] = 0.0f;
                    if (n < 4) g[3] = 0.0f;
                    *reinterpret_cast<float4 *>(o.grads + i0) = make_float4(g[0], g[1], g[2], g[3]);
                } else {
                    const float4 t = *reinterpret_cast<const float4 *>(o.grads + i0);
                    g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
                }
            }
            if (ll_exchange<4>(p.dp, (int)blockIdx.x, (int)threadIdx.x, active, g, tot, 2, &s_timed_out) && active) apply_update(o, d, i0, j0, n, tot);
            return;
        }
        for (int pass = 0; pass < (DP ? 2 : 1); ++pass) {
            // DP: pass 0 folds the partials into this rank's peer-visible gradient buffer, then the handshake, pass 1 sums the
            // ranks and updates; single GPU: one pass does both
            for (int c = (int)rank; c < p.small_chunks; c += S) {
                const int chunk = p.small_chunk_begin + c;
                const int tix = o.chunk_tensor[chunk];
                if (tix < 0) continue;
                const TensorDesc d = o.t[tix];
                const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
                const int64_t j0 = i0 - d.offset;
                if (j0 >= d.count) continue;
                const int n = (int)min((int64_t)4, d.count - j0);
                float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                if (pass == 0) {
                    if (d.nsplit > 0) {
                        sum_partials(o.partials + d.partial_off + j0, d.partial_stride, d.nsplit, g, slots, o.partial_depth);
                        // padding lanes carry zeros: the flat buffers are padded to whole chunks
                        *reinterpret_cast<float4 *>(o.grads + i0) = make_float4(g[0], n > 1 ? g[1] : 0.0f, n > 2 ? g[2] : 0.0f, n > 3 ? g[3] : 0.0f);
                    } else {
                        const float4 t = *reinterpret_cast<const float4 *>(o.grads + i0);
                        g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
                    }
                    if (!DP) apply_update(o, d, i0, j0, n, g);
                } else {
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r < p.dp.world) {
                            const float4 t = __ldcv(reinterpret_cast<const float4 *>(p.dp.grads[r] + i0));
                            g[0] += t.x; g[1] += t.y; g[2] += t.z; g[3] += t.w;
                        }
                    apply_update(o, d, i0, j0, n, g);
                }
            }
            if (DP && pass == 0 && !dp_publish_and_wait(p.dp, (int)threadIdx.x, 2, &s_last, &s_timed_out)) return;
        }
        return;
    }

    uint8_t *tiles = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t a_bytes = BM * BK * 2, b_bytes = TN * BK * 2, stage_bytes = a_bytes + b_bytes;
    const int tm = cluster_id / p.tiles_n, tn = cluster_id - tm * p.tiles_n;
    const int m0 = tm * BM, n0 = tn * TN;
    const int kb_begin = (int)rank * p.kblocks_per_split;
    const int kb_end = min(p.kblocks, kb_begin + p.kblocks_per_split);
    const bool has_k = kb_begin < kb_end;

    if (threadIdx.x == 0) WTRACE(0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 2); mbar_init(&empty_bar[s], 1); }
        mbar_init(&acc_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tma_b)) : "memory");
    }
    if (warp == 1) tmem_alloc(&tmem_base_slot, TN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    pdl_launch_dependents();
    pdl_wait();          // dH and the bf16 inputs were written by the kernel in front
    if (threadIdx.x == 0) WTRACE(1);

    if (warp == 0) {
        // ===================== TMA producer: dH tile [BK batch rows][128 hidden units], two 64-unit boxes =====================
        int s = 0;
        uint32_t phase = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            if (elect_one()) {
                mbar_wait(&empty_bar[s], phase ^ 1);
                uint8_t *sa = tiles + (size_t)s * stage_bytes;
                mbar_expect_tx(&full_bar[s], a_bytes);
                tma_load_2d(&tma_a, &full_bar[s], sa, m0, kb * BK);
                tma_load_2d(&tma_a, &full_bar[s], sa + (size_t)BK * 128, m0 + 64, kb * BK);
            }
            __syncwarp();
            if (++s == kStages) { s = 0; phase ^= 1; }
        }
    } else if (warp == 2) {
        // ===================== TMA producer: X tile [BK batch rows][128 features] =====================
        int s = 0;
        uint32_t phase = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            if (elect_one()) {
                mbar_wait(&empty_bar[s], phase ^ 1);
                uint8_t *sb = tiles + (size_t)s * stage_bytes + a_bytes;
                mbar_expect_tx(&full_bar[s], b_bytes);
                tma_load_2d(&tma_b, &full_bar[s], sb, n0, kb * BK);
                tma_load_2d(&tma_b, &full_bar[s], sb + (size_t)BK * 128, n0 + 64, kb * BK);
            }
            __syncwarp();
            if (++s == kStages) { s = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        int s = 0;
        uint32_t phase = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
            if (elect_one()) {
                mbar_wait(&full_bar[s], phase);
                tc_fence_after();
                const uint32_t sa = smem_u32(tiles + (size_t)s * stage_bytes), sb = sa + a_bytes;
                const uint64_t adesc = make_desc(sa, TC_MAJOR_MN, BK), bdesc = make_desc(sb, TC_MAJOR_MN, BK);
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k)
                    umma_bf16(tmem_base, adesc + (uint64_t)(k * ((UMMA_K * 128) >> 4)), bdesc + (uint64_t)(k * ((UMMA_K * 128) >> 4)), p.idesc,
                              (kb > kb_begin || k > 0) ? 1u : 0u);
                umma_commit(&empty_bar[s]);
                if (kb == kb_end - 1) umma_commit(&acc_full_bar);
                if (kb == kb_begin) WTRACE(2);
                if (kb == kb_end - 1) WTRACE(3);
            }
            __syncwarp();
            if (++s == kStages) { s = 0; phase ^= 1; }
        }
    } else {
        // ===================== epilogue warps: state prefetch, accumulator -> shared memory, cluster reduce, update =====================
        const int et = threadIdx.x - 96;                 // 0..255
        const int q = warp & 3, half = (warp - 3) >> 2;  // TMEM lane quarter, which 64 of the tile's 128 feature columns
        constexpr int rows_own = BM / S;                 // tile rows this CTA reduces: [rank * rows_own, ...)
        constexpr int n_own = rows_own * (TN / 4) / kEpiThreads;   // quads per thread: 2 (S = 8) or 4 (S = 4)
        const bool adam = o.kind != TS_OPT_SGD;
        const bool use_m = adam || o.momentum != 0.0f;
        const bool use_x = adam && o.amsgrad;
        float4 w[n_own], m[n_own], v[n_own], x[n_own];
        int64_t idx[n_own];
#pragma unroll
        for (int i = 0; i < n_own; ++i) {
            idx[i] = -1;
            w[i] = m[i] = v[i] = x[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            {
                const int f = et + i * kEpiThreads;
                const int row = m0 + (int)rank * rows_own + f / (TN / 4), col = n0 + (f % (TN / 4)) * 4;
                if (row < p.H && col < p.D) {             // D is a multiple of 4: a quad is inside or outside as a whole
                    idx[i] = (int64_t)row * p.D + col;
                    w[i] = *reinterpret_cast<const float4 *>(o.params + p.off_w1 + idx[i]);
                    if (use_m) m[i] = *reinterpret_cast<const float4 *>(o.exp_avg + p.off_w1 + idx[i]);
                    if (adam) v[i] = *reinterpret_cast<const float4 *>(o.exp_avg_sq + p.off_w1 + idx[i]);
                    if (use_x) x[i] = *reinterpret_cast<const float4 *>(o.max_exp_avg_sq + p.off_w1 + idx[i]);
                }
            }
        }
        // keep the loads on this side of the wait (the compiler would otherwise sink them next to their use)
#pragma unroll
        for (int i = 0; i < n_own; ++i) asm volatile("" ::"f"(w[i].x), "f"(m[i].x), "f"(v[i].x), "f"(x[i].x) : "memory");

        // ---- this CTA's partial tile: TMEM -> shared memory (the pipeline stages are idle by now)
        if (has_k) {
            if (elect_one()) mbar_wait(&acc_full_bar, 0);
            __syncwarp();
            tc_fence_after();
        }
        if (et == 0) WTRACE(4);
        const int r_local = q * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t acc_s = smem_u32(tiles);
#pragma unroll 1
        for (int c0 = half * 64; c0 < half * 64 + 64; c0 += 32) {
            float vv[32];
            if (has_k) {
                tmem_ld16(t_lane + (uint32_t)c0, *reinterpret_cast<float(*)[16]>(&vv[0]));
                tmem_ld16(t_lane + (uint32_t)(c0 + 16), *reinterpret_cast<float(*)[16]>(&vv[16]));
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) vv[j] = 0.0f;       // no batch rows in this split: it contributes zeros
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
                asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(acc_s + acc_off(r_local, (c0 >> 2) + j)), "f"(vv[4 * j]),
                             "f"(vv[4 * j + 1]), "f"(vv[4 * j + 2]), "f"(vv[4 * j + 3]) : "memory");
        }
        tc_fence_before();
        if (et == 0) WTRACE(5);
        cluster_arrive();                                     // every CTA's tile is in its shared memory
        cluster_wait();
        if (et == 0) WTRACE(6);

        // ---- sum over the cluster in rank order, update, store. All remote loads of a thread are issued before the first
        //      add: a distributed-shared-memory load takes a few hundred nanoseconds, and S of them in a dependent chain
        //      were the longest phase of the kernel.
        float4 part[n_own][S];
#pragma unroll
        for (int i = 0; i < n_own; ++i) {
            const int f = et + i * kEpiThreads;
            const uint32_t addr = acc_s + acc_off((int)rank * rows_own + f / (TN / 4), f % (TN / 4));
#pragma unroll
            for (int r = 0; r < S; ++r) part[i][r] = ld_cluster_v4(addr, (uint32_t)r);
        }
        if (et == 0) WTRACE(7);
        auto update_quad = [&](int i, const float (&g)[4]) {
            float ww[4] = {w[i].x, w[i].y, w[i].z, w[i].w}, mm[4] = {m[i].x, m[i].y, m[i].z, m[i].w};
            float vq[4] = {v[i].x, v[i].y, v[i].z, v[i].w}, xx[4] = {x[i].x, x[i].y, x[i].z, x[i].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) opt_update_element(o, g[e], ww[e], mm[e], vq[e], xx[e]);
            const int64_t gi = p.off_w1 + idx[i];
            *reinterpret_cast<float4 *>(o.params + gi) = make_float4(ww[0], ww[1], ww[2], ww[3]);
            if (!DP) *reinterpret_cast<float4 *>(o.grads + gi) = make_float4(g[0], g[1], g[2], g[3]);   // DP: `grads` keeps this rank's own gradient
            if (use_m) *reinterpret_cast<float4 *>(o.exp_avg + gi) = make_float4(mm[0], mm[1], mm[2], mm[3]);
            if (adam) *reinterpret_cast<float4 *>(o.exp_avg_sq + gi) = make_float4(vq[0], vq[1], vq[2], vq[3]);
            if (use_x) *reinterpret_cast<float4 *>(o.max_exp_avg_sq + gi) = make_float4(xx[0], xx[1], xx[2], xx[3]);
            const __nv_bfloat162 lo = __floats2bfloat162_rn(ww[0], ww[1]), hi = __floats2bfloat162_rn(ww[2], ww[3]);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t *>(&lo);
            pk.y = *reinterpret_cast<const uint32_t *>(&hi);
            *reinterpret_cast<uint2 *>(o.shadow + p.sh_w1 + idx[i]) = pk;
        };
        float4 own[n_own];      // DP: this rank's gradient quads
#pragma unroll
        for (int i = 0; i < n_own; ++i) {
            own[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
            for (int r = 0; r < S; ++r) { g[0] += part[i][r].x; g[1] += part[i][r].y; g[2] += part[i][r].z; g[3] += part[i][r].w; }
            if (idx[i] < 0) continue;
            if (DP) {
                own[i] = make_float4(g[0], g[1], g[2], g[3]);
                *reinterpret_cast<float4 *>(o.grads + p.off_w1 + idx[i]) = own[i];   // peer-visible buffer (and param.grad)
            } else {
                update_quad(i, g);
            }
        }
        if (et == 0) WTRACE(10);
        if (DP && ll) {
            if constexpr (S == 8) {      // n_own = 2: eight floats per thread in three packets
                const float mine[8] = {own[0].x, own[0].y, own[0].z, own[0].w, own[1].x, own[1].y, own[1].z, own[1].w};
                float tot[8];
                if (ll_exchange<8>(p.dp, (int)blockIdx.x, et, idx[0] >= 0 || idx[1] >= 0, mine, tot, 1, &s_timed_out)) {
                    const float g0[4] = {tot[0], tot[1], tot[2], tot[3]}, g1[4] = {tot[4], tot[5], tot[6], tot[7]};
                    if (idx[0] >= 0) update_quad(0, g0);
                    if (idx[1] >= 0) update_quad(1, g1);
                }
            }
        } else if (DP) {
            // ---- every rank's gradient is complete once the handshake returns: this thread's quads from all of them, rank order
            if (dp_publish_and_wait(p.dp, et, 1, &s_last, &s_timed_out)) {
                float4 rv[n_own][8];
#pragma unroll
                for (int i = 0; i < n_own; ++i)
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        rv[i][r] = (idx[i] >= 0 && r < p.dp.world) ? __ldcv(reinterpret_cast<const float4 *>(p.dp.grads[r] + p.off_w1 + idx[i]))
                                                                   : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
                for (int i = 0; i < n_own; ++i) {
                    if (idx[i] < 0) continue;
                    float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r < p.dp.world) { g[0] += rv[i][r].x; g[1] += rv[i][r].y; g[2] += rv[i][r].z; g[3] += rv[i][r].w; }
                    update_quad(i, g);
                }
            }
        }
    }
    if (threadIdx.x == 96) WTRACE(8);   // et == 0
    // the cluster barriers are cluster-wide: the producer / issuer warps take part too. Second barrier: nobody leaves while a
    // peer may still be reading its tile.
    __syncwarp();
    if (warp < 3) { cluster_arrive(); cluster_wait(); }
    cluster_arrive();
    cluster_wait();
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TN);
    }
    if (threadIdx.x == 0) WTRACE(9);
}

}  // namespace

int read_wgrad_trace(unsigned long long *out_host) {
    TS_CUDA_CHECK(cudaMemcpyFromSymbol(out_host, g_wgrad_trace, sizeof(unsigned long long) * 16));
    return TS_OK;
}

bool wgrad_update_supported(int D, int H, int64_t B) { return (H == 128 || H == 256) && D % 8 == 0 && B >= 512; }

int launch_wgrad_update(const WgradUpdateArgs &a, cudaStream_t st) {
    TS_REQUIRE(wgrad_update_supported(a.D, a.H, a.B), TS_E_UNSUPPORTED, "dW1 + update kernel: unsupported shape D=%d H=%d B=%lld", a.D, a.H,
               (long long)a.B);
    WgParams p{};
    p.H = a.H; p.D = a.D; p.B = a.B;
    p.S = a.B >= 2048 ? 8 : 4;
    p.tiles_n = (a.D + TN - 1) / TN;
    p.n_tiles = (a.H / BM) * p.tiles_n;
    p.kblocks = (int)((a.B + BK - 1) / BK);
    p.kblocks_per_split = (p.kblocks + p.S - 1) / p.S;
    p.idesc = make_idesc(BM, TN, TC_MAJOR_MN, TC_MAJOR_MN);
    p.off_w1 = a.off_w1; p.sh_w1 = a.sh_w1;
    p.small_chunk_begin = a.small_chunk_begin; p.small_chunks = a.small_chunks;
    p.opt = a.opt;
    TS_REQUIRE(p.opt.params && p.opt.grads && p.opt.shadow && p.opt.chunk_tensor, TS_E_INVALID, "dW1 + update kernel: buffers missing");
    TS_REQUIRE(p.opt.params_mc == nullptr, TS_E_INVALID, "dW1 + update kernel: not for the nvls exchange");

    CUtensorMap ma, mb;
    int rc = get_map(a.dH, (uint64_t)a.H, (uint64_t)a.B, (uint64_t)a.H, 64, BK, &ma);
    if (rc) return rc;
    rc = get_map(a.X, (uint64_t)a.D, (uint64_t)a.B, (uint64_t)a.D, 64, BK, &mb);
    if (rc) return rc;

    const size_t smem = (size_t)kStages * (BM * BK * 2 + TN * BK * 2) + 1024;
    static bool attr_set[64] = {false};
    int dev = 0;
    TS_CUDA_CHECK(cudaGetDevice(&dev));
    if (!attr_set[dev & 63]) {
        TS_CUDA_CHECK(cudaFuncSetAttribute(wgrad_update_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        TS_CUDA_CHECK(cudaFuncSetAttribute(wgrad_update_kernel<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        TS_CUDA_CHECK(cudaFuncSetAttribute(wgrad_update_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        TS_CUDA_CHECK(cudaFuncSetAttribute(wgrad_update_kernel<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        attr_set[dev & 63] = true;
    }
    const unsigned grid = (unsigned)((p.n_tiles + 1) * p.S);
    const bool dp = a.dp.world > 1;
    p.dp = a.dp;
    // low-latency exchange: clusters of 8 (two quads per thread), one small-tensor chunk per CTA, a slot per CTA in the inbox
    if (p.S != 8 || p.small_chunks > p.S || (int)grid > kLlCtas) p.dp.ll_own = nullptr;
    // DP: every CTA waits for the peers, so the whole grid must be resident: one CTA per SM (160 KiB of shared memory each)
    TS_REQUIRE(!dp || (int)grid <= a.sm_count, TS_E_UNSUPPORTED, "dW1 + update kernel: %u CTAs do not fit on %d SMs at once", grid, a.sm_count);
    if (p.S == 8 && dp) TS_CUDA_CHECK(launch_pdl(wgrad_update_kernel<8, true>, dim3(grid), dim3(kThreads), smem, st, 8u, ma, mb, p));
    else if (p.S == 8) TS_CUDA_CHECK(launch_pdl(wgrad_update_kernel<8, false>, dim3(grid), dim3(kThreads), smem, st, 8u, ma, mb, p));
    else if (dp) TS_CUDA_CHECK(launch_pdl(wgrad_update_kernel<4, true>, dim3(grid), dim3(kThreads), smem, st, 4u, ma, mb, p));
    else TS_CUDA_CHECK(launch_pdl(wgrad_update_kernel<4, false>, dim3(grid), dim3(kThreads), smem, st, 4u, ma, mb, p));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

}  // namespace tsb
