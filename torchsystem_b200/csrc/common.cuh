// Shared device/host helpers for the tsb200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../../include/tsb200.h"

namespace tsb {

// ------------------------------------------------------------------ error plumbing (engine.cu)
void set_error(const char *fmt, ...);
extern long long g_launch_count;

#define TS_CUDA_CHECK(expr)                                                                     \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            tsb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return TS_E_CUDA;                                                                   \
        }                                                                                       \
    } while (0)

#define TS_LAUNCH_CHECK()                                  \
    do {                                                   \
        ++tsb::g_launch_count;                             \
        TS_CUDA_CHECK(cudaGetLastError());                 \
    } while (0)

#define TS_REQUIRE(cond, code, ...)     \
    do {                                \
        if (!(cond)) {                  \
            tsb::set_error(__VA_ARGS__); \
            return (code);              \
        }                               \
    } while (0)

// ------------------------------------------------------------------ small device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ void from_f32(float &dst, float v) { dst = v; }
__device__ __forceinline__ void from_f32(__nv_bfloat16 &dst, float v) { dst = __float2bfloat16_rn(v); }
__device__ __forceinline__ float round_bf16(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// exact (erf) GELU and its derivative, torch.nn.GELU() default
__device__ __forceinline__ float gelu_f(float u) { return 0.5f * u * (1.0f + erff(u * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_grad_f(float u) {
    const float cdf = 0.5f * (1.0f + erff(u * 0.70710678118654752440f));
    const float pdf = expf(-0.5f * u * u) * 0.39894228040143267794f;
    return cdf + u * pdf;
}

// Philox4x32-10 (Salmon et al. 2011): counter (c0..c3), key (k0,k1) -> 4 x 32 random bits.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}

// Keep-decision for element `idx` of hidden layer `layer` at optimizer step `step`: one Philox call
// covers 4 consecutive elements, so neighbouring threads recompute the same block (cheap, stateless).
__device__ __forceinline__ bool philox_keep(uint64_t seed, uint32_t step, uint32_t layer, uint64_t idx, float p) {
    const uint64_t blk = idx >> 2;
    const uint4 r = philox4x32_10(make_uint4((uint32_t)blk, (uint32_t)(blk >> 32), step, layer),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const uint32_t lane = (uint32_t)(idx & 3);
    const uint32_t bits = lane == 0 ? r.x : lane == 1 ? r.y : lane == 2 ? r.z : r.w;
    // uniform in [0,1): keep when u >= p
    return (float)(bits >> 8) * (1.0f / 16777216.0f) >= p;
}

// How the dropout keep-mask of a hidden layer is obtained.
struct DropoutSpec {
    const uint8_t *mask;  // [B, H] 0/1 bytes, or nullptr
    uint64_t seed;
    uint32_t step;
    uint32_t layer;
    float p;      // drop probability; 0 => disabled
    float scale;  // 1/(1-p)
    __device__ __forceinline__ float factor(int64_t row, int col, int H) const {
        if (p <= 0.0f) return 1.0f;
        const int64_t idx = row * (int64_t)H + col;
        const bool keep = mask ? (mask[idx] != 0) : philox_keep(seed, step, layer, (uint64_t)idx, p);
        return keep ? scale : 0.0f;
    }
    // Four consecutive columns starting at a multiple of 4 (one Philox call covers them when H % 4 == 0).
    __device__ __forceinline__ void factor4(int64_t row, int col, int H, float (&f)[4]) const {
        if (p <= 0.0f) { f[0] = f[1] = f[2] = f[3] = 1.0f; return; }
        const int64_t idx = row * (int64_t)H + col;
        if (mask == nullptr && (idx & 3) == 0) {
            const uint64_t blk = (uint64_t)idx >> 2;
            const uint4 r = philox4x32_10(make_uint4((uint32_t)blk, (uint32_t)(blk >> 32), step, layer),
                                          make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
            const float k = 1.0f / 16777216.0f;
            f[0] = ((float)(r.x >> 8) * k >= p) ? scale : 0.0f;
            f[1] = ((float)(r.y >> 8) * k >= p) ? scale : 0.0f;
            f[2] = ((float)(r.z >> 8) * k >= p) ? scale : 0.0f;
            f[3] = ((float)(r.w >> 8) * k >= p) ? scale : 0.0f;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) f[j] = (col + j < H) ? factor(row, col + j, H) : 0.0f;
        }
    }
};

}  // namespace tsb
