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

// ------------------------------------------------------------------ programmatic dependent launch (PDL)
// The kernels of one fit step are launched with programmaticStreamSerialization: a kernel may start while its
// predecessor in the stream is still draining, run its prologue (barrier init, TMEM allocation, descriptor
// prefetch), and must execute pdl_wait() before it touches anything the predecessor wrote or still reads.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, unsigned cluster_x,
                              Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int n = 0;
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
    if (cluster_x > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = cluster_x;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

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

// Counter-based dropout generator. One 32-bit hash (Wellons' "lowbias32" finaliser: xorshift-multiply rounds with
// full avalanche) of (row, column pair, key) yields two 16-bit uniforms, i.e. the keep decisions of two neighbouring
// hidden units; key = f(seed, optimizer step, layer) is mixed on the host. Stateless, so forward and backward (and
// every kernel variant) regenerate the same mask from the element index alone.
__host__ __device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t dropout_bits(uint32_t key, int64_t row, int colpair) {
    return hash32(((uint32_t)row * 0x9E3779B1u) ^ ((uint32_t)((uint64_t)row >> 32) * 0xC2B2AE3Du) ^ ((uint32_t)colpair * 0x85EBCA6Bu) ^ key);
}

// How the dropout keep-mask of a hidden layer is obtained.
struct DropoutSpec {
    const uint8_t *mask;  // [B, H] 0/1 bytes, or nullptr
    uint32_t key;         // hash32-mixed (seed, optimizer step, layer)
    uint32_t thresh;      // drop when the 16-bit uniform is below round(p * 65536)
    float p;              // drop probability; 0 => disabled
    float scale;          // 1/(1-p)
    __device__ __forceinline__ float factor(int64_t row, int col, int H) const {
        if (p <= 0.0f) return 1.0f;
        bool keep;
        if (mask) keep = mask[row * (int64_t)H + col] != 0;
        else keep = ((dropout_bits(key, row, col >> 1) >> ((col & 1) * 16)) & 0xFFFFu) >= thresh;
        return keep ? scale : 0.0f;
    }
    // four consecutive columns starting at a multiple of 4
    __device__ __forceinline__ void factor4(int64_t row, int col, int H, float (&f)[4]) const {
        if (p <= 0.0f) { f[0] = f[1] = f[2] = f[3] = 1.0f; return; }
        if (mask == nullptr && (col & 3) == 0) {
            const uint32_t b0 = dropout_bits(key, row, col >> 1), b1 = dropout_bits(key, row, (col >> 1) + 1);
            f[0] = ((b0 & 0xFFFFu) >= thresh) ? scale : 0.0f;
            f[1] = ((b0 >> 16) >= thresh) ? scale : 0.0f;
            f[2] = ((b1 & 0xFFFFu) >= thresh) ? scale : 0.0f;
            f[3] = ((b1 >> 16) >= thresh) ? scale : 0.0f;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) f[j] = (col + j < H) ? factor(row, col + j, H) : 0.0f;
        }
    }
};

}  // namespace tsb
