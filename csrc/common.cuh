// Shared device/host helpers for the b200dp CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/b200dp.h"

#define LEAKY_SLOPE 0.2f

namespace b200dp {

void set_last_error(const char* fmt, ...);

#define B200DP_CHECK(expr)                                                              \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) {                                                        \
            ::b200dp::set_last_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,       \
                                     cudaGetErrorString(_e));                           \
            return (int)_e;                                                             \
        }                                                                               \
    } while (0)

#define B200DP_REQUIRE(cond, msg)                                                       \
    do {                                                                                \
        if (!(cond)) {                                                                  \
            ::b200dp::set_last_error("%s:%d requirement failed: %s (%s)", __FILE__,     \
                                     __LINE__, #cond, msg);                             \
            return (int)cudaErrorInvalidValue;                                          \
        }                                                                               \
    } while (0)

// ---- programmatic dependent launch (PDL) ----------------------------------------------------------
// The step is a chain of ~10 short kernels; between two of them the GPU idles for the ~1-2 us it takes to
// drain one grid and schedule the next.  Launched with the programmatic-stream-serialization attribute, a
// kernel's CTAs may become resident while its predecessor is still running (as soon as every predecessor CTA
// has executed pdl_trigger); they then block in pdl_wait until the predecessor has completed and its memory
// is visible.  Every kernel calls pdl_wait at its very top, before touching global memory, so the data
// dependencies are exactly those of ordinary stream order; pdl_trigger sits just before a kernel's final
// stores (triggering at the top made the step SLOWER: early dependents hold registers and CTA slots the
// running grid needs), so only the scheduling gap is hidden.  Captured into CUDA graphs as programmatic edges.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE property of a kernel: a process that drives several GPUs
// must set it once on each.  One DeviceOnce per launch site (a function-local static) remembers the devices done.
struct DeviceOnce {
    unsigned long long done[4] = {0ull, 0ull, 0ull, 0ull};     // bit per device ordinal, 0..255
    template <class Kern>
    cudaError_t max_dynamic_smem(Kern kern, int bytes) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        const bool tracked = dev >= 0 && dev < 256;
        if (tracked && ((__atomic_load_n(&done[dev >> 6], __ATOMIC_ACQUIRE) >> (dev & 63)) & 1ull)) return cudaSuccess;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);   // idempotent: a race only repeats it
        if (e == cudaSuccess && tracked) __atomic_fetch_or(&done[dev >> 6], 1ull << (dev & 63), __ATOMIC_RELEASE);
        return e;
    }
};

template <class Kern, class Args>
static inline cudaError_t launch_pdl(Kern kern, dim3 grid, dim3 block, size_t smem, cudaStream_t s, const Args& args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, args);
}

__device__ __forceinline__ float leaky(float v) { return v > 0.f ? v : LEAKY_SLOPE * v; }
// derivative of LeakyReLU evaluated from the *activated* value (sign is preserved)
__device__ __forceinline__ float leaky_grad_from_act(float a) { return a > 0.f ? 1.f : LEAKY_SLOPE; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Input pre-processing shared by conv1 forward and conv1 wgrad (which recomputes
// the 32x32 image instead of storing it): ToTensor+Normalize (reference train.py:85)
// then UpsamplingBilinear2d(size=32), align_corners=True (reference cnn.py:9).
__device__ __forceinline__ float load_pixel(const void* x, int is_u8, long idx, float mean, float std) {
    if (is_u8) {
        float v = (float)((const uint8_t*)x)[idx];
        return __fdiv_rn(__fdiv_rn(v, 255.f) - mean, std);
    }
    return ((const float*)x)[idx];
}

// Fill dst[32*32] from a src[hw*hw] plane held in shared memory.
__device__ __forceinline__ void upsample32_plane(const float* src, int hw, float* dst, int tid, int nthreads) {
    if (hw == 32) {
        for (int i = tid; i < 1024; i += nthreads) dst[i] = src[i];
        return;
    }
    const float scale = (float)(hw - 1) / 31.f;
    for (int i = tid; i < 1024; i += nthreads) {
        int y = i >> 5, x = i & 31;
        float h1r = scale * (float)y, w1r = scale * (float)x;
        int h1 = (int)h1r, w1 = (int)w1r;
        int h1p = (h1 < hw - 1) ? 1 : 0, w1p = (w1 < hw - 1) ? 1 : 0;
        float h1l = h1r - (float)h1, h0l = 1.f - h1l;
        float w1l = w1r - (float)w1, w0l = 1.f - w1l;
        const float* r0 = src + h1 * hw + w1;
        const float* r1 = r0 + h1p * hw;
        dst[i] = h0l * (w0l * r0[0] + w1l * r0[w1p]) + h1l * (w0l * r1[0] + w1l * r1[w1p]);
    }
}

}  // namespace b200dp
