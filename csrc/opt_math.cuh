// Optimizer arithmetic and small device helpers shared by the two reduce+update kernels
// (reduce_update.cu: flag-handshake one-shot / two-shot / NVLS; ll_reduce.cu: push with flagged lines).
//
// Replaces reference train.py:51 (optimizer.step) and the 1/world divide of train.py:39.
#pragma once
#include "common.cuh"

namespace b200dp {

__device__ __forceinline__ uint16_t f32_to_bf16_rn(float f) {
    uint32_t u = __float_as_uint(f);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40u);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
__device__ __forceinline__ float bf16_to_f32(uint32_t h) { return __uint_as_float(h << 16); }

struct OptScalars { float lr, h0, h1, h2, wd, c0, c1; };

// Every operation is an explicit round-to-nearest intrinsic: the compiler may not contract or
// re-associate, so every code path and every rank produce bit-identical parameters from the same
// summed gradient.
__device__ __forceinline__ void opt_apply(int kind, const OptScalars& h, float g, float& p, float& s0, float& s1) {
    if (h.wd != 0.f) g = __fmaf_rn(h.wd, p, g);
    if (kind == 0) {            // Adadelta: h0 = rho, h1 = eps
        const float omr = __fsub_rn(1.f, h.h0);
        s0 = __fmaf_rn(h.h0, s0, __fmul_rn(omr, __fmul_rn(g, g)));
        const float stdv = __fsqrt_rn(__fadd_rn(s0, h.h1));
        const float delta = __fmul_rn(__fdiv_rn(__fsqrt_rn(__fadd_rn(s1, h.h1)), stdv), g);
        s1 = __fmaf_rn(h.h0, s1, __fmul_rn(omr, __fmul_rn(delta, delta)));
        p = __fsub_rn(p, __fmul_rn(h.lr, delta));
    } else if (kind == 1) {     // SGD: h0 = momentum, h1 = nesterov flag
        float d = g;
        if (h.h0 != 0.f) {
            s0 = __fmaf_rn(h.h0, s0, g);
            d = (h.h1 != 0.f) ? __fmaf_rn(h.h0, s0, g) : s0;
        }
        p = __fsub_rn(p, __fmul_rn(h.lr, d));
    } else {                    // Adam: h0 = beta1, h1 = beta2, h2 = eps; c0 = 1/bc1, c1 = 1/sqrt(bc2)
        s0 = __fmaf_rn(h.h0, s0, __fmul_rn(__fsub_rn(1.f, h.h0), g));
        s1 = __fmaf_rn(h.h1, s1, __fmul_rn(__fsub_rn(1.f, h.h1), __fmul_rn(g, g)));
        const float denom = __fmaf_rn(__fsqrt_rn(s1), h.c1, h.h2);
        p = __fsub_rn(p, __fmul_rn(__fmul_rn(h.lr, h.c0), __fdiv_rn(s0, denom)));
    }
}

__device__ __forceinline__ float4 scale4(const float4& v, float k) {
    return make_float4(__fmul_rn(v.x, k), __fmul_rn(v.y, k), __fmul_rn(v.z, k), __fmul_rn(v.w, k));
}
__device__ __forceinline__ float4 add4(const float4& a, const float4& b) {
    return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
}

struct RUCtx {
    OptScalars h;
    int kind;
    float* param; float* grad; float* state0; float* state1; uint16_t* shadow;
};

// hyper (device float[8]) -> scalars; Adam's bias correction comes from the generation number
__device__ __forceinline__ void ru_load_hyper(RUCtx& cx, const float* hyper, uint32_t gen) {
    cx.h.lr = hyper[0]; cx.h.h0 = hyper[1]; cx.h.h1 = hyper[2]; cx.h.h2 = hyper[3]; cx.h.wd = hyper[4];
    cx.h.c0 = cx.h.c1 = 1.f;
    if (cx.kind == 2) {
        const float t = (float)gen;
        cx.h.c0 = 1.f / (1.f - powf(cx.h.h0, t));
        cx.h.c1 = rsqrtf(1.f - powf(cx.h.h1, t));
    }
}

// The update of one float4 is split into its loads and its math+stores so that a loop can put the loads of
// several chunks in flight before the first dependent instruction.
struct RULoaded { float4 p, s0, s1; };
__device__ __forceinline__ RULoaded ru_load(const RUCtx& c, int idx) {
    RULoaded L;
    L.p = *reinterpret_cast<const float4*>(c.param + idx);
    L.s0 = *reinterpret_cast<const float4*>(c.state0 + idx);
    L.s1 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c.kind != 1) L.s1 = *reinterpret_cast<const float4*>(c.state1 + idx);
    return L;
}
__device__ __forceinline__ void ru_apply(const RUCtx& c, int idx, float4 g, RULoaded L) {
    float4 p = L.p, s0 = L.s0, s1 = L.s1;
    const bool two = (c.kind != 1);
    opt_apply(c.kind, c.h, g.x, p.x, s0.x, s1.x);
    opt_apply(c.kind, c.h, g.y, p.y, s0.y, s1.y);
    opt_apply(c.kind, c.h, g.z, p.z, s0.z, s1.z);
    opt_apply(c.kind, c.h, g.w, p.w, s0.w, s1.w);
    *reinterpret_cast<float4*>(c.param + idx) = p;
    *reinterpret_cast<float4*>(c.state0 + idx) = s0;
    if (two) *reinterpret_cast<float4*>(c.state1 + idx) = s1;
    *reinterpret_cast<float4*>(c.grad + idx) = g;     // param.grad == mean over ranks, like train.py:39
    if (c.shadow) {
        uint2 packed;
        packed.x = (uint32_t)f32_to_bf16_rn(p.x) | ((uint32_t)f32_to_bf16_rn(p.y) << 16);
        packed.y = (uint32_t)f32_to_bf16_rn(p.z) | ((uint32_t)f32_to_bf16_rn(p.w) << 16);
        *reinterpret_cast<uint2*>(c.shadow + idx) = packed;
    }
}
__device__ __forceinline__ void ru_update(const RUCtx& c, int idx, float4 g) { ru_apply(c, idx, g, ru_load(c, idx)); }

// Sum the rows of `partials` over one 128-float slice (32 float4 columns x 8 row groups of a 256-thread CTA) into
// sSlice, fixed order.  sRed: [8][128] floats of shared memory.  Ends with a __syncthreads.
__device__ __forceinline__ void ru_sum_partial_rows(const float* partials, int n_partials, int partial_stride,
                                                    int off, int n_here, float (*sRed)[128], float* sSlice, int tid) {
    const int o = tid & 31, pg = tid >> 5;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (o * 4 < n_here) {
        const float* src = partials + off + o * 4;        // off and partial_stride are multiples of 4: 16-byte aligned
        int r = pg;
        for (; r + 56 < n_partials; r += 64) {            // 8 independent loads in flight, added in row order
            float4 t[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) t[u] = __ldcg(reinterpret_cast<const float4*>(src + (long)(r + 8 * u) * partial_stride));
#pragma unroll
            for (int u = 0; u < 8; ++u) s = add4(s, t[u]);
        }
        for (; r < n_partials; r += 8) s = add4(s, __ldcg(reinterpret_cast<const float4*>(src + (long)r * partial_stride)));
    }
    *reinterpret_cast<float4*>(&sRed[pg][o * 4]) = s;
    __syncthreads();
    if (tid < 32) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < 8; ++q) t = add4(t, *reinterpret_cast<const float4*>(&sRed[q][tid * 4]));
        *reinterpret_cast<float4*>(sSlice + tid * 4) = t;
    }
    __syncthreads();
}

// Step bookkeeping shared by the kernels that end a step: the last CTA to finish delivers the loss to the host
// (a 4-byte store into a mapped pinned array: no copy command, nothing for the next step to wait on) and advances
// the generation.  Call with all threads of the CTA; `grid` = CTAs of the launch.
__device__ __forceinline__ void ru_finish_step(uint32_t* step_counter, uint32_t gen, unsigned grid, const float* loss_src,
                                               const unsigned long long* loss_sink, int tid) {
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(step_counter + 1, 1u) == grid - 1) {
            if (loss_sink != nullptr) {
                const unsigned long long ring = loss_sink[0], idx = (unsigned long long)gen - loss_sink[1];
                if (loss_src != nullptr && ring != 0ull && idx < loss_sink[2]) {
                    // no system fence: the host reads the slot only after an event that follows this kernel, and a fence
                    // here kept the step's last kernel alive for a PCIe round trip (5 us per step)
                    reinterpret_cast<volatile float*>(ring)[idx] = __ldcg(loss_src);
                }
                // loss_sink[3]: mapped host word that receives the sticky error word once it is set, so that the host's
                // check at a log point is a plain memory read (every CTA that raised the error did so before it took its
                // ticket).  Nothing is stored while all is well: a store to host memory keeps this kernel alive for a PCIe
                // round trip.
                const unsigned long long err_host = loss_sink[3];
                const uint32_t err = __ldcg(step_counter + 2);
                if (err_host != 0ull && err != 0u) *reinterpret_cast<volatile uint32_t*>(err_host) = err;
            }
            step_counter[1] = 0u;
            __threadfence();
            step_counter[0] = gen;
        }
    }
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

}  // namespace b200dp
