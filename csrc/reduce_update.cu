// The product kernel: gradient reduction across ranks over NVLink peer memory, fused with the
// gradient average and the optimizer update.  One launch per bucket of the flat arena.
//
// Replaces reference train.py:34-39 (ten blocking ncclAllReduce + ten divides), train.py:51
// (optimizer.step: ~7 elementwise kernels x 10 tensors in torch 1.5) and train.py:46 (zero_grad).
// No NCCL call and no separate elementwise kernel runs on this path.
//
// Decomposition: CTA j owns slice j of the bucket, end to end.  A slice never depends on another
// slice, so no grid-wide synchronisation exists anywhere:
//   A. (optional) sum the partial-gradient rows written by the conv backward (fixed order).
//   B. world > 1, one-shot P2P: store the local slice into this rank's staging buffer
//      [parity][arena], __threadfence_system, then write the generation number into flag slot
//      (my_rank, slice) of EVERY peer (st.release.sys over NVLink); spin (ld.acquire.sys, bounded)
//      on the local flag slots (peer, slice); read the slice from every rank's staging buffer with
//      128-bit volatile loads over NVLink and add in rank order -> bitwise identical on all ranks.
//      Staging is double-buffered by step parity: a rank can only be one step ahead of a peer
//      (it needs the peer's flag of step g to finish step g), so buffer g&1 is never overwritten
//      while a peer still reads it.  Flags only ever increase; nothing is reset.
//   C. g *= 1/world; Adadelta / SGD-momentum / Adam on the fp32 master weights and state arenas;
//      optional bf16 shadow copy for the tensor-core kernels; averaged gradient written back.
// The generation counter lives on the device and is advanced by the last CTA of the step's last
// bucket, so the same launch sequence can be replayed from a CUDA graph.
#include "common.cuh"

namespace b200dp {

constexpr int RU_NT = 256;
constexpr int RU_SLICE = 1024;        // floats per CTA for plain buckets (float4 per thread)
constexpr int RU_PSLICE = 32;         // floats per CTA when partial rows are pre-reduced
constexpr int RU_FLAG_GRAIN = 32;     // one flag slot per 32 arena floats
constexpr unsigned RU_SPIN_LIMIT = 1u << 28;

struct RUArgs {
    float* param; float* grad; float* state0; float* state1; uint16_t* shadow;
    int start, count;
    const float* partials; int n_partials, partial_stride;
    int opt_kind;
    const float* hyper;
    uint32_t* step_counter;   // [0] generation, [1] ticket, [2] error flag
    int last_bucket;
    b200dp_comm comm;
    int slice;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint16_t f32_to_bf16_rn(float f) {
    uint32_t u = __float_as_uint(f);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40u);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

struct OptScalars { float lr, h0, h1, h2, wd, c0, c1; };

__device__ __forceinline__ void opt_apply(int kind, const OptScalars& h, float g, float& p, float& s0, float& s1) {
    if (h.wd != 0.f) g = fmaf(h.wd, p, g);
    if (kind == 0) {            // Adadelta: h0 = rho, h1 = eps
        s0 = h.h0 * s0 + (1.f - h.h0) * g * g;
        const float stdv = sqrtf(s0 + h.h1);
        const float delta = sqrtf(s1 + h.h1) / stdv * g;
        s1 = h.h0 * s1 + (1.f - h.h0) * delta * delta;
        p -= h.lr * delta;
    } else if (kind == 1) {     // SGD: h0 = momentum, h1 = nesterov flag
        float d = g;
        if (h.h0 != 0.f) {
            s0 = h.h0 * s0 + g;
            d = (h.h1 != 0.f) ? g + h.h0 * s0 : s0;
        }
        p -= h.lr * d;
    } else {                    // Adam: h0 = beta1, h1 = beta2, h2 = eps; c0 = 1/bc1, c1 = 1/sqrt(bc2)
        s0 = h.h0 * s0 + (1.f - h.h0) * g;
        s1 = h.h1 * s1 + (1.f - h.h1) * g * g;
        const float denom = sqrtf(s1) * h.c1 + h.h2;
        p -= (h.lr * h.c0) * (s0 / denom);
    }
}

__global__ void __launch_bounds__(RU_NT) reduce_update_kernel(RUArgs a) {
    __shared__ float sRed[8][RU_PSLICE];
    __shared__ __align__(16) float sSlice[RU_PSLICE];
    __shared__ int sTimeout;
    const int tid = threadIdx.x;
    const int off = blockIdx.x * a.slice;                 // offset inside the bucket
    const int n_here = min(a.slice, a.count - off);       // multiple of 4 (buckets are 128-aligned)
    const uint32_t gen = a.step_counter[0] + 1u;          // constant for the whole step
    const int world = a.comm.world;
    const bool active = tid * 4 < n_here;
    const int idx = a.start + off + tid * 4;              // arena index of this thread's float4
    if (tid == 0) sTimeout = 0;

    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    // ---- A: local gradient for this slice ----
    if (a.partials) {
        // slice == RU_PSLICE: thread (o = tid%32, pg = tid/32) sums rows pg, pg+8, ... ; then 8 -> 1
        const int o = tid & 31, pg = tid >> 5;
        float s = 0.f;
        if (o < n_here) {
            const float* src = a.partials + off + o;
            int r = pg;
            for (; r + 56 < a.n_partials; r += 64) {          // 8 independent loads in flight, added in row order
                float t[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) t[u] = __ldcg(src + (long)(r + 8 * u) * a.partial_stride);
#pragma unroll
                for (int u = 0; u < 8; ++u) s += t[u];
            }
            for (; r < a.n_partials; r += 8) s += __ldcg(src + (long)r * a.partial_stride);
        }
        sRed[pg][o] = s;
        __syncthreads();
        if (tid < RU_PSLICE) {
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) t += sRed[q][tid];
            sSlice[tid] = t;
        }
        __syncthreads();
        if (active) g = *reinterpret_cast<const float4*>(sSlice + tid * 4);
    } else if (active) {
        g = *reinterpret_cast<const float4*>(a.grad + idx);
    }

    // ---- B: one-shot all-reduce over peer memory ----
    if (world > 1) {
        const int parity = gen & 1u;
        const long sidx = (long)parity * a.comm.arena_len + idx;
        if (active) {
            float* my_stage = nullptr;
#pragma unroll
            for (int r = 0; r < 8; ++r)
                if (r == a.comm.rank) my_stage = a.comm.stage[r];
            *reinterpret_cast<float4*>(my_stage + sidx) = g;
            __threadfence_system();
        }
        __syncthreads();
        const int slot = (a.start + off) / RU_FLAG_GRAIN;
        if (tid < world && tid != a.comm.rank) {
            // publish to peer `tid`, then wait for peer `tid`'s publication of the same slice.
            // (pointer tables are read with compile-time indices: run-time indexing of kernel
            // parameters compiles to slow indexed constant loads)
            uint32_t* peer_flags = nullptr;
            uint32_t* my_flags = nullptr;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                if (r == tid) peer_flags = a.comm.flags[r];
                if (r == a.comm.rank) my_flags = a.comm.flags[r];
            }
            st_release_sys(peer_flags + (long)a.comm.rank * a.comm.max_slices + slot, gen);
            const uint32_t* mine = my_flags + (long)tid * a.comm.max_slices + slot;
            unsigned spins = 0;
            while ((int)(ld_acquire_sys(mine) - gen) < 0) {
                if (++spins > RU_SPIN_LIMIT) { sTimeout = 1; break; }   // dead peer: error, not a hang
                __nanosleep(20);
            }
        }
        __syncthreads();
        if (sTimeout) {
            if (tid == 0) atomicExch(a.step_counter + 2, 1u);
            return;   // leave parameters untouched; the host sees the error flag
        }
        if (active) {
            float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
            float4 v[8];
#pragma unroll
            for (int r = 0; r < 8; ++r)                  // all peer loads in flight together
                if (r < world && r != a.comm.rank) v[r] = ld_volatile_f4(a.comm.stage[r] + sidx);
#pragma unroll
            for (int r = 0; r < 8; ++r) {                // rank order, identical on every rank
                if (r < world) {
                    const float4 t = (r == a.comm.rank) ? g : v[r];
                    sum.x += t.x; sum.y += t.y; sum.z += t.z; sum.w += t.w;
                }
            }
            const float inv = 1.f / (float)world;
            g = make_float4(sum.x * inv, sum.y * inv, sum.z * inv, sum.w * inv);
        }
    }

    // ---- C: optimizer ----
    if (active) {
        OptScalars h;
        h.lr = a.hyper[0]; h.h0 = a.hyper[1]; h.h1 = a.hyper[2]; h.h2 = a.hyper[3]; h.wd = a.hyper[4];
        h.c0 = h.c1 = 1.f;
        if (a.opt_kind == 2) {
            const float t = (float)gen;
            h.c0 = 1.f / (1.f - powf(h.h0, t));
            h.c1 = rsqrtf(1.f - powf(h.h1, t));
        }
        float4 p = *reinterpret_cast<float4*>(a.param + idx);
        float4 s0 = *reinterpret_cast<float4*>(a.state0 + idx);
        float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool two = (a.opt_kind != 1);
        if (two) s1 = *reinterpret_cast<float4*>(a.state1 + idx);
        opt_apply(a.opt_kind, h, g.x, p.x, s0.x, s1.x);
        opt_apply(a.opt_kind, h, g.y, p.y, s0.y, s1.y);
        opt_apply(a.opt_kind, h, g.z, p.z, s0.z, s1.z);
        opt_apply(a.opt_kind, h, g.w, p.w, s0.w, s1.w);
        *reinterpret_cast<float4*>(a.param + idx) = p;
        *reinterpret_cast<float4*>(a.state0 + idx) = s0;
        if (two) *reinterpret_cast<float4*>(a.state1 + idx) = s1;
        *reinterpret_cast<float4*>(a.grad + idx) = g;     // param.grad == mean over ranks, like train.py:39
        if (a.shadow) {
            uint2 packed;
            packed.x = (uint32_t)f32_to_bf16_rn(p.x) | ((uint32_t)f32_to_bf16_rn(p.y) << 16);
            packed.y = (uint32_t)f32_to_bf16_rn(p.z) | ((uint32_t)f32_to_bf16_rn(p.w) << 16);
            *reinterpret_cast<uint2*>(a.shadow + idx) = packed;
        }
    }

    // ---- generation counter: advanced once per step by the last CTA of the last bucket ----
    if (a.last_bucket) {
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            if (atomicAdd(a.step_counter + 1, 1u) == gridDim.x - 1) {
                a.step_counter[1] = 0u;
                __threadfence();
                a.step_counter[0] = gen;
            }
        }
    }
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_reduce_update(float* param, float* grad, float* state0, float* state1,
                                    uint16_t* shadow_bf16, int start, int count, const float* partials,
                                    int n_partials, int partial_stride, int opt_kind, const float* hyper,
                                    uint32_t* step_counter, int last_bucket, const b200dp_comm* comm,
                                    b200dp_stream_t s) {
    B200DP_REQUIRE(start % 128 == 0 && count % 4 == 0 && count > 0, "bucket must be 128-aligned");
    B200DP_REQUIRE(opt_kind >= 0 && opt_kind <= 2, "opt_kind 0 adadelta, 1 sgd, 2 adam");
    RUArgs a{};
    a.param = param; a.grad = grad; a.state0 = state0; a.state1 = state1; a.shadow = shadow_bf16;
    a.start = start; a.count = count;
    a.partials = partials; a.n_partials = n_partials; a.partial_stride = partial_stride;
    a.opt_kind = opt_kind; a.hyper = hyper; a.step_counter = step_counter; a.last_bucket = last_bucket;
    if (comm) {
        a.comm = *comm;
        B200DP_REQUIRE(comm->world >= 1 && comm->world <= 8, "world size 1..8");
        B200DP_REQUIRE(comm->algo == 0, "only the one-shot algorithm is built so far");
        B200DP_REQUIRE(comm->world == 1 || (start + count) / RU_FLAG_GRAIN <= comm->max_slices, "flag block too small");
    } else {
        a.comm.world = 1; a.comm.rank = 0;
    }
    a.slice = partials ? RU_PSLICE : RU_SLICE;
    const int grid = (count + a.slice - 1) / a.slice;
    reduce_update_kernel<<<grid, RU_NT, 0, (cudaStream_t)s>>>(a);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
