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
//   B. world > 1 (one-shot shown; two-shot and NVLS below): store the local slice into this rank's staging buffer
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
#include "opt_math.cuh"

namespace b200dp {

constexpr int RU_NT = 256;
constexpr int RU_CHUNK = 1024;        // floats per CTA iteration (float4 per thread)
constexpr int RU_MAX_CTAS = 148 * 2;  // = resident CTAs (2 per SM at this register count): one handshake wave,
                                      // larger buckets loop over chunks inside each CTA
constexpr int RU_UNROLL = 4;          // chunks whose loads are issued together
constexpr int RU_PSLICE = 128;        // floats per CTA when partial rows are pre-reduced (32 float4 columns x 8 row groups)
constexpr int RU_FLAG_GRAIN = 32;     // one flag slot per 32 arena floats
constexpr unsigned long long RU_DEFAULT_TIMEOUT_NS = 20ull * 1000ull * 1000ull * 1000ull;

struct RUArgs {
    float* param; float* grad; float* state0; float* state1; uint16_t* shadow;
    int start, count;
    const float* partials; int n_partials, partial_stride;
    int opt_kind;
    const float* hyper;
    uint32_t* step_counter;   // [0] generation, [1] ticket, [2] error flag
    int last_bucket;
    const float* loss_src; const unsigned long long* loss_sink;
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
// NVLS: a load from / store to the multicast alias of the staging buffers is executed by the
// NVSwitch on every rank's copy (sm_90+; the alias comes from cuMulticast* / torch symmetric memory).
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// spin until *flag >= gen (generation numbers only grow); bounded so a dead peer is an error, not a hang
__device__ __forceinline__ bool ru_wait(const uint32_t* flag, uint32_t gen, unsigned long long timeout_ns) {
    unsigned spins = 0;
    unsigned long long t0 = 0;
    while ((int)(ld_acquire_sys(flag) - gen) < 0) {
        if ((++spins & 255u) == 0u) {
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > timeout_ns) return false;
        }
        __nanosleep(20);
    }
    return true;
}

__global__ void __launch_bounds__(RU_NT, 2) reduce_update_kernel(RUArgs a) {
    __shared__ __align__(16) float sRed[8][RU_PSLICE];
    __shared__ __align__(16) float sSlice[RU_PSLICE];
    __shared__ int sTimeout;
    const int tid = threadIdx.x;
    const int off = blockIdx.x * a.slice;                 // offset inside the bucket
    const int n_here = min(a.slice, a.count - off);       // multiple of 4 (buckets are 128-aligned)
    const int chunks = (n_here + RU_CHUNK - 1) / RU_CHUNK;  // 1024-float chunks, one float4 per thread each
    const uint32_t gen = a.step_counter[0] + 1u;          // constant for the whole step
    const int world = a.comm.world;
    const int base_idx = a.start + off + tid * 4;         // arena index of this thread's float4 in chunk 0
    if (tid == 0) sTimeout = 0;

    RUCtx cx;
    cx.kind = a.opt_kind; cx.param = a.param; cx.grad = a.grad; cx.state0 = a.state0; cx.state1 = a.state1;
    cx.shadow = a.shadow;
    ru_load_hyper(cx, a.hyper, gen);
    const unsigned long long timeout_ns =
        a.comm.timeout_ms ? (unsigned long long)a.comm.timeout_ms * 1000000ull : RU_DEFAULT_TIMEOUT_NS;

    // ---- A: (optional) sum the partial-gradient rows of this 32-float slice, fixed order ----
    const bool from_rows = a.partials != nullptr;
    if (from_rows) {
        ru_sum_partial_rows(a.partials, a.n_partials, a.partial_stride, off, n_here, sRed, sSlice, tid);
    }
    auto active = [&](int c) { return c * RU_CHUNK + tid * 4 < n_here; };
    auto local_g = [&](int c) -> float4 {
        return from_rows ? *reinterpret_cast<const float4*>(sSlice + tid * 4)
                         : *reinterpret_cast<const float4*>(a.grad + base_idx + c * RU_CHUNK);
    };

    // A time-out (here or in an earlier step: the error word is sticky) skips the exchange and the update but not the
    // bookkeeping below, so the generation keeps advancing and the loss-ring indices stay defined.
    do {
    if (a.step_counter[2] != 0u) break;
    if (world == 1) {
        for (int c0 = 0; c0 < chunks; c0 += RU_UNROLL) {
            float4 g[RU_UNROLL];
            RULoaded L[RU_UNROLL];
#pragma unroll
            for (int u = 0; u < RU_UNROLL; ++u)
                if (c0 + u < chunks && active(c0 + u)) { g[u] = local_g(c0 + u); L[u] = ru_load(cx, base_idx + (c0 + u) * RU_CHUNK); }
#pragma unroll
            for (int u = 0; u < RU_UNROLL; ++u)
                if (c0 + u < chunks && active(c0 + u)) ru_apply(cx, base_idx + (c0 + u) * RU_CHUNK, g[u], L[u]);
        }
    } else {
        // ---- B: all-reduce over peer memory ----
        const int parity = gen & 1u;
        const long pbase = (long)parity * a.comm.arena_len;
        const int rank = a.comm.rank;
        // pointer tables are read with compile-time indices (run-time indexing of kernel parameters
        // compiles to slow indexed constant loads)
        float* my_stage = nullptr;
        uint32_t* my_flags = nullptr;
#pragma unroll
        for (int r = 0; r < 8; ++r)
            if (r == rank) { my_stage = a.comm.stage[r]; my_flags = a.comm.flags[r]; }
        for (int c = 0; c < chunks; ++c)
            if (active(c)) *reinterpret_cast<float4*>(my_stage + pbase + base_idx + c * RU_CHUNK) = local_g(c);   // publish
        __threadfence_system();
        __syncthreads();
        const int slot = (a.start + off) / RU_FLAG_GRAIN;
        const float inv = __fdiv_rn(1.f, (float)world);
        auto sum_ranks = [&](int c) -> float4 {           // rank order: identical on every rank
            const long sidx = pbase + base_idx + c * RU_CHUNK;
            const float4 own = local_g(c);
            float4 v[8];
#pragma unroll
            for (int r = 0; r < 8; ++r)                   // all peer loads in flight together
                if (r < world && r != rank) v[r] = ld_volatile_f4(a.comm.stage[r] + sidx);
            float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                if (r < world) sum = add4(sum, (r == rank) ? own : v[r]);
            }
            return sum;
        };

        if (a.comm.algo == 0) {
            // ---- one-shot: everyone reads everyone (latency-optimal; (world-1) x bytes in per rank) ----
            if (tid < world && tid != rank) {
                uint32_t* peer_flags = nullptr;
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if (r == tid) peer_flags = a.comm.flags[r];
                st_release_sys(peer_flags + (long)rank * a.comm.max_slices + slot, gen);
                if (!ru_wait(my_flags + (long)tid * a.comm.max_slices + slot, gen, timeout_ns)) sTimeout = 1;
            }
            __syncthreads();
            if (sTimeout) {
                if (tid == 0) atomicExch(a.step_counter + 2, 1u);
                break;    // parameters stay untouched; the host sees the error flag; the ticket below is still taken
            }
            for (int c = 0; c < chunks; ++c) {
                if (active(c)) {
                    const float4 s = sum_ranks(c);
                    ru_update(cx, base_idx + c * RU_CHUNK, scale4(s, inv));
                }
            }
        } else {
            // ---- two-shot (algo 1) / NVLS (algo 2): slice j is reduced by ONE owner rank, which then
            //      hands the sum to every peer.  Bytes per rank ~2(world-1)/world x bucket; with NVLS the
            //      switch does the reduction (multimem.ld_reduce) and the broadcast (multimem.st).
            //      One producer per slice also makes the result identical on all ranks by construction.
            const int owner = (int)(blockIdx.x % (unsigned)world);
            uint32_t* flags2_mine = my_flags + 8l * a.comm.max_slices;        // [slot], written by the owner
            const long landing = 2l * a.comm.arena_len;                        // reduced slices land here
            if (rank != owner) {
                if (tid == 0) {
                    uint32_t* owner_flags = nullptr;
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r == owner) owner_flags = a.comm.flags[r];
                    st_release_sys(owner_flags + (long)rank * a.comm.max_slices + slot, gen);   // "my slice is up"
                    if (!ru_wait(flags2_mine + slot, gen, timeout_ns)) sTimeout = 1;                          // "the sum is here"
                }
                __syncthreads();
                if (sTimeout) {
                    if (tid == 0) atomicExch(a.step_counter + 2, 1u);
                    break;
                }
                for (int c0 = 0; c0 < chunks; c0 += RU_UNROLL) {
                    float4 t[RU_UNROLL];
                    RULoaded L[RU_UNROLL];
#pragma unroll
                    for (int u = 0; u < RU_UNROLL; ++u)
                        if (c0 + u < chunks && active(c0 + u)) {
                            t[u] = ld_volatile_f4(my_stage + landing + base_idx + (c0 + u) * RU_CHUNK);
                            L[u] = ru_load(cx, base_idx + (c0 + u) * RU_CHUNK);
                        }
#pragma unroll
                    for (int u = 0; u < RU_UNROLL; ++u)
                        if (c0 + u < chunks && active(c0 + u)) ru_apply(cx, base_idx + (c0 + u) * RU_CHUNK, scale4(t[u], inv), L[u]);
                }
            } else {
                if (tid < world && tid != rank)
                    if (!ru_wait(my_flags + (long)tid * a.comm.max_slices + slot, gen, timeout_ns)) sTimeout = 1;
                __syncthreads();
                if (sTimeout) {
                    if (tid == 0) atomicExch(a.step_counter + 2, 1u);
                    break;
                }
                if (a.comm.algo == 2) {
                    // NVLS: the in-switch reductions of several chunks are in flight together
                    for (int c0 = 0; c0 < chunks; c0 += RU_UNROLL) {
                        float4 s[RU_UNROLL];
                        RULoaded L[RU_UNROLL];
#pragma unroll
                        for (int u = 0; u < RU_UNROLL; ++u)
                            if (c0 + u < chunks && active(c0 + u)) {
                                const int idx = base_idx + (c0 + u) * RU_CHUNK;
                                s[u] = multimem_ld_reduce_f4(a.comm.mc_stage + pbase + idx);   // in-switch sum of all ranks
                                L[u] = ru_load(cx, idx);
                            }
#pragma unroll
                        for (int u = 0; u < RU_UNROLL; ++u)
                            if (c0 + u < chunks && active(c0 + u)) {
                                const int idx = base_idx + (c0 + u) * RU_CHUNK;
                                multimem_st_f4(a.comm.mc_stage + landing + idx, s[u]);         // in-switch broadcast
                                ru_apply(cx, idx, scale4(s[u], inv), L[u]);
                            }
                    }
                } else {
                    for (int c = 0; c < chunks; ++c) {
                        if (active(c)) {
                            const int idx = base_idx + c * RU_CHUNK;
                            const RULoaded L = ru_load(cx, idx);
                            const float4 s = sum_ranks(c);
#pragma unroll
                            for (int r = 0; r < 8; ++r)          // push the sum into every peer's landing area
                                if (r < world && r != rank)
                                    *reinterpret_cast<float4*>(a.comm.stage[r] + landing + idx) = s;
                            ru_apply(cx, idx, scale4(s, inv), L);
                        }
                    }
                }
                __threadfence_system();
                __syncthreads();
                if (tid < world && tid != rank) {
                    uint32_t* peer_flags = nullptr;
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r == tid) peer_flags = a.comm.flags[r];
                    st_release_sys(peer_flags + 8l * a.comm.max_slices + slot, gen);
                }
            }
        }
    }

    } while (false);

    // ---- generation counter: advanced once per step by the last CTA of the last bucket ----
    if (a.last_bucket) ru_finish_step(a.step_counter, gen, gridDim.x, a.loss_src, a.loss_sink, tid);
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_reduce_update(float* param, float* grad, float* state0, float* state1,
                                    uint16_t* shadow_bf16, int start, int count, const float* partials,
                                    int n_partials, int partial_stride, int opt_kind, const float* hyper,
                                    uint32_t* step_counter, int last_bucket, const b200dp_comm* comm,
                                    const float* loss_src, const uint64_t* loss_sink, b200dp_stream_t s) {
    B200DP_REQUIRE(start % 128 == 0 && count % 4 == 0 && count > 0, "bucket must be 128-aligned");
    B200DP_REQUIRE(partials == nullptr || (partial_stride % 4 == 0 && ((uintptr_t)partials & 15) == 0),
                   "partial rows must be 16-byte aligned with a row pitch that is a multiple of 4 floats");
    B200DP_REQUIRE(opt_kind >= 0 && opt_kind <= 2, "opt_kind 0 adadelta, 1 sgd, 2 adam");
    RUArgs a{};
    a.param = param; a.grad = grad; a.state0 = state0; a.state1 = state1; a.shadow = shadow_bf16;
    a.start = start; a.count = count;
    a.partials = partials; a.n_partials = n_partials; a.partial_stride = partial_stride;
    a.opt_kind = opt_kind; a.hyper = hyper; a.step_counter = step_counter; a.last_bucket = last_bucket;
    a.loss_src = loss_src; a.loss_sink = last_bucket ? (const unsigned long long*)loss_sink : nullptr;
    if (comm) {
        a.comm = *comm;
        B200DP_REQUIRE(comm->world >= 1 && comm->world <= 8, "world size 1..8");
        B200DP_REQUIRE(comm->algo >= 0 && comm->algo <= 2, "algo: 0 one-shot, 1 two-shot, 2 NVLS multimem");
        B200DP_REQUIRE(comm->algo != 2 || comm->mc_stage != nullptr, "NVLS needs the multicast alias of the staging block");
        B200DP_REQUIRE(comm->world == 1 || (start + count) / RU_FLAG_GRAIN <= comm->max_slices, "flag block too small");
    } else {
        a.comm.world = 1; a.comm.rank = 0;
    }
    if (partials) {
        a.slice = RU_PSLICE;
    } else {
        const int cap = (comm && comm->max_ctas > 0 && comm->max_ctas < RU_MAX_CTAS) ? comm->max_ctas : RU_MAX_CTAS;
        const int per_cta = (count + cap - 1) / cap;
        a.slice = (per_cta + RU_CHUNK - 1) / RU_CHUNK * RU_CHUNK;
        if (a.slice < RU_CHUNK) a.slice = RU_CHUNK;
    }
    const int grid = (count + a.slice - 1) / a.slice;
    reduce_update_kernel<<<grid, RU_NT, 0, (cudaStream_t)s>>>(a);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
