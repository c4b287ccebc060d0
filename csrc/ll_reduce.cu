// Push-based gradient all-reduce fused with the average and the optimizer update: the latency-regime
// form of the product kernel (whole-model gradients up to ~1 MB: the reference LeNet is 247 KB).
//
// Replaces reference train.py:34-39 (ten blocking ncclAllReduce + ten divides), train.py:51
// (optimizer.step) and train.py:46 (zero_grad).  No NCCL call and no separate elementwise kernel.
//
// Why a second protocol next to reduce_update.cu: the flag-handshake one-shot there is a PULL —
// publish locally, fence, raise a flag in every peer, wait for theirs, then read their buffers over
// NVLink: two fabric round trips per bucket, and a step paid three of them back to back (measured:
// 27 us exposed per step at 8 GPUs).  Here every rank PUSHES its gradient into every peer's
// receive area as soon as a bucket exists, in 8-byte lines that carry their own flag
// {payload, generation} (8-byte stores are single-copy atomic, so a line is either old or complete;
// the layout NCCL calls LL).  Nothing is fenced and nobody waits: the receiver polls LOCAL memory,
// element by element, and the data of the early buckets has long arrived when the step's single
// wait+sum+update kernel runs.  One fabric traversal, one exposed wait per step.
//
//   b200dp_grad_push         fire-and-forget: grad[start, start+count) -> every peer's receive area
//   b200dp_reduce_update_ll  for an arena range: (a) sums the conv-backward partial rows of one sub-range
//                            and pushes that sub-range itself (and, push_rest, the rest of the range),
//                            (b) for every float4 of the range waits for each peer's line(s), adds in
//                            rank order (bit-identical on all ranks), (c) x 1/world, Adadelta / SGD / Adam
//                            on the fp32 master weights, averaged gradient written back, optional bf16
//                            shadow weights, (d) when it is the step's LAST launch: advances the generation
//                            (and can deliver the loss to the host).
//   b200dp_post_loss         one thread: the loss into the host's mapped array, beside the backward pass
// The default LeNet step uses two b200dp_reduce_update_ll launches: the dense buckets (95 % of the bytes:
// push + wait + update in one launch on a side stream, hidden behind the convolution backward) and, last,
// the 10 KB convolution bucket on the main stream.  A step's one-way NVLink latency is paid once.
//
// Receive area of a rank: [2 parities][8 source ranks][arena_len x 2 words].  Parity = generation & 1:
// a rank can be at most one step ahead of a peer (it needs the peer's lines of step g to finish step
// g), so the lines of step g are never overwritten before the peer has consumed them.  The flag IS
// the generation number, which never repeats, so nothing is ever cleared.
// Wire formats: fp32 (line = {float, gen}) or bf16 (line = {2 x bf16, gen}: half the NVLink bytes;
// every rank, the sender included, sums the same bf16-rounded values in fp32, in rank order).
//
// A peer that never delivers: the poll is bounded by wall time (comm.timeout_ms); the CTA then raises
// the sticky error word step_counter[2], skips its update, and STILL takes its completion ticket, so
// the generation keeps advancing and every later launch sees the error word and does nothing but
// bookkeeping.  The host reads the word at every log point and at epoch end (Trainer.check_peers, a 4-byte copy on
// a stream of its own) and raises; ranks that are still alive time out on the raiser's next step.
#include "opt_math.cuh"

namespace b200dp {

constexpr int LL_NT = 256;
constexpr int LL_CHUNK = 1024;         // floats per CTA iteration (one float4 per thread)
constexpr int LL_PSLICE = 128;         // floats per CTA on the partial-row sub-range
constexpr int LL_MAX_CTAS = 148 * 2;
constexpr unsigned long long LL_DEFAULT_TIMEOUT_NS = 20ull * 1000ull * 1000ull * 1000ull;

struct LLArgs {
    float* param; float* grad; float* state0; float* state1; uint16_t* shadow;
    int start, count;                  // arena range this launch reduces and updates
    const float* partials; int n_partials, partial_stride;
    int p_start, p_count;              // sub-range whose local gradient is the sum of the partial rows (p_count 0: none)
    int push_rest;                     // 1: this launch also pushes the rest of the range (nobody called grad_push)
    int last;                          // 1: the step's last launch (advances the generation, takes the completion ticket)
    int opt_kind;
    const float* hyper;
    uint32_t* step_counter;
    const float* loss_src; const unsigned long long* loss_sink;
    b200dp_comm comm;
    int n_pcta, n_chunks, n_chunks_a;  // CTAs on the partial sub-range; 1024-float chunks of the rest; of which before p_start
};

__device__ __forceinline__ void st_volatile_u4(uint32_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_u4(const uint32_t* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    return (uint32_t)f32_to_bf16_rn(lo) | ((uint32_t)f32_to_bf16_rn(hi) << 16);
}
// what every rank adds for a contribution that travelled as bf16 (the sender rounds its own copy the same way)
__device__ __forceinline__ float4 round_bf16x4(const float4& g) {
    return make_float4(bf16_to_f32(f32_to_bf16_rn(g.x)), bf16_to_f32(f32_to_bf16_rn(g.y)),
                       bf16_to_f32(f32_to_bf16_rn(g.z)), bf16_to_f32(f32_to_bf16_rn(g.w)));
}

// word offset of the lines of arena float4 `idx` sent by `src`, inside a rank's receive area
__device__ __forceinline__ long ll_offset(const b200dp_comm& c, int parity, int src, int idx) {
    return ((long)(parity * 8 + src) * c.arena_len + (c.wire ? (idx >> 1) : idx)) * 2;
}

// write the float4 at arena index idx into every peer's receive area
__device__ __forceinline__ void ll_push4(const b200dp_comm& c, int parity, uint32_t gen, int idx, const float4& g) {
    const long off = ll_offset(c, parity, c.rank, idx);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        if (r < c.world && r != c.rank) {
            uint32_t* dst = c.ll[r] + off;
            if (c.wire) {
                st_volatile_u4(dst, pack_bf16x2(g.x, g.y), gen, pack_bf16x2(g.z, g.w), gen);
            } else {
                st_volatile_u4(dst, __float_as_uint(g.x), gen, __float_as_uint(g.y), gen);
                st_volatile_u4(dst + 4, __float_as_uint(g.z), gen, __float_as_uint(g.w), gen);
            }
        }
    }
}

// Wait for every peer's lines of arena float4 idx and add all contributions in rank order.  Returns false on time-out.
__device__ __forceinline__ bool ll_gather4(const b200dp_comm& c, int parity, uint32_t gen, int idx, const float4& own,
                                           unsigned long long timeout_ns, float4& sum) {
    const uint32_t* mine = nullptr;
#pragma unroll
    for (int r = 0; r < 8; ++r)
        if (r == c.rank) mine = c.ll[r];
    float4 v[8];
    unsigned pending = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r)
        if (r < c.world && r != c.rank) pending |= 1u << r;
    unsigned spins = 0;
    unsigned long long t0 = 0;
    while (pending) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (pending & (1u << r)) {
                const uint32_t* src = mine + ll_offset(c, parity, r, idx);
                if (c.wire) {
                    const uint4 a = ld_volatile_u4(src);
                    if (a.y == gen && a.w == gen) {
                        v[r] = make_float4(bf16_to_f32(a.x & 0xffffu), bf16_to_f32(a.x >> 16),
                                           bf16_to_f32(a.z & 0xffffu), bf16_to_f32(a.z >> 16));
                        pending &= ~(1u << r);
                    }
                } else {
                    const uint4 a = ld_volatile_u4(src), b = ld_volatile_u4(src + 4);
                    if (a.y == gen && a.w == gen && b.y == gen && b.w == gen) {
                        v[r] = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
                        pending &= ~(1u << r);
                    }
                }
            }
        }
        if (pending && (++spins & 255u) == 0u) {
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > timeout_ns) return false;
        }
    }
    sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < 8; ++r)
        if (r < c.world) sum = add4(sum, (r == c.rank) ? own : v[r]);
    return true;
}

__global__ void __launch_bounds__(LL_NT) grad_push_kernel(const float* grad, int start, int count,
                                                          const uint32_t* step_counter, b200dp_comm comm) {
    if (step_counter[2] != 0u) return;                   // the job is already in error: stay silent
    const uint32_t gen = step_counter[0] + 1u;
    const int parity = gen & 1u;
    for (int i = (blockIdx.x * LL_NT + threadIdx.x) * 4; i < count; i += gridDim.x * LL_NT * 4)
        ll_push4(comm, parity, gen, start + i, *reinterpret_cast<const float4*>(grad + start + i));
}

__global__ void __launch_bounds__(LL_NT, 2) reduce_update_ll_kernel(LLArgs a) {
    pdl_wait();      // launched with programmatic serialization behind the last backward kernel (no-op otherwise)
    __shared__ __align__(16) float sRed[8][128];
    __shared__ __align__(16) float sSlice[LL_PSLICE];
    const int tid = threadIdx.x;
    const uint32_t gen = a.step_counter[0] + 1u;
    const int parity = gen & 1u;
    const int world = a.comm.world;
    bool dead = a.step_counter[2] != 0u;                 // sticky: an earlier step timed out
    const unsigned long long timeout_ns =
        a.comm.timeout_ms ? (unsigned long long)a.comm.timeout_ms * 1000000ull : LL_DEFAULT_TIMEOUT_NS;
    const float inv = __fdiv_rn(1.f, (float)world);

    RUCtx cx;
    cx.kind = a.opt_kind; cx.param = a.param; cx.grad = a.grad; cx.state0 = a.state0; cx.state1 = a.state1;
    cx.shadow = a.shadow;
    ru_load_hyper(cx, a.hyper, gen);

    if (!dead) {
        if ((int)blockIdx.x < a.n_pcta) {
            // ---- the sub-range that arrives as partial rows: sum, push, gather, update (one float4 per lane of warp 0) ----
            const int off = blockIdx.x * LL_PSLICE;                      // inside the sub-range
            const int n_here = min(LL_PSLICE, a.p_count - off);
            ru_sum_partial_rows(a.partials, a.n_partials, a.partial_stride, off, n_here, sRed, sSlice, tid);   // row column i == arena element p_start + i
            bool bad = false;
            if (tid < 32 && tid * 4 < n_here) {
                const int idx = a.p_start + off + tid * 4;
                float4 g = *reinterpret_cast<const float4*>(sSlice + tid * 4);
                if (world > 1) {
                    ll_push4(a.comm, parity, gen, idx, g);
                    const float4 own = a.comm.wire ? round_bf16x4(g) : g;
                    // observability: how long the step's LAST bucket (pushed just above by every rank) kept CTA 0 waiting
                    // for the peers = the exposed part of the exchange, in ns, readable at step_counter[3]
                    const bool probe = blockIdx.x == 0 && tid == 0;
                    const unsigned long long t0 = probe ? globaltimer_ns() : 0ull;
                    bad = !ll_gather4(a.comm, parity, gen, idx, own, timeout_ns, g);
                    if (probe) a.step_counter[3] = (uint32_t)min(globaltimer_ns() - t0, 0xffffffffull);
                }
                if (!bad) ru_update(cx, idx, scale4(g, inv));
            }
            if (__syncthreads_or(bad)) {
                if (tid == 0) atomicExch(a.step_counter + 2, 1u);
            }
        } else {
            // ---- the rest of the range: gradients are in the arena (and, unless push_rest, already on their way) ----
            const int len_a = a.p_count > 0 ? a.p_start - a.start : a.count;     // floats before the sub-range
            const int stride = gridDim.x - a.n_pcta;
            for (int c = blockIdx.x - a.n_pcta; c < a.n_chunks; c += stride) {
                int idx, end;
                if (c < a.n_chunks_a) { idx = a.start + c * LL_CHUNK; end = a.start + len_a; }
                else { idx = a.p_start + a.p_count + (c - a.n_chunks_a) * LL_CHUNK; end = a.start + a.count; }
                idx += tid * 4;
                bool bad = false;
                if (idx < end) {
                    float4 g = *reinterpret_cast<const float4*>(a.grad + idx);
                    const RULoaded L = ru_load(cx, idx);
                    if (world > 1) {
                        if (a.push_rest) ll_push4(a.comm, parity, gen, idx, g);
                        const float4 own = a.comm.wire ? round_bf16x4(g) : g;
                        bad = !ll_gather4(a.comm, parity, gen, idx, own, timeout_ns, g);
                    }
                    if (!bad) ru_apply(cx, idx, scale4(g, inv), L);
                }
                if (__syncthreads_or(bad)) {
                    if (tid == 0) atomicExch(a.step_counter + 2, 1u);
                    break;
                }
            }
        }
    }
    if (a.last) ru_finish_step(a.step_counter, gen, gridDim.x, a.loss_src, a.loss_sink, tid);
}

}  // namespace b200dp

using namespace b200dp;

// The step's loss goes to the host as a 4-byte store into a mapped pinned array (no copy command).  A kernel that has
// stored to host memory stays alive until the store has crossed PCIe (~5 us), so the store gets a one-thread kernel of
// its own on a stream beside the backward pass instead of riding in a kernel of the step's dependency chain.
__global__ void post_loss_kernel(const float* loss_src, const uint32_t* step_counter, const unsigned long long* sink) {
    const unsigned long long ring = sink[0], idx = (unsigned long long)(step_counter[0] + 1u) - sink[1];
    if (ring != 0ull && idx < sink[2]) reinterpret_cast<volatile float*>(ring)[idx] = __ldcg(loss_src);
}

extern "C" int b200dp_post_loss(const float* loss_src, const uint32_t* step_counter, const uint64_t* loss_sink,
                                b200dp_stream_t s) {
    B200DP_REQUIRE(loss_src && step_counter && loss_sink, "null argument");
    post_loss_kernel<<<1, 1, 0, (cudaStream_t)s>>>(loss_src, step_counter, (const unsigned long long*)loss_sink);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

static int ll_check_comm(const b200dp_comm* comm, int start, int count) {
    B200DP_REQUIRE(comm->world >= 2 && comm->world <= 8, "world size 2..8");
    B200DP_REQUIRE(comm->rank >= 0 && comm->rank < comm->world, "rank out of range");
    B200DP_REQUIRE(start + count <= comm->arena_len, "range beyond the receive area");
    for (int r = 0; r < comm->world; ++r) B200DP_REQUIRE(comm->ll[r] != nullptr, "receive area of a rank is missing (comm.ll)");
    return 0;
}

extern "C" int b200dp_grad_push(const float* grad, int start, int count, const uint32_t* step_counter,
                                const b200dp_comm* comm, b200dp_stream_t s) {
    B200DP_REQUIRE(grad && step_counter && comm, "null argument");
    B200DP_REQUIRE(start % 4 == 0 && count % 4 == 0 && count > 0, "range must be float4-aligned");
    int rc = ll_check_comm(comm, start, count);
    if (rc != 0) return rc;
    int grid = (count / 4 + LL_NT - 1) / LL_NT;
    if (grid > LL_MAX_CTAS) grid = LL_MAX_CTAS;
    grad_push_kernel<<<grid, LL_NT, 0, (cudaStream_t)s>>>(grad, start, count, step_counter, *comm);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int b200dp_reduce_update_ll(float* param, float* grad, float* state0, float* state1, uint16_t* shadow_bf16,
                                       int start, int count, const float* partials, int n_partials,
                                       int partial_stride, int p_start, int p_count, int push_rest, int last, int opt_kind,
                                       const float* hyper, uint32_t* step_counter, const b200dp_comm* comm,
                                       const float* loss_src, const uint64_t* loss_sink, b200dp_stream_t s) {
    B200DP_REQUIRE(start % 4 == 0 && count % 4 == 0 && count > 0, "range must be float4-aligned");
    B200DP_REQUIRE(opt_kind >= 0 && opt_kind <= 2, "opt_kind 0 adadelta, 1 sgd, 2 adam");
    if (partials == nullptr) p_count = 0;
    if (p_count > 0) {
        B200DP_REQUIRE(p_start >= start && p_start + p_count <= start + count && p_start % 4 == 0 && p_count % 4 == 0,
                       "partial-row sub-range must lie inside the range, float4-aligned");
        B200DP_REQUIRE(partial_stride % 4 == 0 && ((uintptr_t)partials & 15) == 0 && n_partials > 0,
                       "partial rows must be 16-byte aligned with a row pitch that is a multiple of 4 floats");
    } else {
        p_start = start;
    }
    LLArgs a{};
    a.param = param; a.grad = grad; a.state0 = state0; a.state1 = state1; a.shadow = shadow_bf16;
    a.start = start; a.count = count;
    a.partials = partials; a.n_partials = n_partials; a.partial_stride = partial_stride;
    a.p_start = p_start; a.p_count = p_count; a.push_rest = push_rest; a.last = last;
    a.opt_kind = opt_kind; a.hyper = hyper; a.step_counter = step_counter;
    a.loss_src = loss_src; a.loss_sink = (const unsigned long long*)loss_sink;   // sink without src: error mirror only
    if (comm && comm->world > 1) {
        int rc = ll_check_comm(comm, start, count);
        if (rc != 0) return rc;
        a.comm = *comm;
    } else {
        a.comm.world = 1; a.comm.rank = 0;
    }
    const int len_a = p_count > 0 ? p_start - start : count;
    const int len_b = p_count > 0 ? (start + count) - (p_start + p_count) : 0;
    a.n_pcta = (p_count + LL_PSLICE - 1) / LL_PSLICE;
    a.n_chunks_a = (len_a + LL_CHUNK - 1) / LL_CHUNK;
    a.n_chunks = a.n_chunks_a + (len_b + LL_CHUNK - 1) / LL_CHUNK;
    B200DP_REQUIRE(a.n_pcta <= LL_MAX_CTAS, "partial-row sub-range too long for one launch");
    int rest = a.n_chunks;
    if (rest > LL_MAX_CTAS - a.n_pcta) rest = LL_MAX_CTAS - a.n_pcta;   // co-resident: spinning CTAs never block a CTA they wait for
    if (rest < 1 && a.n_chunks > 0) rest = 1;
    const int grid = a.n_pcta + rest;
    B200DP_REQUIRE(grid > 0, "empty launch");
    B200DP_CHECK(launch_pdl(reduce_update_ll_kernel, dim3(grid), dim3(LL_NT), 0, (cudaStream_t)s, a));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
