// Fused head: conv3 (a GEMM: the map is 5x5) + LeakyReLU -> FC.0 + LeakyReLU -> FC.2 -> softmax -> loss, and the
// part of the backward that is local to a batch row: dlogits, dz1 = (dlogits W2) * leaky'(h1),
// dz3 = (dz1 W1) * leaky'(a3).  ONE launch, exact fp32 on CUDA cores.
//
// Replaces reference cnn.py:16-17, 20-23 + train.py:48 and the row-local half of their autograd backward
// (train.py:49): three cuDNN/cuBLAS calls, five ATen kernels forward and about as many backward.  It also replaces
// three launches of this library's own first-generation step (conv3 GEMM 6.1 us, FC.0 GEMM 2.9 us, head_tail 8.0 us
// warm) and a third of a fourth (dz3).  Every layer after conv2 is a chain over ONE batch row — only the weight
// gradients need the whole batch — so a CTA takes 4 rows through the whole chain with all intermediates in shared
// memory; what is left for the backward GEMM launch is {dW FC.2, dW FC.0, dW conv3, dP2}, all functions of this
// kernel's outputs only.
//
// Work split inside a CTA (512 threads, 4 batch rows): a dense layer out[r][n] = sum_k in[r][k] W(k,n) is computed with
// thread = (two output columns, one of 8 k-groups) holding 2 x 4 row accumulators; the weights stream through shared memory in
// 64-deep k-slabs (a 4-stage ring running three slabs ahead, filled by TMA tensor-map copies — or by cp.async for shapes
// TMA cannot address —, see "weight stream"), inputs sit transposed [k][4 rows] so
// one 128-bit broadcast read serves 4 FMAs.  The k-quarters are added in fixed order: bit-reproducible.  (8 rows x 256
// threads, the first layout, left 2 warps per scheduler to hide every shared-memory latency: 25 us.)
// FC.2 / softmax / loss / dlogits / dz1 are spread over the 16 warps in three short stages (84 x 10: too small for the
// slab machinery).
#include "common.cuh"
#include "tc_common.cuh"

namespace b200dp {

constexpr int HF_NMAX_ = 128;
constexpr int HF_R = 4;          // batch rows per CTA
constexpr int HF_CT = 64;         // column threads: thread c owns output columns c and c + 64
constexpr int HF_KG = 8;         // k-groups: a slab's 64 k are split over 8 groups of 64 threads
constexpr int HF_NT = HF_CT * HF_KG;
constexpr int HF_WARPS = HF_NT / 32;
constexpr int HF_BK = 64;        // k-slab depth
constexpr int HF_NMAX = HF_NMAX_; // widest layer handled (conv3 out 120, FC.0 out 84)
constexpr int HF_MAXC = 16;      // classes
constexpr int HF_SLAB_REGS = HF_BK * HF_NMAX / HF_NT;   // weights per thread per slab
constexpr int HF_KPT = HF_BK / HF_KG;                    // k per thread per slab
static_assert(HF_R == 4, "one 128-bit read per k serves the 4 row accumulators");
static_assert(HF_KPT == 8, "a thread's k range of a slab is two 16-byte chunks");
constexpr int HF_STAGES = 4;     // weight-slab ring depth

struct HeadFusedArgs {
    const float* p2; int ld_p2;
    const float* w3; const float* b3; const float* w1; const float* b1; const float* w2; const float* b2;
    const int64_t* target;
    float* a3; int ld_a3; float* h1; int ld_h1;
    float* probs; float* loss_rows; float* loss_out;
    float* dlogits; int ld_dl; float* dz1; float* dz3;      // null in eval mode
    double* eval_acc; int* ticket;
    int B, K3, C3, F1, NC, loss_mode;
    long long* dbg;          // profiling aid: CTA 0 stores clock64() stamps of its phases here (NULL = off)
    // TMA variant: tensor maps over W3 [C3][K3] (box 32 k x 128 n), W1 [F1][C3] as k-major operand (box 32 k x 128 n)
    // and as n-major operand (box 32 n x 64 m); SWIZZLE_128B, out-of-bounds elements arrive as zeros
    CUtensorMap tm3, tm1k, tm1n;
};

// ---- weight stream ------------------------------------------------------------------------------------------
// All dense layers of the kernel consume ONE static sequence of 64-deep weight slabs (conv3's, then FC.0's, then FC.0
// transposed for dz3).  The slabs travel global -> shared by TMA (cp.async.bulk.tensor.2d through SWIZZLE_128B tensor maps,
// one elected thread, completion on an mbarrier per stage: 12.7 us per launch) or, for weight shapes whose rows are not
// 16-byte multiples, by cp.async (15.5 us) through a ring of HF_STAGES buffers that runs HF_STAGES-1 slabs ahead of the math and across layer
// boundaries: a CTA has to stream all 232 KB of weights through one SM, so what matters is that several slabs are
// always in flight (first generation: register prefetch one slab ahead = one L2 latency per slab, 23.6 us).
// Wt(k, n) = W[n * sn + k * sk]; exactly one of sn / sk is 1.
// In shared memory a slab keeps the operand's own orientation, so that 16 bytes that are contiguous in global memory
// stay contiguous: k-major layers (conv3, FC.0: W[n][k]) as [n][36] (a thread reads its column's 8 k as two 128-bit
// loads: 8 lanes x 16 B at a 144-byte pitch cover the 32 banks), the n-major layer (dz3: W1[m][n]) as [kk][132].
// One 16-byte cp.async moves what four 4-byte ones did: the first ring version issued 128 warp-wide LDGSTS.32 per
// slab and spent ~1000 cycles per slab in the load/store unit, whatever the thread count (25-30 us per launch).
struct HfLayer { const float* W; int sn, sk, N, K, nslab; };
constexpr int HF_PK = HF_BK + 4;            // k-major slab pitch (floats)
constexpr int HF_PN = HF_NMAX + 4;          // n-major slab pitch
constexpr int HF_STAGE = HF_NMAX * HF_PK;   // floats per ring stage (>= HF_BK * HF_PN)
constexpr uint32_t HF_TMA_SLAB_BYTES = HF_BK * HF_NMAX * 4;   // a TMA slab is dense: 64 x 128 floats
static_assert(HF_STAGE * 4 % 1024 == 0 && HF_STAGE * 4 >= (int)HF_TMA_SLAB_BYTES, "SWIZZLE_128B stages: 1024-byte aligned");
static_assert(HF_STAGE >= HF_BK * HF_PN, "stage must hold either orientation");

__device__ __forceinline__ void hf_cp(float* dst, const float* src, int bytes, bool valid) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int n = valid ? bytes : 0;                       // src-size 0: the destination is zero-filled
    if (bytes == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void hf_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void hf_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// 16-byte copies need rows that start on 16-byte boundaries and hold a multiple of 4 floats
__device__ __forceinline__ bool hf_vec_ok(const HfLayer& L) {
    const int row = (L.sk == 1) ? L.sn : L.sk, inner = (L.sk == 1) ? L.K : L.N;
    return ((reinterpret_cast<uintptr_t>(L.W) & 15) == 0) && (row % 4 == 0) && (inner % 4 == 0);
}

// Issue slab `t` of layer L into ring stage sW (all threads).  Global reads are coalesced along the operand's
// unit-stride index either way.  The code is deliberately compact (loops, one instance): 16 warps share one SM's
// instruction cache.
__device__ __forceinline__ void hf_issue(const HfLayer& L, int t, float* sW, int tid) {
    const int k0 = t * HF_BK;
    const bool kmajor = L.sk == 1;
    if (hf_vec_ok(L)) {
        // 2048 16-byte pieces per slab, 4 per thread; a thread's position along the unit-stride index is the same for
        // all four (only a row offset changes), so the address arithmetic is a handful of integer operations
        constexpr int PER = HF_BK * HF_NMAX / 4 / HF_NT;
        if (kmajor) {
            constexpr int K4 = HF_BK / 4, ROWS = HF_NT / K4;           // 16 pieces per row, 32 rows per pass
            const int k4 = tid % K4, n0 = tid / K4;
            const bool kok = k0 + k4 * 4 < L.K;
            const float* src = L.W + (long)n0 * L.sn + (k0 + k4 * 4);
            float* dst = sW + n0 * HF_PK + k4 * 4;
#pragma unroll
            for (int j = 0; j < PER; ++j) {
                const bool ok = kok && n0 + j * ROWS < L.N;
                hf_cp(dst + j * ROWS * HF_PK, ok ? src + (long)j * ROWS * L.sn : L.W, 16, ok);
            }
        } else {
            constexpr int N4 = HF_NMAX / 4, ROWS = HF_NT / N4;         // 32 pieces per row, 16 rows per pass
            const int n4 = tid % N4, kk0 = tid / N4;
            const bool nok = n4 * 4 < L.N;
            const float* src = L.W + (long)(k0 + kk0) * L.sk + n4 * 4;
            float* dst = sW + kk0 * HF_PN + n4 * 4;
#pragma unroll
            for (int j = 0; j < PER; ++j) {
                const bool ok = nok && k0 + kk0 + j * ROWS < L.K;
                hf_cp(dst + j * ROWS * HF_PN, ok ? src + (long)j * ROWS * L.sk : L.W, 16, ok);
            }
        }
    } else {
#pragma unroll 2
        for (int e = tid; e < HF_BK * HF_NMAX; e += HF_NT) {
            const int kk = kmajor ? (e % HF_BK) : (e / HF_NMAX);
            const int n = kmajor ? (e / HF_BK) : (e % HF_NMAX);
            const bool ok = n < L.N && k0 + kk < L.K;
            hf_cp(sW + (kmajor ? n * HF_PK + kk : kk * HF_PN + n), L.W + (ok ? (long)n * L.sn + (long)(k0 + kk) * L.sk : 0), 4, ok);
        }
    }
}

struct HfStream {
    HfLayer layer[3];
    int n_layers, total;       // slabs in the whole sequence
    int issued;                // next slab (global index) to issue
    float* ring;               // [HF_STAGES][HF_STAGE]

    __device__ __forceinline__ HfLayer pick(int l) const {      // compile-time indices only: the struct stays in registers
        HfLayer L = layer[0];
        if (l == 1) L = layer[1];
        if (l == 2) L = layer[2];
        return L;
    }
    // TMA variant: ONE thread issues a slab as 2 (k-major: two 32-k boxes of 128 rows) or 4 (n-major: four 32-n boxes of
    // 64 rows) tensor copies of 8-16 KB that complete on the stage's mbarrier — no load/store-unit work at all, where the
    // cp.async ring spends 64 warp-wide LDGSTS per slab.
    __device__ __forceinline__ void issue_next_tma(int tid, const HeadFusedArgs& a, uint32_t bars) {
        if (tid == 0 && issued < total) {
            int g = issued, l = 0;
            if (g >= layer[0].nslab) { g -= layer[0].nslab; l = 1; }
            if (l == 1 && g >= layer[1].nslab) { g -= layer[1].nslab; l = 2; }
            const int st = issued % HF_STAGES;
            const uint32_t dst = smem_u32(ring + st * HF_STAGE), bar = bars + 8u * st;
            fence_proxy_async_smem();                    // the stage was last read through the generic proxy
            mbar_expect_tx(bar, HF_TMA_SLAB_BYTES);
            if (l == 2) {
#pragma unroll
                for (int b = 0; b < 4; ++b) tma_load_2d(dst + b * (HF_BK * 128), &a.tm1n, bar, 32 * b, g * HF_BK);
            } else {
                const CUtensorMap* tm = (l == 0) ? &a.tm3 : &a.tm1k;
#pragma unroll
                for (int b = 0; b < 2; ++b) tma_load_2d(dst + b * (HF_NMAX * 128), tm, bar, g * HF_BK + 32 * b, 0);
            }
        }
        ++issued;
    }
    __device__ __forceinline__ void issue_next(int tid) {
        if (issued < total) {
            int g = issued, l = 0;
            if (g >= layer[0].nslab) { g -= layer[0].nslab; l = 1; }
            if (l == 1 && g >= layer[1].nslab) { g -= layer[1].nslab; l = 2; }
            hf_issue(pick(l), g, ring + (issued % HF_STAGES) * HF_STAGE, tid);
        }
        ++issued;
        hf_commit();           // one group per slot, empty past the end: keeps the wait_group arithmetic uniform
    }
};

// acc[c][r] (thread = columns {cn, cn + 64}, k-group kg) += sum over this thread's k of sIn[k][r] * Wt(k, column) for a
// layer whose slabs are global slabs [g0, g0 + nslab).  sIn: [ceil64(K)][HF_R] floats, rows k >= K zero.  One 128-bit
// broadcast read of the inputs serves both columns: 12 shared-memory reads per 64 FMAs.  The barrier at the top of
// every slab also publishes whatever the caller wrote to shared memory before the call.
template <bool TMA>
__device__ __forceinline__ void hf_gemm(HfStream& st, int g0, int nslab, bool kmajor, const float* sIn,
                                        float (&acc)[2][HF_R], int tid, const HeadFusedArgs& a, uint32_t bars) {
    const int cn = tid & (HF_CT - 1), kg = tid / HF_CT;
    for (int t = 0; t < nslab; ++t) {
        if constexpr (TMA) {
            mbar_wait(bars + 8u * ((g0 + t) % HF_STAGES), ((g0 + t) / HF_STAGES) & 1);   // slab g0+t has landed
            __syncthreads();           // everyone is done reading slab g0+t-1
            st.issue_next_tma(tid, a, bars);
        } else {
            hf_wait<HF_STAGES - 2>();  // this thread's copies of slab g0+t have landed ...
            __syncthreads();           // ... everyone's have, and everyone is done reading slab g0+t-1
            st.issue_next(tid);        // refill the buffer slab g0+t-1 occupied
        }
        const float* in = sIn + (t * HF_BK + kg * HF_KPT) * HF_R;
        const float* stage = st.ring + ((g0 + t) % HF_STAGES) * HF_STAGE;
        float wv[2][HF_KPT];
        if constexpr (TMA) {
            // SWIZZLE_128B box: logical (row r, 16-byte chunk q) sits at r * 128 + ((q ^ (r & 7)) << 4)
            const char* sb = reinterpret_cast<const char*>(stage);
            if (kmajor) {      // boxes = k halves, rows = output columns; this thread's 8 k are chunks 2 kq, 2 kq + 1 of its box
                const int box = kg / 4, kq = kg & 3;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const int n = cn + c * HF_CT;
                    const char* rowp = sb + box * (HF_NMAX * 128) + n * 128;
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const float4 w4 = *reinterpret_cast<const float4*>(rowp + (((2 * kq + q) ^ (n & 7)) << 4));
                        wv[c][q * 4] = w4.x; wv[c][q * 4 + 1] = w4.y; wv[c][q * 4 + 2] = w4.z; wv[c][q * 4 + 3] = w4.w;
                    }
                }
            } else {           // boxes = 32-column groups, rows = k; row index & 7 == kk because kg * 8 is a multiple of 8
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const int n = cn + c * HF_CT;
                    const char* colp = sb + (n >> 5) * (HF_BK * 128) + (n & 3) * 4;
                    const int q = (n & 31) >> 2;
#pragma unroll
                    for (int kk = 0; kk < HF_KPT; ++kk)
                        wv[c][kk] = *reinterpret_cast<const float*>(colp + (kg * HF_KPT + kk) * 128 + ((q ^ kk) << 4));
                }
            }
        } else if (kmajor) {
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
                for (int q = 0; q < HF_KPT / 4; ++q) {
                    const float4 w4 = *reinterpret_cast<const float4*>(stage + (cn + c * HF_CT) * HF_PK + kg * HF_KPT + q * 4);
                    wv[c][q * 4] = w4.x; wv[c][q * 4 + 1] = w4.y; wv[c][q * 4 + 2] = w4.z; wv[c][q * 4 + 3] = w4.w;
                }
        } else {
#pragma unroll
            for (int kk = 0; kk < HF_KPT; ++kk) {
                wv[0][kk] = stage[(kg * HF_KPT + kk) * HF_PN + cn];
                wv[1][kk] = stage[(kg * HF_KPT + kk) * HF_PN + cn + HF_CT];
            }
        }
#pragma unroll
        for (int kk = 0; kk < HF_KPT; ++kk) {
            const float4 x0 = *reinterpret_cast<const float4*>(in + kk * HF_R);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                acc[c][0] = fmaf(wv[c][kk], x0.x, acc[c][0]); acc[c][1] = fmaf(wv[c][kk], x0.y, acc[c][1]);
                acc[c][2] = fmaf(wv[c][kk], x0.z, acc[c][2]); acc[c][3] = fmaf(wv[c][kk], x0.w, acc[c][3]);
            }
        }
    }
}

// Fixed-order sum of the k-groups: every group parks its accumulators; then thread t owns ONE output element
// (row t / 128, column t % 128) and adds the 8 partial sums in group order (the first version let the 64 threads of
// group 0 add everything: 1700 cycles per layer on two warps).  Returns the pre-activation of the caller's element.
__device__ __forceinline__ float hf_combine(float* sRed, const float (&acc)[2][HF_R], int tid) {
    const int cn = tid & (HF_CT - 1), kg = tid / HF_CT;
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int r = 0; r < HF_R; ++r) sRed[(kg * HF_R + r) * HF_NMAX + cn + c * HF_CT] = acc[c][r];
    __syncthreads();
    const int r = tid / HF_NMAX, n = tid & (HF_NMAX - 1);
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < HF_KG; ++q) v += sRed[(q * HF_R + r) * HF_NMAX + n];
    return v;
}
static_assert(HF_NT == HF_R * HF_NMAX, "one output element per thread in the layer epilogues");

template <bool TMA>
__global__ void __launch_bounds__(HF_NT, 1) head_fused_kernel(const __grid_constant__ HeadFusedArgs a) {
    extern __shared__ __align__(1024) float smem[];     // SWIZZLE_128B ring stages need 1024-byte alignment
    __shared__ __align__(8) unsigned long long sBar[HF_STAGES];
    const uint32_t bars = smem_u32(sBar);
    const int K3P = (a.K3 + HF_BK - 1) / HF_BK * HF_BK;
    float* sX = smem;                                   // [K3P][8]   conv3 input rows, transposed
    float* sW = sX + K3P * HF_R;                        // [HF_STAGES][HF_STAGE]  weight-slab ring
    float* sRed = sW + HF_STAGES * HF_STAGE;   // [8][4][128] k-group hand-off
    float* sA3 = sRed + HF_KG * HF_R * HF_NMAX;                 // [128][8]   a3 (activated), transposed; rows >= C3 zero
    float* sH1 = sA3 + HF_NMAX * HF_R;                  // [128][8]   h1 (activated), transposed; rows >= F1 zero
    float* sDz1 = sH1 + HF_NMAX * HF_R;                 // [128][8]   dz1, transposed; rows >= F1 zero
    float* sW2 = sDz1 + HF_NMAX * HF_R;                 // [NC][F1] + bias[NC]
    __shared__ int sLast;
    __shared__ float sZ[HF_R * HF_MAXC], sDz[HF_R * HF_MAXC];
    __shared__ float sSum[HF_WARPS];
    __shared__ int sCnt[HF_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cn = tid & (HF_CT - 1), kg = tid / HF_CT;
    const int row0 = blockIdx.x * HF_R;
    const bool train = a.dlogits != nullptr;
    int stamp = 0;
    auto mark = [&]() { if (a.dbg != nullptr && blockIdx.x == 0 && tid == 0) a.dbg[stamp] = clock64(); ++stamp; };
    mark();   // 0: start

    // Weights were written by the previous step's update, at least two launches back in this stream: they are safe
    // to read before the programmatic-dependency wait (only the predecessor, conv2's forward, may still be running).
    HfStream st;
    st.layer[0] = HfLayer{a.w3, a.K3, 1, a.C3, a.K3, (a.K3 + HF_BK - 1) / HF_BK};      // conv3:  Wt(k, n) = W3[n][k]
    st.layer[1] = HfLayer{a.w1, a.C3, 1, a.F1, a.C3, (a.C3 + HF_BK - 1) / HF_BK};      // FC.0:   Wt(k, n) = W1[n][k]
    st.layer[2] = HfLayer{a.w1, 1, a.C3, a.C3, a.F1, (a.F1 + HF_BK - 1) / HF_BK};      // dz3:    Wt(m, n) = W1[m][n]
    st.n_layers = train ? 3 : 2;
    st.total = st.layer[0].nslab + st.layer[1].nslab + (train ? st.layer[2].nslab : 0);
    st.issued = 0;
    st.ring = sW;
    if constexpr (TMA) {
        if (tid == 0) {
            for (int i = 0; i < HF_STAGES; ++i) mbar_init(bars + 8u * i, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
#pragma unroll 1
        for (int i = 0; i < HF_STAGES - 1; ++i) st.issue_next_tma(tid, a, bars);
        for (int i = tid; i < a.NC * a.F1; i += HF_NT) sW2[i] = __ldg(a.w2 + i);
        if (tid < a.NC) sW2[a.NC * a.F1 + tid] = __ldg(a.b2 + tid);
    } else {
        for (int i = tid; i < a.NC * a.F1; i += HF_NT) hf_cp(sW2 + i, a.w2 + i, 4, true);      // joins the first group
        if (tid < a.NC) hf_cp(sW2 + a.NC * a.F1 + tid, a.b2 + tid, 4, true);
#pragma unroll 1
        for (int i = 0; i < HF_STAGES - 1; ++i) st.issue_next(tid);
    }
    for (int i = tid; i < HF_NMAX * HF_R; i += HF_NT) { sA3[i] = 0.f; sH1[i] = 0.f; sDz1[i] = 0.f; }
    mark();   // 1: ring primed, FC.2 weights staged
    pdl_wait();
    mark();   // 2: predecessor complete
    // this warp's row target, needed only in the row tail: fetched now so that its latency is off the critical path
    const int my_target = (warp < HF_R && row0 + warp < a.B) ? (int)a.target[row0 + warp] : 0;

    // ---- stage the 8 input rows, transposed (rows past the batch and the k padding are zero) ----
    for (int i = tid; i < K3P * HF_R; i += HF_NT) {
        const int r = i / K3P, k = i - r * K3P;                       // consecutive threads: consecutive k of one row
        const int row = row0 + r;
        sX[k * HF_R + r] = (row < a.B && k < a.K3) ? a.p2[(long)row * a.ld_p2 + k] : 0.f;
    }
    // (hf_gemm's first barrier orders these stores before the first slab is consumed)
    mark();   // 3: inputs staged
    float acc[2][HF_R];
    int g0 = 0;
    // One instance of the slab loop serves the three dense layers (l = 0 conv3, 1 FC.0, 2 dz3).
#pragma unroll 1
    for (int l = 0; l < st.n_layers; ++l) {
        const float* sIn = (l == 0) ? sX : ((l == 1) ? sA3 : sDz1);
        const int nslab = st.pick(l).nslab;
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int r = 0; r < HF_R; ++r) acc[c][r] = 0.f;
        hf_gemm<TMA>(st, g0, nslab, l != 2, sIn, acc, tid, a, bars);  // its first barrier publishes the previous layer's outputs
        g0 += nslab;
        mark();   // 4 / 6 / 9: slabs of layer l consumed
        const float pre = hf_combine(sRed, acc, tid);
        const int er = tid / HF_NMAX, en = tid & (HF_NMAX - 1);          // this thread's output element
        const bool live = row0 + er < a.B;
        if (l == 2) {
            // ---- dz3 = (dz1 W1) * leaky'(a3) ----
            if (en < a.C3 && live) a.dz3[(long)(row0 + er) * a.C3 + en] = pre * leaky_grad_from_act(sA3[en * HF_R + er]);
            continue;
        }
        // ---- conv3: a3 = leaky(p2 W3^T + b3);  FC.0: h1 = leaky(a3 W1^T + b1) ----
        {
            const int nout = (l == 0) ? a.C3 : a.F1;
            const float* bias = (l == 0) ? a.b3 : a.b1;
            float* sOut = (l == 0) ? sA3 : sH1;
            float* gOut = (l == 0) ? a.a3 : a.h1;
            const int ldo = (l == 0) ? a.ld_a3 : a.ld_h1;
            if (en < nout) {
                const float v = leaky(pre + __ldg(bias + en));
                sOut[en * HF_R + er] = v;
                if (live) gOut[(long)(row0 + er) * ldo + en] = v;
            }
        }
        mark();   // 5 / 7: layer epilogue done
        if (l == 0) continue;
        __syncthreads();

        // ---- FC.2 + softmax + loss + dlogits + dz1, spread over all 16 warps (one warp per row took 6700 cycles) ----
        {
            const float* bias2 = sW2 + a.NC * a.F1;
            // logits: warp = (row, class group); a warp reduces its <= 4 classes over the F1 inputs
            const int tr = warp & (HF_R - 1), cq = warp / HF_R;
            constexpr int KJ = HF_NMAX / 32;
            float h[KJ];
#pragma unroll
            for (int j = 0; j < KJ; ++j) h[j] = sH1[(lane + 32 * j) * HF_R + tr];       // rows >= F1 are zero
#pragma unroll
            for (int i = 0; i < HF_MAXC / (HF_WARPS / HF_R); ++i) {
                const int c = cq + i * (HF_WARPS / HF_R);
                if (c < a.NC) {                                                       // warp-uniform
                    float part = 0.f;
#pragma unroll
                    for (int j = 0; j < KJ; ++j) {
                        const int k = lane + 32 * j;
                        if (k < a.F1) part = fmaf(h[j], sW2[c * a.F1 + k], part);
                    }
                    part = warp_sum(part);
                    if (lane == 0) sZ[tr * HF_MAXC + c] = part + bias2[c];
                }
            }
            __syncthreads();
            // softmax / loss / dlogits: warp r < 4 owns row r, lane c owns class c
            if (warp < HF_R) {
                const int row = row0 + warp;
                const bool rlive = row < a.B;
                const bool cls = lane < a.NC;
                const float z = cls ? sZ[warp * HF_MAXC + lane] : -INFINITY;
                const float mx = warp_max(z);
                const float ex = cls ? expf(z - mx) : 0.f;
                const float se = warp_sum(ex);
                const float p = ex / se;
                const int t = my_target;
                const float pt = __shfl_sync(0xffffffffu, p, t), zt = __shfl_sync(0xffffffffu, z, t);
                if (rlive) {
                    if (lane == 0) a.loss_rows[row] = (a.loss_mode == 0) ? -pt : -(zt - mx - logf(se));
                    if (a.probs && cls) a.probs[(long)row * a.NC + lane] = p;
                }
                if (train) {
                    const float invB = 1.f / (float)a.B, d = (lane == t) ? 1.f : 0.f;
                    float dz = (a.loss_mode == 0) ? -invB * pt * (d - p) : invB * (p - d);
                    if (!cls || !rlive) dz = 0.f;
                    if (lane < HF_MAXC) sDz[warp * HF_MAXC + lane] = dz;
                    if (cls && rlive) a.dlogits[(long)row * a.ld_dl + lane] = dz;
                }
            }
            if (train) {
                __syncthreads();
                // dz1 = (dlogits W2) * leaky'(h1): one element per thread
                const int er = tid / HF_NMAX, k = tid & (HF_NMAX - 1);
                if (k < a.F1) {
                    float g = 0.f;
                    for (int c = 0; c < a.NC; ++c) g = fmaf(sDz[er * HF_MAXC + c], sW2[c * a.F1 + k], g);
                    g *= leaky_grad_from_act(sH1[k * HF_R + er]);
                    sDz1[k * HF_R + er] = g;
                    if (row0 + er < a.B) a.dz1[(long)(row0 + er) * a.F1 + k] = g;
                }
            }
        }
        mark();   // 8: row tail (FC.2, softmax, loss, dlogits, dz1) done
    }   // layers
    mark();       // 10 (training) / 9 (evaluation): dz3 written

    // ---- last CTA: fixed-order batch reduction of the loss (deterministic), eval counters ----
    pdl_trigger();
    __syncthreads();                  // every thread's global stores happen-before thread 0's fence ...
    if (tid == 0) {
        __threadfence();              // ... which publishes them (cumulativity) before the ticket
        sLast = (atomicAdd(a.ticket, 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (!sLast) return;
    __threadfence();
    float v = 0.f;
    int ok = 0;
    for (int r = tid; r < a.B; r += HF_NT) {
        v += __ldcg(a.loss_rows + r);
        if (a.eval_acc) {
            int am = 0; float bst = -INFINITY;
            for (int c = 0; c < a.NC; ++c) {
                const float pv = __ldcg(a.probs + (long)r * a.NC + c);
                if (pv > bst) { bst = pv; am = c; }
            }
            ok += (am == (int)a.target[r]);
        }
    }
    v = warp_sum(v);
    ok = __reduce_add_sync(0xffffffffu, ok);
    if (lane == 0) { sSum[warp] = v; sCnt[warp] = ok; }
    __syncthreads();
    if (tid == 0) {
        float s = 0.f;
        int correct = 0;
#pragma unroll
        for (int w = 0; w < HF_WARPS; ++w) { s += sSum[w]; correct += sCnt[w]; }
        a.loss_out[0] = s / (float)a.B;
        if (a.eval_acc) { a.eval_acc[0] += (double)s; a.eval_acc[1] += (double)correct; }
        *a.ticket = 0;   // re-arm for the next launch
    }
    if (a.dbg != nullptr && blockIdx.x == 0 && tid == 0) a.dbg[15] = clock64();
}

}  // namespace b200dp

using namespace b200dp;

typedef CUresult (*HfEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static HfEncodeFn hf_encoder() {
    static HfEncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            fn = (HfEncodeFn)p;
    }
    return fn;
}
// [rows][cols] fp32, dense rows; box = {32 floats (128 bytes: the SWIZZLE_128B span), box_rows}; zeros out of bounds
static bool hf_encode(CUtensorMap* tm, const float* ptr, int rows, int cols, int box_rows) {
    HfEncodeFn enc = hf_encoder();
    if (enc == nullptr || ((uintptr_t)ptr & 15) != 0 || cols % 4 != 0) return false;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
    cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static long long* g_hf_dbg = nullptr;
static int g_hf_tma = 1;
extern "C" void b200dp_head_fused_set_tma(int on) { g_hf_tma = on; }
extern "C" void b200dp_head_fused_set_debug(long long* dbg) { g_hf_dbg = dbg; }

extern "C" int b200dp_head_fused_supported(int K3, int C3, int F1, int ncls) {
    return K3 >= 1 && K3 <= 4096 && C3 >= 1 && C3 <= HF_NMAX && F1 >= 1 && F1 <= HF_NMAX && ncls >= 1 && ncls <= HF_MAXC;
}

extern "C" int b200dp_head_fused(const float* p2, int ld_p2, const float* w3, const float* b3, const float* w1,
                                 const float* b1, const float* w2, const float* b2, const int64_t* target,
                                 int loss_mode, float* a3, int ld_a3, float* h1, int ld_h1, float* probs,
                                 float* loss_rows, float* loss_out, float* dlogits, int ld_dlogits, float* dz1,
                                 float* dz3, double* eval_acc, int* ticket, int B, int K3, int C3, int F1, int ncls,
                                 b200dp_stream_t s) {
    B200DP_REQUIRE(b200dp_head_fused_supported(K3, C3, F1, ncls), "fused head covers C3, F1 <= 128, ncls <= 16, K3 <= 4096");
    B200DP_REQUIRE(p2 && w3 && b3 && w1 && b1 && w2 && b2 && target && a3 && h1 && loss_rows && loss_out && ticket,
                   "null argument");
    B200DP_REQUIRE(probs != nullptr || eval_acc == nullptr, "eval metrics need the probs buffer");
    B200DP_REQUIRE((dlogits == nullptr) == (dz1 == nullptr) && (dz1 == nullptr) == (dz3 == nullptr),
                   "dlogits, dz1, dz3: all three (training) or none (evaluation)");
    HeadFusedArgs a{p2, ld_p2 > 0 ? ld_p2 : K3, w3, b3, w1, b1, w2, b2, target, a3, ld_a3 > 0 ? ld_a3 : C3, h1,
                    ld_h1 > 0 ? ld_h1 : F1, probs, loss_rows, loss_out, dlogits, ld_dlogits > 0 ? ld_dlogits : ncls, dz1,
                    dz3, eval_acc, ticket, B, K3, C3, F1, ncls, loss_mode, g_hf_dbg};
    const int K3P = (K3 + HF_BK - 1) / HF_BK * HF_BK;
    const size_t smem = (size_t)(K3P * HF_R + HF_STAGES * HF_STAGE + HF_KG * HF_R * HF_NMAX + 3 * HF_NMAX * HF_R + ncls * F1 + ncls) * sizeof(float);
    const int grid = (B + HF_R - 1) / HF_R;
    // weights by TMA tensor map when the rows are 16-byte multiples (the reference geometry: 400 and 120 floats); the
    // cp.async ring serves every other shape
    const bool tma = g_hf_tma && hf_encode(&a.tm3, w3, C3, K3, HF_NMAX) && hf_encode(&a.tm1k, w1, F1, C3, HF_NMAX) &&
                     hf_encode(&a.tm1n, w1, F1, C3, HF_BK);
    // the attribute is set once per device, so to the kernel's ceiling, not to this call's need (K3 varies between calls)
    constexpr int kSmemCeiling = 224 * 1024;    // 227 KB per CTA minus room for the static part
    B200DP_REQUIRE(smem <= (size_t)kSmemCeiling, "K3 too large for the fused head's shared-memory budget");
    if (tma) {
        static DeviceOnce attr;
        B200DP_CHECK(attr.max_dynamic_smem(head_fused_kernel<true>, kSmemCeiling));
        B200DP_CHECK(launch_pdl(head_fused_kernel<true>, dim3(grid), dim3(HF_NT), smem, (cudaStream_t)s, a));
    } else {
        static DeviceOnce attr;
        B200DP_CHECK(attr.max_dynamic_smem(head_fused_kernel<false>, kSmemCeiling));
        B200DP_CHECK(launch_pdl(head_fused_kernel<false>, dim3(grid), dim3(HF_NT), smem, (cudaStream_t)s, a));
    }
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
