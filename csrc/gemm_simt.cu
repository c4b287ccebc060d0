// Small dense layers on CUDA cores: forward y = act(x W^T + b) and the two backward products
// (data gradient with the LeakyReLU derivative folded in, weight gradient with the bias gradient
// folded in), up to four independent problems per launch.
//
// Replaces reference cnn.py:16-17 (conv3 on a 5x5 map == GEMM), cnn.py:20-21 (FC.0) and their
// autograd backward (train.py:49).  The batch dimension here is 256, so these GEMMs are a few
// MFLOP each: the design goal is few launches and enough CTAs, not MMA rate.  A CTA owns a 16x16
// output tile (with 32x32 tiles the K=400 layer ran on 32 of the 148 SMs, each CTA grinding through ~3K
// dependent instructions per thread; 16x16 tiles give 128 CTAs with a third of the work per thread); its 256
// threads are four k-groups of 64 (2x2 micro-tile each) that split every 32-deep k-slab and are summed in fixed
// order at the end (deterministic, no atomics).
// The tensor-core path for the wide model is tc_gemm.cu.
#include "common.cuh"

namespace b200dp {

constexpr int BM = 16, BN = 16, BK = 32, KG = 4, GEMM_NT = 256;
constexpr int TM = BM / 8, TN = BN / 8;            // micro-tile of a thread (8x8 threads per k-group)
constexpr int LT = BM * BK / GEMM_NT;               // operand-tile elements per thread
constexpr int OT = (BM * BN + GEMM_NT - 1) / GEMM_NT;   // output elements per thread in the epilogue
static_assert(BM == BN && BM * BK % GEMM_NT == 0, "tile shape");

struct GemmProblem {
    const float* A; const float* B; float* C;
    const float* bias;      // [N] added before act, or null
    const float* mask_src;  // [M][ldc]: C *= leaky'(mask_src), or null
    float* colsum;          // [M]: sum_k A(m,k), or null
    int M, N, K;
    long sAm, sAk, sBn, sBk;
    int ldc, act;
    int tiles_n, tile_start;
};
struct GemmGroup {
    GemmProblem p[4];
    int count;
};

__device__ __forceinline__ void load_tile(const float* __restrict__ P, long s_mn, long s_k, int mn0, int k0,
                                          int MN, int K, float (&r)[LT], int tid) {
    // BM x BK tile, LT elements per thread.  Pick the thread->element map that makes the unit-stride
    // dimension the fastest-varying one across a warp (coalesced either way).
    if (s_k == 1) {
#pragma unroll
        for (int j = 0; j < LT; ++j) {
            int e = tid + j * GEMM_NT, mn = e / BK, k = e % BK;
            r[j] = (mn0 + mn < MN && k0 + k < K) ? P[(long)(mn0 + mn) * s_mn + (k0 + k)] : 0.f;
        }
    } else {
#pragma unroll
        for (int j = 0; j < LT; ++j) {
            int e = tid + j * GEMM_NT, k = e / BM, mn = e % BM;
            r[j] = (mn0 + mn < MN && k0 + k < K) ? P[(long)(mn0 + mn) * s_mn + (long)(k0 + k) * s_k] : 0.f;
        }
    }
}
__device__ __forceinline__ void store_tile(float (*S)[BM + 1], long s_k, const float (&r)[LT], int tid) {
    if (s_k == 1) {
#pragma unroll
        for (int j = 0; j < LT; ++j) { int e = tid + j * GEMM_NT; S[e % BK][e / BK] = r[j]; }
    } else {
#pragma unroll
        for (int j = 0; j < LT; ++j) { int e = tid + j * GEMM_NT; S[e / BM][e % BM] = r[j]; }
    }
}

__global__ void __launch_bounds__(GEMM_NT) gemm_group_kernel(GemmGroup g) {
    pdl_wait();
    __shared__ float As[BK][BM + 1];
    __shared__ float Bs[BK][BN + 1];
    __shared__ float Red[KG][BM][BN + 1];
    __shared__ float RedCs[KG][BM];

    // Select the problem with compile-time indices: a run-time index into the kernel parameter turns
    // every field access into a slow indexed constant load inside the loops.
    int pi = 0;
    if (g.count > 1 && (int)blockIdx.x >= g.p[1].tile_start) pi = 1;
    if (g.count > 2 && (int)blockIdx.x >= g.p[2].tile_start) pi = 2;
    if (g.count > 3 && (int)blockIdx.x >= g.p[3].tile_start) pi = 3;
    GemmProblem p;
    switch (pi) {
        case 1: p = g.p[1]; break;
        case 2: p = g.p[2]; break;
        case 3: p = g.p[3]; break;
        default: p = g.p[0]; break;
    }
    const int tile = blockIdx.x - p.tile_start;
    const int tm = tile / p.tiles_n, tn = tile - tm * p.tiles_n;
    const int m0 = tm * BM, n0 = tn * BN;

    const int tid = threadIdx.x;
    const int kg = tid >> 6;           // k-group 0..3
    const int t64 = tid & 63;
    const int tx = t64 & 7, ty = t64 >> 3;   // 8x8 threads, TM x TN outputs each
    const bool want_cs = (p.colsum != nullptr) && (tn == 0);

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    float cs[TM];
#pragma unroll
    for (int i = 0; i < TM; ++i) cs[i] = 0.f;

    float ra[LT], rb[LT];
    const int nslab = (p.K + BK - 1) / BK;
    load_tile(p.A, p.sAm, p.sAk, m0, 0, p.M, p.K, ra, tid);
    load_tile(p.B, p.sBn, p.sBk, n0, 0, p.N, p.K, rb, tid);
    for (int t = 0; t < nslab; ++t) {
        store_tile(As, p.sAk, ra, tid);
        store_tile(Bs, p.sBk, rb, tid);
        __syncthreads();
        if (t + 1 < nslab) {
            load_tile(p.A, p.sAm, p.sAk, m0, (t + 1) * BK, p.M, p.K, ra, tid);
            load_tile(p.B, p.sBn, p.sBk, n0, (t + 1) * BK, p.N, p.K, rb, tid);
        }
#pragma unroll
        for (int kk = 0; kk < BK / KG; ++kk) {
            const int k = kg * (BK / KG) + kk;
            float av[TM], bv[TN];
#pragma unroll
            for (int i = 0; i < TM; ++i) av[i] = As[k][ty * TM + i];
#pragma unroll
            for (int j = 0; j < TN; ++j) bv[j] = Bs[k][tx * TN + j];
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            if (want_cs && tx == 0) {
#pragma unroll
                for (int i = 0; i < TM; ++i) cs[i] += av[i];
            }
        }
        __syncthreads();
    }

    pdl_trigger();   // late: see common.cuh
    // ---- fixed-order reduction over the four k-groups + epilogue ----
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) Red[kg][ty * TM + i][tx * TN + j] = acc[i][j];
    if (want_cs && tx == 0) {
#pragma unroll
        for (int i = 0; i < TM; ++i) RedCs[kg][ty * TM + i] = cs[i];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < OT; ++j) {
        const int e = tid + j * GEMM_NT, m = e / BN, n = e % BN;
        if (e < BM * BN && m0 + m < p.M && n0 + n < p.N) {
            float v = ((Red[0][m][n] + Red[1][m][n]) + Red[2][m][n]) + Red[3][m][n];
            if (p.bias) v += p.bias[n0 + n];
            if (p.act == 1) v = leaky(v);
            const long o = (long)(m0 + m) * p.ldc + (n0 + n);
            if (p.mask_src) v *= leaky_grad_from_act(p.mask_src[o]);
            p.C[o] = v;
        }
    }
    if (want_cs && tid < BM && m0 + tid < p.M)
        p.colsum[m0 + tid] = ((RedCs[0][tid] + RedCs[1][tid]) + RedCs[2][tid]) + RedCs[3][tid];
}

static int launch_group(GemmGroup& g, cudaStream_t s) {
    int total = 0;
    for (int i = 0; i < g.count; ++i) {
        GemmProblem& p = g.p[i];
        p.tiles_n = (p.N + BN - 1) / BN;
        p.tile_start = total;
        total += ((p.M + BM - 1) / BM) * p.tiles_n;
    }
    if (total == 0) return 0;
    B200DP_CHECK(launch_pdl(gemm_group_kernel, dim3(total), dim3(GEMM_NT), 0, s, g));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_linear_fwd(const float* x, const float* w, const float* b, float* y, int M, int N, int K,
                                 int act, b200dp_stream_t s) {
    GemmGroup g{};
    g.count = 1;
    g.p[0] = GemmProblem{x, w, y, b, nullptr, nullptr, M, N, K, (long)K, 1, (long)K, 1, N, act, 0, 0};
    return launch_group(g, (cudaStream_t)s);
}

extern "C" int b200dp_linear_bwd_group(const b200dp_bwd_problem* probs, int count, b200dp_stream_t s) {
    B200DP_REQUIRE(count >= 1 && count <= 4, "1..4 problems per launch");
    GemmGroup g{};
    g.count = count;
    for (int i = 0; i < count; ++i) {
        const b200dp_bwd_problem& q = probs[i];
        if (q.kind == 0) {
            // dX[M][K] = dY[M][N] . W[N][K]  -> C[m][k] = sum_n A(m,n) B(k,n), B(k,n) = W[n*K + k]
            g.p[i] = GemmProblem{q.dy, q.w_or_x, q.out, nullptr, q.act_prev, nullptr,
                                 q.M, q.K, q.N, (long)q.N, 1, 1, (long)q.K, q.K, 0, 0, 0};
        } else if (q.kind == 1) {
            // dW[N][K] = sum_m dY[m][n] X[m][k] -> C[n][k] = sum_m A(n,m) B(k,m); colsum_m A(n,m) = db[n]
            g.p[i] = GemmProblem{q.dy, q.w_or_x, q.out, nullptr, nullptr, q.db,
                                 q.N, q.K, q.M, 1, (long)q.N, 1, (long)q.K, q.K, 0, 0, 0};
        } else {
            B200DP_REQUIRE(false, "kind must be 0 (data grad) or 1 (weight grad)");
        }
    }
    return launch_group(g, (cudaStream_t)s);
}
