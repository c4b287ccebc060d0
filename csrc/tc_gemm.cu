// tcgen05 / TMEM / TMA GEMM for the dense layers, directly on the fp32 tensors (kind::tf32).
//
//   C[M][N] = epilogue( sum_k A(m,k) * B(n,k) ),  fp32 accumulation in tensor memory
//
// Each operand is described by how it sits in memory, so no transposed or down-converted copies
// are ever made:
//   K-major  : memory is [MN rows][K cols]   (x for y = x W^T; W for y = x W^T; dY for dX = dY W)
//   MN-major : memory is [K rows][MN cols]   (W for dX = dY W; dY and X for dW = dY^T X)
// Replaces reference cnn.py:16-17, 20-21 (conv3-as-GEMM, FC.0) and their autograd backward
// (train.py:49), up to four problems per launch.
//
// Structure (one CTA = one 128x128 output tile; 192 threads):
//   warp 0     TMA producer: cp.async.bulk.tensor.2d (SWIZZLE_128B, zero fill out of bounds) into a
//              6-stage ring of {A 16 KB, B 16 KB} slabs, 32 k per slab, mbarrier complete_tx
//   warp 1     allocates 128 TMEM columns, then one elected thread issues 4 tcgen05.mma
//              (M128 N128 K8, tf32) per slab and tcgen05.commit's the slab's "empty" barrier
//   warps 2-5  epilogue: tcgen05.ld 32x32b.x16 (thread = output row), bias / LeakyReLU /
//              LeakyReLU' mask, fp32 stores; the last logical column can be routed to a separate
//              vector (bias gradient through a ones column in the activation buffers)
#include <cuda.h>
#include <string.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace b200dp {

constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 32;
constexpr int TC_STAGES = 6;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4;   // 16 KB
constexpr int TC_B_BYTES = TC_BN * TC_BK * 4;   // 16 KB
constexpr int TC_STAGE_BYTES = TC_A_BYTES + TC_B_BYTES;
constexpr int TC_SMEM = TC_STAGES * TC_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TC_THREADS = 192;
constexpr int TC_TMEM_COLS = 128;
constexpr int TC_MAX_PROBLEMS = 4;

struct TcProblem {
    CUtensorMap tmA, tmB;
    const float* bias;       // [N], added before act
    const float* mask; int ld_mask;   // C *= leaky'(mask[m][n])
    float* out; int ld_out;
    float* colext;           // if set: logical column N-1 goes to colext[m] (not to out)
    int M, N, K;
    int a_mn, b_mn, act;
    int tiles_n, tile_start;
};
struct TcGroup {
    TcProblem p[TC_MAX_PROBLEMS];
    int count;
    long long* dbg;   // optional: CTA 0 writes clock64() stamps of its pipeline phases (profiling aid)
};
#define TC_STAMP(i) do { if (g.dbg && blockIdx.x == 0) g.dbg[i] = clock64(); } while (0)

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor), version 1.
//   K-major : SWIZZLE_128B (16 B chunks); 8-row groups 1024 B apart (SBO); LBO unused (=1)
//   MN-major: 32-bit operands must use SWIZZLE_128B_BASE32B (32 B chunks, 4-k-row atoms; the TMA map
//             uses CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B): 32-float MN blocks TC_BK*128 B apart (LBO),
//             4-k-row groups 512 B apart (SBO)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, int mn_major) {
    const uint64_t lbo = mn_major ? (uint64_t)((TC_BK * 128) >> 4) : 1ull;
    const uint64_t sbo = mn_major ? (512 >> 4) : (1024 >> 4);
    const uint64_t layout = mn_major ? 1ull : 2ull;
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

// Per-problem scalars pulled into registers once.  Indexing the kernel parameter with a run-time
// problem index makes every later field access an indexed constant load (LDC c[0][R+off], ~100
// cycles each, re-issued inside the loops); a switch over compile-time indices uses immediate
// constant addresses and keeps the values in registers.
struct TcScalars {
    const CUtensorMap* tmA; const CUtensorMap* tmB;
    const float* bias; const float* mask; float* out; float* colext;
    int ld_mask, ld_out, M, N, K, a_mn, b_mn, act, tiles_n, tile_start;
};
__device__ __forceinline__ TcScalars tc_load_scalars(const TcGroup& g, int pi) {
    TcScalars r;
#define TC_COPY(i)                                                                                          \
    case i: r.tmA = &g.p[i].tmA; r.tmB = &g.p[i].tmB; r.bias = g.p[i].bias; r.mask = g.p[i].mask;          \
            r.out = g.p[i].out; r.colext = g.p[i].colext; r.ld_mask = g.p[i].ld_mask; r.ld_out = g.p[i].ld_out; \
            r.M = g.p[i].M; r.N = g.p[i].N; r.K = g.p[i].K; r.a_mn = g.p[i].a_mn; r.b_mn = g.p[i].b_mn;      \
            r.act = g.p[i].act; r.tiles_n = g.p[i].tiles_n; r.tile_start = g.p[i].tile_start; break;
    switch (pi) {
        TC_COPY(1) TC_COPY(2) TC_COPY(3)
        default: TC_COPY(0)
    }
#undef TC_COPY
    return r;
}

__global__ void __launch_bounds__(TC_THREADS, 1) tc_gemm_kernel(const __grid_constant__ TcGroup g) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;          // SWIZZLE_128B needs 1024 B alignment
    const uint32_t bar_base = base + TC_STAGES * TC_STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (TC_STAGES + s); };
    const uint32_t tmem_full_bar = bar_base + 8u * (2 * TC_STAGES);
    const uint32_t tmem_slot = bar_base + 8u * (2 * TC_STAGES + 1);
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) TC_STAMP(0);
    int pi = 0;
    if (g.count > 1 && (int)blockIdx.x >= g.p[1].tile_start) pi = 1;
    if (g.count > 2 && (int)blockIdx.x >= g.p[2].tile_start) pi = 2;
    if (g.count > 3 && (int)blockIdx.x >= g.p[3].tile_start) pi = 3;
    const TcScalars P = tc_load_scalars(g, pi);
    const int tile = blockIdx.x - P.tile_start;
    const int tm = tile / P.tiles_n, tn = tile - tm * P.tiles_n;
    const int m0 = tm * TC_BM, n0 = tn * TC_BN;
    const int nslab = (P.K + TC_BK - 1) / TC_BK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(tmem_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(tmem_slot, TC_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    if (threadIdx.x == 0) TC_STAMP(1);

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            for (int s = 0; s < nslab; ++s) {
                const int st = s % TC_STAGES;
                if (s >= TC_STAGES) mbar_wait(empty_bar(st), ((s / TC_STAGES) - 1) & 1);
                const uint32_t sa = base + st * TC_STAGE_BYTES, sb = sa + TC_A_BYTES;
                mbar_expect_tx(full_bar(st), TC_STAGE_BYTES);
                const int k0 = s * TC_BK;
                if (P.a_mn) {
#pragma unroll
                    for (int i = 0; i < TC_BM / 32; ++i) tma_load_2d(sa + i * (TC_BK * 128), P.tmA, full_bar(st), m0 + 32 * i, k0);
                } else {
                    tma_load_2d(sa, P.tmA, full_bar(st), k0, m0);
                }
                if (P.b_mn) {
#pragma unroll
                    for (int i = 0; i < TC_BN / 32; ++i) tma_load_2d(sb + i * (TC_BK * 128), P.tmB, full_bar(st), n0 + 32 * i, k0);
                } else {
                    tma_load_2d(sb, P.tmB, full_bar(st), k0, n0);
                }
                if (s < 8) TC_STAMP(8 + s);        // producer issued slab s
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        // instruction descriptor: D=f32, A=B=tf32, majorness per operand, N=128, M=128
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)P.a_mn << 15) | ((uint32_t)P.b_mn << 16) |
                               ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
        const uint32_t a_step = P.a_mn ? 1024u : 32u;     // bytes per 8-k step inside a slab
        const uint32_t b_step = P.b_mn ? 1024u : 32u;
        for (int s = 0; s < nslab; ++s) {
            const int st = s % TC_STAGES;
            mbar_wait(full_bar(st), (s / TC_STAGES) & 1);
            tc_fence_after();
            if (lane == 0 && s < 8) TC_STAMP(16 + s);   // slab s landed
            if (lane == 0) {
                const uint32_t sa = base + st * TC_STAGE_BYTES, sb = sa + TC_A_BYTES;
#pragma unroll
                for (int kk = 0; kk < TC_BK / 8; ++kk) {
                    umma_tf32(tmem_base, make_desc(sa + kk * a_step, P.a_mn), make_desc(sb + kk * b_step, P.b_mn),
                              idesc, (s | kk) != 0);
                }
                umma_commit(empty_bar(st));                    // slab free once these MMAs retire
                if (s == nslab - 1) umma_commit(tmem_full_bar);  // accumulator complete
            }
            __syncwarp();
        }
    } else {
        // ================= epilogue (warps 2..5) =================
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        if (threadIdx.x == 64) TC_STAMP(2);          // accumulator complete
        const int q = warp & 3;                           // TMEM lane quarter this warp may touch
        const int n_out = P.colext ? P.N - 1 : P.N;
        // Stage the warp's 32x128 fp32 sub-tile through shared memory (the operand ring is idle now)
        // so that global stores are row-contiguous: lane = column.  Row pitch 129 floats: both the
        // row-wise writes (lane = row) and the column-wise reads (lane = column) are conflict-free.
        constexpr int TP = TC_BN + 1;
        float* sT = reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw))) + q * 32 * TP;
        const int ncols = min(TC_BN, P.N - n0);           // warp-uniform
#pragma unroll 1
        for (int c = 0; c * 16 < ncols; ++c) {
            uint32_t r[16];
            tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 16), r);
#pragma unroll
            for (int j = 0; j < 16; ++j) sT[lane * TP + c * 16 + j] = __uint_as_float(r[j]);
        }
        __syncwarp();
        float bias_r[TC_BN / 32];
#pragma unroll
        for (int cc = 0; cc < TC_BN / 32; ++cc) {
            const int n = n0 + cc * 32 + lane;
            bias_r[cc] = (P.bias && n < n_out) ? P.bias[n] : 0.f;
        }
        const int rows = min(32, P.M - (m0 + q * 32));    // warp-uniform, may be <= 0
#pragma unroll 1
        for (int rr = 0; rr < rows; ++rr) {
            const int m = m0 + q * 32 + rr;
#pragma unroll
            for (int cc = 0; cc < TC_BN / 32; ++cc) {
                const int n = n0 + cc * 32 + lane;
                if (cc * 32 < ncols && n < P.N) {
                    float v = sT[rr * TP + cc * 32 + lane];
                    if (n < n_out) {
                        v += bias_r[cc];
                        if (P.act == 1) v = leaky(v);
                        if (P.mask) v *= leaky_grad_from_act(P.mask[(long)m * P.ld_mask + n]);
                        P.out[(long)m * P.ld_out + n] = v;
                    } else {
                        P.colext[m] = v;
                    }
                }
            }
        }
    }
    if (threadIdx.x == 64) TC_STAMP(3);              // epilogue of warp 2 done
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TC_TMEM_COLS);
    }
    if (threadIdx.x == 0) TC_STAMP(4);
}

// ---- host side ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// memory matrix [rows][cols] fp32 with leading dimension ld; box = {box_cols (inner), box_rows}
static int encode_map(CUtensorMap* tm, const float* ptr, int rows, int cols, int ld, int box_cols, int box_rows,
                      CUtensorMapSwizzle swz) {
    EncodeTiledFn enc = get_encode();
    B200DP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled not available from the driver");
    B200DP_REQUIRE(((uintptr_t)ptr & 15) == 0 && (ld % 4) == 0, "TMA needs 16-byte aligned base and row pitch");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_last_error("cuTensorMapEncodeTiled failed with %d (rows %d cols %d ld %d)", (int)r, rows, cols, ld);
        return (int)cudaErrorInvalidValue;
    }
    return 0;
}

}  // namespace b200dp

using namespace b200dp;

static int tc_launch(const b200dp_tc_problem* probs, int count, long long* dbg, b200dp_stream_t s);

extern "C" int b200dp_tc_gemm_group(const b200dp_tc_problem* probs, int count, b200dp_stream_t s) {
    return tc_launch(probs, count, nullptr, s);
}
// Profiling aid: same launch, CTA 0 writes clock64() stamps into dbg[0..32) (device memory):
// [0] start, [1] setup done, [2] accumulator complete, [3] epilogue done, [4] end,
// [8+s] producer issued slab s, [16+s] slab s landed (s < 8).
extern "C" int b200dp_tc_gemm_group_timed(const b200dp_tc_problem* probs, int count, long long* dbg, b200dp_stream_t s) {
    return tc_launch(probs, count, dbg, s);
}

static int tc_launch(const b200dp_tc_problem* probs, int count, long long* dbg, b200dp_stream_t s) {
    B200DP_REQUIRE(count >= 1 && count <= TC_MAX_PROBLEMS, "1..4 problems per launch");
    TcGroup g;
    memset(&g, 0, sizeof(g));
    g.count = count;
    g.dbg = dbg;
    int total = 0;
    for (int i = 0; i < count; ++i) {
        const b200dp_tc_problem& q = probs[i];
        TcProblem& p = g.p[i];
        B200DP_REQUIRE(q.M > 0 && q.N > 0 && q.K > 0, "empty problem");
        int rc;
        // K-major: memory [MN][K] -> box {32 k, 128 mn};  MN-major: memory [K][MN] -> box {32 mn, 32 k}
        rc = q.a_mn_major ? encode_map(&p.tmA, q.a, q.K, q.M, q.lda, 32, TC_BK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
                          : encode_map(&p.tmA, q.a, q.M, q.K, q.lda, TC_BK, TC_BM, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        rc = q.b_mn_major ? encode_map(&p.tmB, q.b, q.K, q.N, q.ldb, 32, TC_BK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
                          : encode_map(&p.tmB, q.b, q.N, q.K, q.ldb, TC_BK, TC_BN, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        p.bias = q.bias; p.mask = q.mask; p.ld_mask = q.ld_mask; p.out = q.out; p.ld_out = q.ld_out;
        p.colext = q.colext; p.M = q.M; p.N = q.N; p.K = q.K;
        p.a_mn = q.a_mn_major ? 1 : 0; p.b_mn = q.b_mn_major ? 1 : 0; p.act = q.act;
        p.tiles_n = (q.N + TC_BN - 1) / TC_BN;
        p.tile_start = total;
        total += ((q.M + TC_BM - 1) / TC_BM) * p.tiles_n;
    }
    static bool attr_done = false;
    if (!attr_done) {
        B200DP_CHECK(cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));
        attr_done = true;
    }
    tc_gemm_kernel<<<total, TC_THREADS, TC_SMEM, (cudaStream_t)s>>>(g);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
