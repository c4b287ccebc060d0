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
// Structure (one CTA = one 128 x bn output tile, or one K-split of it; 320 threads):
//   warp 0     TMA producer: cp.async.bulk.tensor.2d (SWIZZLE_128B, zero fill out of bounds) into a
//              6-stage ring of {A 16 KB, B 16 KB} slabs, 32 k per slab, mbarrier complete_tx
//   warp 1     allocates 128 TMEM columns, then one elected thread issues 4 tcgen05.mma
//              (M128 N128 K8, tf32) per slab and tcgen05.commit's the slab's "empty" barrier
//   warps 2-9  epilogue: tcgen05.ld 32x32b.x16 (thread = output row, 64 contiguous bytes per load), bias /
//              LeakyReLU / LeakyReLU' mask, stores straight from registers; the last logical column can be
//              routed to a separate vector (bias gradient through a ones column in the activation buffers)
// With a workspace the launch is re-tiled (32/64/128-wide tiles) and split along K so that about one CTA per
// SM works: a CTA is bound by the rate at which TMA delivers 128-byte rows (~43 B/clk, measured 765 cycles per
// 32 KB slab), so a 50-slab chain on 8 CTAs costs 20 us while 120 CTAs with 12 slabs each cost ~3 us; the CTAs
// of a tile park raw partials and the last one (ticket) sums them in split order: deterministic.
#include <cuda.h>
#include <string.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace b200dp {

constexpr int TC_BM = 128, TC_BN_MAX = 128, TC_BK = 32;
constexpr int TC_STAGES = 6;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4;       // 16 KB
constexpr int TC_B_BYTES = TC_BN_MAX * TC_BK * 4;   // up to 16 KB (bn * 128 B are used)
constexpr int TC_STAGE_BYTES = TC_A_BYTES + TC_B_BYTES;
constexpr int TC_SMEM = TC_STAGES * TC_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TC_EPI_WARPS = 8;
constexpr int TC_THREADS = 32 * (2 + TC_EPI_WARPS);
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
    int bn;                  // tile width (32 / 64 / 128)
    int splits;              // split-K factor: CTAs per output tile
    int tiles_n, cta_start;  // first CTA of this problem
    float* ws;               // split-K partial tiles [tile][split][128][bn]
    unsigned* tickets;       // one arrival counter per tile (zero between launches)
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
    const float* bias; const float* mask; float* out; float* colext; float* ws; unsigned* tickets;
    int ld_mask, ld_out, M, N, K, a_mn, b_mn, act, bn, splits, tiles_n, cta_start;
};
__device__ __forceinline__ TcScalars tc_load_scalars(const TcGroup& g, int pi) {
    TcScalars r;
#define TC_COPY(i)                                                                                          \
    case i: r.tmA = &g.p[i].tmA; r.tmB = &g.p[i].tmB; r.bias = g.p[i].bias; r.mask = g.p[i].mask;          \
            r.out = g.p[i].out; r.colext = g.p[i].colext; r.ld_mask = g.p[i].ld_mask; r.ld_out = g.p[i].ld_out; \
            r.M = g.p[i].M; r.N = g.p[i].N; r.K = g.p[i].K; r.a_mn = g.p[i].a_mn; r.b_mn = g.p[i].b_mn;      \
            r.act = g.p[i].act; r.bn = g.p[i].bn; r.splits = g.p[i].splits; r.tiles_n = g.p[i].tiles_n;     \
            r.cta_start = g.p[i].cta_start; r.ws = g.p[i].ws; r.tickets = g.p[i].tickets; break;
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
    int* s_last = reinterpret_cast<int*>(tmem_slot_ptr + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) TC_STAMP(0);
    int pi = 0;
    if (g.count > 1 && (int)blockIdx.x >= g.p[1].cta_start) pi = 1;
    if (g.count > 2 && (int)blockIdx.x >= g.p[2].cta_start) pi = 2;
    if (g.count > 3 && (int)blockIdx.x >= g.p[3].cta_start) pi = 3;
    const TcScalars P = tc_load_scalars(g, pi);
    const int local = blockIdx.x - P.cta_start;
    const int tile = local / P.splits, split = local - tile * P.splits;
    const int tm = tile / P.tiles_n, tn = tile - tm * P.tiles_n;
    const int m0 = tm * TC_BM, n0 = tn * P.bn;
    // this CTA's K range: slabs [slab0, slab0 + nslab)
    const int nslab_all = (P.K + TC_BK - 1) / TC_BK;
    const int per = nslab_all / P.splits, rem = nslab_all - per * P.splits;
    const int nslab = per + (split < rem ? 1 : 0);
    const int slab0 = split * per + min(split, rem);

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
    pdl_wait();      // set-up above overlaps the previous kernel's tail (common.cuh)

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            const uint32_t stage_tx = TC_A_BYTES + (uint32_t)P.bn * 128u;
            for (int s = 0; s < nslab; ++s) {
                const int st = s % TC_STAGES;
                if (s >= TC_STAGES) mbar_wait(empty_bar(st), ((s / TC_STAGES) - 1) & 1);
                const uint32_t sa = base + st * TC_STAGE_BYTES, sb = sa + TC_A_BYTES;
                mbar_expect_tx(full_bar(st), stage_tx);
                const int k0 = (slab0 + s) * TC_BK;
                if (P.a_mn) {
#pragma unroll
                    for (int i = 0; i < TC_BM / 32; ++i) tma_load_2d(sa + i * (TC_BK * 128), P.tmA, full_bar(st), m0 + 32 * i, k0);
                } else {
                    tma_load_2d(sa, P.tmA, full_bar(st), k0, m0);
                }
                if (P.b_mn) {
                    for (int i = 0; i < P.bn / 32; ++i) tma_load_2d(sb + i * (TC_BK * 128), P.tmB, full_bar(st), n0 + 32 * i, k0);
                } else {
                    tma_load_2d(sb, P.tmB, full_bar(st), k0, n0);
                }
                if (s < 8) TC_STAMP(8 + s);        // producer issued slab s
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        // instruction descriptor: D=f32, A=B=tf32, majorness per operand, N=bn, M=128
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)P.a_mn << 15) | ((uint32_t)P.b_mn << 16) |
                               ((uint32_t)(P.bn >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
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
        // ================= epilogue (warps 2..9): warp reads TMEM lane quarter warp%4, column half (warp-2)/4.
        // tcgen05.ld gives a thread 16 columns of ITS row; storing those directly costs one L1 wavefront per
        // element (32 lanes -> 32 different lines; measured: 16 K cycles for a 128x128 tile).  So the warp's 32 x half
        // sub-tile is transposed through shared memory (the operand ring is idle now) and handled with
        // lane = (row group, 4 consecutive columns): 16-byte shared loads, 16-byte coalesced global accesses, and
        // only half/4 iterations per warp (a lone warp per scheduler pays ~5 cycles per dependent instruction,
        // so the instruction count is what matters). =====
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        if (threadIdx.x == 64) TC_STAMP(2);          // accumulator complete
        pdl_trigger();
        const int q = warp & 3;
        const int h = (warp - 2) >> 2;
        const int half = P.bn >> 1, c_lo = h * half;          // this warp's columns [c_lo, c_lo + half) of the tile
        const int n_out = P.colext ? P.N - 1 : P.N;
        constexpr int TP = TC_BN_MAX / 2 + 4;                  // 68 floats: 16-byte aligned rows, conflict-free 128-bit accesses
        float* sT = reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw))) + (warp - 2) * 32 * TP;
        const int ncols = max(0, min(half, P.N - n0 - c_lo));  // warp-uniform
#pragma unroll 1
        for (int c = 0; c * 16 < ncols; ++c) {
            uint32_t r[16];
            tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c_lo + c * 16), r);
#pragma unroll
            for (int j = 0; j < 16; j += 4)
                *reinterpret_cast<uint4*>(sT + lane * TP + c * 16 + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
        }
        __syncwarp();
        if (threadIdx.x == 64) TC_STAMP(5);
        const int rows = max(0, min(32, P.M - (m0 + q * 32)));    // warp-uniform
        const int vpr = half >> 2;                                 // float4 per row of the sub-tile: 4 / 8 / 16
        const int rpi = 32 / vpr;                                  // rows covered by one warp-wide access: 8 / 4 / 2
        const int c4 = (lane % vpr) * 4, r_in = lane / vpr;        // this lane's columns / row inside an iteration
        const int mrow0 = m0 + q * 32, ncol = n0 + c_lo + c4;
        const bool col_ok = c4 < ncols;
        const bool vec_ok = col_ok && c4 + 3 < ncols && ncol + 3 < n_out && ((P.ld_out & 3) == 0) &&
                            ((reinterpret_cast<uintptr_t>(P.out) & 15) == 0) &&
                            (!P.mask || (((P.ld_mask & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.mask) & 15) == 0)));
        float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (P.bias && col_ok) {
            if (ncol + 0 < n_out) bias4.x = __ldg(P.bias + ncol + 0);
            if (ncol + 1 < n_out) bias4.y = __ldg(P.bias + ncol + 1);
            if (ncol + 2 < n_out) bias4.z = __ldg(P.bias + ncol + 2);
            if (ncol + 3 < n_out) bias4.w = __ldg(P.bias + ncol + 3);
        }
        auto finish4 = [&](int rr, float4 v) {            // fused epilogue of 4 consecutive columns of row rr
            const int m = mrow0 + rr;
            if (vec_ok) {
                v.x += bias4.x; v.y += bias4.y; v.z += bias4.z; v.w += bias4.w;
                if (P.act == 1) { v.x = leaky(v.x); v.y = leaky(v.y); v.z = leaky(v.z); v.w = leaky(v.w); }
                if (P.mask) {
                    const float4 k = *reinterpret_cast<const float4*>(P.mask + (long)m * P.ld_mask + ncol);
                    v.x *= leaky_grad_from_act(k.x); v.y *= leaky_grad_from_act(k.y);
                    v.z *= leaky_grad_from_act(k.z); v.w *= leaky_grad_from_act(k.w);
                }
                *reinterpret_cast<float4*>(P.out + (long)m * P.ld_out + ncol) = v;
            } else {
                const float e[4] = {v.x, v.y, v.z, v.w}, bb[4] = {bias4.x, bias4.y, bias4.z, bias4.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int n = ncol + j;
                    if (c4 + j < ncols && n < P.N) {
                        float x = e[j];
                        if (n < n_out) {
                            x += bb[j];
                            if (P.act == 1) x = leaky(x);
                            if (P.mask) x *= leaky_grad_from_act(P.mask[(long)m * P.ld_mask + n]);
                            P.out[(long)m * P.ld_out + n] = x;
                        } else {
                            P.colext[m] = x;
                        }
                    }
                }
            }
        };

        if (P.splits == 1) {
#pragma unroll 4
            for (int rr = r_in; rr < rows; rr += rpi)
                if (col_ok) finish4(rr, *reinterpret_cast<const float4*>(sT + rr * TP + c4));
        } else {
            // ---- split-K: park the raw partial tile ([128][bn], coalesced 16-byte rows), take a ticket; the last CTA of
            //      the tile sums the partials in split order (deterministic) and runs the epilogue ----
            const long sstride = (long)TC_BM * P.bn;
            float* wt = P.ws + ((long)tile * P.splits + split) * sstride + (q * 32) * P.bn + c_lo + c4;
#pragma unroll 4
            for (int rr = r_in; rr < rows; rr += rpi)
                if (col_ok) *reinterpret_cast<float4*>(wt + rr * P.bn) = *reinterpret_cast<const float4*>(sT + rr * TP + c4);
            __threadfence();
            asm volatile("bar.sync 2, %0;" ::"n"(32 * TC_EPI_WARPS) : "memory");
            if (threadIdx.x == 64) {
                const unsigned t = atomicAdd(P.tickets + tile, 1u);
                const int last = (t == (unsigned)P.splits - 1u);
                if (last) P.tickets[tile] = 0u;               // ready for the next launch / graph replay
                *s_last = last;
            }
            asm volatile("bar.sync 2, %0;" ::"n"(32 * TC_EPI_WARPS) : "memory");
            if (*s_last) {
                __threadfence();
                const float* w0 = P.ws + (long)tile * P.splits * sstride + (q * 32) * P.bn + c_lo + c4;
#pragma unroll 2
                for (int rr = r_in; rr < rows; rr += rpi) {
                    if (col_ok) {
                        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                        for (int sp0 = 0; sp0 < P.splits; sp0 += 4) {        // 4 loads in flight, added in split order
                            float4 x[4];
#pragma unroll
                            for (int d = 0; d < 4; ++d)
                                x[d] = (sp0 + d < P.splits) ? __ldcg(reinterpret_cast<const float4*>(w0 + (sp0 + d) * sstride + rr * P.bn))
                                                            : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                            for (int d = 0; d < 4; ++d) { acc.x += x[d].x; acc.y += x[d].y; acc.z += x[d].z; acc.w += x[d].w; }
                        }
                        finish4(rr, acc);
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

// Tiling plan of a group: tile width and split-K factor per problem, chosen so that the launch has about
// one CTA per SM (these GEMMs are a few hundred MFLOP: a 128x128 tile per CTA leaves most of the GPU idle
// behind a 50-slab dependency chain).  Without a workspace every problem runs unsplit on 128-wide tiles.
struct TcPlan { int bn[TC_MAX_PROBLEMS], splits[TC_MAX_PROBLEMS], tiles[TC_MAX_PROBLEMS]; long ws_off[TC_MAX_PROBLEMS]; int tk_off[TC_MAX_PROBLEMS]; long ws_floats; int n_tickets; int ctas; };
static void tc_plan(const b200dp_tc_problem* probs, int count, bool can_split, TcPlan& pl) {
    const int target = 148;
    auto tiles_for = [&](int bn) {
        int t = 0;
        for (int i = 0; i < count; ++i) t += ((probs[i].M + TC_BM - 1) / TC_BM) * ((probs[i].N + bn - 1) / bn);
        return t;
    };
    int bn = 128;
    if (can_split) {
        if (tiles_for(128) < 48) bn = 64;
        if (bn == 64 && tiles_for(64) < 48) bn = 32;
    }
    const int total_tiles = tiles_for(bn);
    pl.ws_floats = 0; pl.n_tickets = 0; pl.ctas = 0;
    for (int i = 0; i < count; ++i) {
        const int nslab = (probs[i].K + TC_BK - 1) / TC_BK;
        int sp = 1;
        if (can_split && total_tiles < target) {
            sp = target / total_tiles;
            const int max_sp = nslab / 6 > 0 ? nslab / 6 : 1;          // at least 6 slabs per split: a split costs a
                                                                       // partial-tile round trip through L2
            if (sp > max_sp) sp = max_sp;
            if (sp > 16) sp = 16;
            if (sp < 1) sp = 1;
        }
        pl.bn[i] = bn; pl.splits[i] = sp;
        pl.tiles[i] = ((probs[i].M + TC_BM - 1) / TC_BM) * ((probs[i].N + bn - 1) / bn);
        pl.ws_off[i] = pl.ws_floats; pl.tk_off[i] = pl.n_tickets;
        if (sp > 1) { pl.ws_floats += (long)pl.tiles[i] * sp * TC_BM * bn; pl.n_tickets += pl.tiles[i]; }
        pl.ctas += pl.tiles[i] * sp;
    }
}

static int tc_launch(const b200dp_tc_problem* probs, int count, bool retile, float* ws, long ws_floats, unsigned* tickets,
                     int n_tickets, long long* dbg, b200dp_stream_t s);
static long long* g_tc_dbg = nullptr;
// Profiling aid: while set, CTA 0 of every tensor-core GEMM launch writes its clock64() stamps to dbg[0..32).
extern "C" void b200dp_tc_gemm_set_debug(long long* dbg) { g_tc_dbg = dbg; }

extern "C" int b200dp_tc_gemm_group(const b200dp_tc_problem* probs, int count, b200dp_stream_t s) {
    return tc_launch(probs, count, false, nullptr, 0, nullptr, 0, nullptr, s);
}
extern "C" int b200dp_tc_gemm_group_splitk(const b200dp_tc_problem* probs, int count, float* ws, long ws_floats,
                                           uint32_t* tickets, int n_tickets, b200dp_stream_t s) {
    return tc_launch(probs, count, true, ws, ws_floats, tickets, n_tickets, nullptr, s);
}
extern "C" long b200dp_tc_gemm_ws_floats(const b200dp_tc_problem* probs, int count, int* n_tickets) {
    if (count < 1 || count > TC_MAX_PROBLEMS) return -1;
    TcPlan pl;
    tc_plan(probs, count, true, pl);
    if (n_tickets) *n_tickets = pl.n_tickets;
    return pl.ws_floats;
}
// Profiling aid: same launch, CTA 0 writes clock64() stamps into dbg[0..32) (device memory):
// [0] start, [1] setup done, [2] accumulator complete, [3] epilogue done, [4] end,
// [8+s] producer issued slab s, [16+s] slab s landed (s < 8).
extern "C" int b200dp_tc_gemm_group_timed(const b200dp_tc_problem* probs, int count, long long* dbg, b200dp_stream_t s) {
    return tc_launch(probs, count, false, nullptr, 0, nullptr, 0, dbg, s);
}

static int tc_launch(const b200dp_tc_problem* probs, int count, bool retile, float* ws, long ws_floats, unsigned* tickets,
                     int n_tickets, long long* dbg, b200dp_stream_t s) {
    B200DP_REQUIRE(count >= 1 && count <= TC_MAX_PROBLEMS, "1..4 problems per launch");
    TcPlan pl;
    tc_plan(probs, count, retile, pl);
    B200DP_REQUIRE(pl.ws_floats <= ws_floats && pl.n_tickets <= n_tickets && (pl.ws_floats == 0 || (ws && tickets)),
                   "split-K workspace too small: size it with b200dp_tc_gemm_ws_floats");
    TcGroup g;
    memset(&g, 0, sizeof(g));
    g.count = count;
    g.dbg = dbg ? dbg : g_tc_dbg;
    int total = 0;
    for (int i = 0; i < count; ++i) {
        const b200dp_tc_problem& q = probs[i];
        TcProblem& p = g.p[i];
        B200DP_REQUIRE(q.M > 0 && q.N > 0 && q.K > 0, "empty problem");
        const int bn = pl.bn[i];
        int rc;
        // K-major: memory [MN][K] -> box {32 k, rows mn};  MN-major: memory [K][MN] -> box {32 mn, 32 k}
        rc = q.a_mn_major ? encode_map(&p.tmA, q.a, q.K, q.M, q.lda, 32, TC_BK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
                          : encode_map(&p.tmA, q.a, q.M, q.K, q.lda, TC_BK, TC_BM, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        rc = q.b_mn_major ? encode_map(&p.tmB, q.b, q.K, q.N, q.ldb, 32, TC_BK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
                          : encode_map(&p.tmB, q.b, q.N, q.K, q.ldb, TC_BK, bn, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        p.bias = q.bias; p.mask = q.mask; p.ld_mask = q.ld_mask; p.out = q.out; p.ld_out = q.ld_out;
        p.colext = q.colext; p.M = q.M; p.N = q.N; p.K = q.K;
        p.a_mn = q.a_mn_major ? 1 : 0; p.b_mn = q.b_mn_major ? 1 : 0; p.act = q.act;
        p.bn = bn; p.splits = pl.splits[i];
        p.tiles_n = (q.N + bn - 1) / bn;
        p.cta_start = total;
        p.ws = ws ? ws + pl.ws_off[i] : nullptr;
        p.tickets = tickets ? tickets + pl.tk_off[i] : nullptr;
        total += pl.tiles[i] * pl.splits[i];
    }
    static DeviceOnce attr;
    B200DP_CHECK(attr.max_dynamic_smem(tc_gemm_kernel, TC_SMEM));
    B200DP_CHECK(launch_pdl(tc_gemm_kernel, dim3(total), dim3(TC_THREADS), TC_SMEM, (cudaStream_t)s, g));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
