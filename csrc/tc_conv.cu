// Convolutions as implicit GEMMs on the 5th-generation tensor cores (tcgen05.mma kind::tf32,
// accumulators in tensor memory), for the layers where MMA rate matters (the 4x-wide CIFAR-shaped
// model; also selectable for the reference LeNet's conv2).
//
//   fprop : D[(img,cell,win)][oc] = sum_(ic,ky,kx) X[img][ic][oy+ky][ox+kx] W[oc][ic][ky][kx]
//           epilogue: + bias, LeakyReLU(0.2), 2x2 max-pool + argmax across the 4 rows of a pool window
//   dgrad : D[(y,x)][ic]      = sum_(oc,u,v) Zpad[oc][y+u][x+v] W[oc][ic][4-u][4-v]   (Zpad = un-pooled,
//           zero-padded output gradient); epilogue: * LeakyReLU'(P1)
//   wgrad : D[(ic,ky,kx)|1][oc] = sum_(img,pos) X[img][ic][oy+ky][ox+kx] dZ[img][oc][pos]
//           (the extra all-ones row yields the bias gradient); split over image groups into the
//           partial-gradient rows that b200dp_reduce_update sums in fixed order
//
// Replaces reference cnn.py:10-15 (Conv2d/LeakyReLU/MaxPool2d, cuDNN + ATen kernels) and their autograd
// backward (train.py:49).
//
// One CTA = one 128-row accumulator tile, 160 threads:
//   warps 0-3  producers, then epilogue.  The input planes of the CTA's image(s) are staged in shared memory
//              (one TMA bulk copy, cp.async.bulk, when they are contiguous fp32; the first layer normalises
//              and up-samples the raw image instead); each 32-deep K slab of the im2col operand
//              is gathered from there and written straight into the SWIZZLE_128B K-major layout the
//              tensor core reads (lane = k, so both the gather and the swizzled store are conflict-free),
//              followed by fence.proxy.async + mbarrier arrive.  3-stage ring.
//   warp 4     allocates TMEM, one elected lane issues 4 x tcgen05.mma (M128 x NP x K8) per slab and
//              tcgen05.commit's the stage back to the producers.
// Epilogue: tcgen05.ld 32x32b (thread = accumulator row), fused math, coalesced stores.
#include "common.cuh"
#include "tc_common.cuh"

namespace b200dp {

constexpr int CV_STAGES = 3;
constexpr int CV_PRODUCERS = 128;
constexpr int CV_THREADS = 160;
constexpr int CV_A_BYTES = 128 * 128;   // [128 rows][32 floats]

template <int NP>
struct CvRing {
    static constexpr int B_BYTES = NP * 128;
    static constexpr int STAGE_BYTES = CV_A_BYTES + B_BYTES;
    static constexpr int BYTES = CV_STAGES * STAGE_BYTES;
    static constexpr int TMEM_COLS = NP <= 32 ? 32 : (NP <= 64 ? 64 : (NP <= 128 ? 128 : 256));
    static_assert(NP % 16 == 0 && NP <= 256, "MMA N must be a multiple of 16 (M = 128)");
    static_assert(B_BYTES % 1024 == 0, "stage bases must stay 1024-byte aligned");
};

// Everything a role needs to find the ring, the barriers and the accumulator.
struct CvCtl {
    uint8_t* ring_gen;       // generic pointer to stage 0 (1024-byte aligned)
    uint32_t ring, bars, tmem;
    __device__ uint32_t full(int s) const { return bars + 8u * s; }
    __device__ uint32_t empty(int s) const { return bars + 8u * (CV_STAGES + s); }
    __device__ uint32_t accum() const { return bars + 8u * (2 * CV_STAGES); }
    __device__ uint32_t staged() const { return bars + 8u * (2 * CV_STAGES + 1); }   // TMA bulk staging of the input planes
};

// Common prologue: barrier init, TMEM allocation by warp 4.  `bar_mem` points at >= 64 bytes.
template <int NP>
__device__ __forceinline__ void cv_setup(CvCtl& c, uint8_t* ring_gen, uint64_t* bar_mem, uint32_t* tmem_slot) {
    c.ring_gen = ring_gen;
    c.ring = smem_u32(ring_gen);
    c.bars = smem_u32(bar_mem);
    if (threadIdx.x == 0) {
        for (int s = 0; s < CV_STAGES; ++s) { mbar_init(c.full(s), CV_PRODUCERS); mbar_init(c.empty(s), 1); }
        mbar_init(c.accum(), 1);
        mbar_init(c.staged(), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if ((threadIdx.x >> 5) == 4) tmem_alloc(smem_u32(tmem_slot), CvRing<NP>::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    c.tmem = *tmem_slot;
}
template <int NP>
__device__ __forceinline__ void cv_teardown(const CvCtl& c) {
    tc_fence_before();
    __syncthreads();
    if ((threadIdx.x >> 5) == 4) {
        tc_fence_after();
        tmem_dealloc(c.tmem, CvRing<NP>::TMEM_COLS);
    }
}
// producer side of slab `it` (counted over the CTA's whole life)
__device__ __forceinline__ int cv_acquire(const CvCtl& c, int it) {
    const int st = it % CV_STAGES;
    if (it >= CV_STAGES) mbar_wait(c.empty(st), ((it / CV_STAGES) - 1) & 1);
    return st;
}
__device__ __forceinline__ void cv_publish(const CvCtl& c, int st) {
    fence_proxy_async_smem();
    mbar_arrive(c.full(st));
}
// MMA side of slab `it`; `last` also signals the accumulator barrier
template <int NP>
__device__ __forceinline__ void cv_mma(const CvCtl& c, int it, bool last) {
    const int st = it % CV_STAGES;
    mbar_wait(c.full(st), (it / CV_STAGES) & 1);
    tc_fence_after();
    if ((threadIdx.x & 31) == 0) {
        const uint32_t sa = c.ring + st * CvRing<NP>::STAGE_BYTES, sb = sa + CV_A_BYTES;
        const uint32_t idesc = idesc_tf32_m128(NP);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
            umma_tf32(c.tmem, kmajor_sw128_desc(sa + kk * 32), kmajor_sw128_desc(sb + kk * 32), idesc, (it | kk) != 0);
        umma_commit(c.empty(st));
        if (last) umma_commit(c.accum());
    }
    __syncwarp();
}

// ================================================================================================
// fprop + bias + LeakyReLU + 2x2 max-pool (+argmax)
// ================================================================================================
struct TcConvFwdArgs {
    const void* x;        // FIRST: raw [B][CIN][hw][hw] u8|f32; else fp32 [B][CIN][IN_W][IN_W]
    const float* w;       // [COUT][CIN*25]
    const float* bias;
    float* out;           // [B][out_img_stride] holding [COUT][PW][PW]
    uint8_t* idx;         // [B][COUT][PW][PW], nullable
    int B, in_is_u8, in_hw, out_img_stride;
    float mean, std;
};

template <int CIN, int COUT, int IN_W, bool FIRST>
struct TcConvFwdCfg {
    static constexpr int PW = (IN_W - 4) / 2, PP = PW * PW, PLANE = IN_W * IN_W;
    static constexpr int K = CIN * 25, NSLAB = (K + 31) / 32;
    static constexpr int NP = (COUT + 15) / 16 * 16;
    static constexpr int TILES = (PP + 31) / 32;           // 32 pool cells = 128 accumulator rows per tile
    static constexpr int IN_FLOATS = CIN * PLANE + (FIRST ? CIN * PLANE : 0);
    static constexpr size_t SMEM = 1024 + CvRing<NP>::BYTES + (size_t)IN_FLOATS * 4 + 128 * 4 + 128;
};

template <int CIN, int COUT, int IN_W, bool FIRST>
__global__ void __launch_bounds__(CV_THREADS) tc_conv_fwd_kernel(TcConvFwdArgs a) {
    using C = TcConvFwdCfg<CIN, COUT, IN_W, FIRST>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    float* sIn = reinterpret_cast<float*>(ring + CvRing<C::NP>::BYTES);       // [CIN][PLANE]
    float* sRaw = sIn + CIN * C::PLANE;                                        // FIRST only
    int* sRowBase = reinterpret_cast<int*>(sIn + C::IN_FLOATS);                // [128]
    uint64_t* bar_mem = reinterpret_cast<uint64_t*>(sRowBase + 128);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mem + 2 * CV_STAGES + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int img = blockIdx.x / C::TILES, tile = blockIdx.x - img * C::TILES;

    CvCtl ctl;
    cv_setup<C::NP>(ctl, ring, bar_mem, tmem_slot);

    // ---- stage the image and the per-row gather bases (all 160 threads) ----
    if constexpr (FIRST) {
        const int hw = a.in_hw, n_in = hw * hw;
        for (int i = tid; i < CIN * n_in; i += CV_THREADS)
            sRaw[i] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * n_in + i, a.mean, a.std);
        __syncthreads();
        for (int c = 0; c < CIN; ++c) upsample32_plane(sRaw + c * n_in, hw, sIn + c * C::PLANE, tid, CV_THREADS);
    } else {
        // the image's CIN planes are contiguous: one TMA bulk copy, completion on an mbarrier
        constexpr uint32_t bytes = CIN * C::PLANE * 4;
        static_assert(bytes % 16 == 0, "bulk copies move multiples of 16 bytes");
        if (tid == 0) {
            mbar_expect_tx(ctl.staged(), bytes);
            tma_bulk_load(smem_u32(sIn), (const float*)a.x + (long)img * CIN * C::PLANE, bytes, ctl.staged());
        }
        mbar_wait(ctl.staged(), 0);
    }
    if (tid < 128) {
        const int cell = tile * 32 + (tid >> 2), win = tid & 3;
        const int py = cell / C::PW, px = cell - py * C::PW;
        sRowBase[tid] = (cell < C::PP) ? (2 * py + (win >> 1)) * IN_W + 2 * px + (win & 1) : -1;
    }
    __syncthreads();

    if (warp < 4) {
        // ================= producers =================
        for (int s = 0; s < C::NSLAB; ++s) {
            const int st = cv_acquire(ctl, s);
            uint8_t* sa = ctl.ring_gen + st * CvRing<C::NP>::STAGE_BYTES;
            uint8_t* sb = sa + CV_A_BYTES;
            const int k = s * 32 + lane;
            const bool kv = k < C::K;
            const int ic = k / 25, t = k - ic * 25, ky = t / 5, kx = t - ky * 5;
            const int koff = ic * C::PLANE + ky * IN_W + kx;
#pragma unroll 4
            for (int rr = 0; rr < 32; ++rr) {
                const int r = warp * 32 + rr;
                const int rb = sRowBase[r];
                const float v = (kv && rb >= 0) ? sIn[rb + koff] : 0.f;
                *reinterpret_cast<float*>(sa + sw128_off(r, lane)) = v;
            }
            for (int n = warp; n < C::NP; n += 4) {
                const float v = (kv && n < COUT) ? a.w[(long)n * C::K + k] : 0.f;
                *reinterpret_cast<float*>(sb + sw128_off(n, lane)) = v;
            }
            cv_publish(ctl, st);
        }
        // ================= epilogue =================
        mbar_wait(ctl.accum(), 0);
        tc_fence_after();
        const int cell = tile * 32 + warp * 8 + (lane >> 2), win = lane & 3;
#pragma unroll 1
        for (int c0 = 0; c0 < C::NP; c0 += 16) {
            uint32_t r[16];
            tmem_ld16(ctl.tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int n = c0 + j;
                if (n < COUT) {                                   // warp-uniform
                    float best = leaky(__uint_as_float(r[j]) + a.bias[n]);
                    int bi = win;
#pragma unroll
                    for (int o = 1; o <= 2; o <<= 1) {            // max over the 4 rows of the pool window,
                        const float ov = __shfl_xor_sync(0xffffffffu, best, o);   // first maximum wins (torch)
                        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
                    }
                    if (win == 0 && cell < C::PP) {
                        a.out[(long)img * a.out_img_stride + n * C::PP + cell] = best;
                        if (a.idx) a.idx[((long)img * COUT + n) * C::PP + cell] = (uint8_t)bi;
                    }
                }
            }
        }
    } else {
        // ================= MMA issuer =================
        for (int s = 0; s < C::NSLAB; ++s) cv_mma<C::NP>(ctl, s, s == C::NSLAB - 1);
    }
    cv_teardown<C::NP>(ctl);
}

// ================================================================================================
// dgrad of the second convolution (+ LeakyReLU' of the first layer's pooled output)
// ================================================================================================
struct TcConvDgradArgs {
    const float* dp2; const float* p2; int ld_p2; const uint8_t* idx2;   // [B][COUT][25]
    const float* w2;                                                   // [COUT][CIN][25]
    const float* p1;                                                   // [B][CIN][196]
    float* g1;                                                         // [B][CIN][196]
    int B;
};
template <int CIN, int COUT>
struct TcConvDgradCfg {
    static constexpr int ZW = 18, ZPLANE = 324;
    static constexpr int K = COUT * 25, NSLAB = (K + 31) / 32;
    static constexpr int NP = (CIN + 15) / 16 * 16;
    static constexpr int TILES = 2;                                    // 196 rows = 128 + 68
    static constexpr size_t SMEM = 1024 + CvRing<NP>::BYTES + (size_t)COUT * ZPLANE * 4 + 128;
};

template <int CIN, int COUT>
__global__ void __launch_bounds__(CV_THREADS) tc_conv_dgrad_kernel(TcConvDgradArgs a) {
    using C = TcConvDgradCfg<CIN, COUT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    float* sZ = reinterpret_cast<float*>(ring + CvRing<C::NP>::BYTES);        // [COUT][18][18], zero padded
    uint64_t* bar_mem = reinterpret_cast<uint64_t*>(sZ + COUT * C::ZPLANE);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mem + 2 * CV_STAGES + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int img = blockIdx.x / C::TILES, tile = blockIdx.x - img * C::TILES;

    CvCtl ctl;
    cv_setup<C::NP>(ctl, ring, bar_mem, tmem_slot);

    // un-pool dP2 into the zero-padded map (pool2-bwd + LeakyReLU-bwd of the second layer)
    for (int i = tid; i < COUT * C::ZPLANE; i += CV_THREADS) sZ[i] = 0.f;
    __syncthreads();
    for (int i = tid; i < COUT * 25; i += CV_THREADS) {
        const int oc = i / 25, pp = i - oc * 25, py = pp / 5, px = pp - py * 5;
        const int k = a.idx2[(long)img * COUT * 25 + i];
        const float g = a.dp2[(long)img * COUT * 25 + i] * leaky_grad_from_act(a.p2[(long)img * a.ld_p2 + i]);
        sZ[oc * C::ZPLANE + (2 * py + (k >> 1) + 4) * C::ZW + 2 * px + (k & 1) + 4] = g;
    }
    __syncthreads();

    if (warp < 4) {
        for (int s = 0; s < C::NSLAB; ++s) {
            const int st = cv_acquire(ctl, s);
            uint8_t* sa = ctl.ring_gen + st * CvRing<C::NP>::STAGE_BYTES;
            uint8_t* sb = sa + CV_A_BYTES;
            const int k = s * 32 + lane;
            const bool kv = k < C::K;
            const int oc = k / 25, t = k - oc * 25, u = t / 5, v = t - u * 5;
            const int koff = oc * C::ZPLANE + u * C::ZW + v;
#pragma unroll 4
            for (int rr = 0; rr < 32; ++rr) {
                const int r = warp * 32 + rr;
                const int m = tile * 128 + r;                       // output position y*14 + x
                const int y = m / 14, x = m - y * 14;
                const float val = (kv && m < 196) ? sZ[koff + y * C::ZW + x] : 0.f;
                *reinterpret_cast<float*>(sa + sw128_off(r, lane)) = val;
            }
            for (int n = warp; n < C::NP; n += 4) {                 // B(n = ic, k) = W2[oc][ic][4-u][4-v]
                const float val = (kv && n < CIN) ? a.w2[((long)oc * CIN + n) * 25 + (4 - u) * 5 + (4 - v)] : 0.f;
                *reinterpret_cast<float*>(sb + sw128_off(n, lane)) = val;
            }
            cv_publish(ctl, st);
        }
        mbar_wait(ctl.accum(), 0);
        tc_fence_after();
        const int m = tile * 128 + warp * 32 + lane;
#pragma unroll 1
        for (int c0 = 0; c0 < C::NP; c0 += 16) {
            uint32_t r[16];
            tmem_ld16(ctl.tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
            if (m < 196) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int n = c0 + j;
                    if (n < CIN) {
                        const long o = ((long)img * CIN + n) * 196 + m;
                        a.g1[o] = __uint_as_float(r[j]) * leaky_grad_from_act(a.p1[o]);
                    }
                }
            }
        }
    } else {
        for (int s = 0; s < C::NSLAB; ++s) cv_mma<C::NP>(ctl, s, s == C::NSLAB - 1);
    }
    cv_teardown<C::NP>(ctl);
}

// ================================================================================================
// wgrad (+ bias gradient), split over image groups into partial rows
// ================================================================================================
struct TcConvWgradArgs {
    const void* x;            // FIRST: raw images; else fp32 [B][CIN][IN_W][IN_W]
    int in_is_u8, in_hw; float mean, std;
    const float* g;           // [B][COUT][PP] gradient at the pooled cells
    const float* act; int ld_act;   // pooled activation for LeakyReLU' (nullable: g already carries it)
    const uint8_t* idx;       // [B][COUT][PP] argmax inside the pool window
    float* partials; int partial_stride, off_w, off_b;
    int B, img_per_cta;
};
template <int CIN, int COUT, int IN_W, bool FIRST>
struct TcConvWgradCfg {
    static constexpr int OW = IN_W - 4, NPOS = OW * OW, PW = OW / 2, PP = PW * PW, PLANE = IN_W * IN_W;
    static constexpr int MROWS = CIN * 25 + 1;                    // + the all-ones row (bias gradient)
    static constexpr int MTILES = (MROWS + 127) / 128;
    static constexpr int NSLAB = (NPOS + 31) / 32;                // position slabs per image
    static constexpr int NP = (COUT + 15) / 16 * 16;
    static constexpr int NPOS_PAD = NSLAB * 32;
    static constexpr int IN_FLOATS = CIN * PLANE + (FIRST ? CIN * PLANE : 0);
    static constexpr size_t SMEM = 1024 + CvRing<NP>::BYTES + (size_t)(IN_FLOATS + COUT * NPOS_PAD) * 4 + 128 * 4 + 128;
};

template <int CIN, int COUT, int IN_W, bool FIRST>
__global__ void __launch_bounds__(CV_THREADS) tc_conv_wgrad_kernel(TcConvWgradArgs a) {
    using C = TcConvWgradCfg<CIN, COUT, IN_W, FIRST>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    float* sIn = reinterpret_cast<float*>(ring + CvRing<C::NP>::BYTES);       // [CIN][PLANE]
    float* sRaw = sIn + CIN * C::PLANE;                                        // FIRST only
    float* sdZ = sIn + C::IN_FLOATS;                                           // [COUT][NPOS_PAD] dense, zero filled
    int* sRowOff = reinterpret_cast<int*>(sdZ + COUT * C::NPOS_PAD);           // [128]
    uint64_t* bar_mem = reinterpret_cast<uint64_t*>(sRowOff + 128);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mem + 2 * CV_STAGES + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int mtile = blockIdx.y;
    const int img0 = blockIdx.x * a.img_per_cta;
    const int nimg = min(a.img_per_cta, a.B - img0);

    CvCtl ctl;
    cv_setup<C::NP>(ctl, ring, bar_mem, tmem_slot);
    if (tid < 128) {      // gather offset of accumulator row m = (ic,ky,kx); -1 = all-ones row; -2 = unused
        const int m = mtile * 128 + tid;
        const int ic = m / 25, t = m - ic * 25, ky = t / 5, kx = t - ky * 5;
        sRowOff[tid] = (m < CIN * 25) ? ic * C::PLANE + ky * IN_W + kx : (m == CIN * 25 ? -1 : -2);
    }

    for (int ii = 0; ii < nimg; ++ii) {
        const int img = img0 + ii;
        // the staging buffers are read by the producer warps only: once all of them have finished the
        // previous image's slabs (CTA barrier) the buffers can be refilled
        __syncthreads();
        if constexpr (FIRST) {
            const int hw = a.in_hw, n_in = hw * hw;
            for (int i = tid; i < CIN * n_in; i += CV_THREADS)
                sRaw[i] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * n_in + i, a.mean, a.std);
        } else {
            constexpr uint32_t bytes = CIN * C::PLANE * 4;
            static_assert(bytes % 16 == 0, "bulk copies move multiples of 16 bytes");
            if (tid == 0) {
                mbar_expect_tx(ctl.staged(), bytes);
                tma_bulk_load(smem_u32(sIn), (const float*)a.x + (long)img * CIN * C::PLANE, bytes, ctl.staged());
            }
        }
        for (int i = tid; i < COUT * C::NPOS_PAD; i += CV_THREADS) sdZ[i] = 0.f;
        if constexpr (!FIRST) mbar_wait(ctl.staged(), ii & 1);
        __syncthreads();
        if constexpr (FIRST) {
            const int hw = a.in_hw, n_in = hw * hw;
            for (int c = 0; c < CIN; ++c) upsample32_plane(sRaw + c * n_in, hw, sIn + c * C::PLANE, tid, CV_THREADS);
        }
        for (int i = tid; i < COUT * C::PP; i += CV_THREADS) {        // pool-bwd (+ LeakyReLU-bwd) scatter
            const int oc = i / C::PP, pp = i - oc * C::PP, py = pp / C::PW, px = pp - py * C::PW;
            const long o = (long)img * COUT * C::PP + i;
            const int k = a.idx[o];
            float g = a.g[o];
            if (a.act) g *= leaky_grad_from_act(a.act[(long)img * a.ld_act + i]);
            sdZ[oc * C::NPOS_PAD + (2 * py + (k >> 1)) * C::OW + 2 * px + (k & 1)] = g;
        }
        __syncthreads();

        if (warp < 4) {
            for (int s = 0; s < C::NSLAB; ++s) {
                const int it = ii * C::NSLAB + s;
                const int st = cv_acquire(ctl, it);
                uint8_t* sa = ctl.ring_gen + st * CvRing<C::NP>::STAGE_BYTES;
                uint8_t* sb = sa + CV_A_BYTES;
                const int pos = s * 32 + lane;
                const bool pv = pos < C::NPOS;
                const int oy = pos / C::OW, ox = pos - oy * C::OW;
                const int poff = oy * IN_W + ox;
#pragma unroll 4
                for (int rr = 0; rr < 32; ++rr) {
                    const int r = warp * 32 + rr;
                    const int ro = sRowOff[r];
                    const float v = !pv ? 0.f : (ro >= 0 ? sIn[ro + poff] : (ro == -1 ? 1.f : 0.f));
                    *reinterpret_cast<float*>(sa + sw128_off(r, lane)) = v;
                }
                for (int n = warp; n < C::NP; n += 4) {
                    const float v = (n < COUT) ? sdZ[n * C::NPOS_PAD + pos] : 0.f;
                    *reinterpret_cast<float*>(sb + sw128_off(n, lane)) = v;
                }
                cv_publish(ctl, st);
            }
        } else {
            for (int s = 0; s < C::NSLAB; ++s) {
                const int it = ii * C::NSLAB + s;
                cv_mma<C::NP>(ctl, it, ii == nimg - 1 && s == C::NSLAB - 1);
            }
        }
    }

    if (warp < 4) {
        float* row = a.partials + (long)blockIdx.x * a.partial_stride;
        const int m = mtile * 128 + warp * 32 + lane;
        if (nimg > 0) {
            mbar_wait(ctl.accum(), 0);
            tc_fence_after();
        }
#pragma unroll 1
        for (int c0 = 0; c0 < C::NP; c0 += 16) {
            uint32_t r[16];
            if (nimg > 0) {
                tmem_ld16(ctl.tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) r[j] = 0u;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int n = c0 + j;
                if (n < COUT) {
                    if (m < CIN * 25) row[a.off_w + n * (CIN * 25) + m] = __uint_as_float(r[j]);
                    else if (m == CIN * 25) row[a.off_b + n] = __uint_as_float(r[j]);
                }
            }
        }
    }
    cv_teardown<C::NP>(ctl);
}

template <class K, class A>
static int cv_launch(K kern, dim3 grid, size_t smem, const A& args, cudaStream_t s) {
    B200DP_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, CV_THREADS, smem, s>>>(args);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_tc_conv1_fwd(const void* x, int in_is_u8, float mean, float std, const float* w, const float* b,
                                   float* p1, uint8_t* idx1, b200dp_dims d, b200dp_stream_t s) {
    B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
    TcConvFwdArgs a{x, w, b, p1, idx1, d.batch, in_is_u8, d.in_hw, d.c1 * 196, mean, std};
    if (d.in_ch == 3 && d.c1 == 24) {
        using C = TcConvFwdCfg<3, 24, 32, true>;
        return cv_launch(tc_conv_fwd_kernel<3, 24, 32, true>, dim3(d.batch * C::TILES), C::SMEM, a, (cudaStream_t)s);
    }
    if (d.in_ch == 1 && d.c1 == 6) {
        using C = TcConvFwdCfg<1, 6, 32, true>;
        return cv_launch(tc_conv_fwd_kernel<1, 6, 32, true>, dim3(d.batch * C::TILES), C::SMEM, a, (cudaStream_t)s);
    }
    B200DP_REQUIRE(false, "unsupported (in_ch, c1): expected (1,6) or (3,24)");
}

extern "C" int b200dp_tc_conv2_fwd(const float* p1, const float* w, const float* b, float* p2, uint8_t* idx2,
                                   b200dp_dims d, b200dp_stream_t s) {
    const int stride = d.ld_p2 > 0 ? d.ld_p2 : d.c2 * 25;
    TcConvFwdArgs a{p1, w, b, p2, idx2, d.batch, 0, 14, stride, 0.f, 1.f};
    if (d.c1 == 24 && d.c2 == 64) {
        using C = TcConvFwdCfg<24, 64, 14, false>;
        return cv_launch(tc_conv_fwd_kernel<24, 64, 14, false>, dim3(d.batch * C::TILES), C::SMEM, a, (cudaStream_t)s);
    }
    if (d.c1 == 6 && d.c2 == 16) {
        using C = TcConvFwdCfg<6, 16, 14, false>;
        return cv_launch(tc_conv_fwd_kernel<6, 16, 14, false>, dim3(d.batch * C::TILES), C::SMEM, a, (cudaStream_t)s);
    }
    B200DP_REQUIRE(false, "unsupported (c1, c2): expected (6,16) or (24,64)");
}

extern "C" int b200dp_tc_conv2_dgrad(const float* dp2, const float* p2, const uint8_t* idx2, const float* w2,
                                     const float* p1, float* g1, b200dp_dims d, b200dp_stream_t s) {
    TcConvDgradArgs a{dp2, p2, d.ld_p2 > 0 ? d.ld_p2 : d.c2 * 25, idx2, w2, p1, g1, d.batch};
    if (d.c1 == 24 && d.c2 == 64) {
        using C = TcConvDgradCfg<24, 64>;
        return cv_launch(tc_conv_dgrad_kernel<24, 64>, dim3(d.batch * C::TILES), C::SMEM, a, (cudaStream_t)s);
    }
    if (d.c1 == 6 && d.c2 == 16) {
        using C = TcConvDgradCfg<6, 16>;
        return cv_launch(tc_conv_dgrad_kernel<6, 16>, dim3(d.batch * C::TILES), C::SMEM, a, (cudaStream_t)s);
    }
    B200DP_REQUIRE(false, "unsupported (c1, c2): expected (6,16) or (24,64)");
}

extern "C" int b200dp_tc_conv_wgrad_num_partials(b200dp_dims d) {
    int n = (d.batch + 7) / 8;
    if (n > 64) n = 64;
    return n < 1 ? 1 : n;
}

extern "C" int b200dp_tc_conv_wgrad(int layer, const void* x, int in_is_u8, float mean, float std, const float* g,
                                    const float* act, int ld_act, const uint8_t* idx, float* partials,
                                    int partial_stride, int off_w, int off_b, b200dp_dims d, b200dp_stream_t s) {
    const int rows = b200dp_tc_conv_wgrad_num_partials(d);
    const int ipc = (d.batch + rows - 1) / rows;
    TcConvWgradArgs a{x, in_is_u8, d.in_hw, mean, std, g, act, ld_act, idx, partials, partial_stride, off_w, off_b,
                      d.batch, ipc};
    if (layer == 2) {
        a.in_hw = 14;
        if (d.c1 == 24 && d.c2 == 64) {
            using C = TcConvWgradCfg<24, 64, 14, false>;
            return cv_launch(tc_conv_wgrad_kernel<24, 64, 14, false>, dim3(rows, C::MTILES), C::SMEM, a, (cudaStream_t)s);
        }
        if (d.c1 == 6 && d.c2 == 16) {
            using C = TcConvWgradCfg<6, 16, 14, false>;
            return cv_launch(tc_conv_wgrad_kernel<6, 16, 14, false>, dim3(rows, C::MTILES), C::SMEM, a, (cudaStream_t)s);
        }
    } else if (layer == 1) {
        B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
        if (d.in_ch == 3 && d.c1 == 24) {
            using C = TcConvWgradCfg<3, 24, 32, true>;
            return cv_launch(tc_conv_wgrad_kernel<3, 24, 32, true>, dim3(rows, C::MTILES), C::SMEM, a, (cudaStream_t)s);
        }
        if (d.in_ch == 1 && d.c1 == 6) {
            using C = TcConvWgradCfg<1, 6, 32, true>;
            return cv_launch(tc_conv_wgrad_kernel<1, 6, 32, true>, dim3(rows, C::MTILES), C::SMEM, a, (cudaStream_t)s);
        }
    }
    B200DP_REQUIRE(false, "unsupported layer / channel combination");
}
