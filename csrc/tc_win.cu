// "Window-tile" convolutions on tcgen05: implicit GEMM with NO im2col materialisation.
//
// The activations of an image are staged once in shared memory as [channel group of 4][pixel][4 floats]
// (16 bytes per pixel per group).  In that layout a tcgen05 K-major, no-swizzle operand descriptor can
// address an accumulator tile of 128 output pixels = 16 image rows x 8 columns DIRECTLY:
//     core matrix (8 rows x 16 B)  = 8 consecutive pixels of one image row, 4 channels   (contiguous 128 B)
//     SBO (next 8 rows of the tile) = one image row further   (PITCH * 16 B)
//     LBO (next 4 channels of K)    = next channel group      (NPIX * 16 B)
// and a filter tap (ky,kx) is nothing but a different start address (+ (ky*PITCH+kx) * 16 B, always
// 16-byte aligned).  So a 5x5 convolution is 25 x (C/8) tcgen05.mma instructions per tile that read
// the SAME staged image; nothing is gathered per K slab (the previous generation of these kernels spent
// >95% of its time writing im2col slabs with scalar shared-memory stores: tensor pipe 1-3% busy).
// The per-tap weight panels [kgroup][n][4] come from a pre-packed copy (b200dp_tcw_pack_weights) through
// a TMA bulk-copy ring (cp.async.bulk + mbarrier complete_tx).
//
//   kind FWD1 : first layer. normalise + bilinear 28->32 on the fly; the staged "channels" are
//               (ic,kx) pairs (kx-expanded), so only the 5 ky taps remain; epilogue bias+LeakyReLU+2x2 pool+argmax
//   kind FWD2 : second layer, 25 taps over the pooled first-layer map; same epilogue
//   kind DGRAD: second layer's data gradient: 25 flipped taps over the zero-padded, un-pooled output
//               gradient (pool-bwd and LeakyReLU-bwd fused into the staging); epilogue * LeakyReLU'(p1)
//
// Replaces reference cnn.py:10-15 (Conv2d/LeakyReLU/MaxPool2d: cuDNN + ATen) and their backward (train.py:49).
//
// CTA = G images, 320 threads: warps 0-7 stage, warp 8 issues tcgen05.mma (one elected lane), warp 9 runs
// the weight ring, warps 0-3 run the epilogue (tcgen05.ld 32x32b, thread = pixel).  Within a tile lane l of
// warp w is pixel (row 4w + l/8, column l%8), so a 2x2 pool window is lanes {l, l^1, l^8, l^9}: two shuffles.
#include "common.cuh"
#include "tc_common.cuh"

namespace b200dp {

constexpr int WIN_THREADS = 320;
constexpr int WIN_STAGERS = 256;

// K-major / MN-major operand descriptor without swizzle (layout type 0), strides in bytes.
__device__ __forceinline__ uint64_t nosw_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
__device__ __forceinline__ uint32_t idesc_tf32(int n, int a_mn, int b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

enum { WIN_FWD1 = 0, WIN_FWD2 = 1, WIN_DGRAD = 2 };

template <int KIND, int CIN, int COUT>
struct WinCfg {
    // geometry of the staged map
    static constexpr int PITCH = KIND == WIN_FWD1 ? 32 : (KIND == WIN_FWD2 ? 14 : 18);
    static constexpr int NTAP = KIND == WIN_FWD1 ? 5 : 25;
    // K channels per tap (padded to a multiple of 8) and N columns (padded to a multiple of 16)
    static constexpr int KREAL = KIND == WIN_FWD1 ? CIN * 5 : (KIND == WIN_FWD2 ? CIN : COUT);
    static constexpr int KC = (KREAL + 7) / 8 * 8, KG = KC / 4, KSTEPS = KC / 8;
    static constexpr int NREAL = KIND == WIN_DGRAD ? CIN : COUT;
    static constexpr int NP = (NREAL + 15) / 16 * 16;
    // tiles of 16 rows x 8 columns covering the valid outputs
    static constexpr int OUT_W = KIND == WIN_FWD1 ? 28 : (KIND == WIN_FWD2 ? 10 : 14);
    static constexpr int NTR = (OUT_W + 15) / 16, NTC = (OUT_W + 7) / 8, NT = NTR * NTC;
    // rows touched: last tile row 16*NTR-1, plus 4 tap rows; columns run over into the next row
    static constexpr int NPIX = (16 * NTR + 4) * PITCH + 16;
    static constexpr int IMG_BYTES = KG * NPIX * 16;
    static constexpr int WB = KC * NP * 4;                       // weight panel of one tap
    static constexpr bool WIDE = COUT > 16;
    static constexpr int G = KIND == WIN_FWD1 ? (WIDE ? 2 : 1) : 2;     // images per CTA
    static constexpr int RING = KIND == WIN_FWD1 ? 5 : (KIND == WIN_FWD2 ? (WIDE ? 8 : 25) : (WIDE ? 4 : 25));
    static constexpr int ACC_COLS = G * NT * NP;
    static constexpr int TMEM_COLS = ACC_COLS <= 32 ? 32 : (ACC_COLS <= 64 ? 64 : (ACC_COLS <= 128 ? 128 : (ACC_COLS <= 256 ? 256 : 512)));
    static_assert(ACC_COLS <= 512, "accumulators exceed tensor memory");
    static_assert(WB % 16 == 0 && IMG_BYTES % 16 == 0, "alignment");
    // FWD1 scratch: raw planes + upsampled planes (fp32), one image at a time
    static constexpr int PLANE_PAD = 1024 + 8;
    static constexpr int SCRATCH = KIND == WIN_FWD1 ? CIN * (1024 + PLANE_PAD) * 4 : 0;
    static constexpr size_t SMEM = 1024 + (size_t)G * IMG_BYTES + (size_t)RING * WB + SCRATCH + 8 * (2 * RING + 2) + 16;
    static_assert(SMEM <= 232448, "shared memory");
    __host__ __device__ static constexpr int tap_shift(int t) {
        return KIND == WIN_FWD1 ? t * PITCH : (t / 5) * PITCH + (t % 5);
    }
};

struct WinArgs {
    const void* x;            // FWD1: raw [B][CIN][hw][hw] u8|f32;  FWD2: p1 [B][CIN][196];  DGRAD: dp2 [B][COUT][25]
    const float* wpack;       // packed weight panels [NTAP][KG][NP][4]
    const float* bias;        // FWD*: [COUT]
    float* out;               // FWD*: pooled output [B][out_img_stride];  DGRAD: g1 [B][CIN][196]
    uint8_t* idx;             // FWD*: argmax [B][COUT][PW*PW] (nullable)
    const float* act;         // DGRAD: p2 [B][ld_act] (LeakyReLU' of layer 2);  unused otherwise
    const uint8_t* pidx;      // DGRAD: idx2 [B][COUT][25]
    const float* mask;        // DGRAD: p1 [B][CIN][196] (LeakyReLU' of layer 1)
    int B, in_is_u8, in_hw, out_img_stride, ld_act;
    float mean, std;
};

// ---------------------------------------------------------------------------------------------------
// staging (warps 0-7)
// ---------------------------------------------------------------------------------------------------
template <class C, int CIN>
__device__ __forceinline__ void win_stage_fwd2(const WinArgs& a, int img, float* sImg, int tid) {
    // NCHW planes -> [cg][pixel][4]: one float4 store per (pixel, channel group), 4 coalesced plane reads
    const float* src = (const float*)a.x + (long)img * CIN * 196;
    for (int i = tid; i < C::KG * C::NPIX; i += WIN_STAGERS) {
        const int cg = i / C::NPIX, p = i - cg * C::NPIX;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p < 196) {
            const int c = cg * 4;
            if (c + 0 < CIN) v.x = __ldg(src + (c + 0) * 196 + p);
            if (c + 1 < CIN) v.y = __ldg(src + (c + 1) * 196 + p);
            if (c + 2 < CIN) v.z = __ldg(src + (c + 2) * 196 + p);
            if (c + 3 < CIN) v.w = __ldg(src + (c + 3) * 196 + p);
        }
        reinterpret_cast<float4*>(sImg)[i] = v;
    }
}

template <class C, int CIN>
__device__ __forceinline__ void win_stage_fwd1(const WinArgs& a, int img, float* sImg, float* sRaw, float* sUp, int tid) {
    const int hw = a.in_hw, n_in = hw * hw;
    for (int i = tid; i < CIN * n_in; i += WIN_STAGERS)
        sRaw[i] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * n_in + i, a.mean, a.std);
    asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");
    for (int c = 0; c < CIN; ++c) {
        upsample32_plane(sRaw + c * n_in, hw, sUp + c * C::PLANE_PAD, tid, WIN_STAGERS);
        if (tid < 8) sUp[c * C::PLANE_PAD + 1024 + tid] = 0.f;          // kx overrun of the last pixel
    }
    asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");
    // kx-expansion: staged channel c' = ic*5 + kx holds X[ic][p + kx]
    for (int i = tid; i < C::KG * C::NPIX; i += WIN_STAGERS) {
        const int cg = i / C::NPIX, p = i - cg * C::NPIX;
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        if (p < 1024) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int cc = cg * 4 + q;
                if (cc < CIN * 5) { const int ic = cc / 5, kx = cc - ic * 5; v[q] = sUp[ic * C::PLANE_PAD + p + kx]; }
            }
        }
        reinterpret_cast<float4*>(sImg)[i] = make_float4(v[0], v[1], v[2], v[3]);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");     // sRaw / sUp are reused by the next image
}

template <class C, int COUT>
__device__ __forceinline__ void win_stage_dgrad(const WinArgs& a, int img, float* sImg, int tid) {
    // zero-padded 18x18 map of the un-pooled gradient: pool-bwd (scatter to the argmax) + LeakyReLU-bwd
    for (int i = tid; i < C::KG * C::NPIX; i += WIN_STAGERS)
        reinterpret_cast<float4*>(sImg)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");
    for (int i = tid; i < COUT * 25; i += WIN_STAGERS) {
        const int oc = i / 25, pp = i - oc * 25, py = pp / 5, px = pp - py * 5;
        const int k = a.pidx[(long)img * COUT * 25 + i];
        const float g = ((const float*)a.x)[(long)img * COUT * 25 + i] * leaky_grad_from_act(a.act[(long)img * a.ld_act + i]);
        const int p = (2 * py + (k >> 1) + 4) * C::PITCH + 2 * px + (k & 1) + 4;
        sImg[((oc >> 2) * C::NPIX + p) * 4 + (oc & 3)] = g;
    }
}

// ---------------------------------------------------------------------------------------------------
template <int KIND, int CIN, int COUT>
__global__ void __launch_bounds__(WIN_THREADS, 1) win_conv_kernel(const __grid_constant__ WinArgs a) {
    using C = WinCfg<KIND, CIN, COUT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    float* sImg = reinterpret_cast<float*>(base);                                   // [G][KG][NPIX][4]
    uint8_t* sW = base + (size_t)C::G * C::IMG_BYTES;                               // [RING][WB]
    float* sScratch = reinterpret_cast<float*>(sW + (size_t)C::RING * C::WB);
    uint64_t* bar_mem = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sScratch) + C::SCRATCH);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mem + 2 * C::RING + 1);
    const uint32_t bars = smem_u32(bar_mem);
    auto w_full = [&](int s) { return bars + 8u * s; };
    auto w_empty = [&](int s) { return bars + 8u * (C::RING + s); };
    const uint32_t acc_bar = bars + 8u * (2 * C::RING);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int img0 = blockIdx.x * C::G;
    const int nimg = min(C::G, a.B - img0);

    if (tid == 0) {
        for (int s = 0; s < C::RING; ++s) { mbar_init(w_full(s), 1); mbar_init(w_empty(s), 1); }
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) tmem_alloc(smem_u32(tmem_slot), C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    auto load_tap = [&](int t) {
        const int s = t % C::RING;
        mbar_expect_tx(w_full(s), C::WB);
        tma_bulk_load(smem_u32(sW + (size_t)s * C::WB), a.wpack + (size_t)t * (C::WB / 4), C::WB, w_full(s));
    };
    if (warp == 9) {
        // ================= weight ring producer (one lane): fill the ring while the images are staged =================
        if (lane == 0)
            for (int t = 0; t < C::RING && t < C::NTAP; ++t) load_tap(t);
    } else if (warp < 8) {
        // ================= stage the CTA's images =================
        for (int g = 0; g < nimg; ++g) {
            float* dst = sImg + (size_t)g * (C::IMG_BYTES / 4);
            if constexpr (KIND == WIN_FWD1) win_stage_fwd1<C, CIN>(a, img0 + g, dst, sScratch, sScratch + CIN * 1024, tid);
            else if constexpr (KIND == WIN_FWD2) win_stage_fwd2<C, CIN>(a, img0 + g, dst, tid);
            else win_stage_dgrad<C, COUT>(a, img0 + g, dst, tid);
        }
        fence_proxy_async_smem();          // generic-proxy stores -> visible to the tensor core's reads
    }
    __syncthreads();

    if (warp == 9) {
        if (lane == 0)
            for (int t = C::RING; t < C::NTAP; ++t) {          // refill a stage once its MMAs have completed
                mbar_wait(w_empty(t % C::RING), ((t / C::RING) - 1) & 1);
                load_tap(t);
            }
    } else if (warp == 8) {
        // ================= MMA issuer =================
        tc_fence_after();
        const uint32_t idesc = idesc_tf32(C::NP, 0, 0);
        const uint32_t img_base = smem_u32(sImg);
        for (int t = 0; t < C::NTAP; ++t) {
            const int s = t % C::RING;
            mbar_wait(w_full(s), (t / C::RING) & 1);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t wb = smem_u32(sW + (size_t)s * C::WB);
                const uint32_t shift = (uint32_t)C::tap_shift(t) * 16u;
                for (int g = 0; g < nimg; ++g) {
#pragma unroll
                    for (int tl = 0; tl < C::NT; ++tl) {
                        const int r0 = (tl / C::NTC) * 16, c0 = (tl % C::NTC) * 8;
                        const uint32_t a0 = img_base + (uint32_t)g * C::IMG_BYTES + (uint32_t)(r0 * C::PITCH + c0) * 16u + shift;
                        const uint32_t d = tmem + (uint32_t)((g * C::NT + tl) * C::NP);
#pragma unroll
                        for (int ks = 0; ks < C::KSTEPS; ++ks) {
                            const uint32_t aa = a0 + (uint32_t)ks * 2u * C::NPIX * 16u, bb = wb + (uint32_t)ks * 2u * C::NP * 16u;
                            umma_tf32(d, nosw_desc(aa, C::NPIX * 16u, C::PITCH * 16u), nosw_desc(bb, C::NP * 16u, 128u), idesc, (t | ks) != 0);
                        }
                    }
                }
                umma_commit(w_empty(s));
                if (t == C::NTAP - 1) umma_commit(acc_bar);
            }
            __syncwarp();
        }
    } else if (warp < 4) {
        // ================= epilogue: thread = pixel (row 4*warp + lane/8, column lane%8 of the tile) =================
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        const int lr = 4 * warp + (lane >> 3), lc = lane & 7;
        for (int g = 0; g < nimg; ++g) {
            const int img = img0 + g;
#pragma unroll 1
            for (int tl = 0; tl < C::NT; ++tl) {
                const int r = (tl / C::NTC) * 16 + lr, c = (tl % C::NTC) * 8 + lc;
                const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)((g * C::NT + tl) * C::NP);
#pragma unroll 1
                for (int n0 = 0; n0 < C::NP; n0 += 16) {
                    uint32_t v[16];
                    tmem_ld16(taddr + (uint32_t)n0, v);
                    if constexpr (KIND == WIN_DGRAD) {
                        if (r < 14 && c < 14) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int n = n0 + j;
                                if (n < CIN) {
                                    const long o = ((long)img * CIN + n) * 196 + r * 14 + c;
                                    a.out[o] = __uint_as_float(v[j]) * leaky_grad_from_act(__ldg(a.mask + o));
                                }
                            }
                        }
                    } else {
                        constexpr int PW = C::OUT_W / 2, PP = PW * PW;
                        const int win = ((r & 1) << 1) | (c & 1);
                        const int py = r >> 1, px = c >> 1;
                        const bool writer = win == 0 && py < PW && px < PW;
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const int n = n0 + j;
                            if (n < COUT) {                                   // warp-uniform
                                float best = leaky(__uint_as_float(v[j]) + __ldg(a.bias + n));
                                int bi = win;
                                {   // columns: lane ^ 1; rows: lane ^ 8.  First maximum wins (torch's tie rule)
                                    float ov = __shfl_xor_sync(0xffffffffu, best, 1);
                                    int oi = __shfl_xor_sync(0xffffffffu, bi, 1);
                                    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
                                    ov = __shfl_xor_sync(0xffffffffu, best, 8);
                                    oi = __shfl_xor_sync(0xffffffffu, bi, 8);
                                    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
                                }
                                if (writer) {
                                    a.out[(long)img * a.out_img_stride + n * PP + py * PW + px] = best;
                                    if (a.idx) a.idx[((long)img * COUT + n) * PP + py * PW + px] = (uint8_t)bi;
                                }
                            }
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tmem, C::TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------------
// weight packing: OIHW -> per-tap K-major panels [tap][kgroup][n][4] (zero padded), all three in one launch
// ---------------------------------------------------------------------------------------------------
struct PackArgs {
    const float* w1; const float* w2;
    float* p_fwd1; float* p_fwd2; float* p_dgrad;
    int in_ch, c1, c2;
    int kc1, np1, kc2, np2, kcd, npd;
};
__global__ void __launch_bounds__(256) win_pack_kernel(PackArgs a) {
    const int n1 = 5 * a.kc1 * a.np1, n2 = 25 * a.kc2 * a.np2, nd = 25 * a.kcd * a.npd;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n1 + n2 + nd; i += gridDim.x * blockDim.x) {
        if (i < n1) {                                   // [ky][kg][oc][4], k = ic*5 + kx
            const int q = i & 3, n = (i >> 2) % a.np1, kg = (i >> 2) / a.np1 % (a.kc1 / 4), ky = (i >> 2) / a.np1 / (a.kc1 / 4);
            const int k = kg * 4 + q, ic = k / 5, kx = k - ic * 5;
            a.p_fwd1[i] = (n < a.c1 && k < a.in_ch * 5) ? a.w1[((n * a.in_ch + ic) * 5 + ky) * 5 + kx] : 0.f;
        } else if (i < n1 + n2) {                       // [tap][kg][oc][4], k = ic
            const int e = i - n1;
            const int q = e & 3, n = (e >> 2) % a.np2, kg = (e >> 2) / a.np2 % (a.kc2 / 4), t = (e >> 2) / a.np2 / (a.kc2 / 4);
            const int k = kg * 4 + q;
            a.p_fwd2[e] = (n < a.c2 && k < a.c1) ? a.w2[(n * a.c1 + k) * 25 + t] : 0.f;
        } else {                                        // [tap=(u,v)][kg][ic][4], k = oc, flipped filter
            const int e = i - n1 - n2;
            const int q = e & 3, n = (e >> 2) % a.npd, kg = (e >> 2) / a.npd % (a.kcd / 4), t = (e >> 2) / a.npd / (a.kcd / 4);
            const int k = kg * 4 + q;
            a.p_dgrad[e] = (n < a.c1 && k < a.c2) ? a.w2[(k * a.c1 + n) * 25 + (24 - t)] : 0.f;
        }
    }
}

template <int KIND, int CIN, int COUT>
static int win_launch(const WinArgs& a, cudaStream_t s) {
    using C = WinCfg<KIND, CIN, COUT>;
    auto kern = win_conv_kernel<KIND, CIN, COUT>;
    static bool attr_set = false;      // per kernel instantiation; setting it again is harmless
    if (!attr_set) {
        B200DP_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
        attr_set = true;
    }
    kern<<<(a.B + C::G - 1) / C::G, WIN_THREADS, C::SMEM, s>>>(a);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

struct PackLayout { int kc1, np1, kc2, np2, kcd, npd; long o1, o2, od, total; };
static bool pack_layout(b200dp_dims d, PackLayout& L) {
    if (d.in_ch == 3 && d.c1 == 24 && d.c2 == 64) {
        using A = WinCfg<WIN_FWD1, 3, 24>; using B_ = WinCfg<WIN_FWD2, 24, 64>; using D = WinCfg<WIN_DGRAD, 24, 64>;
        L = {A::KC, A::NP, B_::KC, B_::NP, D::KC, D::NP, 0, 0, 0, 0};
    } else if (d.in_ch == 1 && d.c1 == 6 && d.c2 == 16) {
        using A = WinCfg<WIN_FWD1, 1, 6>; using B_ = WinCfg<WIN_FWD2, 6, 16>; using D = WinCfg<WIN_DGRAD, 6, 16>;
        L = {A::KC, A::NP, B_::KC, B_::NP, D::KC, D::NP, 0, 0, 0, 0};
    } else {
        return false;
    }
    L.o1 = 0;
    L.o2 = L.o1 + 5l * L.kc1 * L.np1;
    L.od = L.o2 + 25l * L.kc2 * L.np2;
    L.total = L.od + 25l * L.kcd * L.npd;
    return true;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" long b200dp_tcw_pack_floats(b200dp_dims d) {
    PackLayout L;
    return pack_layout(d, L) ? L.total : -1;
}

extern "C" int b200dp_tcw_pack_weights(const float* w1, const float* w2, float* wpack, b200dp_dims d, b200dp_stream_t s) {
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    PackArgs a{w1, w2, wpack + L.o1, wpack + L.o2, wpack + L.od, d.in_ch, d.c1, d.c2, L.kc1, L.np1, L.kc2, L.np2, L.kcd, L.npd};
    const int grid = (int)((L.total + 255) / 256);
    win_pack_kernel<<<grid < 296 ? grid : 296, 256, 0, (cudaStream_t)s>>>(a);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int b200dp_tcw_conv1_fwd(const void* x, int in_is_u8, float mean, float std, const float* wpack, const float* b,
                                    float* p1, uint8_t* idx1, b200dp_dims d, b200dp_stream_t s) {
    B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    WinArgs a{};
    a.x = x; a.wpack = wpack + L.o1; a.bias = b; a.out = p1; a.idx = idx1;
    a.B = d.batch; a.in_is_u8 = in_is_u8; a.in_hw = d.in_hw; a.out_img_stride = d.c1 * 196; a.mean = mean; a.std = std;
    if (d.in_ch == 3) return win_launch<WIN_FWD1, 3, 24>(a, (cudaStream_t)s);
    return win_launch<WIN_FWD1, 1, 6>(a, (cudaStream_t)s);
}

extern "C" int b200dp_tcw_conv2_fwd(const float* p1, const float* wpack, const float* b, float* p2, uint8_t* idx2,
                                    b200dp_dims d, b200dp_stream_t s) {
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    WinArgs a{};
    a.x = p1; a.wpack = wpack + L.o2; a.bias = b; a.out = p2; a.idx = idx2;
    a.B = d.batch; a.out_img_stride = d.ld_p2 > 0 ? d.ld_p2 : d.c2 * 25;
    if (d.c1 == 24) return win_launch<WIN_FWD2, 24, 64>(a, (cudaStream_t)s);
    return win_launch<WIN_FWD2, 6, 16>(a, (cudaStream_t)s);
}

extern "C" int b200dp_tcw_conv2_dgrad(const float* dp2, const float* p2, const uint8_t* idx2, const float* wpack,
                                      const float* p1, float* g1, b200dp_dims d, b200dp_stream_t s) {
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    WinArgs a{};
    a.x = dp2; a.wpack = wpack + L.od; a.out = g1; a.act = p2; a.pidx = idx2; a.mask = p1;
    a.B = d.batch; a.ld_act = d.ld_p2 > 0 ? d.ld_p2 : d.c2 * 25;
    if (d.c1 == 24) return win_launch<WIN_DGRAD, 24, 64>(a, (cudaStream_t)s);
    return win_launch<WIN_DGRAD, 6, 16>(a, (cudaStream_t)s);
}
