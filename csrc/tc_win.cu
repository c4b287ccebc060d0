// "Window-tile" convolutions on tcgen05 (kind::f16: bf16 operands, fp32 accumulation in tensor memory): implicit GEMM
// with NO im2col materialisation.
//
// The activations of an image are staged once in shared memory as [channel group of 8][pixel][8 bf16]
// (16 bytes per pixel per group).  In that layout a tcgen05 K-major, no-swizzle operand descriptor can
// address an accumulator tile of 128 output pixels = 16 image rows x 8 columns DIRECTLY:
//     core matrix (8 rows x 16 B)  = 8 consecutive pixels of one image row, 8 channels   (contiguous 128 B)
//     SBO (next 8 rows of the tile) = one image row further   (PITCH * 16 B)
//     LBO (next 8 channels of K)    = next channel group      (NPIX * 16 B)
// and a filter tap (ky,kx) is nothing but a different start address (+ (ky*PITCH+kx) * 16 B, always
// 16-byte aligned).  So a 5x5 convolution is 25 x (C/16) tcgen05.mma instructions per tile that read
// the SAME staged image; nothing is gathered per K slab (the first generation of these kernels spent
// >95% of its time writing im2col slabs with scalar shared-memory stores: tensor pipe 1-3% busy).
// The per-tap weight panels [kgroup][n][8] come from a pre-packed bf16 copy (b200dp_tcw_pack_weights) through
// a TMA bulk-copy ring (cp.async.bulk + mbarrier complete_tx).  bf16 rather than tf32 because the MMA phases are
// bound by shared-memory operand bandwidth: half the bytes and half the instructions (measured: the dgrad MMA
// phase 34K -> 19K cycles per CTA, the wide step 190 -> 177 us).
//
//   kind FWD1 : first layer. normalise + bilinear 28->32 on the fly; the staged "channels" are
//               (ic,kx) pairs (kx-expanded), so only the 5 ky taps remain; epilogue bias+LeakyReLU+2x2 pool+argmax
//   kind FWD2 : second layer, 25 taps over the pooled first-layer map; same epilogue
//   kind DGRAD: second layer's data gradient: 25 flipped taps over the zero-padded, un-pooled output
//               gradient (pool-bwd and LeakyReLU-bwd fused into the staging); epilogue * LeakyReLU'(p1)
//
// Replaces reference cnn.py:10-15 (Conv2d/LeakyReLU/MaxPool2d: cuDNN + ATen) and their backward (train.py:49).
//
// CTA = G images, 672 threads: warps 0-15 stage the images and later run the epilogue (four warps per tensor-memory
// lane quadrant: a lone warp per scheduler issues a dependent instruction only every ~5 cycles), warps 16-19 issue
// tcgen05.mma (each owns a subset of the accumulators; uniform control flow keeps the descriptors in uniform
// registers, one elected lane issues), warp 20 runs the weight ring.  Epilogue: tcgen05.ld 32x32b (thread = pixel:
// lane l of quadrant q is row 4q + l/8, column l%8 of the tile), a 32 x 16 transpose through shared memory so that
// a lane owns whole 2x2 pool windows, max + first-max argmax, then bias + LeakyReLU on the pooled value.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace b200dp {

constexpr int WIN_NMMA = 4;                          // MMA-issuing warps, each owns the accumulators ai % 4 == m
constexpr int WIN_WORKERS = 16;                      // staging + epilogue warps (4 per tensor-memory lane quadrant)
constexpr int WIN_THREADS = 32 * (WIN_WORKERS + WIN_NMMA + 1);
constexpr int WIN_STAGERS = 32 * 16;

enum { WIN_FWD1 = 0, WIN_FWD2 = 1, WIN_DGRAD = 2 };

template <int KIND, int CIN, int COUT>
struct WinCfg {
    // geometry of the staged map
    static constexpr int PITCH = KIND == WIN_FWD1 ? 32 : (KIND == WIN_FWD2 ? 14 : 18);
    static constexpr int NTAP = KIND == WIN_FWD1 ? 5 : 25;
    // K channels per tap (padded to a multiple of 16) and N columns (padded to a multiple of 16)
    static constexpr int KREAL = KIND == WIN_FWD1 ? CIN * 5 : (KIND == WIN_FWD2 ? CIN : COUT);
    static constexpr int KC = (KREAL + 15) / 16 * 16, KG = KC / 8, KSTEPS = KC / 16;   // bf16: groups of 8 channels, K = 16 per MMA
    static constexpr int NREAL = KIND == WIN_DGRAD ? CIN : COUT;
    static constexpr int NP = (NREAL + 15) / 16 * 16;
    // tiles of 16 rows x 8 columns covering the valid outputs
    static constexpr int OUT_W = KIND == WIN_FWD1 ? 28 : (KIND == WIN_FWD2 ? 10 : 14);
    static constexpr int NTR = (OUT_W + 15) / 16, NTC = (OUT_W + 7) / 8, NT = NTR * NTC;
    // rows touched: last tile row 16*NTR-1, plus 4 tap rows; columns run over into the next row
    static constexpr int NPIX = (16 * NTR + 4) * PITCH + 16;
    static constexpr int IMG_BYTES = KG * NPIX * 16;
    static constexpr int WB = KC * NP * 2;                       // weight panel of one tap (bf16)
    static constexpr bool WIDE = COUT > 16;
    static constexpr int G = KIND == WIN_FWD1 ? (WIDE ? 2 : 1) : 2;     // images per CTA
    static constexpr int RING = KIND == WIN_FWD1 ? 5 : (KIND == WIN_FWD2 ? (WIDE ? 8 : 25) : (WIDE ? 4 : 25));
    // all taps resident: the panels arrive with ONE bulk copy (25 copies of 512 bytes each paced the MMA issue loop of the
    // reference geometry: the copy unit handles a small copy in ~300 cycles, one after the other)
    static constexpr bool W_ONESHOT = RING >= NTAP;
    static constexpr int ACC_COLS = G * NT * NP;
    static constexpr int TMEM_COLS = ACC_COLS <= 32 ? 32 : (ACC_COLS <= 64 ? 64 : (ACC_COLS <= 128 ? 128 : (ACC_COLS <= 256 ? 256 : 512)));
    static_assert(ACC_COLS <= 512, "accumulators exceed tensor memory");
    static_assert(WB % 16 == 0 && IMG_BYTES % 16 == 0, "alignment");
    // FWD1 scratch: raw planes + upsampled planes (fp32), one image at a time
    static constexpr int PLANE_PAD = 1024 + 8;
    static constexpr int SCRATCH = KIND == WIN_FWD1 ? CIN * (1024 + PLANE_PAD) * 4 : 0;
    static constexpr int POOL_BYTES = KIND == WIN_DGRAD ? 0 : WIN_WORKERS * 16 * 33 * 4;   // epilogue transposes, one per warp
    static constexpr size_t SMEM = 1024 + (size_t)G * IMG_BYTES + (size_t)RING * WB + SCRATCH + POOL_BYTES + 8 * (2 * RING + 2) + 16;
    static_assert(SMEM <= 232448, "shared memory");
    __host__ __device__ static constexpr int tap_shift(int t) {
        return KIND == WIN_FWD1 ? t * PITCH : (t / 5) * PITCH + (t % 5);
    }
};

struct WinArgs {
    const void* x;            // FWD1: raw [B][CIN][hw][hw] u8|f32;  FWD2: p1 [B][CIN][196];  DGRAD: dp2 [B][COUT][25]
    const void* wpack;        // packed bf16 weight panels [NTAP][KG][NP][8]
    const float* bias;        // FWD*: [COUT]
    float* out;               // FWD*: pooled output [B][out_img_stride];  DGRAD: g1 [B][CIN][196]
    uint8_t* idx;             // FWD*: argmax [B][COUT][PW*PW] (nullable)
    const float* act;         // DGRAD: p2 [B][ld_act] (LeakyReLU' of layer 2);  unused otherwise
    const uint8_t* pidx;      // DGRAD: idx2 [B][COUT][25]
    const float* mask;        // DGRAD: p1 [B][CIN][196] (LeakyReLU' of layer 1)
    int B, in_is_u8, in_hw, out_img_stride, ld_act;
    float mean, std;
    long long* dbg;           // profiling aid (nullable): CTA 0 writes clock64() stamps of its phases to dbg[0..16)
};
static long long* g_win_dbg = nullptr;
#define WIN_STAMP(i) do { if (a.dbg && blockIdx.x == 0 && (threadIdx.x & 31) == 0) a.dbg[i] = clock64(); } while (0)

// ---------------------------------------------------------------------------------------------------
// staging (warps 0-15)
// ---------------------------------------------------------------------------------------------------
template <class C, int CIN>
__device__ __forceinline__ void win_stage_fwd2(const WinArgs& a, int img, float* sImg, int tid) {
    // NCHW planes -> [cg][pixel][8 bf16]: one 16-byte store per (pixel, channel group), 8 coalesced plane reads
    const float* src = (const float*)a.x + (long)img * CIN * 196;
#pragma unroll 2
    for (int i = tid; i < C::KG * C::NPIX; i += WIN_STAGERS) {
        const int cg = i / C::NPIX, p = i - cg * C::NPIX;
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = (p < 196 && cg * 8 + q < CIN) ? __ldg(src + (cg * 8 + q) * 196 + p) : 0.f;
        reinterpret_cast<uint4*>(sImg)[i] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                       pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    }
}

template <class C, int CIN>
__device__ __forceinline__ void win_stage_fwd1(const WinArgs& a, int img, float* sImg, float* sRaw, float* sUp, int tid) {
    const int hw = a.in_hw, n_in = hw * hw;
    if (hw == 32) {                      // no resampling: normalise straight into the padded planes
        for (int i = tid; i < CIN * 1024; i += WIN_STAGERS)
            sUp[(i >> 10) * C::PLANE_PAD + (i & 1023)] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * 1024 + i, a.mean, a.std);
    } else {
        for (int i = tid; i < CIN * n_in; i += WIN_STAGERS)
            sRaw[i] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * n_in + i, a.mean, a.std);
        asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");
        for (int c = 0; c < CIN; ++c) upsample32_plane(sRaw + c * n_in, hw, sUp + c * C::PLANE_PAD, tid, WIN_STAGERS);
    }
    if (tid < 8 * CIN) sUp[(tid >> 3) * C::PLANE_PAD + 1024 + (tid & 7)] = 0.f;      // kx overrun of the last pixels
    asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");
    // kx-expansion: staged channel c' = ic*5 + kx holds X[ic][p + kx].  A thread keeps one channel group of 8, so the
    // plane offsets are loop invariants; consecutive lanes take consecutive pixels (conflict-free 16-byte stores).
    constexpr int TPG = WIN_STAGERS / C::KG;                 // threads per channel group
    static_assert(WIN_STAGERS % C::KG == 0, "stagers must split evenly over the channel groups");
    const int cg = tid / TPG, p0 = tid - cg * TPG;
    int off[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const int cc = cg * 8 + q;
        off[q] = cc < CIN * 5 ? (cc / 5) * C::PLANE_PAD + (cc % 5) : -1;
    }
    uint4* dst = reinterpret_cast<uint4*>(sImg) + (size_t)cg * C::NPIX;
#pragma unroll 2
    for (int p = p0; p < C::NPIX; p += TPG) {
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = (p < 1024 && off[q] >= 0) ? sUp[off[q] + p] : 0.f;
        dst[p] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    }
    asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");     // sRaw / sUp are reused by the next image
}

template <class C, int COUT>
__device__ __forceinline__ void win_stage_dgrad(const WinArgs& a, int img, float* sImg, int tid) {
    // zero-padded 18x18 map of the un-pooled gradient: pool-bwd (scatter to the argmax) + LeakyReLU-bwd
    for (int i = tid; i < C::KG * C::NPIX; i += WIN_STAGERS)
        reinterpret_cast<uint4*>(sImg)[i] = make_uint4(0u, 0u, 0u, 0u);
    asm volatile("bar.sync 1, %0;" ::"n"(WIN_STAGERS) : "memory");
    for (int i = tid; i < COUT * 25; i += WIN_STAGERS) {
        const int oc = i / 25, pp = i - oc * 25, py = pp / 5, px = pp - py * 5;
        const int k = a.pidx[(long)img * COUT * 25 + i];
        const float g = ((const float*)a.x)[(long)img * COUT * 25 + i] * leaky_grad_from_act(a.act[(long)img * a.ld_act + i]);
        const int p = (2 * py + (k >> 1) + 4) * C::PITCH + 2 * px + (k & 1) + 4;
        reinterpret_cast<__nv_bfloat16*>(sImg)[((oc >> 3) * C::NPIX + p) * 8 + (oc & 7)] = __float2bfloat16_rn(g);
    }
}

// ---------------------------------------------------------------------------------------------------
template <class C, int M>
__device__ __forceinline__ void win_issue(uint32_t tmem, uint32_t s_img, uint32_t s_w, uint32_t bars, int nimg) {
    const bool leader = elect_one();
    const uint32_t idesc = idesc_bf16_m128(C::NP, 0, 0);
    const uint64_t adesc0 = nosw_desc(s_img, C::NPIX * 16u, C::PITCH * 16u);
    const uint64_t bdesc0 = nosw_desc(s_w, C::NP * 16u, 128u);
    const uint32_t acc_bar = bars + 8u * (2 * C::RING);
#pragma unroll 1
    for (int t = 0; t < C::NTAP; ++t) {
        const int s = t % C::RING;
        if (!C::W_ONESHOT || t == 0) mbar_wait(bars + 8u * s, (t / C::RING) & 1);
        tc_fence_after();
        const uint64_t at = adesc0 + (uint64_t)C::tap_shift(t);                 // the address field counts 16-byte units
        const uint64_t bt = bdesc0 + (uint64_t)(s * (C::WB >> 4));
        const uint32_t acc = t != 0;
        if (leader) {
#pragma unroll
            for (int ai = M; ai < C::G * C::NT; ai += WIN_NMMA) {
                const int g = ai / C::NT, tl = ai % C::NT;
                if (g < nimg) {
                    const int r0 = (tl / C::NTC) * 16, c0 = (tl % C::NTC) * 8;
                    const uint64_t ag = at + (uint64_t)((g * C::IMG_BYTES + (r0 * C::PITCH + c0) * 16) >> 4);
#pragma unroll
                    for (int ks = 0; ks < C::KSTEPS; ++ks)
                        umma_bf16(tmem + (uint32_t)(ai * C::NP), ag + (uint64_t)(ks * 2 * C::NPIX),      // K = 16: two groups of 8
                                  bt + (uint64_t)(ks * 2 * C::NP), idesc, ks != 0 ? 1u : acc);
                }
            }
            umma_commit(bars + 8u * (C::RING + s));
            if (t == C::NTAP - 1) umma_commit(acc_bar);
        }
        __syncwarp();
    }
}

template <int KIND, int CIN, int COUT>
__global__ void __launch_bounds__(WIN_THREADS, 1) win_conv_kernel(const __grid_constant__ WinArgs a) {
    using C = WinCfg<KIND, CIN, COUT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    float* sImg = reinterpret_cast<float*>(base);                                   // [G][KG][NPIX][4]
    uint8_t* sW = base + (size_t)C::G * C::IMG_BYTES;                               // [RING][WB]
    float* sScratch = reinterpret_cast<float*>(sW + (size_t)C::RING * C::WB);
    float* sPool = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(sScratch) + C::SCRATCH);
    uint64_t* bar_mem = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sPool) + C::POOL_BYTES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mem + 2 * C::RING + 1);
    const uint32_t bars = smem_u32(bar_mem);
    auto w_full = [&](int s) { return bars + 8u * s; };
    auto w_empty = [&](int s) { return bars + 8u * (C::RING + s); };
    const uint32_t acc_bar = bars + 8u * (2 * C::RING);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int img0 = blockIdx.x * C::G;
    const int nimg = min(C::G, a.B - img0);
    if (tid == 0) WIN_STAMP(0);

    if (tid == 0) {
        for (int s = 0; s < C::RING; ++s) { mbar_init(w_full(s), 1); mbar_init(w_empty(s), WIN_NMMA); }
        mbar_init(acc_bar, WIN_NMMA);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == WIN_WORKERS) tmem_alloc(smem_u32(tmem_slot), C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (tid == 0) WIN_STAMP(1);
    pdl_wait();      // barrier set-up and TMEM allocation above overlap the previous kernel's tail (common.cuh)

    auto load_tap = [&](int t) {
        const int s = t % C::RING;
        mbar_expect_tx(w_full(s), C::WB);
        tma_bulk_load(smem_u32(sW + (size_t)s * C::WB), (const uint8_t*)a.wpack + (size_t)t * C::WB, C::WB, w_full(s));
    };
    if (warp == WIN_WORKERS + WIN_NMMA) {
        // ================= weight ring producer (one lane): fill the ring while the images are staged =================
        if (lane == 0) {
            if constexpr (C::W_ONESHOT) {
                mbar_expect_tx(w_full(0), C::NTAP * C::WB);
                tma_bulk_load(smem_u32(sW), a.wpack, C::NTAP * C::WB, w_full(0));
            } else {
                for (int t = 0; t < C::RING && t < C::NTAP; ++t) load_tap(t);
            }
        }
    } else if (warp < WIN_WORKERS) {
        // ================= stage the CTA's images =================
        for (int g = 0; g < nimg; ++g) {
            float* dst = sImg + (size_t)g * (C::IMG_BYTES / 4);
            if constexpr (KIND == WIN_FWD1) win_stage_fwd1<C, CIN>(a, img0 + g, dst, sScratch, sScratch + CIN * 1024, tid);
            else if constexpr (KIND == WIN_FWD2) win_stage_fwd2<C, CIN>(a, img0 + g, dst, tid);
            else win_stage_dgrad<C, COUT>(a, img0 + g, dst, tid);
        }
        fence_proxy_async_smem();          // generic-proxy stores -> visible to the tensor core's reads
        if (tid == 0) WIN_STAMP(2);
    }
    __syncthreads();
    if (tid == 0) WIN_STAMP(3);

    if (warp == WIN_WORKERS + WIN_NMMA) {
        if (lane == 0)
            for (int t = C::RING; t < C::NTAP; ++t) {          // refill a stage once its MMAs have completed
                mbar_wait(w_empty(t % C::RING), ((t / C::RING) - 1) & 1);
                load_tap(t);
            }
    } else if (warp >= WIN_WORKERS) {
        // ================= MMA issuers: warp WIN_WORKERS+m owns the accumulators (image g, tile tl) with (g*NT+tl) % 4 == m.
        // The loop is warp-uniform (descriptors live in uniform registers); one elected lane issues. =================
        tc_fence_after();
        const uint32_t sw = smem_u32(sW), si = smem_u32(sImg);
        switch (warp - WIN_WORKERS) {
            case 0: win_issue<C, 0>(tmem, si, sw, bars, nimg); break;
            case 1: win_issue<C, 1>(tmem, si, sw, bars, nimg); break;
            case 2: win_issue<C, 2>(tmem, si, sw, bars, nimg); break;
            default: win_issue<C, 3>(tmem, si, sw, bars, nimg); break;
        }
        if (warp == WIN_WORKERS) WIN_STAMP(4);
    } else {
        // ================= epilogue (16 warps; a lone warp issues a dependent instruction only every ~5 cycles, so the work is spread
        // over four warps per scheduler): warp w reads TMEM lane quadrant w%4 (rows 4q..4q+3 of a tile; thread =
        // pixel: row 4q + lane/8, column lane%8); items are dealt round-robin to the four warps of a quadrant =================
        mbar_wait(acc_bar, 0);
        tc_fence_after();
        pdl_trigger();
        if (tid == 0) WIN_STAMP(5);
        const int q = warp & 3;
        const int lr = 4 * q + (lane >> 3), lc = lane & 7;
        constexpr int NCH = C::NP / 16;                     // work item = (accumulator, 16-column chunk)
#pragma unroll 1
        for (int item = warp >> 2; item < nimg * C::NT * NCH; item += WIN_WORKERS / 4) {
            const int ai = item / NCH, n0 = (item - ai * NCH) * 16;
            const int g = ai / C::NT, tl = ai - g * C::NT;
            const int img = img0 + g;
            {
                const int r = (tl / C::NTC) * 16 + lr, c = (tl % C::NTC) * 8 + lc;
                const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(ai * C::NP);
                {
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
                        // 2x2 max-pool (+ first-max argmax), bias and LeakyReLU.  The warp's 32 pixels x 16 columns go through
                        // shared memory so that a lane owns whole pool windows: lane = (cell e = lane%8 of the 2 x 4 cells of the
                        // warp's 4 x 8 pixels, column group lane/8), 4 columns each.  bias + LeakyReLU are monotonic, so they are
                        // applied after the max (same result, a quarter of the work).  ~4x fewer instructions than pooling
                        // with shuffles in the TMEM register layout, which was what bounded these kernels.
                        constexpr int PW = C::OUT_W / 2, PP = PW * PW;
                        float* sT = sPool + warp * (16 * 33);
#pragma unroll
                        for (int j = 0; j < 16; ++j) sT[j * 33 + lane] = __uint_as_float(v[j]);
                        __syncwarp();
                        const int e = lane & 7, pr = e >> 2, pc = e & 3, cgp = lane >> 3;
                        const int py = (tl / C::NTC) * 8 + 2 * q + pr, px = (tl % C::NTC) * 4 + pc;
                        const bool cell_ok = py < PW && px < PW;
                        const float* src = sT + 16 * pr + 2 * pc;
#pragma unroll
                        for (int jj = 0; jj < 4; ++jj) {
                            const int j = cgp * 4 + jj, n = n0 + j;
                            const float x00 = src[j * 33], x01 = src[j * 33 + 1], x10 = src[j * 33 + 8], x11 = src[j * 33 + 9];
                            float best = x00;
                            int bi = 0;
                            if (x01 > best) { best = x01; bi = 1; }
                            if (x10 > best) { best = x10; bi = 2; }
                            if (x11 > best) { best = x11; bi = 3; }
                            if (cell_ok && n < COUT) {
                                a.out[(long)img * a.out_img_stride + n * PP + py * PW + px] = leaky(best + __ldg(a.bias + n));
                                if (a.idx) a.idx[((long)img * COUT + n) * PP + py * PW + px] = (uint8_t)bi;
                            }
                        }
                        __syncwarp();
                    }
                }
            }
        }
    }
    if (tid == 0) WIN_STAMP(6);
    tc_fence_before();
    __syncthreads();
    if (warp == WIN_WORKERS) {
        tc_fence_after();
        tmem_dealloc(tmem, C::TMEM_COLS);
        WIN_STAMP(7);
    }
}

// ---------------------------------------------------------------------------------------------------
// weight packing: OIHW -> per-tap K-major panels [tap][kgroup][n][4] (zero padded), all three in one launch
// ---------------------------------------------------------------------------------------------------
struct PackArgs {
    const float* w1; const float* w2;
    __nv_bfloat16* pack;      // fwd1 panels, then fwd2, then dgrad
    int in_ch, c1, c2;
    int kc1, np1, kc2, np2, kcd, npd;
};
__global__ void __launch_bounds__(256) win_pack_kernel(PackArgs a) {
    pdl_wait();
    pdl_trigger();
    const int n1 = 5 * a.kc1 * a.np1, n2 = 25 * a.kc2 * a.np2, nd = 25 * a.kcd * a.npd;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n1 + n2 + nd; i += gridDim.x * blockDim.x) {
        float v;
        if (i < n1) {                                   // [ky][kg][oc][8], k = ic*5 + kx
            const int q = i & 7, n = (i >> 3) % a.np1, kg = (i >> 3) / a.np1 % (a.kc1 / 8), ky = (i >> 3) / a.np1 / (a.kc1 / 8);
            const int k = kg * 8 + q, ic = k / 5, kx = k - ic * 5;
            v = (n < a.c1 && k < a.in_ch * 5) ? a.w1[((n * a.in_ch + ic) * 5 + ky) * 5 + kx] : 0.f;
        } else if (i < n1 + n2) {                       // [tap][kg][oc][8], k = ic
            const int e = i - n1;
            const int q = e & 7, n = (e >> 3) % a.np2, kg = (e >> 3) / a.np2 % (a.kc2 / 8), t = (e >> 3) / a.np2 / (a.kc2 / 8);
            const int k = kg * 8 + q;
            v = (n < a.c2 && k < a.c1) ? a.w2[(n * a.c1 + k) * 25 + t] : 0.f;
        } else {                                        // [tap=(u,v)][kg][ic][8], k = oc, flipped filter
            const int e = i - n1 - n2;
            const int q = e & 7, n = (e >> 3) % a.npd, kg = (e >> 3) / a.npd % (a.kcd / 8), t = (e >> 3) / a.npd / (a.kcd / 8);
            const int k = kg * 8 + q;
            v = (n < a.c1 && k < a.c2) ? a.w2[(k * a.c1 + n) * 25 + (24 - t)] : 0.f;
        }
        a.pack[i] = __float2bfloat16_rn(v);             // the three panel sets are contiguous
    }
}

template <int KIND, int CIN, int COUT>
static int win_launch(const WinArgs& a, cudaStream_t s) {
    using C = WinCfg<KIND, CIN, COUT>;
    auto kern = win_conv_kernel<KIND, CIN, COUT>;
    static DeviceOnce attr;            // one per kernel instantiation, tracks the devices already set up
    B200DP_CHECK(attr.max_dynamic_smem(kern, (int)C::SMEM));
    WinArgs b = a;
    b.dbg = g_win_dbg;
    B200DP_CHECK(launch_pdl(kern, dim3((a.B + C::G - 1) / C::G), dim3(WIN_THREADS), C::SMEM, s, b));
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

// Profiling aid: while set, CTA 0 of every window-tile launch writes clock64() stamps to dbg[0..16)
// (0 start, 1 set-up done, 2 staged, 3 after barrier, 4 MMAs issued, 5 accumulators ready, 6 epilogue done, 7 end).
extern "C" void b200dp_tcw_set_debug(long long* dbg) { g_win_dbg = dbg; }

extern "C" long b200dp_tcw_pack_floats(b200dp_dims d) {
    PackLayout L;
    return pack_layout(d, L) ? (L.total + 1) / 2 : -1;      // bf16 elements, counted in floats
}

extern "C" int b200dp_tcw_pack_weights(const float* w1, const float* w2, float* wpack, b200dp_dims d, b200dp_stream_t s) {
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    PackArgs a{w1, w2, reinterpret_cast<__nv_bfloat16*>(wpack), d.in_ch, d.c1, d.c2, L.kc1, L.np1, L.kc2, L.np2, L.kcd, L.npd};
    const int grid = (int)((L.total + 255) / 256);
    B200DP_CHECK(launch_pdl(win_pack_kernel, dim3(grid < 296 ? grid : 296), dim3(256), 0, (cudaStream_t)s, a));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int b200dp_tcw_conv1_fwd(const void* x, int in_is_u8, float mean, float std, const float* wpack, const float* b,
                                    float* p1, uint8_t* idx1, b200dp_dims d, b200dp_stream_t s) {
    B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    WinArgs a{};
    a.x = x; a.wpack = reinterpret_cast<const __nv_bfloat16*>(wpack) + L.o1; a.bias = b; a.out = p1; a.idx = idx1;
    a.B = d.batch; a.in_is_u8 = in_is_u8; a.in_hw = d.in_hw; a.out_img_stride = d.c1 * 196; a.mean = mean; a.std = std;
    if (d.in_ch == 3) return win_launch<WIN_FWD1, 3, 24>(a, (cudaStream_t)s);
    return win_launch<WIN_FWD1, 1, 6>(a, (cudaStream_t)s);
}

extern "C" int b200dp_tcw_conv2_fwd(const float* p1, const float* wpack, const float* b, float* p2, uint8_t* idx2,
                                    b200dp_dims d, b200dp_stream_t s) {
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    WinArgs a{};
    a.x = p1; a.wpack = reinterpret_cast<const __nv_bfloat16*>(wpack) + L.o2; a.bias = b; a.out = p2; a.idx = idx2;
    a.B = d.batch; a.out_img_stride = d.ld_p2 > 0 ? d.ld_p2 : d.c2 * 25;
    if (d.c1 == 24) return win_launch<WIN_FWD2, 24, 64>(a, (cudaStream_t)s);
    return win_launch<WIN_FWD2, 6, 16>(a, (cudaStream_t)s);
}

extern "C" int b200dp_tcw_conv2_dgrad(const float* dp2, const float* p2, const uint8_t* idx2, const float* wpack,
                                      const float* p1, float* g1, b200dp_dims d, b200dp_stream_t s) {
    PackLayout L;
    B200DP_REQUIRE(pack_layout(d, L), "unsupported (in_ch, c1, c2): expected (1,6,16) or (3,24,64)");
    WinArgs a{};
    a.x = dp2; a.wpack = reinterpret_cast<const __nv_bfloat16*>(wpack) + L.od; a.out = g1; a.act = p2; a.pidx = idx2; a.mask = p1;
    a.B = d.batch; a.ld_act = d.ld_p2 > 0 ? d.ld_p2 : d.c2 * 25;
    if (d.c1 == 24) return win_launch<WIN_DGRAD, 24, 64>(a, (cudaStream_t)s);
    return win_launch<WIN_DGRAD, 6, 16>(a, (cudaStream_t)s);
}
