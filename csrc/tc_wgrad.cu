// Convolution weight gradients on tcgen05 (kind::f16, bf16 operands, fp32 accumulation in tensor memory) with both
// operands read in place (MN-major, no swizzle).  (MN-major without swizzle is a 16-bit layout: with tf32 operands the
// tensor core only accepts the SWIZZLE_128B_BASE32B form, whose 4-row swizzle phase a tap shift would break.)
//
//   dW[oc][ic][ky][kx] = sum_img sum_pos dZ[img][oc][pos] * X[img][ic][pos + ky*PITCH + kx],   db[oc] = sum dZ
//
// is a GEMM whose K dimension is the flat output position.  With the activations staged as
// [channel group of 8][position][8 bf16] an MN-major no-swizzle descriptor reads them directly: the
// 16-byte chunk is 8 channels of one position, 8 consecutive positions are one 8-row K group (LBO = 128 B),
// the next channel group is SBO = (positions in the buffer) * 16 B away — and shifting the K origin by a tap
// is a 16-byte-granular start-address change.  So:
//   A (M = 128 rows)  "expanded" input channels c' -> X[plane(c')][pos + off(c')], plus one all-ones channel whose
//                     accumulator row is the bias gradient
//        layer 2: c' = (ic,kx)      (CIN*5+1 rows), the 5 ky taps are 5 start addresses, 5 accumulators
//        layer 1: c' = (ic,ky,kx)   (CIN*25+1 rows), one accumulator (CIN is 1 or 3: everything fits in M)
//   B (N = COUT cols) the un-pooled output gradient dZ (pool-bwd scatter + LeakyReLU' fused into the staging),
//                     zero where no gradient flows, so the garbage columns of the PITCH-wide map add nothing
// Accumulators stay in tensor memory across all images of the CTA; each CTA then writes ONE partial row that
// b200dp_reduce_update sums in fixed order (deterministic, no atomics).
//
// Replaces the weight/bias-gradient half of the autograd backward of cnn.py:10-15 (train.py:49; cuDNN wgrad).
//
// CTA: warps 0-15 stage (dZ once per image, the expanded input per position chunk, double-buffered), warp 16 issues
// the MMAs (uniform control flow, one elected lane), warps 0-3 run the epilogue.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace b200dp {

constexpr int WG_NSTAGE = 16;                        // stager warps (latency-bound work: more warps = more in flight)
constexpr int WG_STAGERS = 32 * WG_NSTAGE;
constexpr int WG_THREADS = WG_STAGERS + 32;          // + the MMA-issuing warp

template <int LAYER, int CIN, int COUT>
struct WgCfg {
    static constexpr bool FIRST = LAYER == 1;
    static constexpr int PITCH = FIRST ? 32 : 14, PLANE = PITCH * PITCH;
    static constexpr int OW = PITCH - 4, PW = OW / 2, PP = PW * PW;
    static constexpr int NTAP = FIRST ? 1 : 5;                       // accumulators (ky taps kept as address shifts)
    static constexpr int CA = FIRST ? CIN * 25 + 1 : CIN * 5 + 1;    // expanded channels + the ones channel
    static constexpr int KGA = (CA + 7) / 8;                          // groups of 8 bf16 channels (16 bytes)
    static_assert(CA <= 128, "expanded channels must fit the 128 accumulator rows");
    static constexpr int NP = (COUT + 15) / 16 * 16, KGB = NP / 8;
    static constexpr int POS16 = ((OW - 1) * PITCH + OW + 15) / 16 * 16;  // K extent of one image (16 positions per MMA)
    // positions per chunk.  The chunk pipeline is a round trip per chunk (stage -> MMAs -> completion frees the buffer) and
    // with two buffers that round trip, not the staging work, sets the pace: 256-position chunks where shared memory
    // allows two such buffers (single input channel) instead of 128 took the reference LeNet's conv1 wgrad from 11.8 us down
    static constexpr int P = FIRST ? (CIN == 1 ? 256 : 128) : POS16;
    static constexpr int NPOS = (POS16 + P - 1) / P * P;              // B holds whole chunks: no MMA reads past its end
    static constexpr int NCHUNK = NPOS / P;
    static constexpr int TAP_SPAN = FIRST ? 0 : 4 * PITCH;
    static constexpr int PA = P + TAP_SPAN;                           // positions held by the A buffer
    static constexpr int A_BYTES = 16 * PA * 16;                      // M = 128 addresses 16 channel groups of 8
    static constexpr int B_BYTES = KGB * NPOS * 16;                   // the whole image's output gradient (staged once per image)
    static constexpr int MAXOFF = 4 * PITCH + 4;
    static constexpr int PLANE_PAD = (NCHUNK * P + TAP_SPAN + MAXOFF + 8 + 3) / 4 * 4;
    static constexpr int PLANES_BYTES = CIN * PLANE_PAD * 4 + (FIRST ? CIN * 1024 * 4 : 0);
    static constexpr int FIXED = 1024 + PLANES_BYTES + 64 + 128 * 4;   // + offset table of the expanded channels
    static constexpr int NBUF = (FIXED + 2 * (A_BYTES + B_BYTES) <= 232448) ? 2 : 1;
    static constexpr size_t SMEM = FIXED + (size_t)NBUF * (A_BYTES + B_BYTES);
    static_assert(SMEM <= 232448, "shared memory");
    static constexpr int ACC_COLS = NTAP * NP;
    static constexpr int TMEM_COLS = ACC_COLS <= 32 ? 32 : (ACC_COLS <= 64 ? 64 : (ACC_COLS <= 128 ? 128 : (ACC_COLS <= 256 ? 256 : 512)));
    static_assert(ACC_COLS <= 512, "tensor memory");
    // expanded channel -> (plane, offset inside the plane)
    __host__ __device__ static constexpr int plane_of(int c) { return FIRST ? c / 25 : c / 5; }
    __host__ __device__ static constexpr int off_of(int c) {
        return FIRST ? ((c % 25) / 5) * PITCH + (c % 5) : c % 5;
    }
};

struct WgArgs {
    const void* x;            // layer 1: raw images [B][CIN][hw][hw] u8|f32;  layer 2: p1 [B][CIN][196]
    int in_is_u8, in_hw; float mean, std;
    const float* g;           // [B][COUT][PP] gradient at the pooled cells
    const float* act; int ld_act;   // pooled activation for LeakyReLU' (nullable: g already carries it)
    const uint8_t* idx;       // [B][COUT][PP] argmax inside the pool window
    float* partials; int partial_stride, off_w, off_b;
    int B, img_per_cta;
    long long* dbg;           // profiling aid (nullable): CTA 0 writes clock64() stamps to dbg[0..16)
};
static long long* g_wg_dbg = nullptr;
#define WG_STAMP(i) do { if (a.dbg && blockIdx.x == 0 && (threadIdx.x & 31) == 0) a.dbg[i] = clock64(); } while (0)

template <int LAYER, int CIN, int COUT>
__global__ void __launch_bounds__(WG_THREADS, 1) wg_conv_kernel(const __grid_constant__ WgArgs a) {
    using C = WgCfg<LAYER, CIN, COUT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    uint8_t* sAB = base;                                               // [NBUF] A: [16][PA][8 bf16], then [NBUF] B: [KGB][NPOS][8 bf16]
    float* sPl = reinterpret_cast<float*>(base + (size_t)C::NBUF * (C::A_BYTES + C::B_BYTES));   // [CIN][PLANE_PAD]
    float* sRaw = sPl + CIN * C::PLANE_PAD;                            // layer 1 with 28x28 input: raw planes
    uint64_t* bar_mem = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sPl) + C::PLANES_BYTES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_mem + 5);
    int* sOff = reinterpret_cast<int*>(bar_mem + 8);                   // [128] expanded channel -> offset into sPl (-1 ones, -2 zero)
    const uint32_t bars = smem_u32(bar_mem);
    auto staged = [&](int b) { return bars + 8u * b; };                // stagers -> MMA warp (count 256)
    auto drained = [&](int b) { return bars + 8u * (2 + b); };         // MMA warp -> stagers (tcgen05.commit)
    const uint32_t acc_bar = bars + 8u * 4;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int img0 = blockIdx.x * a.img_per_cta;
    const int nimg = max(0, min(a.img_per_cta, a.B - img0));
    const int nchunks = nimg * C::NCHUNK;
    if (tid == 0) WG_STAMP(0);

    if (tid < 128) sOff[tid] = tid < C::CA - 1 ? C::plane_of(tid) * C::PLANE_PAD + C::off_of(tid) : (tid == C::CA - 1 ? -1 : -2);
    if (tid == 0) {
        for (int b = 0; b < 2; ++b) { mbar_init(staged(b), WG_STAGERS); mbar_init(drained(b), 1); }
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == WG_NSTAGE) tmem_alloc(smem_u32(tmem_slot), C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (tid == 0) WG_STAMP(1);
    pdl_wait();      // set-up above overlaps the previous kernel's tail (common.cuh)

    if (warp < WG_NSTAGE) {
        // ================= stagers =================
        for (int ch = 0; ch < nchunks; ++ch) {
            const int ii = ch / C::NCHUNK, ck = ch - ii * C::NCHUNK, img = img0 + ii;
            const int b = C::NBUF == 2 ? (ch & 1) : 0;
            if (ck == 0) {
                // ---- the image's planes (normalised / up-sampled for the first layer), zero tail ----
                asm volatile("bar.sync 1, %0;" ::"n"(WG_STAGERS) : "memory");      // previous image's expansion reads are done
                if constexpr (C::FIRST) {
                    const int hw = a.in_hw, n_in = hw * hw;
                    if (hw == 32) {
                        for (int i = tid; i < CIN * 1024; i += WG_STAGERS)
                            sPl[(i >> 10) * C::PLANE_PAD + (i & 1023)] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * 1024 + i, a.mean, a.std);
                    } else {
                        for (int i = tid; i < CIN * n_in; i += WG_STAGERS)
                            sRaw[i] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * n_in + i, a.mean, a.std);
                        asm volatile("bar.sync 1, %0;" ::"n"(WG_STAGERS) : "memory");
                        for (int c = 0; c < CIN; ++c) upsample32_plane(sRaw + c * n_in, hw, sPl + c * C::PLANE_PAD, tid, WG_STAGERS);
                    }
                } else {
                    const float* src = (const float*)a.x + (long)img * CIN * C::PLANE;
                    for (int i = tid; i < CIN * C::PLANE; i += WG_STAGERS) {
                        const int c = i / C::PLANE;
                        sPl[c * C::PLANE_PAD + (i - c * C::PLANE)] = __ldg(src + i);
                    }
                }
                for (int i = tid; i < CIN * (C::PLANE_PAD - C::PLANE); i += WG_STAGERS) {
                    const int c = i / (C::PLANE_PAD - C::PLANE);
                    sPl[c * C::PLANE_PAD + C::PLANE + (i - c * (C::PLANE_PAD - C::PLANE))] = 0.f;
                }
                asm volatile("bar.sync 1, %0;" ::"n"(WG_STAGERS) : "memory");
            }
            if (tid == 0 && ch == 0) WG_STAMP(2);                       // first image's planes ready
            // ---- wait until the tensor core has finished reading this buffer (chunk ch - NBUF) ----
            if (ch >= C::NBUF) mbar_wait(drained(b), ((ch / C::NBUF) - 1) & 1);
            const int ib = C::NBUF == 2 ? (ii & 1) : 0;                  // B buffer of this image
            uint4* sA = reinterpret_cast<uint4*>(sAB + (size_t)b * C::A_BYTES);
            uint4* sB = reinterpret_cast<uint4*>(sAB + (size_t)C::NBUF * C::A_BYTES + (size_t)ib * C::B_BYTES);
            const int p0 = ck * C::P;
            if (ck == 0) {
                // B, once per image: zero, then scatter every pooled gradient to its argmax position (pool-bwd) with
                // LeakyReLU' applied.  (The wait above covers this buffer too: chunk ch-NBUF is later than every chunk
                // of the image that last used it.)
                for (int i = tid; i < C::KGB * C::NPOS; i += WG_STAGERS) sB[i] = make_uint4(0u, 0u, 0u, 0u);
                asm volatile("bar.sync 1, %0;" ::"n"(WG_STAGERS) : "memory");
                __nv_bfloat16* sBh = reinterpret_cast<__nv_bfloat16*>(sB);
                const long o0 = (long)img * COUT * C::PP;
#pragma unroll 4
                for (int i = tid; i < COUT * C::PP; i += WG_STAGERS) {
                    const int oc = i / C::PP, rem = i - oc * C::PP;
                    const int py = rem / C::PW, px = rem - py * C::PW;
                    const int k = a.idx[o0 + i];
                    float g = a.g[o0 + i];
                    if (a.act) g *= leaky_grad_from_act(a.act[(long)img * a.ld_act + i]);
                    const int pos = (2 * py + (k >> 1)) * C::PITCH + 2 * px + (k & 1);
                    sBh[((oc >> 3) * C::NPOS + pos) * 8 + (oc & 7)] = __float2bfloat16_rn(g);
                }
            }
            if (tid == 0 && ch == 0) WG_STAMP(3);                       // first image's dZ staged
            // A, per chunk: work item = (channel group, block of 32 positions), dealt round-robin to the 8 warps;
            // lanes run over positions (conflict-free 16-byte stores)
            constexpr int PBLK = (C::PA + 31) / 32;
            for (int item = warp; item < C::KGA * PBLK; item += WG_NSTAGE) {
                const int cg = item / PBLK, p = (item - cg * PBLK) * 32 + lane;
                const int4 o0 = reinterpret_cast<const int4*>(sOff)[cg * 2], o1 = reinterpret_cast<const int4*>(sOff)[cg * 2 + 1];
                if (p < C::PA) {
                    const float* src = sPl + p0 + p;
                    auto tap = [&](int o) { return o >= 0 ? src[o] : (o == -1 ? 1.f : 0.f); };
                    sA[(size_t)cg * C::PA + p] = make_uint4(pack_bf16x2(tap(o0.x), tap(o0.y)), pack_bf16x2(tap(o0.z), tap(o0.w)),
                                                            pack_bf16x2(tap(o1.x), tap(o1.y)), pack_bf16x2(tap(o1.z), tap(o1.w)));
                }
            }
            fence_proxy_async_smem();
            mbar_arrive(staged(b));
            if (tid == 0 && ch < 2) WG_STAMP(4 + ch);                  // chunk 0 / 1 staged
        }
        if (tid == 0) WG_STAMP(6);                                     // everything staged
    } else {
        // ================= MMA issuer (warp 8): uniform loop, one elected lane =================
        const bool leader = elect_one();
        const uint32_t idesc = idesc_bf16_m128(C::NP, 1, 1);            // bf16 x bf16 -> f32, A and B MN-major
        for (int ch = 0; ch < nchunks; ++ch) {
            const int b = C::NBUF == 2 ? (ch & 1) : 0;
            mbar_wait(staged(b), (ch / C::NBUF) & 1);
            tc_fence_after();
            const int ii = ch / C::NCHUNK, ck = ch - ii * C::NCHUNK;
            const uint32_t sa = smem_u32(sAB + (size_t)b * C::A_BYTES);
            const uint32_t sb = smem_u32(sAB + (size_t)C::NBUF * C::A_BYTES + (size_t)(C::NBUF == 2 ? (ii & 1) : 0) * C::B_BYTES);
            const uint64_t adesc = nosw_desc(sa, 128u, C::PA * 16u);      // LBO: next 8 positions; SBO: next channel group
            const uint64_t bdesc = nosw_desc(sb, 128u, C::NPOS * 16u) + (uint64_t)(ck * C::P);
            if (leader) {
#pragma unroll 1
                for (int j = 0; j < C::P / 16; ++j) {           // 16 positions (two 8-row K groups, LBO apart) per MMA
#pragma unroll
                    for (int t = 0; t < C::NTAP; ++t)
                        umma_bf16(tmem + (uint32_t)(t * C::NP), adesc + (uint64_t)(t * C::PITCH + 16 * j), bdesc + (uint64_t)(16 * j),
                                  idesc, (ch | j) != 0);
                }
                umma_commit(drained(b));
                if (ch == nchunks - 1) umma_commit(acc_bar);
            }
            __syncwarp();
            if (ch == 0) WG_STAMP(9);
        }
        WG_STAMP(10);
    }

    // ================= epilogue: warps 0-3 read the accumulators (thread = row = expanded channel), 16 columns at a
    // time, into shared memory (the operand buffers are idle now); all 16 warps then write the partial row in the
    // arena's [oc][ic][ky][kx] order with consecutive lanes on consecutive addresses.  (Storing straight from the
    // accumulator layout puts a warp's 32 lanes on ~7 different lines per instruction; that was half of this kernel.) ====
    if (warp < WG_NSTAGE) {
        float* row = a.partials + (long)blockIdx.x * a.partial_stride;
        if (nchunks > 0) {
            mbar_wait(acc_bar, 0);
            tc_fence_after();
        }
        pdl_trigger();
        if (tid == 0) WG_STAMP(7);
        constexpr int KW = CIN * 25, TP = 17;
        float* sT = reinterpret_cast<float*>(sAB);                      // [NTAP][128][TP]
        const int m = (warp & 3) * 32 + lane;
#pragma unroll 1
        for (int n0 = 0; n0 < C::NP; n0 += 16) {
            if (warp < 4) {
#pragma unroll 1
                for (int t = 0; t < C::NTAP; ++t) {
                    uint32_t v[16];
                    if (nchunks > 0) {
                        tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(t * C::NP + n0), v);
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = 0u;
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) sT[(t * 128 + m) * TP + j] = __uint_as_float(v[j]);
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(WG_STAGERS) : "memory");
            for (int i = tid; i < 16 * KW; i += WG_STAGERS) {
                const int n = i / KW, k = i - n * KW;
                if (n0 + n < COUT) {
                    int t, mm;
                    if constexpr (C::FIRST) { t = 0; mm = k; }
                    else { const int ic = k / 25, r = k - ic * 25; t = r / 5; mm = ic * 5 + (r - t * 5); }
                    row[a.off_w + (n0 + n) * KW + k] = sT[(t * 128 + mm) * TP + n];
                }
            }
            if (tid < 16 && n0 + tid < COUT) row[a.off_b + n0 + tid] = sT[(C::CA - 1) * TP + tid];   // ones row, tap 0
            asm volatile("bar.sync 1, %0;" ::"n"(WG_STAGERS) : "memory");
        }
    }
    if (tid == 0) WG_STAMP(8);
    tc_fence_before();
    __syncthreads();
    if (warp == WG_NSTAGE) {
        tc_fence_after();
        tmem_dealloc(tmem, C::TMEM_COLS);
    }
}

template <int LAYER, int CIN, int COUT>
static int wg_launch(const WgArgs& a, int rows, cudaStream_t s) {
    using C = WgCfg<LAYER, CIN, COUT>;
    auto kern = wg_conv_kernel<LAYER, CIN, COUT>;
    static DeviceOnce attr;            // one per kernel instantiation, tracks the devices already set up
    B200DP_CHECK(attr.max_dynamic_smem(kern, (int)C::SMEM));
    WgArgs b = a;
    b.dbg = g_wg_dbg;
    B200DP_CHECK(launch_pdl(kern, dim3(rows), dim3(WG_THREADS), C::SMEM, s, b));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b200dp

using namespace b200dp;

// Profiling aid (see b200dp_tcw_set_debug): stamps 0 start, 1 set-up, 2 planes of image 0, 3 dZ of image 0, 4/5 chunk 0/1
// staged, 6 all staged, 7 accumulators ready, 8 epilogue done, 9 first chunk's MMAs issued, 10 all issued.
extern "C" void b200dp_tcw_wgrad_set_debug(long long* dbg) { g_wg_dbg = dbg; }

extern "C" int b200dp_tcw_conv_wgrad_num_partials(b200dp_dims d) {
    // two images per CTA: measured against one image per CTA (256 rows) on the reference LeNet, round 2: 73.7 vs 79.7
    // us per step (the kernels get slower, 6.8 -> 10.1 and 11.8 -> 13.6 us: twice the set-ups, twice the partial rows)
    int n = (d.batch + 1) / 2;
    if (n > 128) n = 128;
    return n < 1 ? 1 : n;
}

extern "C" int b200dp_tcw_conv_wgrad(int layer, const void* x, int in_is_u8, float mean, float std, const float* g,
                                     const float* act, int ld_act, const uint8_t* idx, float* partials,
                                     int partial_stride, int off_w, int off_b, b200dp_dims d, b200dp_stream_t s) {
    const int rows = b200dp_tcw_conv_wgrad_num_partials(d);
    const int ipc = (d.batch + rows - 1) / rows;
    WgArgs a{x, in_is_u8, d.in_hw, mean, std, g, act, ld_act, idx, partials, partial_stride, off_w, off_b, d.batch, ipc, nullptr};
    if (layer == 2) {
        if (d.c1 == 24 && d.c2 == 64) return wg_launch<2, 24, 64>(a, rows, (cudaStream_t)s);
        if (d.c1 == 6 && d.c2 == 16) return wg_launch<2, 6, 16>(a, rows, (cudaStream_t)s);
    } else if (layer == 1) {
        B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
        if (d.in_ch == 3 && d.c1 == 24) return wg_launch<1, 3, 24>(a, rows, (cudaStream_t)s);
        if (d.in_ch == 1 && d.c1 == 6) return wg_launch<1, 1, 6>(a, rows, (cudaStream_t)s);
    }
    B200DP_REQUIRE(false, "unsupported layer / channel combination");
}
