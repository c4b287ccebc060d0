// conv5x5 + bias + LeakyReLU(0.2) + maxpool2x2 (+argmax) forward, CUDA-core path.
//
// Replaces reference cnn.py:9-15 (UpsamplingBilinear2d, Conv2d, LeakyReLU, MaxPool2d x2) and, for the
// first stage, the host transform of train.py:85.  One launch per stage; the conv output never
// reaches memory un-pooled.  A thread owns one pooled cell (2x2 conv outputs) for OCT output
// channels: it keeps the 6x6 input patch in registers and streams weights from shared memory
// (warp-broadcast reads).  The tcgen05 path for the wide model lives in tc_win.cu / tc_wgrad.cu.
#include "common.cuh"

namespace b200dp {

template <int CIN, int COUT, int IN_W, int OCT, int GPB, int IPB, bool FIRST>
struct ConvFwdCfg {
    static constexpr int PW = (IN_W - 4) / 2;
    static constexpr int PP = PW * PW;
    static constexpr int COUT_BLK = OCT * GPB;
    static constexpr int WORK = IPB * GPB * PP;
    static constexpr int NT = (WORK + 31) / 32 * 32;
    static constexpr int PLANE = IN_W * IN_W;
    static constexpr int SMEM_FLOATS =
        IPB * CIN * PLANE + CIN * 25 * COUT_BLK + COUT_BLK + (FIRST ? CIN * PLANE : 0);
    static_assert(COUT % COUT_BLK == 0, "COUT must be a multiple of OCT*GPB");
    static_assert(!FIRST || IPB == 1, "first stage handles one image per CTA");
};

struct ConvFwdArgs {
    const void* x;      // FIRST: raw [B][CIN][hw][hw] u8|f32 ; else float [B][CIN][IN_W][IN_W]
    const float* w;     // [COUT][CIN][25]
    const float* bias;  // [COUT]
    float* out;         // [B][COUT][PW][PW]
    uint8_t* idx;       // same shape, nullable
    int B, in_is_u8, in_hw;
    float mean, std;
    int out_img_stride;  // floats between consecutive images of `out` (>= COUT*PP)
};

template <int CIN, int COUT, int IN_W, int OCT, int GPB, int IPB, bool FIRST>
__global__ void __launch_bounds__(ConvFwdCfg<CIN, COUT, IN_W, OCT, GPB, IPB, FIRST>::NT)
conv5_pool_fwd_kernel(ConvFwdArgs a) {
    pdl_wait();
    using C = ConvFwdCfg<CIN, COUT, IN_W, OCT, GPB, IPB, FIRST>;
    extern __shared__ __align__(16) float smem[];
    float* sIn = smem;                                   // [IPB][CIN][PLANE]
    float* sW = sIn + IPB * CIN * C::PLANE;              // [CIN*25][COUT_BLK]
    float* sB = sW + CIN * 25 * C::COUT_BLK;             // [COUT_BLK]
    float* sRaw = sB + C::COUT_BLK;                      // FIRST only: [CIN][hw*hw]

    const int tid = threadIdx.x;
    const int img0 = blockIdx.x * IPB;
    const int oc0 = blockIdx.y * C::COUT_BLK;

    // ---- stage weights (transposed so a thread's OCT channels are contiguous) ----
    for (int i = tid; i < CIN * 25 * C::COUT_BLK; i += C::NT) {
        int ocl = i / (CIN * 25), r = i - ocl * (CIN * 25);
        sW[r * C::COUT_BLK + ocl] = a.w[(long)(oc0 + ocl) * (CIN * 25) + r];
    }
    if (tid < C::COUT_BLK) sB[tid] = a.bias[oc0 + tid];

    // ---- stage input ----
    if constexpr (FIRST) {
        const int hw = a.in_hw, n_in = hw * hw;
        const int img = img0;  // IPB == 1
        for (int i = tid; i < CIN * n_in; i += C::NT)
            sRaw[i] = load_pixel(a.x, a.in_is_u8, (long)img * CIN * n_in + i, a.mean, a.std);
        __syncthreads();
        for (int c = 0; c < CIN; ++c) upsample32_plane(sRaw + c * n_in, hw, sIn + c * C::PLANE, tid, C::NT);
    } else {
        const float* xin = (const float*)a.x;
        const int nimg = min(IPB, a.B - img0);
        const float4* src = reinterpret_cast<const float4*>(xin + (long)img0 * CIN * C::PLANE);
        float4* dst = reinterpret_cast<float4*>(sIn);
        static_assert((CIN * C::PLANE) % 4 == 0, "plane must be float4-divisible");
        for (int i = tid; i < nimg * CIN * C::PLANE / 4; i += C::NT) dst[i] = src[i];
    }
    __syncthreads();

    // ---- compute ----
    const int pos = tid % C::PP;
    const int rest = tid / C::PP;
    const int grp = rest % GPB;
    const int il = rest / GPB;
    const int img = img0 + il;
    if (il >= IPB || img >= a.B) return;
    const int py = pos / C::PW, px = pos - py * C::PW;

    float acc[OCT][4];
#pragma unroll
    for (int o = 0; o < OCT; ++o) { acc[o][0] = acc[o][1] = acc[o][2] = acc[o][3] = 0.f; }

#pragma unroll 1
    for (int ic = 0; ic < CIN; ++ic) {
        const float* in = sIn + (il * CIN + ic) * C::PLANE + (2 * py) * IN_W + 2 * px;
        float patch[6][6];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
#pragma unroll
            for (int c2 = 0; c2 < 3; ++c2) {
                float2 v = *reinterpret_cast<const float2*>(in + r * IN_W + 2 * c2);
                patch[r][2 * c2] = v.x;
                patch[r][2 * c2 + 1] = v.y;
            }
        }
        const float* wp = sW + (ic * 25) * C::COUT_BLK + grp * OCT;
#pragma unroll
        for (int ky = 0; ky < 5; ++ky) {
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
#pragma unroll
                for (int o = 0; o < OCT; ++o) {
                    float wv = wp[(ky * 5 + kx) * C::COUT_BLK + o];
                    acc[o][0] = fmaf(wv, patch[ky][kx], acc[o][0]);
                    acc[o][1] = fmaf(wv, patch[ky][kx + 1], acc[o][1]);
                    acc[o][2] = fmaf(wv, patch[ky + 1][kx], acc[o][2]);
                    acc[o][3] = fmaf(wv, patch[ky + 1][kx + 1], acc[o][3]);
                }
            }
        }
    }

    pdl_trigger();   // late: the next kernel's CTAs become resident while this grid stores its results
    // ---- epilogue: bias + LeakyReLU + 2x2 max (first max wins, like torch) ----
#pragma unroll
    for (int o = 0; o < OCT; ++o) {
        const int oc = oc0 + grp * OCT + o;
        const float bv = sB[grp * OCT + o];
        float best = leaky(acc[o][0] + bv);
        int bi = 0;
#pragma unroll
        for (int k = 1; k < 4; ++k) {
            float v = leaky(acc[o][k] + bv);
            if (v > best) { best = v; bi = k; }
        }
        const long off = ((long)img * COUT + oc) * C::PP + pos;
        a.out[(long)img * a.out_img_stride + oc * C::PP + pos] = best;
        if (a.idx) a.idx[off] = (uint8_t)bi;
    }
}

template <int CIN, int COUT, int IN_W, int OCT, int GPB, int IPB, bool FIRST>
static int launch_conv_fwd(const ConvFwdArgs& a, cudaStream_t s) {
    using C = ConvFwdCfg<CIN, COUT, IN_W, OCT, GPB, IPB, FIRST>;
    auto kern = conv5_pool_fwd_kernel<CIN, COUT, IN_W, OCT, GPB, IPB, FIRST>;
    const size_t smem = C::SMEM_FLOATS * sizeof(float);
    static DeviceOnce attr;         // one per kernel instantiation
    if (smem > 48 * 1024) B200DP_CHECK(attr.max_dynamic_smem(kern, (int)smem));
    dim3 grid((a.B + IPB - 1) / IPB, COUT / C::COUT_BLK);
    B200DP_CHECK(launch_pdl(kern, dim3(grid), dim3(C::NT), smem, s, a));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_conv1_fwd(const void* x, int in_is_u8, float mean, float std, const float* w, const float* b,
                                float* p1, uint8_t* idx1, b200dp_dims d, b200dp_stream_t s) {
    B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
    ConvFwdArgs a{x, w, b, p1, idx1, d.batch, in_is_u8, d.in_hw, mean, std, d.c1 * 196};
    if (d.in_ch == 1 && d.c1 == 6) return launch_conv_fwd<1, 6, 32, 3, 2, 1, true>(a, (cudaStream_t)s);
    if (d.in_ch == 3 && d.c1 == 24) return launch_conv_fwd<3, 24, 32, 4, 2, 1, true>(a, (cudaStream_t)s);
    B200DP_REQUIRE(false, "unsupported (in_ch, c1): expected (1,6) or (3,24)");
}

extern "C" int b200dp_conv2_fwd(const float* p1, const float* w, const float* b, float* p2, uint8_t* idx2,
                                b200dp_dims d, b200dp_stream_t s) {
    ConvFwdArgs a{p1, w, b, p2, idx2, d.batch, 0, 14, 0.f, 1.f, d.ld_p2 > 0 ? d.ld_p2 : d.c2 * 25};
    B200DP_REQUIRE(a.out_img_stride >= d.c2 * 25, "ld_p2 smaller than c2*25");
    if (d.c1 == 6 && d.c2 == 16) return launch_conv_fwd<6, 16, 14, 4, 4, 1, false>(a, (cudaStream_t)s);
    if (d.c1 == 24 && d.c2 == 64) return launch_conv_fwd<24, 64, 14, 4, 4, 2, false>(a, (cudaStream_t)s);
    B200DP_REQUIRE(false, "unsupported (c1, c2): expected (6,16) or (24,64)");
}
