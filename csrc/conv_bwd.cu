// Fused convolution backward for the LeNet-sized model, CUDA-core path: ONE launch per step does
//   pool2-bwd + LeakyReLU-bwd -> conv2 dgrad -> pool1-bwd + LeakyReLU-bwd,
//   conv2 wgrad + bgrad (exploiting that only 1 of 4 dZ2 entries per pool window is non-zero),
//   conv1 wgrad + bgrad (the 32x32 input is recomputed from the raw image, never stored).
// conv1 dgrad is never needed (the input requires no gradient).
//
// Replaces the autograd backward of reference cnn.py:9-15 (train.py:49): ~12 ATen/cuDNN kernels.
//
// A CTA owns images cta, cta+grid, ... and keeps its weight-gradient partial sums in registers;
// at the end it writes ONE row of `partials` (the scratch buffer; layout = the arena layout of
// the conv bucket, so column i of a row is arena element bucket_start+i).  Rows are summed in
// fixed order by b200dp_reduce_update -> bit-reproducible gradients, no atomics.
#include "common.cuh"

namespace b200dp {

constexpr int CB_NT = 256;
constexpr int CB_MAX_PARTIALS = 256;
constexpr int ZP = 18;  // 10x10 conv2-output map zero-padded by 4 on each side

struct ConvBwdArgs {
    const void* x; int in_is_u8; int in_hw; float mean, std;
    const float* p1; const uint8_t* idx1;
    const float* p2; int ld_p2; const uint8_t* idx2; const float* dp2;
    const float* w2;
    float* partials; int partial_stride;
    int off_w1, off_b1, off_w2, off_b2;
    int B;
};

template <int C0, int C1, int C2>
struct ConvBwdCfg {
    static constexpr int W2N = C2 * C1 * 25;
    static constexpr int W1N = C1 * C0 * 25;
    static constexpr int W2_PER_THREAD = (W2N + CB_NT - 1) / CB_NT;
    static constexpr int W1_PER_THREAD = (W1N + CB_NT - 1) / CB_NT;
    static constexpr int OG = 4;                       // oc groups in the dgrad phase
    static constexpr int C1P = (C1 + 1) / 2 * 2;        // ic padded to even for float2 weight reads
    static_assert(C2 % OG == 0, "C2 must split into 4 groups");
    static_assert(49 * OG <= CB_NT, "dgrad mapping needs 196 threads");
    static_assert(C1 <= CB_NT / 32, "one warp per conv1 channel for the bias gradient");
    // shared memory carve-up (floats unless noted)
    static constexpr int F_X0 = C0 * 1024, F_RAW = C0 * 1024, F_P1 = C1 * 196, F_DZ2 = C2 * ZP * ZP,
                         F_W2F = C2 * 25 * C1P, F_PART = OG * C1 * 196, F_G1 = C1 * 196, F_G2 = C2 * 25,
                         I_POS2 = C2 * 25;
    static constexpr size_t SMEM = (size_t)(F_X0 + F_RAW + F_P1 + F_DZ2 + F_W2F + F_PART + F_G1 + F_G2 + I_POS2) * 4
                                   + (size_t)(C1 * 196 * 2 + 15) / 16 * 16;
};

template <int C0, int C1, int C2>
__global__ void __launch_bounds__(CB_NT, 2) conv_bwd_kernel(ConvBwdArgs a) {
    pdl_wait();
    using C = ConvBwdCfg<C0, C1, C2>;
    extern __shared__ __align__(16) float smem[];
    float* sX0 = smem;
    float* sRaw = sX0 + C::F_X0;
    float* sP1 = sRaw + C::F_RAW;
    float* sdZ2 = sP1 + C::F_P1;
    float* sW2f = sdZ2 + C::F_DZ2;     // [oc][u*5+v][C1P], flipped taps: W2[oc][ic][(4-u)*5+(4-v)]
    float* sPart = sW2f + C::F_W2F;    // [OG][C1][196]
    float* sG1 = sPart + C::F_PART;    // dL/d(conv1 pre-activation) at the pooled argmax, [C1][196]
    float* sG2 = sG1 + C::F_G1;        // dL/d(conv2 pre-activation) at the pooled argmax, [C2][25]
    int* sPos2 = reinterpret_cast<int*>(sG2 + C::F_G2);            // y*,x* packed: y*<<8 | x*
    uint16_t* sOff1 = reinterpret_cast<uint16_t*>(sPos2 + C::I_POS2);   // argmax position in the 32x32 input, per (c, cell)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int hw = a.in_hw, n_in = hw * hw;

    // flipped / transposed conv2 weights, staged once per CTA
    for (int i = tid; i < C2 * 25 * C::C1P; i += CB_NT) {
        int ic = i % C::C1P, r = i / C::C1P, t = r % 25, oc = r / 25;
        int u = t / 5, v = t - u * 5;
        sW2f[i] = (ic < C1) ? a.w2[((long)oc * C1 + ic) * 25 + (4 - u) * 5 + (4 - v)] : 0.f;
    }

    float acc_w2[C::W2_PER_THREAD];
    float acc_w1[C::W1_PER_THREAD];
#pragma unroll
    for (int j = 0; j < C::W2_PER_THREAD; ++j) acc_w2[j] = 0.f;
#pragma unroll
    for (int j = 0; j < C::W1_PER_THREAD; ++j) acc_w1[j] = 0.f;
    float acc_b2 = 0.f;   // thread tid < C2
    float acc_b1 = 0.f;   // lane 0 of warp w < C1

    for (int img = blockIdx.x; img < a.B; img += gridDim.x) {
        __syncthreads();  // previous image's readers are done with the shared buffers
        // ---- phase A: loads ----
        for (int i = tid; i < C0 * n_in; i += CB_NT)
            sRaw[i] = load_pixel(a.x, a.in_is_u8, (long)img * C0 * n_in + i, a.mean, a.std);
        for (int i = tid; i < C1 * 196; i += CB_NT) {
            sP1[i] = a.p1[(long)img * C1 * 196 + i];
            const int pp = i % 196, py = pp / 14, px = pp - py * 14, k = a.idx1[(long)img * C1 * 196 + i];
            sOff1[i] = (uint16_t)((2 * py + (k >> 1)) * 32 + 2 * px + (k & 1));
        }
        for (int i = tid; i < C2 * 25; i += CB_NT) {
            const long o = (long)img * C2 * 25 + i;
            const int pp = i % 25, py = pp / 5, px = pp - py * 5, k = a.idx2[o];
            sG2[i] = a.dp2[o] * leaky_grad_from_act(a.p2[(long)img * a.ld_p2 + i]);
            sPos2[i] = ((2 * py + (k >> 1)) << 8) | (2 * px + (k & 1));   // (y*, x*) of the conv2 output
        }
        for (int i = tid; i < C::F_DZ2; i += CB_NT) sdZ2[i] = 0.f;
        __syncthreads();
        // ---- phase B: input recompute, scatter dZ2 ----
        for (int c = 0; c < C0; ++c) upsample32_plane(sRaw + c * n_in, hw, sX0 + c * 1024, tid, CB_NT);
        for (int i = tid; i < C2 * 25; i += CB_NT) {
            const int oc = i / 25, ps = sPos2[i];
            sdZ2[oc * ZP * ZP + ((ps >> 8) + 4) * ZP + (ps & 255) + 4] = sG2[i];
        }
        __syncthreads();
        // ---- phase C: conv2 dgrad, dense correlation of padded dZ2 with flipped weights ----
        if (tid < 49 * C::OG) {
            const int blk = tid % 49, og = tid / 49;
            const int by = blk / 7, bx = blk - by * 7;
            float acc[C1][4];
#pragma unroll
            for (int ic = 0; ic < C1; ++ic) acc[ic][0] = acc[ic][1] = acc[ic][2] = acc[ic][3] = 0.f;
#pragma unroll 1
            for (int oo = 0; oo < C2 / C::OG; ++oo) {
                const int oc = og * (C2 / C::OG) + oo;
                const float* zp = sdZ2 + oc * ZP * ZP + (2 * by) * ZP + 2 * bx;
                float patch[6][6];
#pragma unroll
                for (int r = 0; r < 6; ++r)
#pragma unroll
                    for (int c2 = 0; c2 < 3; ++c2) {
                        float2 v = *reinterpret_cast<const float2*>(zp + r * ZP + 2 * c2);
                        patch[r][2 * c2] = v.x; patch[r][2 * c2 + 1] = v.y;
                    }
                const float* wp = sW2f + oc * 25 * C::C1P;
#pragma unroll
                for (int u = 0; u < 5; ++u)
#pragma unroll
                    for (int v = 0; v < 5; ++v) {
                        float wv[C::C1P];
#pragma unroll
                        for (int q = 0; q < C::C1P / 2; ++q) {
                            float2 t2 = *reinterpret_cast<const float2*>(wp + (u * 5 + v) * C::C1P + 2 * q);
                            wv[2 * q] = t2.x; wv[2 * q + 1] = t2.y;
                        }
#pragma unroll
                        for (int ic = 0; ic < C1; ++ic) {
                            acc[ic][0] = fmaf(wv[ic], patch[u][v], acc[ic][0]);
                            acc[ic][1] = fmaf(wv[ic], patch[u][v + 1], acc[ic][1]);
                            acc[ic][2] = fmaf(wv[ic], patch[u + 1][v], acc[ic][2]);
                            acc[ic][3] = fmaf(wv[ic], patch[u + 1][v + 1], acc[ic][3]);
                        }
                    }
            }
#pragma unroll
            for (int ic = 0; ic < C1; ++ic) {
                float* dst = sPart + (og * C1 + ic) * 196 + (2 * by) * 14 + 2 * bx;
                dst[0] = acc[ic][0]; dst[1] = acc[ic][1]; dst[14] = acc[ic][2]; dst[15] = acc[ic][3];
            }
        }
        // ---- phase E (independent of C): conv2 wgrad over the 25 non-zero positions per channel ----
#pragma unroll
        for (int j = 0; j < C::W2_PER_THREAD; ++j) {
            const int o = tid + j * CB_NT;
            if (o < C::W2N) {
                const int t = o % 25, r = o / 25, ic = r % C1, oc = r / C1;
                const int ky = t / 5, kx = t - ky * 5;
                const float* pin = sP1 + ic * 196 + ky * 14 + kx;
                float s = 0.f;
#pragma unroll 5
                for (int pp = 0; pp < 25; ++pp) {
                    const int ps = sPos2[oc * 25 + pp];
                    s = fmaf(sG2[oc * 25 + pp], pin[(ps >> 8) * 14 + (ps & 255)], s);
                }
                acc_w2[j] += s;
            }
        }
        if (tid < C2) {
            float s = 0.f;
            for (int pp = 0; pp < 25; ++pp) s += sG2[tid * 25 + pp];
            acc_b2 += s;
        }
        __syncthreads();
        // ---- phase D: sum the dgrad groups (fixed order), apply pool1-bwd mask -> g1 ----
        for (int i = tid; i < C1 * 196; i += CB_NT) {
            float dp1 = 0.f;
#pragma unroll
            for (int g = 0; g < C::OG; ++g) dp1 += sPart[g * C1 * 196 + i];
            sG1[i] = dp1 * leaky_grad_from_act(sP1[i]);
        }
        __syncthreads();
        // ---- phase F: conv1 wgrad over the 196 non-zero positions per channel ----
#pragma unroll
        for (int j = 0; j < C::W1_PER_THREAD; ++j) {
            const int o = tid + j * CB_NT;
            if (o < C::W1N) {
                const int t = o % 25, r = o / 25, i0 = r % C0, c = r / C0;
                const int ky = t / 5, kx = t - ky * 5;
                const float* xin = sX0 + i0 * 1024 + ky * 32 + kx;
                float s = 0.f;
#pragma unroll 7
                for (int pp = 0; pp < 196; ++pp) s = fmaf(sG1[c * 196 + pp], xin[sOff1[c * 196 + pp]], s);
                acc_w1[j] += s;
            }
        }
        if (warp < C1) {
            float s = 0.f;
            for (int pp = lane; pp < 196; pp += 32) s += sG1[warp * 196 + pp];
            s = warp_sum(s);
            acc_b1 += s;
        }
    }

    // ---- one partial row per CTA ----
    pdl_trigger();   // late: see common.cuh
    float* row = a.partials + (long)blockIdx.x * a.partial_stride;
#pragma unroll
    for (int j = 0; j < C::W2_PER_THREAD; ++j) {
        const int o = tid + j * CB_NT;
        if (o < C::W2N) row[a.off_w2 + o] = acc_w2[j];
    }
#pragma unroll
    for (int j = 0; j < C::W1_PER_THREAD; ++j) {
        const int o = tid + j * CB_NT;
        if (o < C::W1N) row[a.off_w1 + o] = acc_w1[j];
    }
    if (tid < C2) row[a.off_b2 + tid] = acc_b2;
    if (warp < C1 && lane == 0) row[a.off_b1 + warp] = acc_b1;
}

static int num_partials_for(int B) {
    int n = B;     // one image per CTA while rows last: two or three CTAs share an SM and overlap their phases
    if (n > CB_MAX_PARTIALS) n = CB_MAX_PARTIALS;
    return n < 1 ? 1 : n;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_conv_bwd_num_partials(b200dp_dims d) { return num_partials_for(d.batch); }

extern "C" int b200dp_conv_bwd(const void* x, int in_is_u8, float mean, float std, const float* p1,
                               const uint8_t* idx1, const float* p2, const uint8_t* idx2, const float* dp2,
                               const float* w2, float* partials, int partial_stride, int off_w1, int off_b1,
                               int off_w2, int off_b2, b200dp_dims d, b200dp_stream_t s) {
    B200DP_REQUIRE(d.in_ch == 1 && d.c1 == 6 && d.c2 == 16,
                   "the fused CUDA-core conv backward covers the reference LeNet (1,6,16); "
                   "the wide model goes through the tensor-core kernels");
    B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
    using C = ConvBwdCfg<1, 6, 16>;
    B200DP_REQUIRE(off_w1 + C::W1N <= partial_stride && off_w2 + C::W2N <= partial_stride &&
                   off_b1 + 6 <= partial_stride && off_b2 + 16 <= partial_stride, "partial row too short");
    ConvBwdArgs a{x, in_is_u8, d.in_hw, mean, std, p1, idx1, p2, d.ld_p2 > 0 ? d.ld_p2 : 400, idx2, dp2, w2, partials, partial_stride,
                  off_w1, off_b1, off_w2, off_b2, d.batch};
    auto kern = conv_bwd_kernel<1, 6, 16>;
    static DeviceOnce attr;
    B200DP_CHECK(attr.max_dynamic_smem(kern, (int)C::SMEM));
    B200DP_CHECK(launch_pdl(kern, dim3(num_partials_for(d.batch)), dim3(CB_NT), C::SMEM, (cudaStream_t)s, a));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

// =================================================================================================
// Allocation-free variant of the fused convolution backward.
//
// Who needs it: a host program that may not call cudaMalloc where it calls us.  Typical cases are
// a process whose device memory has all been handed to one pool allocator up front (a raw
// cudaMalloc then fails or fragments the pool's view of the GPU), and a control loop with a
// latency budget that cannot absorb the driver lock and implicit synchronisation an allocation
// may take.  b200dp_conv_bwd makes such a caller find room for the partial-gradient rows itself;
// this entry point keeps those rows in a fixed array that is part of the library image instead,
// sums them, and hands back only the four finished gradients.
//
// The array is B200DP_NOALLOC_MAX_ROWS x B200DP_NOALLOC_MAX_STRIDE floats (2,752,512 bytes).  Every
// GPU that loads this library carries it from load until the process exits, used or not.
namespace b200dp {

constexpr int NA_MAX_ROWS = CB_MAX_PARTIALS;   // 256 == B200DP_NOALLOC_MAX_ROWS
constexpr int NA_MAX_STRIDE = 2688;            //        B200DP_NOALLOC_MAX_STRIDE
constexpr int NA_MAX_DEVICES = 64;

static __device__ __align__(256) float g_conv_bwd_rows[NA_MAX_ROWS * NA_MAX_STRIDE];

// Sum rows 0..n_rows-1 of the four owned column ranges, row 0 first (fixed order => repeatable).
__global__ void __launch_bounds__(256) conv_bwd_rows_sum_kernel(const float* __restrict__ rows, int n_rows,
                                                                int stride, float* __restrict__ out,
                                                                int off_w1, int n_w1, int off_b1, int n_b1,
                                                                int off_w2, int n_w2, int off_b2, int n_b2) {
    const int total = n_w1 + n_b1 + n_w2 + n_b2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int col;
        if (i < n_w1) col = off_w1 + i;
        else if (i < n_w1 + n_b1) col = off_b1 + (i - n_w1);
        else if (i < n_w1 + n_b1 + n_w2) col = off_w2 + (i - n_w1 - n_b1);
        else col = off_b2 + (i - n_w1 - n_b1 - n_w2);
        float s = 0.f;
        for (int r = 0; r < n_rows; ++r) s += rows[(long)r * stride + col];
        out[col] = s;
    }
}

// The name g_conv_bwd_rows is only a handle in host code; the address a kernel can use has to be
// asked for, per device, with that device current.  Done once per device, on first use.
static float* g_rows_addr[NA_MAX_DEVICES] = {nullptr};

static int noalloc_device_setup(float** rows_out) {
    int dev = -1;
    B200DP_CHECK(cudaGetDevice(&dev));
    B200DP_REQUIRE(dev >= 0 && dev < NA_MAX_DEVICES, "device ordinal beyond the per-device table");
    float* cached = __atomic_load_n(&g_rows_addr[dev], __ATOMIC_ACQUIRE);
    if (cached == nullptr) {
        float* p = nullptr;
        B200DP_CHECK(cudaGetSymbolAddress((void**)&p, g_conv_bwd_rows));
        B200DP_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
        B200DP_CHECK(cudaFuncSetAttribute(conv_bwd_kernel<1, 6, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)ConvBwdCfg<1, 6, 16>::SMEM));
        __atomic_store_n(&g_rows_addr[dev], p, __ATOMIC_RELEASE);   // two racing first calls store the same address
        cached = p;
    }
    *rows_out = cached;
    return 0;
}

}  // namespace b200dp

extern "C" int b200dp_conv_bwd_noalloc(const void* x, int in_is_u8, float mean, float std, const float* p1,
                                       const uint8_t* idx1, const float* p2, const uint8_t* idx2,
                                       const float* dp2, const float* w2, float* grad_out, int grad_len,
                                       int off_w1, int off_b1, int off_w2, int off_b2, b200dp_dims d,
                                       b200dp_stream_t s) {
    using C = ConvBwdCfg<1, 6, 16>;
    B200DP_REQUIRE(d.in_ch == 1 && d.c1 == 6 && d.c2 == 16, "covers the reference LeNet geometry (1,6,16) only");
    B200DP_REQUIRE(d.in_hw == 28 || d.in_hw == 32, "in_hw must be 28 or 32");
    B200DP_REQUIRE(d.batch >= 1, "empty batch");
    if (grad_len > NA_MAX_STRIDE || grad_len < 1) {
        set_last_error("b200dp_conv_bwd_noalloc: grad_len %d is outside 1..%d floats, the row length the "
                       "built-in scratch array was sized for; use b200dp_conv_bwd with your own buffer",
                       grad_len, NA_MAX_STRIDE);
        return (int)cudaErrorInvalidValue;
    }
    const int n_rows = num_partials_for(d.batch);
    if (n_rows > NA_MAX_ROWS) {
        set_last_error("b200dp_conv_bwd_noalloc: %d partial rows exceed the built-in maximum %d", n_rows, NA_MAX_ROWS);
        return (int)cudaErrorInvalidValue;
    }
    B200DP_REQUIRE(off_w1 >= 0 && off_b1 >= 0 && off_w2 >= 0 && off_b2 >= 0 &&
                   off_w1 + C::W1N <= grad_len && off_w2 + C::W2N <= grad_len &&
                   off_b1 + 6 <= grad_len && off_b2 + 16 <= grad_len, "gradient ranges do not fit in grad_len");
    float* rows = nullptr;
    int rc = noalloc_device_setup(&rows);
    if (rc != 0) return rc;
    ConvBwdArgs a{x, in_is_u8, d.in_hw, mean, std, p1, idx1, p2, d.ld_p2 > 0 ? d.ld_p2 : 400, idx2, dp2, w2, rows, grad_len,
                  off_w1, off_b1, off_w2, off_b2, d.batch};
    // Nothing in the array has to start at zero: each of the n_rows CTAs overwrites all four column
    // ranges of its own row, and the summing pass reads exactly those cells of rows 0..n_rows-1.
    // Whatever an earlier, larger call left in further rows or in other columns is never read.
    conv_bwd_kernel<1, 6, 16><<<n_rows, CB_NT, C::SMEM, (cudaStream_t)s>>>(a);
    B200DP_CHECK(cudaGetLastError());
    const int total = C::W1N + 6 + C::W2N + 16;
    conv_bwd_rows_sum_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)s>>>(
        rows, n_rows, grad_len, grad_out, off_w1, C::W1N, off_b1, 6, off_w2, C::W2N, off_b2, 16);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
