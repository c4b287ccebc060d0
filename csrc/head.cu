// Head tail: FC.2 + softmax + loss (+ eval metrics) + the backward through softmax/loss/FC.2's
// data path, one warp per batch row, one launch.
//
// Replaces reference cnn.py:22-23 (Linear(84,10), Softmax) and train.py:48 (nll_loss on the
// probabilities) and their backward (train.py:49); in eval mode it replaces the per-batch
// `.item()` pair of train.py:68-70 with device-side accumulation.
//
//   logits = h1 W^T + b ; p = softmax(logits)
//   ref :  L = -(1/B) sum_i p_i[t_i]        dL/dz_j = -(1/B) p_t (delta_jt - p_j)
//   xent:  L = -(1/B) sum_i log p_i[t_i]    dL/dz_j =  (1/B) (p_j - delta_jt)
//   dz1 = (dlogits W) * leaky'(h1)          (gradient w.r.t. FC.0's pre-activation)
#include "common.cuh"

namespace b200dp {

constexpr int HT_ROWS = 8;      // rows (warps) per CTA
constexpr int HT_MAXC = 16;     // max classes
constexpr int HT_MAXK = 12;     // per-lane k elements: F1 <= 384

struct HeadTailArgs {
    const float* h1; const float* w; const float* b; const int64_t* target;
    float* probs; float* loss_rows; float* loss_out; float* dlogits; float* dz1;
    double* eval_acc; int* ticket;
    int B, F1, ncls, loss_mode;
    int ld_h1, ld_dl;
};

__global__ void __launch_bounds__(HT_ROWS * 32) head_tail_kernel(HeadTailArgs a) {
    pdl_wait();
    extern __shared__ float sW[];          // [ncls][F1] then bias[ncls]
    float* sBias = sW + a.ncls * a.F1;
    __shared__ int sLast;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < a.ncls * a.F1; i += blockDim.x) sW[i] = a.w[i];
    if (tid < a.ncls) sBias[tid] = a.b[tid];
    __syncthreads();

    const int row = blockIdx.x * HT_ROWS + warp;
    if (row < a.B) {
        float h[HT_MAXK];
#pragma unroll
        for (int j = 0; j < HT_MAXK; ++j) {
            int k = lane + 32 * j;
            h[j] = (k < a.F1) ? a.h1[(long)row * a.ld_h1 + k] : 0.f;
        }
        float z[HT_MAXC];
#pragma unroll
        for (int c = 0; c < HT_MAXC; ++c) {
            float part = 0.f;
            if (c < a.ncls) {
#pragma unroll
                for (int j = 0; j < HT_MAXK; ++j) {
                    int k = lane + 32 * j;
                    if (k < a.F1) part = fmaf(h[j], sW[c * a.F1 + k], part);
                }
            }
            z[c] = (c < a.ncls) ? warp_sum(part) + sBias[c] : -INFINITY;
        }
        float mx = z[0];
#pragma unroll
        for (int c = 1; c < HT_MAXC; ++c) mx = fmaxf(mx, z[c]);
        float e[HT_MAXC], se = 0.f;
#pragma unroll
        for (int c = 0; c < HT_MAXC; ++c) { e[c] = (c < a.ncls) ? expf(z[c] - mx) : 0.f; se += e[c]; }
        const float inv = 1.f / se;
        const int t = (int)a.target[row];
        float pt = 0.f, zt = 0.f;
#pragma unroll
        for (int c = 0; c < HT_MAXC; ++c) {
            e[c] *= inv;                                  // probabilities
            if (c == t) { pt = e[c]; zt = z[c]; }
        }
        const float lrow = (a.loss_mode == 0) ? -pt : -(zt - mx - logf(se));
        if (lane == 0) a.loss_rows[row] = lrow;
        if (a.probs && lane < a.ncls) {
            float pv = 0.f;
#pragma unroll
            for (int c = 0; c < HT_MAXC; ++c) if (c == lane) pv = e[c];
            a.probs[(long)row * a.ncls + lane] = pv;
        }
        if (a.dlogits) {
            const float invB = 1.f / (float)a.B;
            float dz[HT_MAXC];
#pragma unroll
            for (int c = 0; c < HT_MAXC; ++c) {
                float d = (c == t) ? 1.f : 0.f;
                dz[c] = (a.loss_mode == 0) ? -invB * pt * (d - e[c]) : invB * (e[c] - d);
                if (c >= a.ncls) dz[c] = 0.f;
            }
            if (lane < a.ncls) {
                float dv = 0.f;
#pragma unroll
                for (int c = 0; c < HT_MAXC; ++c) if (c == lane) dv = dz[c];
                a.dlogits[(long)row * a.ld_dl + lane] = dv;
            }
            if (a.dz1) {
#pragma unroll
                for (int j = 0; j < HT_MAXK; ++j) {
                    int k = lane + 32 * j;
                    if (k < a.F1) {
                        float g = 0.f;
#pragma unroll
                        for (int c = 0; c < HT_MAXC; ++c)
                            if (c < a.ncls) g = fmaf(dz[c], sW[c * a.F1 + k], g);
                        a.dz1[(long)row * a.F1 + k] = g * leaky_grad_from_act(h[j]);
                    }
                }
            }
        }
        // (the last CTA derives argmax == target from the stored probabilities when eval_acc is set)
    }

    // ---- last CTA: fixed-order batch reduction (deterministic) ----
    __threadfence();
    pdl_trigger();   // late: see common.cuh
    __syncthreads();
    if (tid == 0) sLast = (atomicAdd(a.ticket, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (!sLast) return;
    __threadfence();
    // all 256 threads of the last CTA: thread t sums rows t, t+256, ... ; warp butterflies; warp 0 adds the
    // eight warp sums in order.  Fixed order => bit-reproducible; ~3 L2 round trips instead of B/32.
    __shared__ float sSum[HT_ROWS];
    __shared__ int sCnt[HT_ROWS];
    float v = 0.f;
    int ok = 0;
    for (int r = tid; r < a.B; r += HT_ROWS * 32) {
        v += __ldcg(a.loss_rows + r);
        if (a.eval_acc) {
            int am = 0; float bst = -INFINITY;
            for (int c = 0; c < a.ncls; ++c) {
                float pv = __ldcg(a.probs + (long)r * a.ncls + c);
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
        for (int w = 0; w < HT_ROWS; ++w) { s += sSum[w]; correct += sCnt[w]; }
        a.loss_out[0] = s / (float)a.B;
        if (a.eval_acc) { a.eval_acc[0] += (double)s; a.eval_acc[1] += (double)correct; }
        *a.ticket = 0;   // re-arm for the next launch
    }
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_head_tail(const float* h1, const float* w, const float* b, const int64_t* target,
                                int loss_mode, float* probs, float* loss_rows, float* loss_out, float* dlogits,
                                float* dz1, double* eval_acc, int* ticket, int B, int F1, int ncls, int ld_h1,
                                int ld_dlogits, b200dp_stream_t s) {
    B200DP_REQUIRE(ncls >= 1 && ncls <= HT_MAXC, "ncls <= 16");
    B200DP_REQUIRE(F1 >= 1 && F1 <= 32 * HT_MAXK, "F1 <= 384");
    B200DP_REQUIRE(probs != nullptr || eval_acc == nullptr, "eval metrics need the probs buffer");
    B200DP_REQUIRE(loss_rows && loss_out && ticket, "loss_rows, loss_out and ticket are required");
    HeadTailArgs a{h1, w, b, target, probs, loss_rows, loss_out, dlogits, dz1, eval_acc, ticket, B, F1, ncls, loss_mode,
                   ld_h1 > 0 ? ld_h1 : F1, ld_dlogits > 0 ? ld_dlogits : ncls};
    const int grid = (B + HT_ROWS - 1) / HT_ROWS;
    const size_t smem = (size_t)(ncls * F1 + ncls) * sizeof(float);
    B200DP_CHECK(launch_pdl(head_tail_kernel, dim3(grid), dim3(HT_ROWS * 32), smem, (cudaStream_t)s, a));
    B200DP_CHECK(cudaGetLastError());
    return 0;
}
