/*
 * b200dp — C ABI of the in-tree CUDA library (csrc/*.cu, built for sm_100a only).
 *
 * Every entry point is `extern "C"`, takes raw device pointers plus the CUDA
 * stream to launch on, allocates nothing unless its name says so, and returns
 * a cudaError_t-compatible int (0 = success).  Python binds these with ctypes
 * (b200dp/ops/native.py); a C/C++ host program can link the .so directly.
 *
 * Each declaration cites the reference code it replaces (file:line in
 * ownzonefeng/pytorch-distributed-training-example).
 *
 * Tensor layouts are PyTorch's: activations NCHW fp32, conv weights OIHW,
 * linear weights (out,in), all contiguous.
 */
#ifndef B200DP_H_
#define B200DP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200dp_stream_t; /* cudaStream_t */

/* Model geometry.  LeNet family: conv5(in_ch->c1)+pool, conv5(c1->c2)+pool,
 * conv5(c2->c3) on a 5x5 map (== GEMM), fc(c3->f1), fc(f1->ncls).
 * Supported: {in_ch=1,c1=6,c2=16} (reference cnn.py:9-23) and the 4x widened
 * CIFAR-shaped {in_ch=3,c1=24,c2=64}; c3, f1 are free; ncls <= 16. */
typedef struct {
    int batch;
    int in_ch;
    int in_hw;   /* 28 (bilinear-upsampled to 32 on the fly) or 32 */
    int c1, c2, c3, f1, ncls;
    int ld_p2;   /* row pitch (floats) of the pooled conv2 output p2 [B][ld_p2]; 0 = dense c2*25 */
} b200dp_dims;

int b200dp_version(void);
const char* b200dp_last_error(void);

/* ---- forward -------------------------------------------------------------
 * conv1 stage: [normalise (x/255-mean)/std when in_is_u8] -> bilinear 28->32
 * (align_corners) -> conv5x5 + bias -> LeakyReLU(0.2) -> maxpool 2x2 (+argmax).
 * Replaces train.py:85 (transform), cnn.py:9-12.   x: [B][in_ch][hw][hw] u8|f32,
 * out p1: [B][c1][14][14], idx1 (nullable): uint8 argmax 0..3 per pooled cell. */
int b200dp_conv1_fwd(const void* x, int in_is_u8, float mean, float std, const float* w, const float* b,
                     float* p1, uint8_t* idx1, b200dp_dims d, b200dp_stream_t s);

/* conv2 stage: conv5x5 + bias -> LeakyReLU -> maxpool (+argmax).  cnn.py:13-15.
 * p1: [B][c1][14][14] -> p2: [B][c2][5][5] */
int b200dp_conv2_fwd(const float* p1, const float* w, const float* b, float* p2, uint8_t* idx2,
                     b200dp_dims d, b200dp_stream_t s);

/* Dense layer y = act(x W^T + b) on CUDA cores; act: 0 none, 1 LeakyReLU(0.2).
 * conv3 (cnn.py:16-17, a GEMM because the map is 5x5) and FC.0 (cnn.py:20-21). */
int b200dp_linear_fwd(const float* x, const float* w, const float* b, float* y, int M, int N, int K, int act,
                      b200dp_stream_t s);

/* Head tail, one warp per row: logits = h1 W^T + b -> softmax -> loss -> dlogits
 * -> dz1 = (dlogits W) * leaky'(h1).   cnn.py:22-23 + train.py:48 (+ backward).
 * loss_mode 0 = "ref": nll_loss on probabilities, L = -mean p[target];
 *           1 = "xent": cross-entropy.
 * loss_out[0] = batch mean loss (deterministic fixed-order sum).
 * eval_acc (nullable): [0] += sum-reduced loss, [1] += #correct (double[2]) —
 * the eval loop of train.py:60-72 without per-batch host syncs.
 * dlogits/dz1 may be null in eval mode.  ticket: int[1] zero-initialised scratch.
 * ld_h1 / ld_dlogits: row pitches of h1 and dlogits in floats (0 = dense F1 / ncls). */
int b200dp_head_tail(const float* h1, const float* w, const float* b, const int64_t* target, int loss_mode,
                     float* probs, float* loss_rows, float* loss_out, float* dlogits, float* dz1,
                     double* eval_acc, int* ticket, int B, int F1, int ncls, int ld_h1, int ld_dlogits,
                     b200dp_stream_t s);

/* Fused head (csrc/head_fused.cu), one launch, exact fp32: conv3 (GEMM on the 5x5 map) + LeakyReLU -> FC.0 + LeakyReLU
 * -> FC.2 -> softmax -> loss (+ eval counters), and in training the row-local backward: dlogits,
 * dz1 = (dlogits W2) * leaky'(h1), dz3 = (dz1 W1) * leaky'(a3).  A CTA carries 4 batch rows through the whole chain
 * in shared memory.  Replaces cnn.py:16-17, 20-23, train.py:48 and the row-local half of their backward (train.py:49);
 * equivalent to b200dp_linear_fwd x2 + b200dp_head_tail + the kind-0 problem of b200dp_linear_bwd_group that produces
 * dz3.  What remains for the backward GEMM launch: dW FC.2, dW FC.0, dW conv3, dP2.
 * p2: [B][ld_p2] (K3 = c2*25 valid columns), a3: [B][ld_a3], h1: [B][ld_h1], dlogits: [B][ld_dlogits] (pitches 0 =
 * dense), dz1: [B][F1], dz3: [B][C3] dense.  dlogits / dz1 / dz3 all NULL = evaluation.  Other arguments as
 * b200dp_head_tail.  Limits: C3, F1 <= 128, ncls <= 16 (b200dp_head_fused_supported). */
int b200dp_head_fused_supported(int K3, int C3, int F1, int ncls);
/* Profiling aid: while dbg != NULL, CTA 0 of every fused-head launch stores clock64() stamps of its phases in dbg[0..16). */
void b200dp_head_fused_set_debug(long long* dbg);
/* The weights reach shared memory by TMA tensor maps (cp.async.bulk.tensor, SWIZZLE_128B; default, needs 16-byte-multiple
 * rows) or by a cp.async ring (any shape).  on = 0 forces the cp.async ring (tests, comparisons). */
void b200dp_head_fused_set_tma(int on);
int b200dp_head_fused(const float* p2, int ld_p2, const float* w3, const float* b3, const float* w1, const float* b1,
                      const float* w2, const float* b2, const int64_t* target, int loss_mode, float* a3, int ld_a3,
                      float* h1, int ld_h1, float* probs, float* loss_rows, float* loss_out, float* dlogits,
                      int ld_dlogits, float* dz1, float* dz3, double* eval_acc, int* ticket, int B, int K3, int C3,
                      int F1, int ncls, b200dp_stream_t s);

/* ---- backward ------------------------------------------------------------
 * Up to four independent GEMM-shaped backward problems in ONE launch.
 * kind 0: dX[M][K] = (dY[M][N] W[N][K]) * leaky'(act_prev[M][K])   (act_prev nullable)
 * kind 1: dW[N][K] = dY[M][N]^T X[M][K],  db[N] = colsum(dY)       (db nullable)
 * Replaces the autograd backward of cnn.py:16-22 (train.py:49). */
typedef struct {
    int kind;
    const float* dy; const float* w_or_x; const float* act_prev;
    float* out; float* db;
    int M, N, K;
} b200dp_bwd_problem;
int b200dp_linear_bwd_group(const b200dp_bwd_problem* probs, int count, b200dp_stream_t s);

/* Tensor-core GEMMs (tcgen05.mma kind::tf32, TMEM accumulators, TMA-fed), up to four problems
 * per launch, directly on fp32 tensors:  C[M][N] = epi( sum_k A(m,k) B(n,k) ).
 * An operand is K-major when its memory is [MN rows][K cols] and MN-major when it is
 * [K rows][MN cols] (lda/ldb = row pitch in floats, multiple of 4; base 16-byte aligned), so
 *   y  = x W^T   : a = x (K-major),   b = W (K-major)
 *   dX = dY W    : a = dY (K-major),  b = W (MN-major)
 *   dW = dY^T X  : a = dY (MN-major), b = X (MN-major)
 * need no transposed copies.  epi: + bias[n], LeakyReLU (act = 1), * LeakyReLU'(mask[m][n]).
 * colext (nullable): logical column N-1 is written to colext[m] instead of out — with a column
 * of ones in X this yields the bias gradient from the same GEMM.
 * Replaces cnn.py:16-17, 20-22 and their backward (train.py:49); cuBLAS in the reference. */
typedef struct {
    const float* a; int lda; int a_mn_major;
    const float* b; int ldb; int b_mn_major;
    int M, N, K;
    const float* bias; int act;
    const float* mask; int ld_mask;
    float* out; int ld_out;
    float* colext;
} b200dp_tc_problem;
int b200dp_tc_gemm_group(const b200dp_tc_problem* probs, int count, b200dp_stream_t s);
/* Same, with a workspace: the launch is re-tiled (32/64/128-wide tiles) and split along K so that about one CTA per
 * SM works on it; the CTAs of a tile park raw partial tiles in `ws`, and the last one to arrive (a ticket per tile in
 * `tickets`, zero-initialised, left zero again) sums them in split order — deterministic — and runs the epilogue.
 * Sizes from b200dp_tc_gemm_ws_floats (returns floats needed, *n_tickets = counters needed).  One launch at a time
 * per workspace. */
int b200dp_tc_gemm_group_splitk(const b200dp_tc_problem* probs, int count, float* ws, long ws_floats,
                                uint32_t* tickets, int n_tickets, b200dp_stream_t s);
long b200dp_tc_gemm_ws_floats(const b200dp_tc_problem* probs, int count, int* n_tickets);
/* Profiling aid: same launch; CTA 0 records clock64() stamps of its pipeline phases in dbg[0..32). */
int b200dp_tc_gemm_group_timed(const b200dp_tc_problem* probs, int count, long long* dbg, b200dp_stream_t s);

/* "Window-tile" convolutions (csrc/tc_win.cu): tcgen05 implicit GEMM that reads the staged image in place —
 * an accumulator tile is a 16x8-pixel window addressed by a no-swizzle K-major descriptor, a filter tap is a
 * start-address shift, so no im2col operand is ever written (kind::f16 on bf16 operands, fp32 accumulation in
 * tensor memory; activations and weights stay fp32 in global memory and are converted while staging / packing).
 * Same tensors and semantics as b200dp_conv1_fwd / b200dp_conv2_fwd, except that the weights come
 * pre-packed: call b200dp_tcw_pack_weights once per step (after the optimizer update) into a scratch of
 * b200dp_tcw_pack_floats(d) floats; it packs the first layer's forward panels, the second layer's forward
 * panels and its flipped data-gradient panels.
 *   b200dp_tcw_conv2_dgrad : g1[B][c1][196] = (pool2-bwd, LeakyReLU-bwd, conv2 dgrad) * LeakyReLU'(p1)
 *   b200dp_tcw_conv_wgrad  (csrc/tc_wgrad.cu; kind::f16 on bf16 operands, both MN-major and read in place):
 *                           layer 2: dW2/db2 from (x = p1, g = dp2, act = p2, idx = idx2)
 *                           layer 1: dW1/db1 from (x = raw images, g = g1, act = NULL, idx = idx1)
 *     written as partial rows [b200dp_tcw_conv_wgrad_num_partials(d)][partial_stride] at columns
 *     off_w / off_b (summed in fixed order by b200dp_reduce_update).
 * Supported geometries: (in_ch,c1,c2) = (1,6,16) and (3,24,64).
 * Replaces cnn.py:10-15 and their backward (train.py:49); cuDNN in the reference. */
long b200dp_tcw_pack_floats(b200dp_dims d);
int b200dp_tcw_pack_weights(const float* w1, const float* w2, float* wpack, b200dp_dims d, b200dp_stream_t s);
int b200dp_tcw_conv1_fwd(const void* x, int in_is_u8, float mean, float std, const float* wpack, const float* b,
                         float* p1, uint8_t* idx1, b200dp_dims d, b200dp_stream_t s);
int b200dp_tcw_conv2_fwd(const float* p1, const float* wpack, const float* b, float* p2, uint8_t* idx2,
                         b200dp_dims d, b200dp_stream_t s);
int b200dp_tcw_conv2_dgrad(const float* dp2, const float* p2, const uint8_t* idx2, const float* wpack,
                           const float* p1, float* g1, b200dp_dims d, b200dp_stream_t s);
int b200dp_tcw_conv_wgrad_num_partials(b200dp_dims d);
int b200dp_tcw_conv_wgrad(int layer, const void* x, int in_is_u8, float mean, float std, const float* g,
                          const float* act, int ld_act, const uint8_t* idx, float* partials,
                          int partial_stride, int off_w, int off_b, b200dp_dims d, b200dp_stream_t s);
/* Profiling aid: while dbg != NULL, CTA 0 of every window-tile launch writes clock64() stamps of its phases
 * (start, set-up, staged, barrier, MMAs issued, accumulators ready, epilogue done, end) to dbg[0..8). */
void b200dp_tcw_set_debug(long long* dbg);
/* Same for the weight-gradient kernels (stamps: start, set-up, planes of image 0, dZ of image 0, chunk 0 / 1 staged,
 * MMAs done, end; bench/wg_probe.py) and for the tf32 GEMM (pipeline stamps; bench/tc_probe.py). */
void b200dp_tcw_wgrad_set_debug(long long* dbg);
void b200dp_tc_gemm_set_debug(long long* dbg);

/* Fused convolution backward for one step: pool2-bwd + leaky-bwd + conv2 dgrad
 * + conv2 wgrad/bgrad + pool1-bwd + leaky-bwd + conv1 wgrad/bgrad (no conv1
 * dgrad: the input needs no gradient).   train.py:49 / cnn.py:9-15.
 * Covers the reference geometry (in_ch=1, c1=6, c2=16).  A CTA owns images cta, cta+grid, ...
 * and writes ONE row of the scratch buffer `partials` ([n_partials][partial_stride] floats,
 * n_partials = b200dp_conv_bwd_num_partials(d)); within a row the four gradients sit at the
 * column offsets off_w1/off_b1/off_w2/off_b2 (the arena layout of the conv bucket, so a row can
 * be summed straight into the gradient arena).  Columns the kernel does not own are never
 * written: allocate the buffer zeroed.  Rows are summed in fixed order by b200dp_reduce_update
 * (deterministic; no atomics). */
int b200dp_conv_bwd_num_partials(b200dp_dims d);
int b200dp_conv_bwd(const void* x, int in_is_u8, float mean, float std, const float* p1, const uint8_t* idx1,
                    const float* p2, const uint8_t* idx2, const float* dp2, const float* w2, float* partials,
                    int partial_stride, int off_w1, int off_b1, int off_w2, int off_b2,
                    b200dp_dims d, b200dp_stream_t s);

/* Allocation-free form of b200dp_conv_bwd (no counterpart in the reference; an addition of this
 * library).  For hosts that cannot call cudaMalloc at the point where they call us — for example
 * because one pool allocator already owns all device memory, or because the call sits in a loop
 * whose latency budget has no room for an allocation.  The partial-gradient rows live in an array
 * built into the library (B200DP_NOALLOC_MAX_ROWS x B200DP_NOALLOC_MAX_STRIDE floats = 2,752,512
 * bytes, resident on every GPU that loads the library until the process exits); the call sums
 * them in fixed order and writes only the four finished gradients:
 *     grad_out[off_w1 .. +150)  dL/dW conv1      grad_out[off_b1 .. +6)   dL/db conv1
 *     grad_out[off_w2 .. +2400) dL/dW conv2      grad_out[off_b2 .. +16)  dL/db conv2
 * Other cells of grad_out are left alone.  grad_len (floats in grad_out, also the row length used
 * inside) must be <= B200DP_NOALLOC_MAX_STRIDE, else the call fails with cudaErrorInvalidValue
 * and b200dp_last_error() says why.  It calls no cudaMalloc / cudaFree and no allocating library
 * routine.  The first call on each device looks up the array's device address and sets a kernel
 * attribute; make that first call before any stream capture begins.  All calls on a device share
 * the one array: two calls whose kernels can overlap on the GPU (different streams, no ordering
 * between them) race on it — order them.  Nothing has to be cleared between calls. */
#define B200DP_NOALLOC_MAX_ROWS 256
#define B200DP_NOALLOC_MAX_STRIDE 2688
int b200dp_conv_bwd_noalloc(const void* x, int in_is_u8, float mean, float std, const float* p1,
                            const uint8_t* idx1, const float* p2, const uint8_t* idx2, const float* dp2,
                            const float* w2, float* grad_out, int grad_len, int off_w1, int off_b1,
                            int off_w2, int off_b2, b200dp_dims d, b200dp_stream_t s);

/* ---- gradient reduction + optimizer update (the product kernel) -----------
 * One launch per bucket [start, start+count) of the flat arenas:
 *   1. (optional) sum `n_partials` partial-gradient rows into grad[]
 *   2. (world > 1) publish the bucket into this rank's symmetric staging
 *      buffer, signal per-slice flags in every peer, wait for the peers'
 *      flags, read every rank's slice over NVLink (P2P loads) and sum in rank
 *      order — identical result on all ranks
 *   3. scale by 1/world, apply Adadelta / SGD-momentum / Adam to param[] and
 *      the state arenas (fp32), optionally write bf16 shadow weights.
 * Replaces train.py:34-39 (10x all_reduce + divide) and train.py:51
 * (optimizer.step), and train.py:46 (zero_grad: nothing accumulates).
 *
 * partials (nullable): [n_partials][partial_stride] rows whose column i is arena element start+i.
 * hyper (device, float[8]): {lr, h0, h1, h2, weight_decay, -, -, -}; Adadelta h0=rho h1=eps,
 * SGD h0=momentum h1=nesterov, Adam h0=beta1 h1=beta2 h2=eps (bias correction is computed on the
 * device from the generation number).  Read on the device so a captured CUDA graph sees
 * scheduler changes.
 * step_counter (device, uint32[4], zero-initialised): [0] generation (= optimizer steps done),
 * [1] ticket, [2] error word (1 = a peer did not answer within comm->timeout_ms; sticky: later launches
 * skip the exchange and the update but keep the bookkeeping going), [3] (push protocol) nanoseconds the step's last
 * bucket kept CTA 0 waiting for the peers' lines — the exposed part of the exchange.  The LAST
 * bucket launch of a step (last_bucket != 0) advances the generation.
 * loss_src / loss_sink (both nullable; used by the last bucket launch only): loss_sink is a device
 * uint64[4] = {address of a mapped pinned float array (0 = off), generation stored at index 0,
 * array length, address of a mapped pinned uint32 (0 = off)}; the step with generation g stores *loss_src
 * (when loss_src is given) into array[g - sink[1]] — how the loss reaches the host without a copy command
 * (train.py:55 loss.item()) — and, once the error word step_counter[2] is set, copies it to the uint32: the
 * host's check for a dead peer is then a plain memory read. */
typedef struct {
    int world, rank;
    int algo;                 /* 0 one-shot: every rank reads every peer's slice (latency regime);
                                 1 two-shot: slice j is summed by rank j%world, which pushes the sum to all;
                                 2 NVLS: as two-shot, but the NVSwitch reduces (multimem.ld_reduce) and
                                   broadcasts (multimem.st) through the multicast alias */
    float* stage[8];          /* every rank's staging base (peer-mapped):
                                 [2 parities][arena_len] published slices, then [arena_len] landing area
                                 for reduced slices (two-shot / NVLS) => 3*arena_len floats */
    uint32_t* flags[8];       /* every rank's flag block (peer-mapped): [8][max_slices] "slice published by
                                 rank r", then [max_slices] "reduced slice delivered" => 9*max_slices words */
    float* mc_stage;          /* multicast alias of the staging block (NVLS) or NULL */
    int arena_len;            /* floats per parity buffer */
    int max_slices;           /* flag slots per source rank (arena_len / 32) */
    int max_ctas;             /* 0 = fill the GPU (2 CTAs per SM).  A bucket that is reduced WHILE backward kernels still
                                 run should use few CTAs (32): CTAs that spin on a slow peer's flags would otherwise hold
                                 the SMs the backward kernels need, and with several chunks in flight per thread a few
                                 CTAs already cover the NVLink latency-bandwidth product */
    /* ---- push protocol (b200dp_grad_push / b200dp_reduce_update_ll; ignored by b200dp_reduce_update) ---- */
    uint32_t* ll[8];          /* every rank's receive area (peer-mapped, zero-initialised):
                                 [2 parities][8 source ranks][arena_len x 2 words] of flagged 8-byte lines */
    int wire;                 /* gradient wire format of the push protocol: 0 = fp32 lines {float, gen},
                                 1 = bf16 lines {2 x bf16, gen} (half the NVLink bytes; fp32 accumulation) */
    uint32_t timeout_ms;      /* bound on every wait for a peer (0 = 20 s): a peer that does not deliver raises
                                 step_counter[2] instead of hanging the GPU (both protocols) */
} b200dp_comm;

int b200dp_reduce_update(float* param, float* grad, float* state0, float* state1, uint16_t* shadow_bf16,
                         int start, int count, const float* partials, int n_partials, int partial_stride,
                         int opt_kind, const float* hyper, uint32_t* step_counter, int last_bucket,
                         const b200dp_comm* comm, const float* loss_src, const uint64_t* loss_sink,
                         b200dp_stream_t s);

/* Push form of the same product kernel, for whole-model gradients in the latency regime (<= ~1 MB; the reference
 * LeNet is 247 KB).  Replaces train.py:34-39 + train.py:51 like b200dp_reduce_update, with ONE exposed wait per
 * step instead of one handshake per bucket:
 *   b200dp_grad_push         as soon as backward has produced grad[start, start+count), write it into every
 *                            peer's receive area (comm->ll) as 8-byte lines {payload, generation} over NVLink;
 *                            returns without waiting for anybody.
 *   b200dp_reduce_update_ll  ends the step for the arena range [start, start+count): sums the partial rows of the
 *                            sub-range [p_start, p_start+p_count) (rows' column i == arena element p_start+i;
 *                            partials NULL / p_count 0: none) and pushes that sub-range itself, pushes the rest
 *                            too when push_rest != 0 (nobody called b200dp_grad_push this step), then for every
 *                            element waits for the peers' lines in LOCAL memory, adds in rank order, scales by
 *                            1/world, applies the optimizer and, when last != 0 (the step's last launch), advances the
 *                            generation and delivers the loss (loss_src / loss_sink as in b200dp_reduce_update).  A step
 *                            may split its arena over several launches on disjoint ranges (e.g. the dense layers'
 *                            buckets beside the convolution backward, the convolution bucket at the end).
 * comm may be NULL (or world 1): plain local update.  A wait that exceeds comm->timeout_ms raises the sticky
 * error word step_counter[2]; from then on every launch of either protocol only does its bookkeeping. */
int b200dp_grad_push(const float* grad, int start, int count, const uint32_t* step_counter,
                     const b200dp_comm* comm, b200dp_stream_t s);
int b200dp_reduce_update_ll(float* param, float* grad, float* state0, float* state1, uint16_t* shadow_bf16,
                            int start, int count, const float* partials, int n_partials, int partial_stride,
                            int p_start, int p_count, int push_rest, int last, int opt_kind, const float* hyper,
                            uint32_t* step_counter, const b200dp_comm* comm, const float* loss_src,
                            const uint64_t* loss_sink, b200dp_stream_t s);

/* The step's loss to the host without a copy command and off the step's dependency chain: a one-thread kernel stores
 * *loss_src into the mapped pinned array described by loss_sink (slot = step_counter[0] + 1 - sink[1]; see
 * b200dp_reduce_update).  Launch it on a stream beside the backward pass, after the kernel that produced the loss and
 * before the step's last reduce+update launch (which advances step_counter[0]); pass that launch loss_src = NULL.
 * (A kernel that stores to host memory stays alive for a PCIe round trip; inside the last kernel of the step that was
 * 5 us per step.)  Replaces train.py:55 loss.item(). */
int b200dp_post_loss(const float* loss_src, const uint32_t* step_counter, const uint64_t* loss_sink, b200dp_stream_t s);

/* ---- symmetric memory plumbing (replaces the NCCL communicator of train.py:26-31) ----
 * Allocation of peer-mappable device memory and exchange via CUDA IPC handles
 * (64 opaque bytes each, shipped between ranks by torch.distributed at start-up). */
int b200dp_symm_alloc(void** ptr, uint64_t bytes);           /* cudaMalloc + zero */
int b200dp_symm_zero(void* ptr, uint64_t bytes);             /* cudaMemset + synchronize: forget all generations
                                                                (callers that move the step counter back; collective) */
int b200dp_symm_free(void* ptr);
int b200dp_symm_export(void* ptr, uint8_t handle_out[64]);
int b200dp_symm_import(const uint8_t handle[64], void** ptr_out);
int b200dp_symm_close(void* imported_ptr);

/* ---- native step pump (csrc/pump.cu) ---------------------------------------
 * The host side of the hot loop in C++: replaces the DataLoader iteration, the two .cuda()
 * copies and loss.item() of train.py:44-45,55.  Step k of a run uses ring slot k%depth:
 * graph_exec[k%depth] is an instantiated CUDA graph with two parallel branches — the whole
 * training step reading dev_x/dev_y[k%depth], and the host->device copies of pinned slot
 * (k+1)%depth into device slot (k+1)%depth (the next step's batch, uploaded off the critical
 * path).  The pump gathers the samples (order[k*batch .. +batch) indexes the base arrays) into
 * the pinned slots from a helper thread, uploads the first batch of a run itself, and launches
 * the graphs.  Only ring-slot reuse is synchronised (one event per pinned slot), so the host
 * runs up to depth-1 steps ahead of the GPU.  The loss of every step is delivered by the step's
 * last b200dp_reduce_update launch (its loss_sink argument).
 * b200dp_pump_run returns once everything is queued; b200dp_pump_drain waits for the queued steps. */
#define B200DP_PUMP_MAX_DEPTH 8
typedef struct b200dp_pump b200dp_pump;
typedef struct {
    const void* images;        /* host base array, one sample = image_bytes contiguous bytes */
    const int64_t* targets;    /* host base labels */
    int64_t image_bytes;
    int batch, depth;
    void* host_x[B200DP_PUMP_MAX_DEPTH];      /* pinned ring: batch*image_bytes each */
    void* host_y[B200DP_PUMP_MAX_DEPTH];      /* pinned ring: batch int64 each */
    void* dev_x[B200DP_PUMP_MAX_DEPTH];       /* device ring, same sizes */
    void* dev_y[B200DP_PUMP_MAX_DEPTH];
    void* graph_exec[B200DP_PUMP_MAX_DEPTH];  /* cudaGraphExec_t per slot */
} b200dp_pump_desc;
/* Upload of one ring slot by a small kernel that reads the pinned host slot through its mapped address (bytes: multiple
 * of 16, both pointers 16-byte aligned).  Meant as the prefetch branch of a step's graph: a kernel node there costs
 * nothing on the step's critical path, a memcpy node cost ~4 us per step.  Replaces train.py:45 (.cuda() copies). */
int b200dp_fetch_slot(void* dst_device, const void* src_pinned_host, int64_t bytes, b200dp_stream_t s);
int b200dp_pump_create(b200dp_pump** out, const b200dp_pump_desc* desc);
/* sink_dev (nullable): the stepper's loss-sink descriptor on the device; the run points it at loss_ring_host (a mapped
 * pinned float array, element k = loss of the run's step k; 0 = off), gen0 = generation of the run's first step. */
int b200dp_pump_run(b200dp_pump* p, const int64_t* order, int n_steps, uint64_t* sink_dev, uint64_t loss_ring_host,
                    uint64_t gen0, b200dp_stream_t s);
/* Host-side prefetch across runs: gather the first (two) batches of the NEXT run into their ring slots now, while the
 * GPU works through the queued steps.  The next b200dp_pump_run uses them if its order starts with the same sample
 * indices, and gathers as usual otherwise.  The upload stays inside the run. */
int b200dp_pump_pregather(b200dp_pump* p, const int64_t* order, int n_steps);
int b200dp_pump_wait_first(b200dp_pump* p);   /* until step 0 of the latest run finished */
int b200dp_pump_drain(b200dp_pump* p);
int b200dp_pump_destroy(b200dp_pump* p);

/* NVTX ranges for host-side phases (visible in Nsight tools; no-ops otherwise). */
void b200dp_nvtx_push(const char* name);
void b200dp_nvtx_pop(void);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H_ */
