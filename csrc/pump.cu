// Native step pump: the per-batch host work of an epoch, done by a C++ loop instead of Python.
//
// Replaces the reference's hot loop plumbing (train.py:44-45 DataLoader iteration + .cuda() copies,
// train.py:55 loss.item()).  Step k of a run uses ring slot (base+k)%depth.  Its CUDA graph has two branches:
// the whole training step (forward, backward, fused reduce+update) on device slot k%depth, and, beside
// it, the host->device copies of pinned slot (k+1)%depth — the NEXT step's batch — so the upload never
// sits on the step's critical path.  The step's loss reaches the host without a copy command: the
// last reduce+update kernel of the step stores it into a mapped pinned array (see b200dp_reduce_update).
// The only host<->GPU synchronisation is ring-slot reuse: a helper thread gathers step j's (shuffled)
// samples into pinned slot j%depth as soon as the graph that last read the slot has finished, while
// the calling thread launches; the host therefore runs up to depth-1 steps ahead of the GPU and the
// launching thread's per-step cost is one cudaGraphLaunch and one event record.
#include <nvtx3/nvToolsExt.h>
#include <string.h>

#include <stdlib.h>
#include <time.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace b200dp {

// B200DP_PUMP_TRACE=1: host timestamps (CLOCK_MONOTONIC, the clock of Python's time.perf_counter_ns) and device times
// of every run, printed to stderr by b200dp_pump_drain — where does a short epoch's fixed cost go.
static inline long now_ns() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (long)ts.tv_sec * 1000000000L + ts.tv_nsec;
}
struct RunTrace {
    static constexpr int kMax = 48;
    long t_entry = 0, t_inline = 0, t_copy = 0, t_exit = 0, t_launch[kMax] = {}, t_gather0[kMax] = {}, t_gather1[kMax] = {};
    int n = 0;
    cudaEvent_t ev_start = nullptr, ev_step[kMax] = {};
};

struct Pump {
    b200dp_pump_desc d;
    bool trace = false;
    std::vector<RunTrace*> runs;
    RunTrace* cur = nullptr;
    // slot_ev[s]: recorded after the last queued operation that READS pinned slot s (the graph whose prefetch
    // branch copies it to the device, or the explicit copy of a run's first step)
    cudaEvent_t slot_ev[B200DP_PUMP_MAX_DEPTH];
    bool slot_pending[B200DP_PUMP_MAX_DEPTH];
    cudaEvent_t first_done;                  // recorded after step 0 of the latest run(): lets the caller log its loss
    bool first_valid = false;
    // launcher <-> helper hand-off
    std::thread helper;
    std::mutex mu;
    std::condition_variable cv;
    const int64_t* order = nullptr;
    long req = 0, filled = 0, launched = 0;  // steps to gather / gathered / graphs launched in the current run()
    // Ring position of the current run's step 0.  A run does not restart at slot 0: it continues behind the slots the
    // previous run used last (+1 for the slot its last graph pre-fetched into), so that its first batches go into the
    // slots whose readers finished longest ago — a log-interval boundary then waits for at most one step, not for most
    // of the queue, before it can gather.
    long base = 0;
    int slot_of(long step) const { return (int)((base + step) % d.depth); }
    bool quit = false, busy = false, failed = false;
    int device = 0;
    // A run's first batch is uploaded on a stream of its own, beside the steps that are still queued (in the compute
    // stream the two copies sat between the last graph of one run and the first graph of the next: a ~20 us bubble at
    // every log-interval boundary).
    cudaStream_t copy_st = nullptr;
    cudaEvent_t copy_done = nullptr;
    // b200dp_pump_pregather: batches of the NEXT run that already sit in their ring slots, and the sample indices they
    // were gathered from (the run checks them against its own order)
    int pre_n = 0;
    std::vector<int64_t> pre_idx;

    // ---- gather: the batch rows are split over a small pool (a 788 KB CIFAR batch gathered row by row by one thread
    // took as long as the GPU step); the helper thread does share 0 and waits for the others ----
    static constexpr int kWorkers = 3;
    std::thread workers[kWorkers];
    std::mutex wmu;
    std::condition_variable wcv;
    long job_step = -1;                      // step being gathered (-1: none)
    unsigned job_seq = 0;                    // bumped per job
    int job_left = 0;                        // worker shares still running
    bool wquit = false;

    void gather_rows(long step, int j0, int j1) {
        const int s = slot_of(step);
        const int64_t* idx = order + step * (long)d.batch;
        uint8_t* hx = (uint8_t*)d.host_x[s];
        int64_t* hy = (int64_t*)d.host_y[s];
        const uint8_t* base = (const uint8_t*)d.images;
        const size_t ib = (size_t)d.image_bytes;
        for (int j = j0; j < j1; ++j) {
            memcpy(hx + (size_t)j * ib, base + (size_t)idx[j] * ib, ib);
            hy[j] = d.targets[idx[j]];
        }
    }
    void worker_main(int w) {
        unsigned seen = 0;
        std::unique_lock<std::mutex> lk(wmu);
        for (;;) {
            wcv.wait(lk, [&] { return wquit || job_seq != seen; });
            if (wquit) return;
            seen = job_seq;
            const long step = job_step;
            lk.unlock();
            const int per = (d.batch + kWorkers) / (kWorkers + 1);
            gather_rows(step, std::min(d.batch, (w + 1) * per), std::min(d.batch, (w + 2) * per));
            lk.lock();
            if (--job_left == 0) wcv.notify_all();
        }
    }
    void gather(long step) {
        const bool split = (size_t)d.batch * (size_t)d.image_bytes >= (256u << 10);   // small batches: not worth a hand-off
        if (!split) { gather_rows(step, 0, d.batch); return; }
        {
            std::lock_guard<std::mutex> lk(wmu);
            job_step = step; job_left = kWorkers; ++job_seq;
        }
        wcv.notify_all();
        const int per = (d.batch + kWorkers) / (kWorkers + 1);
        gather_rows(step, 0, std::min(d.batch, per));
        std::unique_lock<std::mutex> lk(wmu);
        wcv.wait(lk, [&] { return job_left == 0; });
    }

    // Pinned slot j%depth was last read on behalf of step j-depth: by graph j-depth-1 (its prefetch branch),
    // whose event exists once `launched >= j-depth`, or, for the run's first step, by the explicit copy
    // queued before graph 0.  Readers queued by earlier runs still have slot_pending set.
    void helper_main() {
        cudaSetDevice(device);
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] {
                if (quit) return true;
                if (filled >= req) return false;
                const long j = filled;                         // reader of slot j%depth before us: see above
                return j < d.depth || (j == d.depth ? launched >= 1 : launched >= j - d.depth);
            });
            if (quit) return;
            const long k = filled;
            const int s = slot_of(k);
            busy = true;
            const bool wait_ev = slot_pending[s];
            lk.unlock();
            if (wait_ev && cudaEventSynchronize(slot_ev[s]) != cudaSuccess) {
                lk.lock();
                failed = true; busy = false; filled = req;
                cv.notify_all();
                continue;
            }
            RunTrace* tr = cur;
            if (tr && k < RunTrace::kMax) tr->t_gather0[k] = now_ns();
            gather(k);
            if (tr && k < RunTrace::kMax) tr->t_gather1[k] = now_ns();
            lk.lock();
            slot_pending[s] = false;
            busy = false;
            filled = k + 1;
            cv.notify_all();
        }
    }
};

}  // namespace b200dp

using namespace b200dp;

__global__ void set_sink_kernel(unsigned long long* sink, unsigned long long ring, unsigned long long gen0,
                                unsigned long long len) {
    sink[0] = ring; sink[1] = gen0; sink[2] = len;
}

extern "C" int b200dp_pump_create(b200dp_pump** out, const b200dp_pump_desc* desc) {
    B200DP_REQUIRE(out && desc, "null argument");
    B200DP_REQUIRE(desc->depth >= 2 && desc->depth <= B200DP_PUMP_MAX_DEPTH, "depth 2..8");
    B200DP_REQUIRE(desc->batch > 0 && desc->image_bytes > 0 && desc->images && desc->targets, "bad dataset");
    for (int s = 0; s < desc->depth; ++s)
        B200DP_REQUIRE(desc->host_x[s] && desc->host_y[s] && desc->dev_x[s] && desc->dev_y[s] && desc->graph_exec[s],
                       "ring slot incomplete");
    Pump* p = new Pump();
    p->d = *desc;
    B200DP_CHECK(cudaGetDevice(&p->device));
    { const char* t = getenv("B200DP_PUMP_TRACE"); p->trace = t && t[0] == '1'; }
    for (int s = 0; s < desc->depth; ++s) {
        B200DP_CHECK(cudaEventCreateWithFlags(&p->slot_ev[s], cudaEventDisableTiming));
        p->slot_pending[s] = false;
    }
    B200DP_CHECK(cudaEventCreateWithFlags(&p->first_done, cudaEventDisableTiming));
    B200DP_CHECK(cudaStreamCreateWithFlags(&p->copy_st, cudaStreamNonBlocking));
    B200DP_CHECK(cudaEventCreateWithFlags(&p->copy_done, cudaEventDisableTiming));
    for (int w = 0; w < Pump::kWorkers; ++w) p->workers[w] = std::thread([p, w] { p->worker_main(w); });
    p->helper = std::thread([p] { p->helper_main(); });
    *out = (b200dp_pump*)p;
    return 0;
}

// Gather the first batches of the NEXT run into their ring slots now (call it while the GPU still works through the
// previous run): the next b200dp_pump_run with the same leading sample indices starts with its upload instead of ~15 us
// of gathering per batch with the GPU idle.  Host work only; the upload stays with the run.
extern "C" int b200dp_pump_pregather(b200dp_pump* handle, const int64_t* order, int n_steps) {
    Pump* p = (Pump*)handle;
    B200DP_REQUIRE(p && order && n_steps >= 0, "bad arguments");
    const int n = n_steps < 2 ? n_steps : 2;
    p->pre_n = 0;
    p->order = order;
    for (int k = 0; k < n; ++k) {
        const int sl = p->slot_of(k);
        bool pend;
        { std::lock_guard<std::mutex> lk(p->mu); pend = p->slot_pending[sl]; }
        if (pend) B200DP_CHECK(cudaEventSynchronize(p->slot_ev[sl]));
        p->gather(k);
        std::lock_guard<std::mutex> lk(p->mu);
        p->slot_pending[sl] = false;
    }
    p->pre_idx.assign(order, order + (size_t)n * p->d.batch);
    p->pre_n = n;
    p->order = nullptr;
    return 0;
}

extern "C" int b200dp_pump_run(b200dp_pump* handle, const int64_t* order, int n_steps, uint64_t* sink_dev,
                               uint64_t loss_ring_host, uint64_t gen0, b200dp_stream_t stream) {
    Pump* p = (Pump*)handle;
    B200DP_REQUIRE(p && order && n_steps >= 0, "bad arguments");
    if (n_steps == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int depth = p->d.depth;
    nvtxRangePushA("b200dp.pump_run");
    // The first two batches of a run are gathered HERE, on the launching thread: handing them to the helper costs two
    // thread wake-ups (~30-50 us each) with the GPU idle at the start of every epoch / log interval.
    const int inline_steps = n_steps < 2 ? n_steps : 2;
    p->order = order;
    RunTrace* tr = nullptr;
    if (p->trace) {
        tr = new RunTrace();
        tr->n = std::min(n_steps, (int)RunTrace::kMax);
        B200DP_CHECK(cudaEventCreate(&tr->ev_start));
        for (int k = 0; k < tr->n; ++k) B200DP_CHECK(cudaEventCreate(&tr->ev_step[k]));
        tr->t_entry = now_ns();
        B200DP_CHECK(cudaEventRecord(tr->ev_start, st));
        std::lock_guard<std::mutex> lk(p->mu);
        p->cur = tr;
        p->runs.push_back(tr);
    }
    int pre = std::min(p->pre_n, inline_steps);       // batches that b200dp_pump_pregather has put in place already
    if (pre > 0 && memcmp(order, p->pre_idx.data(), (size_t)pre * p->d.batch * sizeof(int64_t)) != 0) pre = 0;
    p->pre_n = 0;
    if (sink_dev != nullptr) {
        // where this run's losses go: a one-thread kernel with the three words as arguments, queued behind the previous
        // run's last graph (whose steps still use the old descriptor) and ahead of this run's first (a 24-byte copy here
        // is a copy-engine command between two graphs: ~5 us of bubble per run)
        set_sink_kernel<<<1, 1, 0, st>>>((unsigned long long*)sink_dev, loss_ring_host, gen0,
                                         loss_ring_host ? (unsigned long long)n_steps : 0ull);
        B200DP_CHECK(cudaGetLastError());
    }
    for (int k = pre; k < inline_steps; ++k) {
        const int sl = p->slot_of(k);
        bool pend;
        { std::lock_guard<std::mutex> lk(p->mu); pend = p->slot_pending[sl]; }
        if (pend) B200DP_CHECK(cudaEventSynchronize(p->slot_ev[sl]));
        p->gather(k);
        std::lock_guard<std::mutex> lk(p->mu);
        p->slot_pending[sl] = false;
    }
    if (tr) tr->t_inline = now_ns();
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->req = n_steps;
        p->filled = inline_steps;
        p->launched = 0;
        p->failed = false;
    }
    p->cv.notify_all();
    auto wait_filled = [&](long k) {
        std::unique_lock<std::mutex> lk(p->mu);
        p->cv.wait(lk, [&] { return p->filled > k; });
        return !p->failed;
    };
    auto mark = [&](int s) -> cudaError_t {      // the operation just queued reads pinned slot s
        cudaError_t e = cudaEventRecord(p->slot_ev[s], st);
        std::lock_guard<std::mutex> lk(p->mu);
        p->slot_pending[s] = true;
        return e;
    };
    int rc = 0;
    cudaError_t e = cudaSuccess;
    // the first step's batch goes up with two explicit copies; every later batch is prefetched by the
    // graph of the step before it
    if (!wait_filled(0)) { set_last_error("pump: waiting for a ring slot failed"); rc = (int)cudaErrorUnknown; }
    if (rc == 0) {
        const int s0 = p->slot_of(0), s1 = (s0 + 1) % depth;
        // Device slot s0 is free once the last graph that read it has finished; that graph recorded slot_ev[s1] (every
        // graph marks the slot after its own), and the last writer of the slot (a prefetch branch) recorded slot_ev[s0],
        // which the gather above has waited for.  An event that was never recorded does not block.
        e = cudaStreamWaitEvent(p->copy_st, p->slot_ev[s1], 0);
        const size_t xb = (size_t)p->d.batch * p->d.image_bytes, yb = (size_t)p->d.batch * sizeof(int64_t);
        const ptrdiff_t hoff = (const char*)p->d.host_y[s0] - (const char*)p->d.host_x[s0];
        const ptrdiff_t doff = (const char*)p->d.dev_y[s0] - (const char*)p->d.dev_x[s0];
        if (e == cudaSuccess && hoff == doff && hoff >= (ptrdiff_t)xb && hoff < (ptrdiff_t)xb + 64) {
            // images and labels share one allocation on both sides: one copy
            e = cudaMemcpyAsync(p->d.dev_x[s0], p->d.host_x[s0], (size_t)hoff + yb, cudaMemcpyHostToDevice, p->copy_st);
        } else if (e == cudaSuccess) {
            e = cudaMemcpyAsync(p->d.dev_x[s0], p->d.host_x[s0], xb, cudaMemcpyHostToDevice, p->copy_st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(p->d.dev_y[s0], p->d.host_y[s0], yb, cudaMemcpyHostToDevice, p->copy_st);
        }
        if (e == cudaSuccess) e = cudaEventRecord(p->copy_done, p->copy_st);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(st, p->copy_done, 0);
        if (e == cudaSuccess) e = mark(s0);
        if (e != cudaSuccess) { set_last_error("pump: first copy: %s", cudaGetErrorString(e)); rc = (int)e; }
        if (tr) tr->t_copy = now_ns();
    }
    for (long k = 0; rc == 0 && k < n_steps; ++k) {
        if (k + 1 < n_steps && !wait_filled(k + 1)) {          // graph k's prefetch branch reads pinned slot (k+1)%depth
            set_last_error("pump: waiting for a ring slot failed"); rc = (int)cudaErrorUnknown; break;
        }
        e = cudaGraphLaunch((cudaGraphExec_t)p->d.graph_exec[p->slot_of(k)], st);
        if (e == cudaSuccess) e = mark(p->slot_of(k + 1));
        if (e == cudaSuccess && k == 0) { e = cudaEventRecord(p->first_done, st); p->first_valid = true; }
        if (tr && k < tr->n) { tr->t_launch[k] = now_ns(); if (e == cudaSuccess) e = cudaEventRecord(tr->ev_step[k], st); }
        if (e != cudaSuccess) { set_last_error("pump: step %ld: %s", k, cudaGetErrorString(e)); rc = (int)e; break; }
        {
            std::lock_guard<std::mutex> lk(p->mu);
            p->launched = k + 1;
        }
        p->cv.notify_all();
    }
    {   // stop the helper from touching `order` after we return
        std::unique_lock<std::mutex> lk(p->mu);
        p->req = 0;                                            // nothing new is started ...
        p->cv.wait(lk, [&] { return !p->busy; });              // ... and a gather in flight (error path) finishes
        p->base = (p->base + n_steps + 1) % depth;             // next run: behind this run's slots (see Pump::base)
        p->cur = nullptr;
    }
    if (tr) tr->t_exit = now_ns();
    nvtxRangePop();
    return rc;
}

extern "C" int b200dp_pump_drain(b200dp_pump* handle) {
    Pump* p = (Pump*)handle;
    B200DP_REQUIRE(p, "null pump");
    for (int s = 0; s < p->d.depth; ++s) {
        bool pend;
        { std::lock_guard<std::mutex> lk(p->mu); pend = p->slot_pending[s]; }
        if (pend) {
            B200DP_CHECK(cudaEventSynchronize(p->slot_ev[s]));
            std::lock_guard<std::mutex> lk(p->mu);
            p->slot_pending[s] = false;
        }
    }
    if (p->trace && !p->runs.empty()) {
        const long t_drained = now_ns();
        B200DP_CHECK(cudaDeviceSynchronize());
        for (RunTrace* tr : p->runs) {
            const long t0 = tr->t_entry;
            fprintf(stderr, "[pump trace] run of %d steps: entry %ld ns | inline gathers done +%.1f us | first copy queued +%.1f | "
                    "returned +%.1f | drained +%.1f\n", tr->n, t0, (tr->t_inline - t0) / 1e3, (tr->t_copy - t0) / 1e3,
                    (tr->t_exit - t0) / 1e3, (t_drained - t0) / 1e3);
            for (int k = 0; k < tr->n; ++k) {
                float ms = 0.f;
                cudaEventElapsedTime(&ms, tr->ev_start, tr->ev_step[k]);
                fprintf(stderr, "[pump trace]   step %2d: gathered %+8.1f..%+8.1f us  launched %+8.1f us  finished on the device %+8.1f us\n",
                        k, tr->t_gather0[k] ? (tr->t_gather0[k] - t0) / 1e3 : 0.0, tr->t_gather1[k] ? (tr->t_gather1[k] - t0) / 1e3 : 0.0,
                        (tr->t_launch[k] - t0) / 1e3, ms * 1e3);
                cudaEventDestroy(tr->ev_step[k]);
            }
            cudaEventDestroy(tr->ev_start);
            delete tr;
        }
        p->runs.clear();
    }
    return 0;
}

extern "C" int b200dp_pump_wait_first(b200dp_pump* handle) {
    Pump* p = (Pump*)handle;
    B200DP_REQUIRE(p, "null pump");
    if (p->first_valid) B200DP_CHECK(cudaEventSynchronize(p->first_done));
    return 0;
}

extern "C" int b200dp_pump_destroy(b200dp_pump* handle) {
    Pump* p = (Pump*)handle;
    if (!p) return 0;
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->quit = true;
    }
    p->cv.notify_all();
    if (p->helper.joinable()) p->helper.join();
    {
        std::lock_guard<std::mutex> lk(p->wmu);
        p->wquit = true;
    }
    p->wcv.notify_all();
    for (int w = 0; w < Pump::kWorkers; ++w)
        if (p->workers[w].joinable()) p->workers[w].join();
    for (int s = 0; s < p->d.depth; ++s) {
        if (p->slot_pending[s]) cudaEventSynchronize(p->slot_ev[s]);
        cudaEventDestroy(p->slot_ev[s]);
    }
    cudaEventDestroy(p->first_done);
    cudaEventDestroy(p->copy_done);
    cudaStreamDestroy(p->copy_st);
    delete p;
    return 0;
}

// ---- host -> device upload by a kernel ---------------------------------------------------------------------------
// The next step's batch is uploaded beside the running step.  As a memcpy NODE in the step's graph that upload cost
// ~4 us per step on the critical path (the copy engine's set-up serialises with the graph's kernel launches); as a small
// KERNEL node that reads the pinned slot through its mapped address it is just another parallel branch.  A few CTAs
// are enough to keep PCIe busy and leave the SMs to the step.
__global__ void __launch_bounds__(256) fetch_slot_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src, long n16) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (long)gridDim.x * blockDim.x) dst[i] = src[i];
}

extern "C" int b200dp_fetch_slot(void* dst_device, const void* src_pinned_host, int64_t bytes, b200dp_stream_t s) {
    B200DP_REQUIRE(dst_device && src_pinned_host && bytes > 0 && bytes % 16 == 0, "16-byte multiple required");
    B200DP_REQUIRE((((uintptr_t)dst_device | (uintptr_t)src_pinned_host) & 15) == 0, "16-byte alignment required");
    void* dev_view = nullptr;    // the device-side address of the pinned host slot (identical under UVA, but ask)
    B200DP_CHECK(cudaHostGetDevicePointer(&dev_view, const_cast<void*>(src_pinned_host), 0));
    const long n16 = bytes / 16;
    int grid = (int)((n16 + 255) / 256);
    if (grid > 8) grid = 8;
    fetch_slot_kernel<<<grid, 256, 0, (cudaStream_t)s>>>((uint4*)dst_device, (const uint4*)dev_view, n16);
    B200DP_CHECK(cudaGetLastError());
    return 0;
}

// ---- NVTX ranges for the Python side (profiling aid; no-ops without a tool attached) ----
extern "C" void b200dp_nvtx_push(const char* name) { nvtxRangePushA(name); }
extern "C" void b200dp_nvtx_pop(void) { nvtxRangePop(); }
