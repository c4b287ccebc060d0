// Native step pump: the per-batch host work of an epoch, done by a C++ loop instead of Python.
//
// Replaces the reference's hot loop plumbing (train.py:44-45 DataLoader iteration + .cuda() copies,
// train.py:55 loss.item()):  for every step the pump gathers the step's (shuffled) samples of this
// rank's shard into a pinned ring slot, launches ONE instantiated CUDA graph whose first nodes copy
// that slot to the device and whose remaining nodes are the whole training step (forward, backward,
// fused reduce+update), and queues a 4-byte device->host copy of the step's loss into a pinned array.
// Nothing is synchronised except ring-slot reuse: before slot s is refilled the pump waits for the
// event recorded after the last graph that read it, so the host runs at most `depth-1` steps ahead.
//
// A helper thread pre-gathers the next slot while the calling thread launches, so the per-step host
// cost on the launching thread is a graph launch + one tiny copy + one event record.
#include <nvtx3/nvToolsExt.h>
#include <string.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>

#include "common.cuh"

namespace b200dp {

struct Pump {
    b200dp_pump_desc d;
    cudaEvent_t done[B200DP_PUMP_MAX_DEPTH];
    cudaEvent_t first_done;                  // recorded after step 0 of the latest run(): lets the caller log its loss
    bool first_valid = false;
    bool pending[B200DP_PUMP_MAX_DEPTH];
    // helper-thread hand-off: the launcher posts "gather step k into slot s", the helper answers "filled"
    std::thread helper;
    std::mutex mu;
    std::condition_variable cv;
    const int64_t* order = nullptr;
    long req = 0, filled = 0, total = 0;     // steps requested / gathered so far in the current run()
    bool quit = false, busy = false;
    bool slot_free[B200DP_PUMP_MAX_DEPTH];
    int device = 0;

    void gather(long step) {
        const int s = (int)(step % d.depth);
        const int64_t* idx = order + step * (long)d.batch;
        uint8_t* hx = (uint8_t*)d.host_x[s];
        int64_t* hy = (int64_t*)d.host_y[s];
        const uint8_t* base = (const uint8_t*)d.images;
        const size_t ib = (size_t)d.image_bytes;
        for (int j = 0; j < d.batch; ++j) {
            memcpy(hx + (size_t)j * ib, base + (size_t)idx[j] * ib, ib);
            hy[j] = d.targets[idx[j]];
        }
    }

    void helper_main() {
        cudaSetDevice(device);
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] { return quit || (filled < req && slot_free[filled % d.depth]); });
            if (quit) return;
            const long k = filled;
            slot_free[k % d.depth] = false;
            busy = true;
            lk.unlock();
            gather(k);
            lk.lock();
            busy = false;
            filled = k + 1;
            cv.notify_all();
        }
    }
};

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_pump_create(b200dp_pump** out, const b200dp_pump_desc* desc) {
    B200DP_REQUIRE(out && desc, "null argument");
    B200DP_REQUIRE(desc->depth >= 2 && desc->depth <= B200DP_PUMP_MAX_DEPTH, "depth 2..8");
    B200DP_REQUIRE(desc->batch > 0 && desc->image_bytes > 0 && desc->images && desc->targets, "bad dataset");
    for (int s = 0; s < desc->depth; ++s)
        B200DP_REQUIRE(desc->host_x[s] && desc->host_y[s] && desc->graph_exec[s], "ring slot incomplete");
    Pump* p = new Pump();
    p->d = *desc;
    B200DP_CHECK(cudaGetDevice(&p->device));
    for (int s = 0; s < desc->depth; ++s) {
        B200DP_CHECK(cudaEventCreateWithFlags(&p->done[s], cudaEventDisableTiming));
        p->pending[s] = false;
        p->slot_free[s] = true;
    }
    B200DP_CHECK(cudaEventCreateWithFlags(&p->first_done, cudaEventDisableTiming));
    p->helper = std::thread([p] { p->helper_main(); });
    *out = (b200dp_pump*)p;
    return 0;
}

extern "C" int b200dp_pump_run(b200dp_pump* handle, const int64_t* order, int n_steps, float* loss_host,
                               b200dp_stream_t stream) {
    Pump* p = (Pump*)handle;
    B200DP_REQUIRE(p && order && n_steps >= 0, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    nvtxRangePushA("b200dp.pump_run");
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->order = order;
        p->req = n_steps;
        p->filled = 0;
        for (int s = 0; s < p->d.depth; ++s) p->slot_free[s] = !p->pending[s];
    }
    p->cv.notify_all();
    int rc = 0;
    for (long k = 0; k < n_steps; ++k) {
        const int s = (int)(k % p->d.depth);
        if (p->pending[s]) {                                   // the graph that last read this slot must be done
            cudaError_t e = cudaEventSynchronize(p->done[s]);
            if (e != cudaSuccess) { set_last_error("pump: event sync: %s", cudaGetErrorString(e)); rc = (int)e; break; }
            p->pending[s] = false;
            std::lock_guard<std::mutex> lk(p->mu);
            p->slot_free[s] = true;
            p->cv.notify_all();
        }
        {
            std::unique_lock<std::mutex> lk(p->mu);
            p->cv.wait(lk, [&] { return p->filled > k; });     // helper has gathered step k
        }
        cudaError_t e = cudaGraphLaunch((cudaGraphExec_t)p->d.graph_exec[s], st);
        if (e == cudaSuccess && loss_host)
            e = cudaMemcpyAsync(loss_host + k, p->d.loss_dev, sizeof(float), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaEventRecord(p->done[s], st);
        if (e == cudaSuccess && k == 0) { e = cudaEventRecord(p->first_done, st); p->first_valid = true; }
        if (e != cudaSuccess) { set_last_error("pump: step %ld: %s", k, cudaGetErrorString(e)); rc = (int)e; break; }
        p->pending[s] = true;
    }
    {   // stop the helper from touching `order` after we return
        std::unique_lock<std::mutex> lk(p->mu);
        p->req = 0;                                            // nothing new is started ...
        p->cv.wait(lk, [&] { return !p->busy; });              // ... and a gather in flight (error path) finishes
    }
    nvtxRangePop();
    return rc;
}

extern "C" int b200dp_pump_drain(b200dp_pump* handle) {
    Pump* p = (Pump*)handle;
    B200DP_REQUIRE(p, "null pump");
    for (int s = 0; s < p->d.depth; ++s)
        if (p->pending[s]) {
            B200DP_CHECK(cudaEventSynchronize(p->done[s]));
            p->pending[s] = false;
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
    for (int s = 0; s < p->d.depth; ++s) {
        if (p->pending[s]) cudaEventSynchronize(p->done[s]);
        cudaEventDestroy(p->done[s]);
    }
    cudaEventDestroy(p->first_done);
    delete p;
    return 0;
}

// ---- NVTX ranges for the Python side (profiling aid; no-ops without a tool attached) ----
extern "C" void b200dp_nvtx_push(const char* name) { nvtxRangePushA(name); }
extern "C" void b200dp_nvtx_pop(void) { nvtxRangePop(); }
