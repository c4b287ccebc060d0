// Library runtime: error reporting and symmetric (peer-mapped) memory plumbing.
//
// The reference gets its communicator from dist.init_process_group('nccl') (train.py:26-31).
// Here torch.distributed is only the bootstrap channel: each rank allocates its staging + flag
// block with b200dp_symm_alloc, exports a CUDA IPC handle (64 opaque bytes), the handles are
// all-gathered once, and every rank maps every peer's block.  After that the step path touches
// peer memory directly from kernels (reduce_update.cu) and never calls a communication library.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace b200dp {
static thread_local char g_err[512] = "";
void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_version(void) { return 100; }
extern "C" const char* b200dp_last_error(void) { return g_err; }

extern "C" int b200dp_symm_alloc(void** ptr, uint64_t bytes) {
    B200DP_REQUIRE(ptr != nullptr && bytes > 0, "bad arguments");
    B200DP_CHECK(cudaMalloc(ptr, bytes));
    B200DP_CHECK(cudaMemset(*ptr, 0, bytes));
    B200DP_CHECK(cudaDeviceSynchronize());
    return 0;
}
extern "C" int b200dp_symm_zero(void* ptr, uint64_t bytes) {
    B200DP_REQUIRE(ptr != nullptr && bytes > 0, "bad arguments");
    B200DP_CHECK(cudaMemset(ptr, 0, bytes));
    B200DP_CHECK(cudaDeviceSynchronize());
    return 0;
}
extern "C" int b200dp_symm_free(void* ptr) {
    B200DP_CHECK(cudaFree(ptr));
    return 0;
}
extern "C" int b200dp_symm_export(void* ptr, uint8_t handle_out[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    cudaIpcMemHandle_t h;
    B200DP_CHECK(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle_out, &h, 64);
    return 0;
}
extern "C" int b200dp_symm_import(const uint8_t handle[64], void** ptr_out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    B200DP_CHECK(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int b200dp_symm_close(void* imported_ptr) {
    B200DP_CHECK(cudaIpcCloseMemHandle(imported_ptr));
    return 0;
}
