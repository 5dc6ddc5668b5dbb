// kd_build_gpu_noalloc — the tree build for callers that cannot obtain device memory.
//
// Who needs this: code that runs where calling cudaMalloc (or any allocator that may fall back to
// it, torch's caching allocator included) is not allowed or not safe — a plug-in executing inside
// another framework's pre-planned memory arena, a service that has handed the whole device to a
// fixed pool at start-up and forbids further allocation so that it can never fail or fragment
// later, or a latency-critical loop where an allocation's implicit device synchronisation is
// unacceptable.  kd_build_gpu needs a scratch workspace from such a caller; this variant brings
// its own: a fixed 16 MiB array that is part of the library image on the device.
//
// Consequences of the scratch being one static array:
//   * every GPU on which the library's device code is loaded keeps those 16 MiB reserved until the
//     process exits, whether or not this entry point is ever called;
//   * the problem must fit: kd_build_workspace_bytes(n, dim) <= 16 MiB (about 400 000 points in
//     3-D).  Larger problems are refused with an error, nothing is launched;
//   * all calls on one device use the same array and it keeps whatever the previous call left in
//     it.  The build clears the parts it needs zeroed at the start of every call, so back-to-back
//     calls on one stream are fine; two builds that can overlap on the GPU (different streams,
//     different host threads) would both write the array and corrupt each other — serialise them;
//   * the first call on each device looks the array's device address up (see below).  Make that
//     first call outside of any stream capture.
//
// No counterpart in the reference (it has no device code at all).
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>

#include "../../include/kdtree_b200.h"
#include "kd_error.h"

namespace kd {
int build_gpu(const float* points, int64_t n, int dim, int32_t* perm, float* tree_points, void* workspace,
              size_t workspace_bytes, cudaStream_t stream);
}

namespace {

constexpr size_t kNoAllocScratchBytes = (size_t)16 << 20;
constexpr int kMaxDevices = 64;

// In host code this name is only a handle for the runtime's symbol table, NOT a device address:
// handing it to a kernel would make the kernel fault (or silently touch host memory where the GPU
// can reach it).  The usable address differs per device and comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

std::mutex g_mu;
void* g_scratch_addr[kMaxDevices] = {};

// Device address of the scratch array on the CURRENT device; resolved once per device.
int scratch_for_current_device(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return kd::fail(std::string("kd_build_gpu_noalloc: cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return kd::fail("kd_build_gpu_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return kd::fail(std::string("kd_build_gpu_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

}  // namespace

extern "C" int kd_build_gpu_noalloc(const float* points, int64_t n, int dim, int32_t* perm, float* tree_points,
                                    void* stream) {
    if (!points || !perm || n <= 0 || dim <= 0) return kd::fail("kd_build_gpu_noalloc: bad arguments");
    const size_t need = kd_build_workspace_bytes(n, dim);
    if (need == 0 || need > kNoAllocScratchBytes)
        return kd::fail("kd_build_gpu_noalloc: n=" + std::to_string(n) + ", dim=" + std::to_string(dim) + " needs " +
                        std::to_string(need) + " bytes of scratch but the built-in scratch holds " +
                        std::to_string(kNoAllocScratchBytes) +
                        " bytes (16 MiB); use kd_build_gpu with a caller-provided workspace for larger problems");
    void* scratch = nullptr;
    if (int rc = scratch_for_current_device(&scratch)) return rc;
    // same kernels, same order, same results as kd_build_gpu; it zeroes the histogram and error words itself
    return kd::build_gpu(points, n, dim, perm, tree_points, scratch, kNoAllocScratchBytes, (cudaStream_t)stream);
}

// Companion of kd_build_status for the built-in scratch: the caller of kd_build_gpu_noalloc never sees a
// workspace pointer, so it could not ask whether the device-side select of its last build failed.  Copies one
// int from the scratch array to the host (no allocation) and synchronises the stream.
extern "C" int kd_build_status_noalloc(int64_t n, int dim, void* stream) {
    if (n <= 0 || dim <= 0) return kd::fail("kd_build_status_noalloc: bad arguments");
    const size_t need = kd_build_workspace_bytes(n, dim);
    if (need == 0 || need > kNoAllocScratchBytes)
        return kd::fail("kd_build_status_noalloc: no build of this size can have used the built-in scratch");
    void* scratch = nullptr;
    if (int rc = scratch_for_current_device(&scratch)) return rc;
    return kd_build_status(scratch, n, dim, stream);
}
