// pt_noalloc.cu -- pt_trace_closest_device_noalloc: the binned closest-hit trace of a device ray batch with NO allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). pt_trace_closest_device sizes its scratch
// (work cursor, bin counters, per-ray bin ids, the processing permutation: launch_ray_order + trace_persistent) with cudaMalloc
// on first use; this variant keeps the same scratch in a file-scope __device__ array of a fixed maximum size instead, so the
// call itself issues only memsets and kernel launches on the context's stream.
//
//   * limit: PT_NOALLOC_MAX_RAYS (1 048 576) rays per call; larger batches are refused with PT_ERR_LIMIT;
//   * cost: ~9.4 MB of device memory on EVERY GPU that loads libb200pt.so, held until the process exits;
//   * the array is shared by every call and every context on a device and keeps the last call's contents: what must start at
//     zero (cursor, bin counters) is cleared by each call, everything else is fully rewritten for [0, n) before it is read;
//   * calls that overlap on one GPU (two contexts, two host threads) RACE on it -- serialise them.
// In host code the array's name is only a stand-in: its device address comes from cudaGetSymbolAddress, once per device, with
// that device current, in this translation unit.
#include "pt_context.h"
#include <mutex>

namespace {

constexpr long long kMaxRays = PT_NOALLOC_MAX_RAYS;
// launch_ray_order sizes its counters from the batch: n <= 2^20 rays gives at most 5 bits per axis (+3 octant bits) = 2^18
// counters, whatever PT_RAY_ORDER_BITS / PT_RAY_ORDER_PER_BIN say, followed by up to 2048 scan-tile totals
constexpr size_t kBinWords = (1u << 18) + 2048;
constexpr size_t kCursorWords = 64;                              // the work cursor on a 256-byte line of its own
constexpr size_t kOffCursor = 0, kOffBins = kOffCursor + kCursorWords, kOffRayBin = kOffBins + kBinWords,
                 kOffPerm = kOffRayBin + (size_t)kMaxRays, kWords = kOffPerm + (size_t)kMaxRays;

static __device__ __align__(256) unsigned g_noalloc_scratch[kWords];

constexpr int kMaxDevices = 64;
unsigned* g_scratch_addr[kMaxDevices];                           // per device, resolved once
std::mutex g_scratch_mutex;

// device address of the scratch array on `device`, which must be the current device
int scratch_for_device(int device, unsigned** out) {
    if (device < 0 || device >= kMaxDevices) return fail(PT_ERR_LIMIT, "pt_trace_closest_device_noalloc: device ordinal %d beyond %d", device, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_addr[device]) {
        unsigned* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(PT_ERR_CUDA, "pt_trace_closest_device_noalloc: cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return PT_OK;
}

}  // namespace

extern "C" {

int64_t pt_noalloc_max_rays(void) { return kMaxRays; }

int pt_trace_closest_device_noalloc(pt_context* c, const void* d_rays, int64_t n, void* d_hits) {
    if (!c || !d_rays || !d_hits || n < 0) return fail(PT_ERR_INVALID, "pt_trace_closest_device_noalloc: bad argument");
    if (n > kMaxRays)
        return fail(PT_ERR_LIMIT, "pt_trace_closest_device_noalloc: %lld rays exceed the fixed scratch of %lld rays (PT_NOALLOC_MAX_RAYS); "
                    "split the batch or use pt_trace_closest_device", (long long)n, (long long)kMaxRays);
    if (c->dims.wide) return fail(PT_ERR_UNSUPPORTED, "pt_trace_closest_device_noalloc: binary BVH layout only");
    CK(cudaSetDevice(c->cfg.device));
    int rc = pt_check_scene(c, "pt_trace_closest_device_noalloc", false);
    if (rc == PT_OK) rc = pt_join_slices(c);          // chained frames still on the slice streams (a device-side wait, no allocation)
    if (rc != PT_OK) return rc;
    unsigned* scratch = nullptr;
    if ((rc = scratch_for_device(c->cfg.device, &scratch)) != PT_OK) return rc;
    if (n == 0) return PT_OK;
    int* cursor = reinterpret_cast<int*>(scratch + kOffCursor);
    unsigned* bins = scratch + kOffBins;
    unsigned* ray_bin = scratch + kOffRayBin;
    int* perm = reinterpret_cast<int*>(scratch + kOffPerm);
    const SceneDev s = pt_scene_dev(c);
    CK(cudaEventRecord(c->ev0, c->stream));
    // counting sort of the rays by origin cell + direction octant (clears the counters it uses), then the persistent traversal
    // kernel walking the batch through that permutation (clears its cursor); results land at each ray's own index
    int launches = launch_ray_order(s, (const float4*)d_rays, n, bins, ray_bin, perm, c->stream);
    launch_trace_closest(s, (const float4*)d_rays, n, (float4*)d_hits, cursor, c->dims, c->stream, perm);
    CK(cudaEventRecord(c->ev1, c->stream));
    CK(cudaGetLastError());
    c->counters.kernel_launches += launches + 1;
    c->counters.closest_rays += (uint64_t)n;
    return PT_OK;
}

}  // extern "C"
