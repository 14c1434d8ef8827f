// noalloc.cu — rt_matrix_cells_static_dev: the device-resident matrix build for callers that must not allocate.
//
// Who that is.  A dispatch service usually sizes its GPU memory once at start-up and hands all of it to a pool
// (torch's caching allocator, RMM, its own arena).  On the request path such a process cannot call cudaMalloc:
// the driver has nothing left to give even though the pool has room, and cudaMalloc/cudaFree are device-wide
// synchronisation points that would stall every other request's stream.  The ordinary route through this library
// (rt_unit_vectors_dev into a buffer the caller allocates, then rt_matrix_cells_dev) asks such a caller for one
// more buffer — the per-stop unit vectors — that it has no way to provide without going back to its allocator.
// This entry point keeps that scratch inside the library image instead.
//
// The scratch is the file-scope array below: 3 x kStaticMaxNodes doubles = 384 KiB of device memory that every
// GPU which loads this library holds until the process exits, whether or not the entry point is ever called.
//
// Sharing rules.  There is ONE array per GPU, shared by every engine, stream and thread of the process:
//   * two calls whose kernels can overlap on the same GPU (different streams, or different host threads using
//     different streams) race on it — serialise them on one stream, or order them with events;
//   * the array keeps the previous call's vectors.  Nothing in it has to start at zero: a call overwrites entries
//     [0, 3n) before its matrix kernel reads exactly those entries, so stale data beyond 3n is never looked at.
//
// Set-up.  A __device__ array's name is, in host code, a stand-in symbol and not a device address: handing it to a
// kernel makes the kernel fault (or, where the GPU can reach host memory, quietly use the wrong memory).  The
// first call on each GPU therefore resolves the real address once with cudaGetSymbolAddress, with that GPU current,
// and caches it per device ordinal.  Make that first call outside any stream capture (it is a runtime query, not
// stream work); later calls only enqueue two kernels.
#include "common.cuh"

namespace rt {

namespace {

constexpr int kStaticMaxNodes = 16384;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) double g_static_xyz[3 * kStaticMaxNodes];

std::mutex g_static_mu;
double *g_static_ptr[kMaxDevices] = {nullptr};

// Device address of g_static_xyz on `device` (which must be current); resolved on first use per device.
int static_scratch(rt_engine *e, int device, double **out)
{
    if (device < 0 || device >= kMaxDevices)
        return fail(e, RT_ERR_INVALID_ARG, "rt_matrix_cells_static_dev: device ordinal %d not supported (max %d)", device,
                    kMaxDevices - 1);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[device]) {
        double *ptr = nullptr;
        cudaError_t st = cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_static_xyz);
        if (st != cudaSuccess || !ptr)
            return fail(e, RT_ERR_CUDA, "cudaGetSymbolAddress(static scratch) failed on device %d: %s", device,
                        cudaGetErrorString(st));
        g_static_ptr[device] = ptr;
    }
    *out = g_static_ptr[device];
    return RT_OK;
}

}  // namespace
}  // namespace rt

using namespace rt;

extern "C" {

int rt_matrix_static_max_nodes(void) { return kStaticMaxNodes; }

int rt_matrix_cells_static_dev(rt_engine *e, const double *lat_d, const double *lon_d, int n, double dist_scale,
                               double time_scale, void *cells_d, void *stream)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!lat_d || !lon_d || !cells_d || n <= 0)
        return fail(e, RT_ERR_INVALID_ARG, "rt_matrix_cells_static_dev: bad argument");
    if (n > kStaticMaxNodes)
        return fail(e, RT_ERR_INVALID_ARG,
                    "rt_matrix_cells_static_dev: n = %d exceeds the static scratch (%d nodes); use rt_unit_vectors_dev + "
                    "rt_matrix_cells_dev with a buffer of your own",
                    n, kStaticMaxNodes);
    RT_TRY(check_scales(e, dist_scale, time_scale));
    DeviceGuard g(e->device);
    double *xyz = nullptr;
    RT_TRY(static_scratch(e, e->device, &xyz));
    cudaStream_t st = pick_stream(e, stream);
    RT_TRY(launch_unit_vectors(e, lat_d, lon_d, n, 1, xyz, st));
    return launch_matrix_cells(e, xyz, n, 1, 0, n, dist_scale, time_scale, static_cast<int2 *>(cells_d), st);
}

}  // extern "C"
