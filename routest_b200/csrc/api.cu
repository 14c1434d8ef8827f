// api.cu — the extern "C" surface declared in include/routest.h: engine lifetime, host-buffer entry points and
// the thin _dev wrappers.  No torch types, no exceptions across the boundary, no CPU fallback.
#include <cmath>
#include <new>

#include "common.cuh"

namespace rt {

static thread_local char g_create_err[512] = "";

int fail(rt_engine *e, int code, const char *fmt, ...)
{
    char *dst = e ? e->err : g_create_err;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
    return code;
}

int reserve(rt_engine *e, DevBuf &b, size_t bytes)
{
    if (bytes <= b.cap) return RT_OK;
    if (b.p) {
        RT_CUDA(e, cudaStreamSynchronize(e->stream));
        RT_CUDA(e, cudaStreamSynchronize(e->copy_stream));
        RT_CUDA(e, cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = (bytes + (size_t(1) << 20) - 1) & ~((size_t(1) << 20) - 1);
    RT_CUDA(e, cudaMalloc(&b.p, want));
    b.cap = want;
    return RT_OK;
}

static int validate_coords(rt_engine *e, const double *lat, const double *lon, long long count)
{
    for (long long i = 0; i < count; ++i) {
        if (!(lat[i] >= -90.0 && lat[i] <= 90.0) || !(lon[i] >= -180.0 && lon[i] <= 180.0))
            return fail(e, RT_ERR_INVALID_ARG, "coordinate %lld out of range or not finite (lat=%g lon=%g)", i, lat[i],
                        lon[i]);
    }
    return RT_OK;
}

}  // namespace rt

using namespace rt;

extern "C" {

int rt_version(void) { return 100; }

int rt_matrix_pitch(int n) { return n <= 0 ? 0 : (n + 15) & ~15; }

int rt_create(int device, rt_engine **out)
{
    if (!out) return fail(nullptr, RT_ERR_INVALID_ARG, "rt_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t st = cudaGetDeviceCount(&count);
    if (st != cudaSuccess || count <= 0)
        return fail(nullptr, RT_ERR_NO_DEVICE, "no CUDA device available (%s); this engine has no CPU fallback",
                    st != cudaSuccess ? cudaGetErrorString(st) : "device count is 0");
    if (device < 0 || device >= count)
        return fail(nullptr, RT_ERR_INVALID_ARG, "rt_create: device %d out of range (0..%d)", device, count - 1);
    cudaDeviceProp prop;
    if ((st = cudaGetDeviceProperties(&prop, device)) != cudaSuccess)
        return fail(nullptr, RT_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(st));
    if (prop.major != 10)
        return fail(nullptr, RT_ERR_NO_DEVICE, "device %d is sm_%d%d; this library carries sm_100a code only", device,
                    prop.major, prop.minor);
    rt_engine *e = new (std::nothrow) rt_engine();
    if (!e) return fail(nullptr, RT_ERR_OUT_OF_MEMORY, "rt_create: out of host memory");
    e->device = device;
    DeviceGuard g(device);
    bool ok = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < 4; ++i) ok = cudaEventCreateWithFlags(&e->ev[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        fail(nullptr, RT_ERR_CUDA, "rt_create: stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
        rt_destroy(e);
        return RT_ERR_CUDA;
    }
    *out = e;
    return RT_OK;
}

void rt_destroy(rt_engine *e)
{
    if (!e) return;
    {
        DeviceGuard g(e->device);
        if (e->stream) cudaStreamSynchronize(e->stream);
        if (e->copy_stream) cudaStreamSynchronize(e->copy_stream);
        eta_model_free(e);
        DevBuf *bufs[] = {&e->coords, &e->xyz, &e->plane_a[0], &e->plane_a[1], &e->plane_b[0],
                          &e->plane_b[1], &e->plane_c[0], &e->plane_c[1]};
        for (DevBuf *b : bufs)
            if (b->p) cudaFree(b->p);
        for (cudaEvent_t ev : e->ev)
            if (ev) cudaEventDestroy(ev);
        if (e->stream) cudaStreamDestroy(e->stream);
        if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    }
    delete e;
}

const char *rt_last_error(const rt_engine *e) { return e ? e->err : g_create_err; }

void *rt_engine_stream(rt_engine *e) { return e ? static_cast<void *>(e->stream) : nullptr; }

int rt_host_alloc(rt_engine *e, size_t bytes, void **out)
{
    if (!e || !out) return fail(e, RT_ERR_INVALID_ARG, "rt_host_alloc: NULL argument");
    DeviceGuard g(e->device);
    RT_CUDA(e, cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return RT_OK;
}

int rt_host_free(rt_engine *e, void *p)
{
    if (!e) return fail(e, RT_ERR_INVALID_ARG, "rt_host_free: NULL engine");
    DeviceGuard g(e->device);
    if (p) RT_CUDA(e, cudaFreeHost(p));
    return RT_OK;
}

int rt_selftest_fast_sqrt(rt_engine *e, uint32_t lo_bits, uint32_t hi_bits, uint64_t *mismatches)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!mismatches || lo_bits > hi_bits) return fail(e, RT_ERR_INVALID_ARG, "rt_selftest_fast_sqrt: bad argument");
    if (lo_bits < 0x0D000000u || hi_bits > 0x7F800000u)
        return fail(e, RT_ERR_INVALID_ARG, "rt_selftest_fast_sqrt: range must lie in [2^-101, inf)");
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard g(e->device);
    RT_TRY(reserve(e, e->coords, 8));
    unsigned long long *cnt = static_cast<unsigned long long *>(e->coords.p);
    RT_CUDA(e, cudaMemsetAsync(cnt, 0, 8, e->stream));
    RT_TRY(run_fast_sqrt_check(e, lo_bits, hi_bits, cnt, e->stream));
    unsigned long long h = 0;
    RT_CUDA(e, cudaMemcpyAsync(&h, cnt, 8, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    *mismatches = h;
    return RT_OK;
}

int rt_unit_vectors_dev(rt_engine *e, const double *lat_d, const double *lon_d, int n, int batch, double *xyz_d,
                        void *stream)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!lat_d || !lon_d || !xyz_d || n <= 0 || batch <= 0)
        return fail(e, RT_ERR_INVALID_ARG, "rt_unit_vectors_dev: bad argument");
    DeviceGuard g(e->device);
    return launch_unit_vectors(e, lat_d, lon_d, n, batch, xyz_d, pick_stream(e, stream));
}

int rt_matrix_cells_dev(rt_engine *e, const double *xyz_d, int n, int batch, int row_begin, int row_end,
                        double dist_scale, double time_scale, void *cells_d, void *stream)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!xyz_d || !cells_d || n <= 0 || batch <= 0 || row_begin < 0 || row_end > n || row_begin > row_end)
        return fail(e, RT_ERR_INVALID_ARG, "rt_matrix_cells_dev: bad argument");
    RT_TRY(check_scales(e, dist_scale, time_scale));
    DeviceGuard g(e->device);
    return launch_matrix_cells(e, xyz_d, n, batch, row_begin, row_end, dist_scale, time_scale,
                               static_cast<int2 *>(cells_d), pick_stream(e, stream));
}

int rt_matrix_build(rt_engine *e, const double *lat, const double *lon, int n, int row_begin, int row_end,
                    double dist_scale, double time_scale, int32_t *dist_out, int32_t *time_out, double *metres_out)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!lat || !lon || n <= 0 || row_begin < 0 || row_end > n || row_begin > row_end)
        return fail(e, RT_ERR_INVALID_ARG, "rt_matrix_build: bad argument (n=%d rows=[%d,%d))", n, row_begin, row_end);
    RT_TRY(check_scales(e, dist_scale, time_scale));
    RT_TRY(validate_coords(e, lat, lon, n));
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard g(e->device);

    const size_t nb = static_cast<size_t>(n) * sizeof(double);
    RT_TRY(reserve(e, e->coords, 2 * nb));
    RT_TRY(reserve(e, e->xyz, 3 * nb));
    double *lat_d = static_cast<double *>(e->coords.p), *lon_d = lat_d + n;
    RT_CUDA(e, cudaMemcpyAsync(lat_d, lat, nb, cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(lon_d, lon, nb, cudaMemcpyHostToDevice, e->stream));
    RT_TRY(launch_unit_vectors(e, lat_d, lon_d, n, 1, static_cast<double *>(e->xyz.p), e->stream));

    const int rows = row_end - row_begin;
    if (rows == 0 || (!dist_out && !time_out && !metres_out)) {
        RT_CUDA(e, cudaStreamSynchronize(e->stream));
        return RT_OK;
    }
    // Row chunks are double-buffered: the D2H copy of chunk k (copy_stream) overlaps the kernel of chunk k+1.
    const int pitch_i32 = (n + 31) & ~31, pitch_f64 = (n + 15) & ~15;
    const int chunk_rows = static_cast<int>(std::max<long long>(1, std::min<long long>(rows, (32LL << 20) / pitch_i32)));
    for (int b = 0; b < 2; ++b) {
        if (dist_out) RT_TRY(reserve(e, e->plane_a[b], static_cast<size_t>(chunk_rows) * pitch_i32 * 4));
        if (time_out) RT_TRY(reserve(e, e->plane_b[b], static_cast<size_t>(chunk_rows) * pitch_i32 * 4));
        if (metres_out) RT_TRY(reserve(e, e->plane_c[b], static_cast<size_t>(chunk_rows) * pitch_f64 * 8));
        if (rows <= chunk_rows) break;
    }
    int k = 0;
    for (int r = row_begin; r < row_end; r += chunk_rows, ++k) {
        const int b = k & 1, r1 = std::min(row_end, r + chunk_rows), cr = r1 - r;
        if (k >= 2) RT_CUDA(e, cudaStreamWaitEvent(e->stream, e->ev[2 + b], 0));  // buffer b drained
        int32_t *dd = dist_out ? static_cast<int32_t *>(e->plane_a[b].p) : nullptr;
        int32_t *td = time_out ? static_cast<int32_t *>(e->plane_b[b].p) : nullptr;
        double *md = metres_out ? static_cast<double *>(e->plane_c[b].p) : nullptr;
        RT_TRY(launch_matrix_planes(e, static_cast<double *>(e->xyz.p), n, r, r1, dist_scale, time_scale, dd, td, md,
                                    pitch_i32, pitch_f64, e->stream));
        RT_CUDA(e, cudaEventRecord(e->ev[b], e->stream));
        RT_CUDA(e, cudaStreamWaitEvent(e->copy_stream, e->ev[b], 0));
        const size_t off = static_cast<size_t>(r - row_begin) * n;
        if (dd)
            RT_CUDA(e, cudaMemcpy2DAsync(dist_out + off, static_cast<size_t>(n) * 4, dd, static_cast<size_t>(pitch_i32) * 4,
                                         static_cast<size_t>(n) * 4, cr, cudaMemcpyDeviceToHost, e->copy_stream));
        if (td)
            RT_CUDA(e, cudaMemcpy2DAsync(time_out + off, static_cast<size_t>(n) * 4, td, static_cast<size_t>(pitch_i32) * 4,
                                         static_cast<size_t>(n) * 4, cr, cudaMemcpyDeviceToHost, e->copy_stream));
        if (md)
            RT_CUDA(e, cudaMemcpy2DAsync(metres_out + off, static_cast<size_t>(n) * 8, md, static_cast<size_t>(pitch_f64) * 8,
                                         static_cast<size_t>(n) * 8, cr, cudaMemcpyDeviceToHost, e->copy_stream));
        RT_CUDA(e, cudaEventRecord(e->ev[2 + b], e->copy_stream));
    }
    RT_CUDA(e, cudaStreamSynchronize(e->copy_stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    return RT_OK;
}

}  // extern "C"
