// problem.cu — extern "C" entry points around rt_problem: a batch of device-resident instances on which the
// matrix build, the greedy construction, move evaluation and local search run without the matrix ever leaving HBM.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <new>
#include <vector>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include "problem.cuh"

namespace rt {

int greedy_f64_launch(rt_engine *e, const double *D_d, int n, const double *demand_d, const double *cap_d,
                      const double *maxd_d, unsigned long long *keys_d, int *order_d, int *order_out_d, int *trip_off_d,
                      int *ntrips_d, int *status_d, cudaStream_t st);

namespace {

template <typename T>
int dev_alloc(rt_problem *p, T *&ptr, size_t count)
{
    void *raw = nullptr;
    RT_CUDA(p->e, cudaMalloc(&raw, std::max<size_t>(count, 1) * sizeof(T)));
    ptr = static_cast<T *>(raw);
    p->all_allocs[p->n_allocs++] = raw;
    return RT_OK;
}

int next_pow2(int v)
{
    int r = 1;
    while (r < v) r <<= 1;
    return r;
}

int stage(rt_problem *p, size_t bytes)
{
    if (bytes <= p->host_stage_bytes) return RT_OK;
    if (p->host_stage) cudaFreeHost(p->host_stage);
    p->host_stage = nullptr;
    p->host_stage_bytes = 0;
    RT_CUDA(p->e, cudaHostAlloc(&p->host_stage, bytes, cudaHostAllocDefault));
    p->host_stage_bytes = bytes;
    return RT_OK;
}

// interleaves / de-interleaves dense host-layout planes with the device cell layout
__global__ void planes_to_cells_kernel(const int *__restrict__ dist, const int *__restrict__ tim, int n, int pitch,
                                       int2 *__restrict__ cells)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= pitch) return;
    cells[static_cast<size_t>(i) * pitch + j] =
        j < n ? make_int2(dist[static_cast<size_t>(i) * n + j], tim[static_cast<size_t>(i) * n + j]) : make_int2(0, 0);
}

__global__ void cells_to_planes_kernel(const int2 *__restrict__ cells, int n, int pitch, int *__restrict__ dist,
                                       int *__restrict__ tim)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= n) return;
    const int2 c = cells[static_cast<size_t>(i) * pitch + j];
    dist[static_cast<size_t>(i) * n + j] = c.x;
    tim[static_cast<size_t>(i) * n + j] = c.y;
}

// Host side of rt_problem_plan: every id in [0, n) and exactly n-1 customers among the P positions.  On the call's
// critical path (the GPU is only busy with the input copies meanwhile), hence four ids per step: the sign bit of
// v | (n-1 - v) is set iff v is negative or above n-1.
// dlat^2 + dlon^2 in radians over the bounding box of the coordinates: an upper bound of every pair's squared central
// angle, hence of its chord^2 (the move kernels drop the cell function's range test when it is small enough).  The same
// pass says whether every coordinate is finite and in range (a NaN fails every comparison, so it can neither move the
// box nor pass the range test): *valid.  On a request's critical path, hence two values per step.
double angle2_bound_of(const double *lat, const double *lon, long long count, bool *valid = nullptr)
{
    double la0 = lat[0], la1 = lat[0], lo0 = lon[0], lo1 = lon[0];
    long long i = 0;
    bool ok = true;
#if defined(__SSE2__)
    __m128d vla0 = _mm_set1_pd(la0), vla1 = vla0, vlo0 = _mm_set1_pd(lo0), vlo1 = vlo0;
    __m128d vok = _mm_castsi128_pd(_mm_set1_epi32(-1));
    const __m128d la_lo = _mm_set1_pd(-90.0), la_hi = _mm_set1_pd(90.0), lo_lo = _mm_set1_pd(-180.0), lo_hi = _mm_set1_pd(180.0);
    for (; i + 2 <= count; i += 2) {
        const __m128d a = _mm_loadu_pd(lat + i), o = _mm_loadu_pd(lon + i);
        vla0 = _mm_min_pd(vla0, a);
        vla1 = _mm_max_pd(vla1, a);
        vlo0 = _mm_min_pd(vlo0, o);
        vlo1 = _mm_max_pd(vlo1, o);
        vok = _mm_and_pd(vok, _mm_and_pd(_mm_and_pd(_mm_cmpge_pd(a, la_lo), _mm_cmple_pd(a, la_hi)),
                                         _mm_and_pd(_mm_cmpge_pd(o, lo_lo), _mm_cmple_pd(o, lo_hi))));
    }
    alignas(16) double t[2];
    _mm_store_pd(t, vla0);
    la0 = t[0] < t[1] ? t[0] : t[1];
    _mm_store_pd(t, vla1);
    la1 = t[0] > t[1] ? t[0] : t[1];
    _mm_store_pd(t, vlo0);
    lo0 = t[0] < t[1] ? t[0] : t[1];
    _mm_store_pd(t, vlo1);
    lo1 = t[0] > t[1] ? t[0] : t[1];
    ok = _mm_movemask_pd(vok) == 3;
#endif
    for (; i < count; ++i) {
        la0 = lat[i] < la0 ? lat[i] : la0;
        la1 = lat[i] > la1 ? lat[i] : la1;
        lo0 = lon[i] < lo0 ? lon[i] : lo0;
        lo1 = lon[i] > lo1 ? lon[i] : lo1;
        ok = ok && lat[i] >= -90.0 && lat[i] <= 90.0 && lon[i] >= -180.0 && lon[i] <= 180.0;
    }
    ok = ok && lat[0] >= -90.0 && lat[0] <= 90.0 && lon[0] >= -180.0 && lon[0] <= 180.0;
    if (valid) *valid = ok;
    const double DEG = 0.017453292519943295, dla = (la1 - la0) * DEG, dlo = (lo1 - lo0) * DEG;
    const double b = dla * dla + dlo * dlo;
    return (ok && b == b) ? b : -1.0;  // unknown when a coordinate is unusable (the pass is refused elsewhere)
}

bool positions_well_formed(const int32_t *nd, int P, int n)
{
    const int32_t top = n - 1;
    int32_t acc = 0, zeros = 0;
    int q = 0;
#if defined(__SSE2__)
    __m128i vacc = _mm_setzero_si128(), vz = _mm_setzero_si128();
    const __m128i vtop = _mm_set1_epi32(top), zero = _mm_setzero_si128();
    for (; q + 8 <= P; q += 8) {
        const __m128i v0 = _mm_loadu_si128(reinterpret_cast<const __m128i *>(nd + q));
        const __m128i v1 = _mm_loadu_si128(reinterpret_cast<const __m128i *>(nd + q + 4));
        vacc = _mm_or_si128(vacc, _mm_or_si128(_mm_or_si128(v0, _mm_sub_epi32(vtop, v0)),
                                               _mm_or_si128(v1, _mm_sub_epi32(vtop, v1))));
        vz = _mm_sub_epi32(vz, _mm_add_epi32(_mm_cmpeq_epi32(v0, zero), _mm_cmpeq_epi32(v1, zero)));
    }
    alignas(16) int32_t t[4];
    _mm_store_si128(reinterpret_cast<__m128i *>(t), vacc);
    acc = t[0] | t[1] | t[2] | t[3];
    _mm_store_si128(reinterpret_cast<__m128i *>(t), vz);
    zeros = t[0] + t[1] + t[2] + t[3];
#endif
    for (; q < P; ++q) {
        const int32_t v = nd[q];
        acc |= v | (top - v);
        zeros += v == 0;
    }
    return acc >= 0 && zeros == P - n + 1;
}

__global__ void reset_search_kernel(int *done, int *moves, int batch)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < batch) {
        done[i] = 0;
        moves[i] = 0;
    }
}

}  // namespace
}  // namespace rt

using namespace rt;

#define RT_PROBLEM_GUARD(p)                       \
    if (!(p)) return RT_ERR_INVALID_ARG;          \
    rt_engine *e = (p)->e;                        \
    std::lock_guard<std::mutex> lk(e->mu);        \
    DeviceGuard dg(e->device)

extern "C" {

int rt_problem_create(rt_engine *e, int n, int batch, rt_problem **out)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!out || n < 2 || batch < 1) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_create: need n >= 2, batch >= 1");
    *out = nullptr;
    if (2 * n > kMaxPositions)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_create: n = %d exceeds the %d positions a move key can address", n,
                    kMaxPositions / 2);
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    rt_problem *p = new (std::nothrow) rt_problem();
    if (!p) return fail(e, RT_ERR_OUT_OF_MEMORY, "out of host memory");
    p->e = e;
    p->n = n;
    p->batch = batch;
    p->pitch = rt_matrix_pitch(n);
    p->pstride = (2 * n + 3) & ~3;  // P = n + R <= 2n - 1
    p->rstride = n;
    p->hist_cap = 4 * n;
    const size_t B = batch, N = n, PS = p->pstride, RS = p->rstride;
    const int n2 = next_pow2(n - 1);
    int rc = RT_OK;
    auto A = [&](int r) { if (rc == RT_OK) rc = r; };
    A(dev_alloc(p, p->lat, B * N));
    A(dev_alloc(p, p->lon, B * N));
    A(dev_alloc(p, p->xyz, B * 3 * N));
    A(dev_alloc(p, p->cells, B * N * p->pitch));
    A(dev_alloc(p, p->stops, B * N));
    A(dev_alloc(p, p->cap, B));
    A(dev_alloc(p, p->max_dist, B));
    A(dev_alloc(p, p->npos, B));
    A(dev_alloc(p, p->nroutes, B));
    A(dev_alloc(p, p->node, B * PS));
    A(dev_alloc(p, p->node_tmp, B * PS));
    A(dev_alloc(p, p->RA, B * PS));
    A(dev_alloc(p, p->RB, B * PS));
    A(dev_alloc(p, p->RC, B * PS));
    A(dev_alloc(p, p->RD, B * PS));
    A(dev_alloc(p, p->H, B * PS));
    A(dev_alloc(p, p->PX, B * 3 * PS));
    A(dev_alloc(p, p->E, B * PS));
    A(dev_alloc(p, p->SK, B * PS));
    A(dev_alloc(p, p->ER, B * PS));
    A(dev_alloc(p, p->Bv, B * PS));
    A(dev_alloc(p, p->Zv, B * PS));
    A(dev_alloc(p, p->route, B * PS));
    A(dev_alloc(p, p->rstart, B * (RS + 1)));
    A(dev_alloc(p, p->rload, B * RS));
    A(dev_alloc(p, p->rdist, B * RS));
    A(dev_alloc(p, p->cost, B));
    A(dev_alloc(p, p->feasible, B));
    A(dev_alloc(p, p->best, B + 1));  // + the coordinate flag of rt_problem_plan: keys and flag leave in one copy
    A(dev_alloc(p, p->done, B));
    A(dev_alloc(p, p->moves, B));
    A(dev_alloc(p, p->ticket, 1));
    A(dev_alloc(p, p->hist, B * p->hist_cap));
    A(dev_alloc(p, p->gkeys, B * n2));
    A(dev_alloc(p, p->gorder, B * n2));
    A(dev_alloc(p, p->gdemand, B * N));
    A(dev_alloc(p, p->gcap, B));
    A(dev_alloc(p, p->gmaxd, B));
    A(dev_alloc(p, p->g_order_out, B * N));
    A(dev_alloc(p, p->g_trip_off, B * (N + 1)));
    A(dev_alloc(p, p->g_ntrips, B));
    A(dev_alloc(p, p->g_status, B));
    if (rc == RT_OK) {
        p->coord_flag = reinterpret_cast<int *>(p->best + B);
        if (cudaMemset(p->coord_flag, 0, sizeof(unsigned long long)) != cudaSuccess)
            rc = fail(e, RT_ERR_CUDA, "rt_problem_create: cudaMemset failed");
    }
    if (rc != RT_OK) {
        for (int i = 0; i < p->n_allocs; ++i) cudaFree(p->all_allocs[i]);
        delete p;
        return rc;
    }
    *out = p;
    return RT_OK;
}

void rt_problem_destroy(rt_problem *p)
{
    if (!p) return;
    rt_engine *e = p->e;
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    cudaStreamSynchronize(e->stream);
    for (int i = 0; i < p->n_allocs; ++i) cudaFree(p->all_allocs[i]);
    for (cudaEvent_t ev : {p->ev_in[0], p->ev_in[1], p->ev_done})
        if (ev) cudaEventDestroy(ev);
    if (p->side_stream) {
        cudaStreamSynchronize(p->side_stream);
        cudaStreamDestroy(p->side_stream);
        cudaEventDestroy(p->ev_fork);
        cudaEventDestroy(p->ev_join);
    }
    for (int i = 0; i < p->n_xpeer_mapped; ++i) cudaIpcCloseMemHandle(p->xpeer_mapped[i]);
    if (p->xpeers) cudaFree(p->xpeers);
    if (p->xbuf) cudaFree(p->xbuf);
    for (int i = 0; i < p->n_xtab_mapped; ++i) cudaIpcCloseMemHandle(p->xtab_mapped[i]);
    if (p->xtab_peers) cudaFree(p->xtab_peers);
    if (p->xtab) cudaFree(p->xtab);
    for (rt::DevBuf *bf : {&p->cache_inter, &p->cache_intra, &p->cache_tiles, &p->cache_misc, &p->batch_buf})
        if (bf->p) cudaFree(bf->p);
    if (p->Gbuf.p) cudaFree(p->Gbuf.p);
    if (p->GTbuf.p) cudaFree(p->GTbuf.p);
    if (p->host_stage) cudaFreeHost(p->host_stage);
    delete p;
}

int rt_problem_set_coords(rt_problem *p, const double *lat, const double *lon)
{
    RT_PROBLEM_GUARD(p);
    if (!lat || !lon) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_set_coords: NULL argument");
    const long long total = static_cast<long long>(p->n) * p->batch;
    for (long long i = 0; i < total; ++i)
        if (!(lat[i] >= -90.0 && lat[i] <= 90.0) || !(lon[i] >= -180.0 && lon[i] <= 180.0))
            return fail(e, RT_ERR_INVALID_ARG, "coordinate %lld out of range or not finite", i);
    RT_CUDA(e, cudaMemcpyAsync(p->lat, lat, total * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(p->lon, lon, total * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    RT_TRY(launch_unit_vectors(e, p->lat, p->lon, p->n, p->batch, p->xyz, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    p->has_coords = true;
    p->has_matrix = false;
    p->cells_from_coords = false;
    p->angle2_bound = angle2_bound_of(lat, lon, total);
    return RT_OK;
}

int rt_problem_set_stops(rt_problem *p, const int32_t *stops, const int32_t *cap, const int64_t *max_dist)
{
    RT_PROBLEM_GUARD(p);
    if (!stops || !cap || !max_dist) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_set_stops: NULL argument");
    const long long total = static_cast<long long>(p->n) * p->batch;
    for (long long i = 0; i < total; ++i) {
        const int32_t *s = stops + 4 * i;
        if (s[0] > s[1] || s[2] < 0 || s[3] < 0 || s[0] < 0 || s[1] > (1 << 28) || s[2] > (1 << 28) || s[3] > (1 << 28))
            return fail(e, RT_ERR_INVALID_ARG,
                        "stop %lld: need 0 <= ready <= due <= 2^28, 0 <= service, demand <= 2^28 (got %d,%d,%d,%d)", i,
                        s[0], s[1], s[2], s[3]);
    }
    for (int b = 0; b < p->batch; ++b)
        if (cap[b] < 0 || max_dist[b] < 0) return fail(e, RT_ERR_INVALID_ARG, "negative capacity or distance limit");
    RT_CUDA(e, cudaMemcpyAsync(p->stops, stops, total * sizeof(int4), cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(p->cap, cap, p->batch * sizeof(int), cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(p->max_dist, max_dist, p->batch * sizeof(long long), cudaMemcpyHostToDevice, e->stream));
    p->dist_unlimited = std::all_of(max_dist, max_dist + p->batch, [](int64_t v) { return v >= (1LL << 61); });
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    p->has_stops = true;
    return RT_OK;
}

int rt_problem_build_matrix(rt_problem *p, double dist_scale, double time_scale)
{
    RT_PROBLEM_GUARD(p);
    if (!p->has_coords) return fail(e, RT_ERR_STATE, "rt_problem_build_matrix: set coordinates first");
    RT_TRY(check_scales(e, dist_scale, time_scale));
    RT_TRY(launch_matrix_cells(e, p->xyz, p->n, p->batch, 0, p->n, dist_scale, time_scale, p->cells, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    p->dist_scale = dist_scale;
    p->time_scale = time_scale;
    p->has_matrix = true;
    p->cells_from_coords = true;
    p->symmetric = true;
    p->supplied = 0;
    return RT_OK;
}

int rt_problem_set_matrix(rt_problem *p, int instance, const int32_t *dist, const int32_t *tim)
{
    RT_PROBLEM_GUARD(p);
    if (!dist || !tim || instance < 0 || instance >= p->batch)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_set_matrix: bad argument");
    const size_t cnt = static_cast<size_t>(p->n) * p->n;
    for (size_t i = 0; i < cnt; ++i)
        if (dist[i] < 0 || tim[i] < 0 || dist[i] >= (1 << 29) || tim[i] >= (1 << 29))
            return fail(e, RT_ERR_INVALID_ARG, "matrix entries must lie in [0, 2^29)");
    bool sym = true;
    for (int i = 0; i < p->n && sym; ++i)
        for (int j = 0; j < i; ++j)
            if (dist[static_cast<size_t>(i) * p->n + j] != dist[static_cast<size_t>(j) * p->n + i] ||
                tim[static_cast<size_t>(i) * p->n + j] != tim[static_cast<size_t>(j) * p->n + i]) {
                sym = false;
                break;
            }
    RT_TRY(reserve(e, e->plane_a[0], cnt * 4));
    RT_TRY(reserve(e, e->plane_b[0], cnt * 4));
    int *dd = static_cast<int *>(e->plane_a[0].p), *td = static_cast<int *>(e->plane_b[0].p);
    RT_CUDA(e, cudaMemcpyAsync(dd, dist, cnt * 4, cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(td, tim, cnt * 4, cudaMemcpyHostToDevice, e->stream));
    dim3 grid((p->pitch + 255) / 256, p->n);
    planes_to_cells_kernel<<<grid, 256, 0, e->stream>>>(dd, td, p->n, p->pitch,
                                                        p->cells + static_cast<size_t>(instance) * p->n * p->pitch);
    RT_CUDA(e, cudaGetLastError());
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    p->has_matrix = true;
    // one flag for the whole batch: the first installed matrix decides, later ones may only clear it
    p->symmetric = (p->cells_from_coords || p->supplied == 0) ? sym : (p->symmetric && sym);
    p->cells_from_coords = false;
    ++p->supplied;
    if (p->has_routes) {  // routes installed earlier refer to the old matrix
        RT_TRY(ensure_position_matrix(p));
        RT_TRY(state_refresh(p, e->stream));
        RT_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    return RT_OK;
}

int rt_problem_get_matrix(rt_problem *p, int instance, int32_t *dist, int32_t *tim)
{
    RT_PROBLEM_GUARD(p);
    if (!dist || !tim || instance < 0 || instance >= p->batch)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_get_matrix: bad argument");
    if (!p->has_matrix) return fail(e, RT_ERR_STATE, "rt_problem_get_matrix: no matrix yet");
    const size_t cnt = static_cast<size_t>(p->n) * p->n;
    RT_TRY(reserve(e, e->plane_a[0], cnt * 4));
    RT_TRY(reserve(e, e->plane_b[0], cnt * 4));
    int *dd = static_cast<int *>(e->plane_a[0].p), *td = static_cast<int *>(e->plane_b[0].p);
    dim3 grid((p->n + 255) / 256, p->n);
    cells_to_planes_kernel<<<grid, 256, 0, e->stream>>>(p->cells + static_cast<size_t>(instance) * p->n * p->pitch, p->n,
                                                        p->pitch, dd, td);
    RT_CUDA(e, cudaGetLastError());
    RT_CUDA(e, cudaMemcpyAsync(dist, dd, cnt * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(tim, td, cnt * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    return RT_OK;
}

int rt_problem_greedy(rt_problem *p, const double *demand, const double *cap, const double *max_dist, int use_tw,
                      int32_t *order_out, int32_t *trip_off_out, int32_t *n_trips_out, int32_t *status_out)
{
    RT_PROBLEM_GUARD(p);
    if (!demand || !cap || !max_dist || !order_out || !trip_off_out || !n_trips_out || !status_out)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_greedy: NULL argument");
    if (!p->has_matrix) return fail(e, RT_ERR_STATE, "rt_problem_greedy: no matrix yet");
    if (use_tw && !p->has_stops) return fail(e, RT_ERR_STATE, "rt_problem_greedy: time windows requested but no stops set");
    const size_t B = p->batch, N = p->n;
    RT_CUDA(e, cudaMemcpyAsync(p->gdemand, demand, B * N * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(p->gcap, cap, B * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(p->gmaxd, max_dist, B * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    RT_TRY(greedy_launch(p, use_tw != 0, e->stream));
    // results: order [B][n-1] (device rows have stride n), trip offsets [B][n+1] -> host [B][n]
    std::vector<int> tmp(B * (N + 1));
    RT_CUDA(e, cudaMemcpy2DAsync(order_out, (N - 1) * 4, p->g_order_out, N * 4, (N - 1) * 4, B, cudaMemcpyDeviceToHost,
                                 e->stream));
    RT_CUDA(e, cudaMemcpyAsync(tmp.data(), p->g_trip_off, B * (N + 1) * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(n_trips_out, p->g_ntrips, B * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(status_out, p->g_status, B * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    for (size_t b = 0; b < B; ++b)
        for (size_t k = 0; k < N; ++k) trip_off_out[b * N + k] = k <= static_cast<size_t>(n_trips_out[b]) ? tmp[b * (N + 1) + k] : 0;
    return RT_OK;
}

int rt_greedy_trips(rt_engine *e, const double *D, int n, const double *demand, double cap, double max_dist,
                    int32_t *order_out, int32_t *trip_off_out, int32_t *n_trips_out)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!D || !demand || !order_out || !trip_off_out || !n_trips_out || n < 1)
        return fail(e, RT_ERR_INVALID_ARG, "rt_greedy_trips: bad argument");
    *n_trips_out = 0;
    trip_off_out[0] = 0;
    if (n == 1) return RT_OK;
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    const size_t N = n, n2 = next_pow2(n - 1);
    // one scratch block: D | demand | cap,maxd | keys | order | order_out | trip_off | ntrips,status
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~size_t(255); return o; };
    const size_t oD = take(N * N * 8), oDem = take(N * 8), oLim = take(16), oKeys = take(n2 * 8), oOrd = take(n2 * 4),
                 oOut = take(N * 4), oOff = take((N + 1) * 4), oRes = take(8);
    RT_TRY(reserve(e, e->plane_c[0], off));
    char *base = static_cast<char *>(e->plane_c[0].p);
    const double lim[2] = {cap, max_dist};
    RT_CUDA(e, cudaMemcpyAsync(base + oD, D, N * N * 8, cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(base + oDem, demand, N * 8, cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(base + oLim, lim, 16, cudaMemcpyHostToDevice, e->stream));
    RT_TRY(greedy_f64_launch(e, reinterpret_cast<double *>(base + oD), n, reinterpret_cast<double *>(base + oDem),
                             reinterpret_cast<double *>(base + oLim), reinterpret_cast<double *>(base + oLim) + 1,
                             reinterpret_cast<unsigned long long *>(base + oKeys), reinterpret_cast<int *>(base + oOrd),
                             reinterpret_cast<int *>(base + oOut), reinterpret_cast<int *>(base + oOff),
                             reinterpret_cast<int *>(base + oRes), reinterpret_cast<int *>(base + oRes) + 1, e->stream));
    int res[2] = {0, 0};
    std::vector<int> offs(N + 1);
    RT_CUDA(e, cudaMemcpyAsync(order_out, base + oOut, (N - 1) * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(offs.data(), base + oOff, (N + 1) * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(res, base + oRes, 8, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    *n_trips_out = res[0];
    for (int k = 0; k <= res[0]; ++k) trip_off_out[k] = offs[k];
    if (res[1] != RT_OK)
        return fail(e, res[1], "a stop fits no trip on its own (capacity or maximum_distance); %d trips built", res[0]);
    return RT_OK;
}

int rt_problem_set_routes(rt_problem *p, const int32_t *order, const int32_t *trip_off, const int32_t *n_trips)
{
    RT_PROBLEM_GUARD(p);
    if (!order || !trip_off || !n_trips) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_set_routes: NULL argument");
    if (!p->has_matrix || !p->has_stops) return fail(e, RT_ERR_STATE, "rt_problem_set_routes: matrix and stops first");
    const size_t B = p->batch, N = p->n, PS = p->pstride;
    std::vector<int> node(B * PS, 0), npos(B);
    std::vector<char> seen(N);
    for (size_t b = 0; b < B; ++b) {
        const int R = n_trips[b];
        const int32_t *off = trip_off + b * N, *ord = order + b * (N - 1);
        if (R < 1 || R > static_cast<int>(N) - 1 || off[0] != 0 || off[R] != static_cast<int>(N) - 1)
            return fail(e, RT_ERR_INVALID_ARG, "instance %zu: trips must cover all %zu customers exactly once", b, N - 1);
        std::fill(seen.begin(), seen.end(), 0);
        int *nd = node.data() + b * PS, P = 0;
        nd[P++] = 0;
        for (int k = 0; k < R; ++k) {
            if (off[k + 1] < off[k] || off[k + 1] > static_cast<int>(N) - 1)
                return fail(e, RT_ERR_INVALID_ARG, "instance %zu: trip offsets must not decrease or pass %zu", b, N - 1);
            for (int i = off[k]; i < off[k + 1]; ++i) {
                const int c = ord[i];
                if (c < 1 || c >= static_cast<int>(N) || seen[c])
                    return fail(e, RT_ERR_INVALID_ARG, "instance %zu: customer %d out of range or repeated", b, c);
                seen[c] = 1;
                nd[P++] = c;
            }
            nd[P++] = 0;
        }
        npos[b] = P;
    }
    p->pmax = *std::max_element(npos.begin(), npos.end());
    p->pmin = *std::min_element(npos.begin(), npos.end());
    RT_TRY(ensure_position_matrix(p));
    RT_CUDA(e, cudaMemcpyAsync(p->node, node.data(), B * PS * 4, cudaMemcpyHostToDevice, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(p->npos, npos.data(), B * 4, cudaMemcpyHostToDevice, e->stream));
    reset_search_kernel<<<(p->batch + 255) / 256, 256, 0, e->stream>>>(p->done, p->moves, p->batch);
    RT_CUDA(e, cudaGetLastError());
    RT_TRY(state_refresh(p, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    p->has_routes = true;
    return RT_OK;
}

int rt_problem_get_routes(rt_problem *p, int32_t *order_out, int32_t *trip_off_out, int32_t *n_trips_out)
{
    RT_PROBLEM_GUARD(p);
    if (!order_out || !trip_off_out || !n_trips_out) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_get_routes: NULL argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_get_routes: no routes set");
    const size_t B = p->batch, N = p->n, PS = p->pstride;
    std::vector<int> node(B * PS), npos(B);
    RT_CUDA(e, cudaMemcpyAsync(node.data(), p->node, B * PS * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(npos.data(), p->npos, B * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    for (size_t b = 0; b < B; ++b) {
        const int *nd = node.data() + b * PS;
        int32_t *off = trip_off_out + b * N, *ord = order_out + b * (N - 1);
        int R = 0, c = 0;
        off[0] = 0;
        for (int q = 1; q < npos[b]; ++q) {
            if (nd[q] == 0) off[++R] = c;
            else ord[c++] = nd[q];
        }
        n_trips_out[b] = R;
    }
    return RT_OK;
}

int rt_problem_get_state(rt_problem *p, int instance, int32_t *npos_out, int32_t *node, int32_t *route, int32_t *Bv,
                         int32_t *Zv, int64_t *rload, int64_t *rdist, int64_t *cost, int32_t *feasible)
{
    RT_PROBLEM_GUARD(p);
    if (instance < 0 || instance >= p->batch || !npos_out)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_get_state: bad argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_get_state: no routes set");
    const size_t PS = p->pstride, RS = p->rstride, po = instance * PS, ro = instance * RS;
    auto D2H = [&](void *dst, const void *src, size_t bytes) {
        return dst ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, e->stream) : cudaSuccess;
    };
    RT_CUDA(e, D2H(npos_out, p->npos + instance, 4));
    RT_CUDA(e, D2H(node, p->node + po, PS * 4));
    RT_CUDA(e, D2H(route, p->route + po, PS * 4));
    RT_CUDA(e, D2H(Bv, p->Bv + po, PS * 4));
    RT_CUDA(e, D2H(Zv, p->Zv + po, PS * 4));
    RT_CUDA(e, D2H(rload, p->rload + ro, RS * 8));
    RT_CUDA(e, D2H(rdist, p->rdist + ro, RS * 8));
    RT_CUDA(e, D2H(cost, p->cost + instance, 8));
    RT_CUDA(e, D2H(feasible, p->feasible + instance, 4));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    return RT_OK;
}

int rt_problem_build_matrix_dev(rt_problem *p, double dist_scale, double time_scale, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!p->has_coords) return fail(e, RT_ERR_STATE, "rt_problem_build_matrix_dev: set coordinates first");
    RT_TRY(check_scales(e, dist_scale, time_scale));
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    RT_TRY(launch_unit_vectors(e, p->lat, p->lon, p->n, p->batch, p->xyz, st));
    RT_TRY(launch_matrix_cells(e, p->xyz, p->n, p->batch, 0, p->n, dist_scale, time_scale, p->cells, st));
    p->dist_scale = dist_scale;
    p->time_scale = time_scale;
    p->has_matrix = true;
    p->cells_from_coords = true;
    p->symmetric = true;
    p->supplied = 0;
    return RT_OK;
}

int rt_problem_refresh_dev(rt_problem *p, int what, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_refresh_dev: no routes set");
    if (!(what & 7)) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_refresh_dev: what must have bit 0, 1 or 2 set");
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    if (what & 4) {  // the route-order matrix itself, whether or not the move kernels read one
        RT_TRY(position_matrix_refresh(p, st));
        what &= ~1;
        if (!(what & 3)) return RT_OK;
    }
    if (eval_without_matrix(p)) what &= ~1;  // bit 0: "what the move kernels need of the matrix" — nothing here
    if (!(what & 3)) return RT_OK;
    // both: the state tables run on the side stream underneath the matrix kernel; the tables alone re-derive the
    // position-ordered unit vectors themselves, so bit 1 never depends on an earlier bit-0 refresh
    if ((what & 3) == 3) return state_refresh(p, st);
    if (what & 1) RT_TRY(position_matrix_refresh(p, st));
    if (what & 2) RT_TRY(state_tables_refresh(p, st, /*with_vectors=*/true));
    return RT_OK;
}

int rt_problem_plan_dev(rt_problem *p, double dist_scale, double time_scale, int families, void *best_keys_d,
                        void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!best_keys_d || !(families & 7)) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_plan_dev: bad argument");
    if (!p->has_coords || !p->has_stops || !p->has_routes)
        return fail(e, RT_ERR_STATE, "rt_problem_plan_dev: coordinates, stops and routes must be on the device");
    RT_TRY(check_scales(e, dist_scale, time_scale));
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    p->ext_best = static_cast<unsigned long long *>(best_keys_d);  // reset by the refresh, reduced into directly
    RT_TRY(position_unit_vectors_launch(p, st));  // unit vectors, already in route order (+ the route scan)
    p->dist_scale = dist_scale;
    p->time_scale = time_scale;
    p->cells_from_coords = true;
    p->symmetric = true;  // whatever matrix was installed before, this pass runs on SPEC §2 cells
    p->supplied = 0;
    int rc = state_refresh(p, st, nullptr, /*vectors_ready=*/true);
    // a replica of a multi-GPU search scores the shard set by rt_problem_set_shard (default: everything)
    if (rc == RT_OK)
        rc = eval_moves_launch(p, families, p->shard_index, p->shard_count, static_cast<unsigned long long *>(best_keys_d),
                               true, st);
    p->ext_best = nullptr;
    p->ext_best_clean = false;
    return rc;
}

// Enqueues one pass (inputs -> best keys in the problem's pinned staging buffer).  Synchronous form: what the matrix
// needs travels on the compute stream, the rest on the copy stream.  Asynchronous form (rt_problem_plan_begin): every
// input travels on the copy stream, so that the copies of this request overlap whatever an earlier request — of
// another problem of the same engine — still has running on the compute stream.
static int plan_enqueue(rt_problem *p, const double *lat, const double *lon, const int32_t *stops, const int32_t *cap,
                        const int64_t *max_dist, const int32_t *node, const int32_t *npos, double dist_scale,
                        double time_scale, int families, bool async, const char *who,
                        unsigned long long *keys_dev = nullptr, cudaStream_t caller = nullptr)
{
    // keys_dev != nullptr: the caller's-stream form (rt_problem_plan_upload_dev) — every copy and kernel goes to
    // `caller`, the keys stay on the device in signed-order form and nothing is staged for the host.
    rt_engine *e = p->e;
    if (!lat || !lon || !stops || !cap || !max_dist || !node || !npos || !(families & 7))
        return fail(e, RT_ERR_INVALID_ARG, "%s: bad argument", who);
    if (p->plan_pending) return fail(e, RT_ERR_STATE, "%s: the previous rt_problem_plan_begin of this problem has not been ended", who);
    RT_TRY(check_scales(e, dist_scale, time_scale));
    const size_t B = p->batch, N = p->n, PS = p->pstride;
    for (size_t b = 0; b < B; ++b)
        if (npos[b] < static_cast<int>(N) + 1 || npos[b] > static_cast<int>(PS) || node[b * PS] != 0 ||
            node[b * PS + npos[b] - 1] != 0)
            return fail(e, RT_ERR_INVALID_ARG, "instance %zu: malformed position array", b);
    if (!p->ev_in[0]) {
        for (cudaEvent_t *ev : {&p->ev_in[0], &p->ev_in[1], &p->ev_done})
            RT_CUDA(e, cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
    }
    cudaStream_t st = keys_dev ? caller : e->stream, cs = keys_dev ? caller : e->copy_stream;
    static const bool dbg = getenv("ROUTEST_DEBUG_PLAN") != nullptr;  // prints where the call's host time goes
    auto now = []() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; };
    p->plan_t[0] = dbg ? now() : 0;
    // What the matrix needs (coordinates, routes) goes first; the stop records and limits, which only the state tables
    // read, follow on the copy stream underneath the unit-vector and matrix kernels.
    cudaStream_t first = async ? cs : st;
    RT_CUDA(e, cudaMemcpyAsync(p->lat, lat, B * N * 8, cudaMemcpyHostToDevice, first));
    RT_CUDA(e, cudaMemcpyAsync(p->lon, lon, B * N * 8, cudaMemcpyHostToDevice, first));
    RT_CUDA(e, cudaMemcpyAsync(p->node, node, B * PS * 4, cudaMemcpyHostToDevice, first));
    RT_CUDA(e, cudaMemcpyAsync(p->npos, npos, B * 4, cudaMemcpyHostToDevice, first));
    RT_CUDA(e, cudaEventRecord(p->ev_in[0], first));
    if (async) RT_CUDA(e, cudaStreamWaitEvent(st, p->ev_in[0], 0));      // the kernels wait for their inputs
    else RT_CUDA(e, cudaStreamWaitEvent(cs, p->ev_in[0], 0));            // earlier work on the buffers has been ordered
    RT_CUDA(e, cudaMemcpyAsync(p->stops, stops, B * N * 16, cudaMemcpyHostToDevice, cs));
    RT_CUDA(e, cudaMemcpyAsync(p->cap, cap, B * 4, cudaMemcpyHostToDevice, cs));
    RT_CUDA(e, cudaMemcpyAsync(p->max_dist, max_dist, B * 8, cudaMemcpyHostToDevice, cs));
    p->dist_unlimited = std::all_of(max_dist, max_dist + B, [](int64_t v) { return v >= (1LL << 61); });
    RT_CUDA(e, cudaEventRecord(p->ev_in[1], cs));
    p->plan_t[1] = dbg ? now() : 0;
    // The unit vectors go out first (the kernel itself falls back to the depot for a node id it cannot index), so the
    // GPU has work while the host checks what every other kernel indexes with: each position holds a node id below n
    // and there are exactly n-1 customers among them (so the route count is P - n and every table fits).  That a
    // customer is not repeated is not checked here: a repeat gives a meaningless plan, not a fault.  Coordinates
    // are range-checked by that kernel (off the host's critical path); stop records are taken as they are
    // (rt_problem_set_stops is the validating entry point) — out-of-range values cannot fault a kernel.
    // sizes first (they depend on npos[] only, checked above): the vector kernel carries the route scan of large instances
    p->pmax = 0;
    p->pmin = npos[0];
    for (size_t b = 0; b < B; ++b) {
        p->pmax = std::max(p->pmax, npos[b]);
        p->pmin = std::min(p->pmin, npos[b]);
    }
    p->cells_from_coords = true;
    p->symmetric = true;
    RT_TRY(ensure_position_matrix(p));
    p->plan_flag_seq = ++p->plan_seq;  // the flag is written with this call's number: no reset launch needed
    p->ext_best = keys_dev;  // (may be null) reset by the scan that rides along, reduced into directly by the evaluation
    RT_TRY(position_unit_vectors_launch(p, st, keys_dev ? nullptr : p->coord_flag, p->plan_flag_seq));
    for (size_t b = 0; b < B; ++b) {
        if (!positions_well_formed(node + b * PS, npos[b], static_cast<int>(N))) {
            p->has_routes = false;  // the copy above has already replaced the previous routes
            p->ext_best = nullptr;
            p->scan_done = false;
            cudaStreamSynchronize(st);  // the caller's buffers are still being read
            cudaStreamSynchronize(cs);
            return fail(e, RT_ERR_INVALID_ARG, "instance %zu: position array must hold node ids below %zu with %zu customers", b, N, N - 1);
        }
    }
    // the box of the coordinates, and — for the caller's-stream form, which has no later host step that could report the
    // device-side flag — their validity, while the GPU is busy with the copies and the vector kernel
    bool coords_ok = true;
    p->angle2_bound = angle2_bound_of(lat, lon, static_cast<long long>(B * N), &coords_ok);
    if (keys_dev && !coords_ok) {
        p->has_coords = p->has_routes = false;
        p->ext_best = nullptr;
        p->scan_done = false;
        cudaStreamSynchronize(st);  // the caller's buffers are still being read
        return fail(e, RT_ERR_INVALID_ARG, "%s: a coordinate is out of range or not finite", who);
    }
    p->plan_t[2] = dbg ? now() : 0;
    // the matrix of this pass is built in route order by the refresh below (state.cu); the node-order copy that
    // rt_problem_build_matrix keeps for the greedy and for rt_problem_get_matrix is not needed here
    p->has_coords = p->has_stops = p->has_routes = true;
    p->has_matrix = false;
    p->cells_from_coords = true;
    p->symmetric = true;
    p->supplied = 0;
    p->dist_scale = dist_scale;
    p->time_scale = time_scale;
    RT_TRY(ensure_position_matrix(p));
    int rc_refresh = state_refresh(p, st, keys_dev ? nullptr : p->ev_in[1], /*vectors_ready=*/true);
    // a replica of a multi-GPU search scores its own shard (rt_problem_set_shard); the default is everything
    if (keys_dev) {
        if (rc_refresh == RT_OK)
            rc_refresh = eval_moves_launch(p, families, p->shard_index, p->shard_count, keys_dev, true, st);
        p->ext_best = nullptr;
        p->ext_best_clean = false;
        return rc_refresh;
    }
    RT_TRY(rc_refresh);
    RT_TRY(eval_moves_launch(p, families, p->shard_index, p->shard_count, p->best, false, st));
    RT_TRY(stage(p, (B + 1) * 8));
    RT_CUDA(e, cudaMemcpyAsync(p->host_stage, p->best, (B + 1) * 8, cudaMemcpyDeviceToHost, st));  // keys + flag, one copy
    RT_CUDA(e, cudaEventRecord(p->ev_done, st));
    p->plan_t[3] = dbg ? now() : 0;
    p->plan_pending = true;
    return RT_OK;
}

// Waits for the pass enqueued last and hands its keys out.
// have_lock: the caller holds the engine's lock (rt_problem_plan).  rt_problem_plan_end does not — its wait must not
// keep other threads from enqueueing — so a failure takes the lock just for the moment it writes the engine's error
// text; the pass's own bookkeeping (plan_pending, the staging buffer) belongs to this problem alone, one pass per
// problem being in flight by contract.
static int plan_collect(rt_problem *p, uint64_t *best_keys_out, const char *who, bool have_lock)
{
    rt_engine *e = p->e;
    auto failed = [&](int code, const char *msg) {
        if (have_lock) return fail(e, code, "%s: %s", who, msg);
        std::lock_guard<std::mutex> lk(e->mu);
        return fail(e, code, "%s: %s", who, msg);
    };
    if (!best_keys_out) return failed(RT_ERR_INVALID_ARG, "NULL argument");
    if (!p->plan_pending) return failed(RT_ERR_STATE, "no pass has been begun on this problem");
    p->plan_pending = false;
    if (cudaEventSynchronize(p->ev_done) != cudaSuccess) return failed(RT_ERR_CUDA, "waiting for the pass failed");
    if (p->plan_t[0] != 0) {
        static double acc[4];
        static int calls;
        timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        const double t_end = ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
        acc[0] += p->plan_t[1] - p->plan_t[0], acc[1] += p->plan_t[2] - p->plan_t[1], acc[2] += p->plan_t[3] - p->plan_t[2];
        acc[3] += t_end - p->plan_t[3];
        if (++calls % 50 == 0) {
            fprintf(stderr, "[routest] plan: copies issued %.1f us, first launch + validation %.1f, launches issued %.1f, wait %.1f\n",
                    acc[0] / 50, acc[1] / 50, acc[2] / 50, acc[3] / 50);
            acc[0] = acc[1] = acc[2] = acc[3] = 0;
        }
    }
    const size_t B = p->batch;
    const unsigned long long *res_h = static_cast<const unsigned long long *>(p->host_stage);
    memcpy(best_keys_out, res_h, B * 8);
    if (*reinterpret_cast<const int *>(res_h + B) == p->plan_flag_seq) {
        p->has_coords = p->has_routes = false;
        return failed(RT_ERR_INVALID_ARG, "a coordinate is out of range or not finite");
    }
    return RT_OK;
}

int rt_problem_plan(rt_problem *p, const double *lat, const double *lon, const int32_t *stops, const int32_t *cap,
                    const int64_t *max_dist, const int32_t *node, const int32_t *npos, double dist_scale,
                    double time_scale, int families, uint64_t *best_keys_out)
{
    RT_PROBLEM_GUARD(p);
    if (!best_keys_out) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_plan: bad argument");
    RT_TRY(plan_enqueue(p, lat, lon, stops, cap, max_dist, node, npos, dist_scale, time_scale, families, false,
                        "rt_problem_plan"));
    return plan_collect(p, best_keys_out, "rt_problem_plan", /*have_lock=*/true);
}

int rt_problem_plan_begin(rt_problem *p, const double *lat, const double *lon, const int32_t *stops, const int32_t *cap,
                          const int64_t *max_dist, const int32_t *node, const int32_t *npos, double dist_scale,
                          double time_scale, int families)
{
    RT_PROBLEM_GUARD(p);
    return plan_enqueue(p, lat, lon, stops, cap, max_dist, node, npos, dist_scale, time_scale, families, true,
                        "rt_problem_plan_begin");
}

int rt_problem_plan_end(rt_problem *p, uint64_t *best_keys_out)
{
    if (!p) return RT_ERR_INVALID_ARG;
    DeviceGuard dg(p->e->device);  // no engine lock: the wait must not keep other requests from being enqueued
    return plan_collect(p, best_keys_out, "rt_problem_plan_end", /*have_lock=*/false);
}

int rt_problem_plan_upload_dev(rt_problem *p, const double *lat, const double *lon, const int32_t *stops,
                               const int32_t *cap, const int64_t *max_dist, const int32_t *node, const int32_t *npos,
                               double dist_scale, double time_scale, int families, void *best_keys_d, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!best_keys_d) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_plan_upload_dev: bad argument");
    DeviceGuard dg(e->device);
    return plan_enqueue(p, lat, lon, stops, cap, max_dist, node, npos, dist_scale, time_scale, families, false,
                        "rt_problem_plan_upload_dev", static_cast<unsigned long long *>(best_keys_d),
                        pick_stream(e, stream));
}

int rt_problem_get_position_matrix(rt_problem *p, int instance, int32_t *dist, int32_t *tim)
{
    RT_PROBLEM_GUARD(p);
    if (!dist || !tim || instance < 0 || instance >= p->batch)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_get_position_matrix: bad argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_get_position_matrix: no routes set");
    if (p->shard_count != 1) return fail(e, RT_ERR_STATE, "rt_problem_get_position_matrix needs shard 0/1");
    if (eval_without_matrix(p)) RT_TRY(position_matrix_refresh(p, e->stream));  // not kept fresh when nothing reads it
    int P = 0;
    RT_CUDA(e, cudaMemcpy(&P, p->npos + instance, 4, cudaMemcpyDeviceToHost));
    const size_t cnt = static_cast<size_t>(P) * P;
    RT_TRY(reserve(e, e->plane_a[0], cnt * 4));
    RT_TRY(reserve(e, e->plane_b[0], cnt * 4));
    int *dd = static_cast<int *>(e->plane_a[0].p), *td = static_cast<int *>(e->plane_b[0].p);
    dim3 grid((P + 255) / 256, P);
    cells_to_planes_kernel<<<grid, 256, 0, e->stream>>>(
        position_matrix(p) + static_cast<size_t>(instance) * p->grows * p->gpitch, P, p->gpitch, dd, td);
    RT_CUDA(e, cudaGetLastError());
    RT_CUDA(e, cudaMemcpyAsync(dist, dd, cnt * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(tim, td, cnt * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    return RT_OK;
}

int rt_problem_position_stride(const rt_problem *p) { return p ? p->pstride : 0; }

// true when the rows of G this replica keeps fresh cover shard (index, count)
static bool shard_covered(const rt_problem *p, int index, int count)
{
    return p->shard_count == 1 || (p->shard_count == count && p->shard_index == index);
}

int rt_problem_set_shard(rt_problem *p, int shard_index, int shard_count)
{
    RT_PROBLEM_GUARD(p);
    if (shard_count < 1 || shard_index < 0 || shard_index >= shard_count)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_set_shard: bad argument");
    p->shard_index = shard_index;
    p->shard_count = shard_count;
    if (p->has_routes) {
        RT_TRY(state_refresh(p, e->stream));
        RT_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    return RT_OK;
}

int rt_problem_eval_moves(rt_problem *p, int families, int shard_index, int shard_count, uint64_t *best_keys_out)
{
    RT_PROBLEM_GUARD(p);
    if (!best_keys_out || shard_count < 1 || shard_index < 0 || shard_index >= shard_count || !(families & 7))
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_eval_moves: bad argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_eval_moves: no routes set");
    if (!shard_covered(p, shard_index, shard_count))
        return fail(e, RT_ERR_STATE, "rt_problem_eval_moves: shard %d/%d is not the one set by rt_problem_set_shard", shard_index, shard_count);
    RT_TRY(eval_moves_launch(p, families, shard_index, shard_count, p->best, false, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(best_keys_out, p->best, p->batch * 8, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    return RT_OK;
}

int rt_problem_eval_moves_dev(rt_problem *p, int families, int shard_index, int shard_count, void *best_keys_d,
                              void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!best_keys_d || shard_count < 1 || shard_index < 0 || shard_index >= shard_count || !(families & 7))
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_eval_moves_dev: bad argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_eval_moves_dev: no routes set");
    if (!shard_covered(p, shard_index, shard_count))
        return fail(e, RT_ERR_STATE, "rt_problem_eval_moves_dev: shard %d/%d is not the one set by rt_problem_set_shard",
                    shard_index, shard_count);
    DeviceGuard dg(e->device);
    return eval_moves_launch(p, families, shard_index, shard_count, static_cast<unsigned long long *>(best_keys_d), true,
                             pick_stream(e, stream));
}

int rt_problem_eval_dump(rt_problem *p, int instance, int family, int32_t *delta_out, uint8_t *feas_out)
{
    RT_PROBLEM_GUARD(p);
    if (!delta_out || !feas_out || instance < 0 || instance >= p->batch || family < 0 || family > 2)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_eval_dump: bad argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_eval_dump: no routes set");
    if (p->shard_count != 1) return fail(e, RT_ERR_STATE, "rt_problem_eval_dump needs the whole route-order matrix (shard 0/1)");
    int P = 0;
    RT_CUDA(e, cudaMemcpy(&P, p->npos + instance, 4, cudaMemcpyDeviceToHost));
    const size_t cnt = static_cast<size_t>(P) * P;
    RT_TRY(reserve(e, e->plane_a[0], cnt * 4));
    RT_TRY(reserve(e, e->plane_b[0], cnt));
    int *dd = static_cast<int *>(e->plane_a[0].p);
    unsigned char *fd = static_cast<unsigned char *>(e->plane_b[0].p);
    RT_CUDA(e, cudaMemsetAsync(dd, 0, cnt * 4, e->stream));
    RT_CUDA(e, cudaMemsetAsync(fd, 2, cnt, e->stream));  // 2 = not a candidate
    RT_TRY(eval_dump_launch(p, instance, family, dd, fd, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(delta_out, dd, cnt * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(feas_out, fd, cnt, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    return RT_OK;
}

int rt_problem_apply_moves(rt_problem *p, const uint64_t *keys)
{
    RT_PROBLEM_GUARD(p);
    if (!keys) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_apply_moves: NULL argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_apply_moves: no routes set");
    // the per-instance "search finished" flags belong to the search loops; a direct apply neither obeys nor leaves them
    RT_TRY(stage(p, static_cast<size_t>(p->batch) * 4));
    int *done_h = static_cast<int *>(p->host_stage);
    RT_CUDA(e, cudaMemsetAsync(p->done, 0, p->batch * 4, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(p->best, keys, p->batch * 8, cudaMemcpyHostToDevice, e->stream));
    RT_TRY(apply_moves_launch(p, p->best, false, false, e->stream));
    RT_CUDA(e, cudaMemcpyAsync(done_h, p->done, p->batch * 4, cudaMemcpyDeviceToHost, e->stream));
    RT_CUDA(e, cudaMemsetAsync(p->done, 0, p->batch * 4, e->stream));
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    for (int b = 0; b < p->batch; ++b)
        if (done_h[b] == 2)
            return fail(e, RT_ERR_INVALID_ARG, "rt_problem_apply_moves: the key of instance %d is not a move of its current routes (left unchanged)", b);
    return RT_OK;
}

int rt_problem_apply_moves_dev(rt_problem *p, const void *keys_d, int improving_only, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!keys_d) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_apply_moves_dev: NULL argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_apply_moves_dev: no routes set");
    DeviceGuard dg(e->device);
    return apply_moves_launch(p, static_cast<const unsigned long long *>(keys_d), true, improving_only != 0,
                              pick_stream(e, stream));
}

int64_t rt_problem_table_entries(rt_problem *p)
{
    if (!p) return 0;
    rt_engine *e = p->e;
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    if (!p->has_routes || !batch_usable(p) || batch_ensure(p) != RT_OK) return 0;
    return static_cast<int64_t>(batch_table_entries(p));
}

int rt_problem_eval_table_dev(rt_problem *p, int families, int shard_index, int shard_count, void *table_d, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!table_d || shard_count < 1 || shard_index < 0 || shard_index >= shard_count || !(families & 7))
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_eval_table_dev: bad argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_eval_table_dev: no routes set");
    if (!batch_usable(p)) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_eval_table_dev handles at most 1024 routes per instance");
    if (!shard_covered(p, shard_index, shard_count))
        return fail(e, RT_ERR_STATE, "rt_problem_eval_table_dev: shard %d/%d is not the one set by rt_problem_set_shard",
                    shard_index, shard_count);
    DeviceGuard dg(e->device);
    RT_TRY(batch_ensure(p));  // no-op after rt_problem_table_entries
    return eval_table_launch(p, families, shard_index, shard_count, static_cast<unsigned long long *>(table_d), true,
                             pick_stream(e, stream));
}

int rt_problem_apply_batch_dev(rt_problem *p, const void *table_d, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!table_d) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_apply_batch_dev: NULL argument");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_apply_batch_dev: no routes set");
    if (!batch_usable(p)) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_apply_batch_dev handles at most 1024 routes per instance");
    DeviceGuard dg(e->device);
    RT_TRY(batch_ensure(p));
    return batch_select_apply(p, static_cast<const unsigned long long *>(table_d), true, pick_stream(e, stream));
}

int rt_problem_search_status(rt_problem *p, int32_t *done_out, int32_t *moves_out, int64_t *cost_out)
{
    RT_PROBLEM_GUARD(p);
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_search_status: no routes set");
    RT_CUDA(e, cudaDeviceSynchronize());  // _dev calls may have used any stream
    if (done_out) RT_CUDA(e, cudaMemcpy(done_out, p->done, p->batch * 4, cudaMemcpyDeviceToHost));
    if (moves_out) RT_CUDA(e, cudaMemcpy(moves_out, p->moves, p->batch * 4, cudaMemcpyDeviceToHost));
    if (cost_out) RT_CUDA(e, cudaMemcpy(cost_out, p->cost, p->batch * 8, cudaMemcpyDeviceToHost));
    return RT_OK;
}

static int local_search_impl(rt_problem *p, int families, int max_moves, int mode, int32_t *moves_out,
                             int64_t *cost_out, uint64_t *keys_out, int keys_capacity);

int rt_problem_local_search(rt_problem *p, int families, int max_moves, int32_t *moves_out, int64_t *cost_out,
                            uint64_t *keys_out, int keys_capacity)
{
    return local_search_impl(p, families, max_moves, 0, moves_out, cost_out, keys_out, keys_capacity);
}

int rt_problem_local_search_full(rt_problem *p, int families, int max_moves, int32_t *moves_out, int64_t *cost_out,
                                 uint64_t *keys_out, int keys_capacity)
{
    return local_search_impl(p, families, max_moves, 1, moves_out, cost_out, keys_out, keys_capacity);
}

int rt_problem_local_search_batched(rt_problem *p, int families, int max_iterations, int32_t *moves_out,
                                    int64_t *cost_out, uint64_t *keys_out, int keys_capacity)
{
    return local_search_impl(p, families, max_iterations, 2, moves_out, cost_out, keys_out, keys_capacity);
}

static int local_search_impl(rt_problem *p, int families, int max_moves, int mode, int32_t *moves_out,
                             int64_t *cost_out, uint64_t *keys_out, int keys_capacity)
{
    RT_PROBLEM_GUARD(p);
    if (!(families & 7) || max_moves < 0) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_local_search: bad argument");
    if (p->shard_count != 1) return fail(e, RT_ERR_STATE, "rt_problem_local_search runs on an unsharded replica (shard 0/1)");
    if (!p->has_routes) return fail(e, RT_ERR_STATE, "rt_problem_local_search: no routes set");
    RT_TRY(stage(p, static_cast<size_t>(p->batch) * 4));
    int *done_h = static_cast<int *>(p->host_stage);
    reset_search_kernel<<<(p->batch + 255) / 256, 256, 0, e->stream>>>(p->done, p->moves, p->batch);
    RT_CUDA(e, cudaGetLastError());
    // Incremental search (cache.cu: only what a move touched is re-scored) whenever its table applies; otherwise
    // every iteration re-scores the whole neighbourhood on the route-order matrix.  Same keys either way.
    const bool batched = mode & 2;  // SPEC.md §8.1: max_moves counts iterations, each applies a batch of moves
    if (batched && !batch_usable(p))
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_local_search_batched handles at most 1024 routes per instance");
    // (up to ~5,000 positions re-scoring everything is the faster of the two — us per move, full vs table: 33 vs 55 at
    // 1,500 stops, 36 vs 59 at 2,000, 47 vs 66 at 3,000, 78 vs 79 at 5,000, 124 vs 89 at 7,000: an iteration of the
    // full search is two or three dependent launches, the table's bookkeeping is a longer chain of small ones)
    int cache_from = 4608;  // positions; ROUTEST_LS_INCREMENTAL_FROM overrides (the tests force either driver at any size)
    if (const char *ev = getenv("ROUTEST_LS_INCREMENTAL_FROM")) cache_from = atoi(ev);
    const bool cached = !batched && !(mode & 1) && cache_usable(p) && p->pmax > cache_from;
    if (batched) RT_TRY(batch_begin(p, e->stream));
    else if (cached) RT_TRY(cache_begin(p, e->stream));
    else RT_TRY(state_refresh(p, e->stream));
    // The host only paces the loop: kChunk iterations are enqueued blind (an instance that has reached its local
    // optimum turns its own later iterations into no-ops on the device), then one 4-byte-per-instance read decides
    // whether to go on.
    auto one_iteration = [&]() -> int {
        if (batched) return batch_iteration(p, families, e->stream);
        if (cached) {
            RT_TRY(cache_evaluate(p, families, e->stream));
            return cache_apply(p, true, e->stream);
        }
        RT_TRY(eval_moves_launch(p, families, 0, 1, p->best, false, e->stream));
        return apply_moves_launch(p, p->best, false, true, e->stream);
    };
    // One cooperative launch runs the whole search where that form applies (search_persistent_kernel: a fused-cell
    // problem up to ~4,600 positions); the host only looks at done[] between launches of kPersistChunk iterations.
    if (!batched && !cached && max_moves > 0) {
        const int kPersistChunk = 4096;
        int issued = 0, rc = RT_OK;
        bool all_done = false;
        while (!all_done && issued < max_moves) {
            const int todo = std::min(kPersistChunk, max_moves - issued);
            rc = search_persistent_launch(p, families, todo, true, e->stream);
            if (rc != RT_OK) break;
            issued += todo;
            RT_CUDA(e, cudaMemcpyAsync(done_h, p->done, p->batch * 4, cudaMemcpyDeviceToHost, e->stream));
            RT_CUDA(e, cudaStreamSynchronize(e->stream));
            all_done = true;
            for (int b = 0; b < p->batch; ++b) all_done = all_done && done_h[b] != 0;
        }
        if (rc > 0) return rc;
        if (rc == RT_OK) max_moves = 0;  // done: fall through to the read-back below
        if (getenv("ROUTEST_DEBUG") && rc == RT_OK) fprintf(stderr, "[routest] local search (full): one cooperative launch\n");
    }
    // An iteration is a fixed sequence of a dozen small launches with constant arguments: it is captured once into a
    // CUDA graph and replayed (the loop is launch-bound at small and medium sizes).  If capture is refused for any
    // reason the iterations are launched one kernel at a time instead.
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    if (max_moves > 1 && cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
        const int rc_it = one_iteration();
        const cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
        if (rc_it != RT_OK || ce != cudaSuccess || !graph ||
            cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) {
            if (graph) cudaGraphDestroy(graph);
            graph = nullptr;
            exec = nullptr;
            (void)cudaGetLastError();
        }
    }
    if (getenv("ROUTEST_DEBUG"))
        fprintf(stderr, "[routest] local search (%s): iterations %s\n", batched ? "batched" : cached ? "incremental" : "full",
                exec ? "replayed from a CUDA graph" : "launched kernel by kernel");
    const int kChunk = batched ? 8 : 32;  // a batched iteration is a full evaluation: overshoot less
    int issued = 0;
    bool all_done = false;
    int loop_rc = RT_OK;
    while (!all_done && issued < max_moves && loop_rc == RT_OK) {
        // the blind run grows with the search (a quarter of what has been issued, between 4 and kChunk): a short
        // search overshoots its end by a few iterations instead of kChunk, a long one is not paced more often
        const int todo = std::min(std::max(4, std::min(kChunk, issued / 4)), max_moves - issued);
        for (int i = 0; i < todo && loop_rc == RT_OK; ++i) {
            if (exec) {
                if (cudaGraphLaunch(exec, e->stream) != cudaSuccess) loop_rc = fail(e, RT_ERR_CUDA, "cudaGraphLaunch failed");
            } else {
                loop_rc = one_iteration();
            }
        }
        if (loop_rc != RT_OK) break;
        issued += todo;
        RT_CUDA(e, cudaMemcpyAsync(done_h, p->done, p->batch * 4, cudaMemcpyDeviceToHost, e->stream));
        RT_CUDA(e, cudaStreamSynchronize(e->stream));
        all_done = true;
        for (int b = 0; b < p->batch; ++b) all_done = all_done && done_h[b] != 0;
    }
    if (exec) cudaGraphExecDestroy(exec);
    if (graph) cudaGraphDestroy(graph);
    RT_TRY(loop_rc);
    if (cached) RT_TRY(state_refresh(p, e->stream));  // bring the route-order matrix back in step with the routes
    if (moves_out) RT_CUDA(e, cudaMemcpyAsync(moves_out, p->moves, p->batch * 4, cudaMemcpyDeviceToHost, e->stream));
    if (cost_out) RT_CUDA(e, cudaMemcpyAsync(cost_out, p->cost, p->batch * 8, cudaMemcpyDeviceToHost, e->stream));
    if (keys_out && keys_capacity > 0) {
        const int w = std::min(keys_capacity, p->hist_cap);
        RT_CUDA(e, cudaMemcpy2DAsync(keys_out, static_cast<size_t>(keys_capacity) * 8, p->hist,
                                     static_cast<size_t>(p->hist_cap) * 8, static_cast<size_t>(w) * 8, p->batch,
                                     cudaMemcpyDeviceToHost, e->stream));
    }
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    return RT_OK;
}

}  // extern "C"
