// common.cuh — engine object, error plumbing and small device helpers shared by every translation unit.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/routest.h"

namespace rt {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// A growable device buffer that is allocated once and kept (no cudaMalloc on the steady-state path).
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

}  // namespace rt

struct rt_engine {
    int device = 0;
    cudaStream_t stream = nullptr;   // compute + H2D
    cudaStream_t copy_stream = nullptr;  // D2H of finished row chunks, overlapped with the next chunk's kernel
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::mutex mu;
    char err[512] = {0};
    rt::DevBuf coords, xyz, plane_a[2], plane_b[2], plane_c[2];
    void *eta_model = nullptr;  // rt::EtaModel*, owned (eta.cu)
};

namespace rt {

int fail(rt_engine *e, int code, const char *fmt, ...);

#define RT_CUDA(e, call)                                                                                  \
    do {                                                                                                  \
        cudaError_t _st = (call);                                                                         \
        if (_st != cudaSuccess)                                                                           \
            return rt::fail((e), _st == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA,   \
                            "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_st), __FILE__, __LINE__); \
    } while (0)

#define RT_TRY(expr)                 \
    do {                             \
        int _rc = (expr);            \
        if (_rc != RT_OK) return _rc; \
    } while (0)

int reserve(rt_engine *e, DevBuf &b, size_t bytes);

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

inline cudaStream_t pick_stream(rt_engine *, void *s) { return reinterpret_cast<cudaStream_t>(s); }

// launchers implemented in matrix.cu
// bad_flag_d (optional): set to 1 on the device when a coordinate is out of range or not finite
int launch_unit_vectors(rt_engine *e, const double *lat_d, const double *lon_d, int n, int batch, double *xyz_d,
                        cudaStream_t st, int *bad_flag_d = nullptr);
int launch_matrix_cells(rt_engine *e, const double *xyz_d, int n, int batch, int r0, int r1, double dist_scale,
                        double time_scale, int2 *cells_d, cudaStream_t st);
int launch_matrix_cells_ld(rt_engine *e, const double *xyz_d, int n, int ld, int batch, int r0, int r1,
                           double dist_scale, double time_scale, int2 *cells_d, int pitch, cudaStream_t st);
int launch_matrix_planes(rt_engine *e, const double *xyz_d, int n, int r0, int r1, double dist_scale, double time_scale,
                         int32_t *dist_d, int32_t *time_d, double *metres_d, int pitch_i32, int pitch_f64,
                         cudaStream_t st);
int check_scales(rt_engine *e, double dist_scale, double time_scale);
int eta_model_free(rt_engine *e);
int run_fast_sqrt_check(rt_engine *e, unsigned lo, unsigned hi, unsigned long long *mismatches_d, cudaStream_t st);

}  // namespace rt
