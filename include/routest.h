/*
 * routest.h — C ABI of libroutest_cuda.so, the B200 engine behind the route optimizer's hot path.
 *
 * The reference (PaullyMac/routest, backend/route_optimizer_twx2) is pure Python and has no FFI of its own; the
 * seam this library plugs into is the body of Flaskr/utils.py:_multi_stop (utils.py:85-139) and
 * Flaskr/ml.py:predict_eta_minutes (ml.py:23-58).  Each entry point below names the reference lines whose work it
 * takes over, or says that the reference has no counterpart (the matrix is a remote HTTPS call there).
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns an int status (RT_OK == 0); rt_last_error() gives the message of the last failure
 *     on that engine (or, with a NULL engine, of the last failed rt_create on this thread);
 *   - no exception crosses the boundary, and there is NO CPU fallback: without a usable GPU rt_create fails;
 *   - plain pointers and sizes only.  Functions without a suffix take HOST buffers owned by the caller and do
 *     their own host<->device copies; functions ending in _dev take DEVICE pointers on the engine's GPU plus a
 *     cudaStream_t passed as void* — taken literally, so NULL is CUDA's default stream (which is also what
 *     torch.cuda.current_stream().cuda_stream is unless the caller changed it) — and never synchronise;
 *   - an engine is bound to one GPU and serialises calls made on it from several threads (the reference runs
 *     its handlers on arbitrary Flask worker threads, app.py:8); use one engine per GPU and per process.
 */
#ifndef ROUTEST_H
#define ROUTEST_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_OK 0
#define RT_ERR_INVALID_ARG 1
#define RT_ERR_CUDA 2
#define RT_ERR_NO_DEVICE 3
#define RT_ERR_OUT_OF_MEMORY 4
#define RT_ERR_INFEASIBLE_STOP 5 /* a stop that fits no trip on its own; the reference loops forever (utils.py:118-139) */
#define RT_ERR_STATE 6

#define RT_EARTH_RADIUS_M 6371008.8

typedef struct rt_engine rt_engine;

/* Library version, major*10000 + minor*100 + patch.  No reference counterpart. */
int rt_version(void);

/* Engine lifetime.  `device` is a CUDA ordinal.  Replaces nothing in the reference (it holds no state besides the
 * lazily loaded model global, ml.py:4,11-21, whose unlocked initialisation this handle makes explicit). */
int rt_create(int device, rt_engine **out);
void rt_destroy(rt_engine *e);
const char *rt_last_error(const rt_engine *e);

/* The engine's own non-blocking stream as a cudaStream_t (what the host-buffer entry points run on). */
void *rt_engine_stream(rt_engine *e);

/* Pinned host memory for callers that want asynchronous copies (optional; any host pointer is accepted). */
int rt_host_alloc(rt_engine *e, size_t bytes, void **out);
int rt_host_free(rt_engine *e, void *p);

/* Cells per row of the device-resident interleaved matrix for n nodes (n rounded up to 16 cells = 128 B). */
int rt_matrix_pitch(int n);

/*
 * Distance/time matrix from coordinates — takes over utils.py:94-109, where the reference POSTs the coordinates
 * to OpenRouteService and uses the returned `distances` as is.  Node 0 is the depot (utils.py:94).
 * Builds rows [row_begin,row_end) of the n x n matrix of SPEC.md §2 (great-circle, R = RT_EARTH_RADIUS_M):
 *   dist_out  int32 [rows][n]  distance in units of 1/dist_scale metres            (may be NULL)
 *   time_out  int32 [rows][n]  travel time in units of 1/(time_scale) per metre    (may be NULL)
 *   metres_out double [rows][n] FP64 verification output in metres                 (may be NULL)
 * lat/lon in degrees.  Requires pi*R*scale < 2^29 for both scales.
 */
int rt_matrix_build(rt_engine *e, const double *lat, const double *lon, int n, int row_begin, int row_end,
                    double dist_scale, double time_scale, int32_t *dist_out, int32_t *time_out, double *metres_out);

/* Per-stop unit vectors (SPEC.md §1) for `batch` instances of n nodes: lat_d/lon_d are [batch][n] doubles,
 * xyz_d receives [batch][3][n] doubles.  No reference counterpart. */
int rt_unit_vectors_dev(rt_engine *e, const double *lat_d, const double *lon_d, int n, int batch, double *xyz_d,
                        void *stream);

/* Device-resident matrix for `batch` instances: cells_d is [batch][rows][pitch] of {int32 dist,int32 time} with
 * pitch = rt_matrix_pitch(n) and rows = row_end-row_begin; pad cells are written as zero.  This is the layout the
 * move-evaluation kernels read.  Same reference lines as rt_matrix_build. */
int rt_matrix_cells_dev(rt_engine *e, const double *xyz_d, int n, int batch, int row_begin, int row_end,
                        double dist_scale, double time_scale, void *cells_d, void *stream);

/* Self-test: counts floats with bit pattern in [lo_bits, hi_bits) for which the matrix kernel's straight-line
 * square root differs from IEEE sqrtf (must be 0; SPEC.md §2 relies on it).  No reference counterpart. */
int rt_selftest_fast_sqrt(rt_engine *e, uint32_t lo_bits, uint32_t hi_bits, uint64_t *mismatches);

#ifdef __cplusplus
}
#endif
#endif /* ROUTEST_H */
