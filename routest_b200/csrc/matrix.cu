// matrix.cu — stop-to-stop distance/time matrix from coordinates (SPEC.md §1-§2).
//
// Takes over Flaskr/utils.py:94-109 of the reference, where the matrix is fetched from a remote service.
//
// Kernel shape.  The matrix is pure output: n^2 cells of 8 B against 24 B of input per stop, so the kernel is
// bound by HBM *writes* (algorithmic bytes per cell: 8 = int32 dist + int32 time).  Each thread owns four columns
// and keeps their unit vectors in registers while its block walks down a strip of rows, whose unit vectors sit in
// shared memory (one broadcast LDS per row).  Per cell the straight-line path costs 6 FP64 ops for the chord, one F2F, one
// MUFU and ~14 FP32/INT ops (geometry.cuh), i.e. ~25 issue slots against the ~40 per cell per SM that the HBM
// write rate leaves room for; every store is a full 16-byte STG whose warp footprint is 512 contiguous bytes.
#include <algorithm>

#include "common.cuh"
#include "geometry.cuh"

namespace rt {

namespace {

constexpr int kThreads = 256;
constexpr int kMaxRowsPerBlock = 64;
constexpr int kBlocksPerSM = 4;
constexpr int kMinWaves = 5;

__global__ void __launch_bounds__(256) unit_vectors_kernel(const double *__restrict__ lat,
                                                           const double *__restrict__ lon, int n, int total,
                                                           double *__restrict__ xyz, int *__restrict__ bad_flag)
{
    const double DEG = 0.017453292519943295;
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    int b = idx / n, i = idx - b * n;
    double la = lat[idx], lo = lon[idx];
    if (bad_flag && (!(la >= -90.0 && la <= 90.0) || !(lo >= -180.0 && lo <= 180.0))) {
        *bad_flag = 1;  // out of range or not finite: reported by the host after the pass; the value is neutralised
        la = 0.0;
        lo = 0.0;
    }
    double sp, cp, sl, cl;
    det_sincos(la * DEG, sp, cp);
    det_sincos(lo * DEG, sl, cl);
    double *o = xyz + static_cast<size_t>(b) * 3 * n;
    o[i] = cp * cl;
    o[n + i] = cp * sl;
    o[2 * n + i] = sp;
}

struct MatArgs {
    const double *xyz;
    int n, ld, r0, r1, rows_per_block;  // ld: stride between the x/y/z component arrays (>= n); instance stride 3*ld
    int out_rows;     // rows between two instances in the output (0: r1 - r0, the dense default)
    float fd, ft;
    int lim_hi;       // fast_path_limit_hi(fd, ft)
    int2 *cells;      // MODE 0
    int pitch;        //   cells per row
    int32_t *dist;    // MODE 1 (either may be null)
    int32_t *tim;
    int pitch_i32;
    double *metres;   // MODE 2
    int pitch_f64;
};

enum { kCells = 0, kPlanes = 1, kMetres = 2 };

// grid = (column blocks of 1024, row strips of rows_per_block, instances).  The strip's unit vectors are staged in
// shared memory once (broadcast LDS per row); the four column vectors of a thread stay in registers.
template <int MODE>
__global__ void __launch_bounds__(kThreads, kBlocksPerSM) matrix_kernel(const MatArgs a)
{
    __shared__ double srow[3][kMaxRowsPerBlock];

    const int n = a.n;
    const double *__restrict__ X = a.xyz + static_cast<size_t>(blockIdx.z) * 3 * a.ld;
    const double *__restrict__ Y = X + a.ld;
    const double *__restrict__ Z = Y + a.ld;
    const int lane = threadIdx.x & 31;
    const int wb = blockIdx.x * (blockDim.x * 4) + (threadIdx.x >> 5) * 128;  // a block spans blockDim.x * 4 columns

    // column ownership: 16-byte stores, 512 contiguous bytes per warp-wide store instruction
    int col[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
        col[k] = (MODE == kPlanes) ? wb + 4 * lane + k : wb + (k >> 1) * 64 + 2 * lane + (k & 1);

    double bx[4], by[4], bz[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int c = col[k] < n ? col[k] : 0;
        bx[k] = X[c];
        by[k] = Y[c];
        bz[k] = Z[c];
    }

    const int row0 = a.r0 + blockIdx.y * a.rows_per_block;
    const int nrows = min(a.rows_per_block, a.r1 - row0);
    for (int t = threadIdx.x; t < 3 * nrows; t += blockDim.x) {
        int comp = t / nrows, r = t - comp * nrows;
        srow[comp][r] = (comp == 0 ? X : comp == 1 ? Y : Z)[row0 + r];
    }
    __syncthreads();

    const size_t rows_total = static_cast<size_t>(a.out_rows ? a.out_rows : a.r1 - a.r0);
    const size_t out_row0 = static_cast<size_t>(blockIdx.z) * rows_total + static_cast<size_t>(row0 - a.r0);

    // four cells of one row: chord^2 in FP64, then either the straight-line tail (all four below the fast-path
    // limit, i.e. every cell of a metro instance, the diagonal included) or the fully guarded generic cell
    const bool v0 = col[0] < n, v1 = col[1] < n, v2 = col[2] < n, v3 = col[3] < n;
    auto four_cells = [&](int r, int2 (&c)[4]) {
        const double ax = srow[0][r], ay = srow[1][r], az = srow[2][r];
        double s[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            double dx = ax - bx[k], dy = ay - by[k], dz = az - bz[k];
            s[k] = fma(dx, dx, fma(dy, dy, dz * dz));
        }
        const int hmax = max(max(__double2hiint(s[0]), __double2hiint(s[1])),
                             max(__double2hiint(s[2]), __double2hiint(s[3])));
        if (hmax < a.lim_hi) {
#pragma unroll
            for (int k = 0; k < 4; ++k) c[k] = cell_i32_fast(s[k], a.fd, a.ft);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) c[k] = cell_i32(ax, ay, az, bx[k], by[k], bz[k], a.fd, a.ft);
        }
        if (!v3) {  // pad columns of the last column block are written as zero
            if (!v0) c[0] = make_int2(0, 0);
            if (!v1) c[1] = make_int2(0, 0);
            if (!v2) c[2] = make_int2(0, 0);
            c[3] = make_int2(0, 0);
        }
    };

    if (MODE == kCells) {
        if (col[0] >= a.pitch) return;  // per-thread exit is fine: no barrier follows
        int4 *out0 = reinterpret_cast<int4 *>(a.cells + out_row0 * a.pitch + col[0]);
        int4 *out1 = reinterpret_cast<int4 *>(a.cells + out_row0 * a.pitch + col[2]);
        const bool st1 = col[2] < a.pitch;
        const size_t step = static_cast<size_t>(a.pitch) / 2;  // int4 = 2 cells
#pragma unroll 2
        for (int r = 0; r < nrows; ++r) {
            int2 c[4];
            four_cells(r, c);
            out0[r * step] = make_int4(c[0].x, c[0].y, c[1].x, c[1].y);
            if (st1) out1[r * step] = make_int4(c[2].x, c[2].y, c[3].x, c[3].y);
        }
    } else if (MODE == kPlanes) {
        if (col[0] >= a.pitch_i32) return;
        const size_t base = out_row0 * a.pitch_i32 + col[0];
        int4 *od = a.dist ? reinterpret_cast<int4 *>(a.dist + base) : nullptr;
        int4 *ot = a.tim ? reinterpret_cast<int4 *>(a.tim + base) : nullptr;
        const size_t step = static_cast<size_t>(a.pitch_i32) / 4;
#pragma unroll 2
        for (int r = 0; r < nrows; ++r) {
            int2 c[4];
            four_cells(r, c);
            if (od) od[r * step] = make_int4(c[0].x, c[1].x, c[2].x, c[3].x);
            if (ot) ot[r * step] = make_int4(c[0].y, c[1].y, c[2].y, c[3].y);
        }
    } else {
        if (col[0] >= a.pitch_f64) return;
        double2 *out0 = reinterpret_cast<double2 *>(a.metres + out_row0 * a.pitch_f64 + col[0]);
        double2 *out1 = reinterpret_cast<double2 *>(a.metres + out_row0 * a.pitch_f64 + col[2]);
        const bool st1 = col[2] < a.pitch_f64;
        const size_t step = static_cast<size_t>(a.pitch_f64) / 2;
        for (int r = 0; r < nrows; ++r) {
            const double ax = srow[0][r], ay = srow[1][r], az = srow[2][r];
            double m[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int kk;
                double s = chord2(ax, ay, az, bx[k], by[k], bz[k], kk);
                m[k] = col[k] < n ? theta_f64(s, kk) * RT_EARTH_RADIUS_M : 0.0;
            }
            out0[r * step] = make_double2(m[0], m[1]);
            if (st1) out1[r * step] = make_double2(m[2], m[3]);
        }
    }
}

// Self-test kernel: fast_sqrtf against the IEEE sqrtf for every float whose bit pattern lies in [lo, hi).
__global__ void fast_sqrt_check_kernel(unsigned lo, unsigned hi, unsigned long long *mismatches)
{
    unsigned long long bad = 0;
    const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * blockDim.x;
    for (unsigned long long b = lo + static_cast<unsigned long long>(blockIdx.x) * blockDim.x + threadIdx.x; b < hi;
         b += stride) {
        float x = __uint_as_float(static_cast<unsigned>(b));
        if (__float_as_uint(fast_sqrtf(x)) != __float_as_uint(sqrtf(x))) ++bad;
    }
    if (bad) atomicAdd(mismatches, bad);
}

// Rows per block: the largest strip that still yields >= kMinWaves full waves of blocks, so that the tail wave (the
// hardware dispatches a new block whenever one retires) costs a few percent at most; never below 8 rows, which
// keeps the per-block prologue (12 column loads + one barrier) under ~5% of the block's work.  kMinWaves = 5:
// measured against 10 / 7 / 3 on the bench step (64 instances 1,040 wide): 103 / 98.6 / 97.8 µs, 98.5 at 5; no
// difference at 10,001 nodes.
int pick_rows_per_block(int rows, int col_blocks, int batch, int threads)
{
    const int blocks_per_sm = std::min(16, 65536 / (64 * threads));  // 64 registers per thread
    const long long want_blocks = static_cast<long long>(kMinWaves) * kNumSMs * blocks_per_sm;
    int rpb = kMaxRowsPerBlock;
    while (rpb > 8 && static_cast<long long>((rows + rpb - 1) / rpb) * col_blocks * batch < want_blocks) rpb >>= 1;
    return rpb;
}

template <int MODE>
int launch_matrix(rt_engine *e, MatArgs a, int batch, int pitch_cols, cudaStream_t st)
{
    const int rows = a.r1 - a.r0;
    if (rows <= 0 || batch <= 0) return RT_OK;
    // Column blocks of whole warps (128 columns each), at most eight warps per block: pick the split that leaves the
    // fewest idle warps (1,040 columns = 9 warps -> 3 blocks of 3, not 2 blocks of 5), larger blocks on ties.
    const int warps = (pitch_cols + 127) / 128;
    int col_blocks = (warps + 7) / 8, wpb = (warps + col_blocks - 1) / col_blocks;
    for (int cb = col_blocks + 1; cb <= (warps + 1) / 2; ++cb) {
        const int w = (warps + cb - 1) / cb;
        if (cb * w - warps < col_blocks * wpb - warps) {
            col_blocks = cb;
            wpb = w;
        }
    }
    const int threads = wpb * 32;
    a.rows_per_block = pick_rows_per_block(rows, col_blocks, batch, threads);
    const int row_blocks = (rows + a.rows_per_block - 1) / a.rows_per_block;
    if (row_blocks > 65535 || batch > 65535) return fail(e, RT_ERR_INVALID_ARG, "matrix grid too large");
    dim3 grid(col_blocks, row_blocks, batch);
    matrix_kernel<MODE><<<grid, threads, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

}  // namespace

int check_scales(rt_engine *e, double dist_scale, double time_scale)
{
    const double lim = 536870912.0;  // 2^29: a four-term delta of cells must fit an int32
    const double half_circ = 3.14159265358979323846 * RT_EARTH_RADIUS_M;
    if (!(dist_scale > 0.0) || !(time_scale > 0.0) || !(half_circ * dist_scale < lim) || !(half_circ * time_scale < lim))
        return fail(e, RT_ERR_INVALID_ARG, "dist_scale/time_scale must be > 0 and keep pi*R*scale below 2^29");
    return RT_OK;
}

int run_fast_sqrt_check(rt_engine *e, unsigned lo, unsigned hi, unsigned long long *mismatches_d, cudaStream_t st)
{
    fast_sqrt_check_kernel<<<kNumSMs * 8, 256, 0, st>>>(lo, hi, mismatches_d);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

int launch_unit_vectors(rt_engine *e, const double *lat_d, const double *lon_d, int n, int batch, double *xyz_d,
                        cudaStream_t st, int *bad_flag_d)
{
    const long long total = static_cast<long long>(n) * batch;
    if (total <= 0) return RT_OK;
    if (total > 0x7fffffffLL) return fail(e, RT_ERR_INVALID_ARG, "n*batch too large");
    unit_vectors_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(lat_d, lon_d, n,
                                                                                    static_cast<int>(total), xyz_d, bad_flag_d);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

// The same kernel on unit vectors whose component arrays are ld apart (the position-ordered vectors of a route
// state): n x n cells per instance into rows of `pitch` cells.
int launch_matrix_cells_ld(rt_engine *e, const double *xyz_d, int n, int ld, int batch, int r0, int r1,
                           double dist_scale, double time_scale, int2 *cells_d, int pitch, cudaStream_t st)
{
    // rows [r0, r1) of every instance, written in place: row r of instance b lands at cells_d[(b*n + r) * pitch]
    MatArgs a{};
    a.xyz = xyz_d;
    a.n = n;
    a.ld = ld;
    a.r0 = r0;
    a.r1 = r1;
    a.out_rows = n;
    cells_d += static_cast<size_t>(r0) * pitch;
    a.fd = static_cast<float>(RT_EARTH_RADIUS_M * dist_scale);
    a.ft = static_cast<float>(RT_EARTH_RADIUS_M * time_scale);
    a.lim_hi = fast_path_limit_hi(a.fd, a.ft);
    a.cells = cells_d;
    a.pitch = pitch;
    return launch_matrix<kCells>(e, a, batch, pitch, st);
}

int launch_matrix_cells(rt_engine *e, const double *xyz_d, int n, int batch, int r0, int r1, double dist_scale,
                        double time_scale, int2 *cells_d, cudaStream_t st)
{
    MatArgs a{};
    a.xyz = xyz_d;
    a.n = n;
    a.ld = n;
    a.r0 = r0;
    a.r1 = r1;
    a.fd = static_cast<float>(RT_EARTH_RADIUS_M * dist_scale);
    a.ft = static_cast<float>(RT_EARTH_RADIUS_M * time_scale);
    a.lim_hi = fast_path_limit_hi(a.fd, a.ft);
    a.cells = cells_d;
    a.pitch = rt_matrix_pitch(n);
    return launch_matrix<kCells>(e, a, batch, a.pitch, st);
}

int launch_matrix_planes(rt_engine *e, const double *xyz_d, int n, int r0, int r1, double dist_scale, double time_scale,
                         int32_t *dist_d, int32_t *time_d, double *metres_d, int pitch_i32, int pitch_f64,
                         cudaStream_t st)
{
    MatArgs a{};
    a.xyz = xyz_d;
    a.n = n;
    a.ld = n;
    a.r0 = r0;
    a.r1 = r1;
    a.fd = static_cast<float>(RT_EARTH_RADIUS_M * dist_scale);
    a.ft = static_cast<float>(RT_EARTH_RADIUS_M * time_scale);
    a.lim_hi = fast_path_limit_hi(a.fd, a.ft);
    a.dist = dist_d;
    a.tim = time_d;
    a.pitch_i32 = pitch_i32;
    a.metres = metres_d;
    a.pitch_f64 = pitch_f64;
    if (dist_d || time_d) RT_TRY(launch_matrix<kPlanes>(e, a, 1, pitch_i32, st));
    if (metres_d) RT_TRY(launch_matrix<kMetres>(e, a, 1, pitch_f64, st));
    return RT_OK;
}

}  // namespace rt
