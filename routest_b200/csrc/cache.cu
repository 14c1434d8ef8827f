// cache.cu — local search that re-scores only what a move changed (SPEC.md §8 semantics, same keys, same routes).
//
// A candidate's delta and feasibility depend on the state of at most two routes.  After a move has been applied,
// every candidate between two untouched routes is exactly what it was, so the search keeps, per instance,
//   inter[ra][rb]  the best key among relocate(p in ra -> edge q of rb) and swap(p in ra, q in rb), ra != rb
//   intra[r]       the best key among the 2-opt / relocate / swap candidates inside route r
// with positions stored RELATIVE to the start of their route (a relocate shifts every later route by one position,
// which leaves relative positions — and the order of keys inside a block — untouched).  One iteration is then:
//   apply the winner, note the (at most two) routes it touched  ->  refresh the route state (no matrix rebuild)
//   -> clear the touched rows/columns of the table  ->  re-score the tiles of the (p, q) plane that intersect a
//   touched route, cells rebuilt on the fly from the position-ordered unit vectors (or gathered from a supplied
//   matrix)  ->  re-sweep the touched routes  ->  arg-min over the table with positions made absolute again.
// The winner is the same key the full evaluation (moves.cu) would return, so the applied sequence and the final
// routes are identical; tests/test_moves_gpu.py runs both and compares them with the oracle.
// No reference counterpart (the reference has no improvement phase, utils.py:111-139).
#include <algorithm>
#include <climits>

#include "geometry.cuh"
#include "problem.cuh"

namespace rt {

namespace {

constexpr int kTileRows = 8, kTileCols = 128;
constexpr int kTallRows = 16;     // column-band tiles are tall: their row loop is cheap and there are P/8 of them otherwise
constexpr int kTallFlag = 1 << 30;  // in a packed tile: rows are counted in units of kTallRows
constexpr int kBandThreads = kTileCols;
constexpr int kAllDirty = -2;

struct CacheArgs {
    // state (same arrays the move kernels read)
    const int *node, *npos, *route, *rstart;
    const int4 *RA, *RB, *RC, *RD, *H;
    const int2 *E, *SK;
    const double *PX;      // position-ordered unit vectors (null: gather from cells)
    const int2 *cells;     // node-order matrix (supplied-matrix path)
    float fd, ft;
    int lim_hi;
    int n, pitch, pstride, rstride, families;
    // table
    unsigned long long *inter, *intra;  // [batch][cr][cr], [batch][cr]
    int cr;
    const int *dirty;                   // [batch][2]
    int *tiles, *ntiles;                // [batch][tile_cap] packed (row_tile << 16 | col_tile), [batch]
    int tile_cap;
    int *cursor;                        // [batch] work counter
    unsigned long long *best;           // [batch]
};

__device__ __forceinline__ unsigned long long rel_key(int delta, int family, int pr, int qr)
{
    return (static_cast<unsigned long long>(static_cast<unsigned>(delta) + 0x80000000u) << 32) |
           (static_cast<unsigned long long>(family) << 30) | (static_cast<unsigned long long>(pr) << 15) |
           static_cast<unsigned long long>(qr);
}

__device__ __forceinline__ bool is_dirty(int r, int d0, int d1) { return d0 == kAllDirty || r == d0 || r == d1; }

// cell(position a, position b) without the route-order matrix
__device__ __forceinline__ int2 pos_cell(const CacheArgs &a, int b, int pa, int pb)
{
    if (a.PX) {
        const double *px = a.PX + static_cast<size_t>(b) * 3 * a.pstride, *py = px + a.pstride, *pz = py + a.pstride;
        const double ax = px[pa], ay = py[pa], az = pz[pa], bx = px[pb], by = py[pb], bz = pz[pb];
        const double dx = ax - bx, dy = ay - by, dz = az - bz;
        const double s2 = fma(dx, dx, fma(dy, dy, dz * dz));
        return __double2hiint(s2) < a.lim_hi ? cell_i32_fast(s2, a.fd, a.ft) : cell_i32(ax, ay, az, bx, by, bz, a.fd, a.ft);
    }
    const int *nd = a.node + static_cast<size_t>(b) * a.pstride;
    return __ldg(&a.cells[static_cast<size_t>(b) * a.n * a.pitch + static_cast<size_t>(nd[pa]) * a.pitch + nd[pb]]);
}

// Clears what the last move invalidated and lists the tiles of the (p, q) plane that must be re-scored.
__global__ void __launch_bounds__(256) cache_prepare_kernel(const CacheArgs a)
{
    const int b = blockIdx.x, tid = threadIdx.x;
    const int P = a.npos[b], R = P - a.n;
    const int d0 = a.dirty[2 * b], d1 = a.dirty[2 * b + 1];
    unsigned long long *inter = a.inter + static_cast<size_t>(b) * a.cr * a.cr;
    unsigned long long *intra = a.intra + static_cast<size_t>(b) * a.cr;
    const int *rstart = a.rstart + static_cast<size_t>(b) * (a.rstride + 1);
    __shared__ int s_n;
    if (tid == 0) {
        s_n = 0;
        a.cursor[b] = 0;
    }
    __syncthreads();
    const int nrt = (P + kTileRows - 1) / kTileRows, nct = (P + kTileCols - 1) / kTileCols;
    int *tiles = a.tiles + static_cast<size_t>(b) * a.tile_cap;
    if (d0 == kAllDirty) {
        for (int i = tid; i < R * a.cr; i += blockDim.x) inter[static_cast<size_t>(i / a.cr) * a.cr + i % a.cr] = kNoMove;
        for (int i = tid; i < R; i += blockDim.x) intra[i] = kNoMove;
        const int ntall = (P + kTallRows - 1) / kTallRows;
        for (int i = tid; i < ntall * nct; i += blockDim.x) tiles[i] = kTallFlag | ((i / nct) << 16) | (i % nct);
        if (tid == 0) a.ntiles[b] = ntall * nct;
        return;
    }
    for (int k = 0; k < 2; ++k) {
        const int d = k == 0 ? d0 : d1;
        if (d < 0 || (k == 1 && d1 == d0)) continue;
        for (int i = tid; i < R; i += blockDim.x) {
            inter[static_cast<size_t>(d) * a.cr + i] = kNoMove;
            inter[static_cast<size_t>(i) * a.cr + d] = kNoMove;
        }
        if (tid == 0) intra[d] = kNoMove;
        // the route's positions, its starting separator included (it is the insertion edge "before the first stop")
        const int lo = rstart[d], hi = rstart[d + 1] - 1;
        const int rt0 = lo / kTileRows, rt1 = hi / kTileRows, ct0 = lo / kTileCols, ct1 = hi / kTileCols;
        // column band: short tiles while there are fewer of them than blocks to run them (latency), tall ones beyond
        const bool tall = P > 4096;
        const int ntall = tall ? (P + kTallRows - 1) / kTallRows : nrt;
        const int nrows = (rt1 - rt0 + 1) * nct, ncols = (ct1 - ct0 + 1) * ntall;
        __shared__ int s_base;
        if (tid == 0) {
            s_base = s_n;
            s_n += nrows + ncols;
        }
        __syncthreads();
        const int base = s_base;
        for (int i = tid; i < nrows; i += blockDim.x) tiles[base + i] = ((rt0 + i / nct) << 16) | (i % nct);
        for (int i = tid; i < ncols; i += blockDim.x)
            tiles[base + nrows + i] = (tall ? kTallFlag : 0) | ((i / (ct1 - ct0 + 1)) << 16) | (ct0 + i % (ct1 - ct0 + 1));
        __syncthreads();
    }
    if (tid == 0) a.ntiles[b] = s_n;
}

// Re-scores listed tiles (8 rows x 128 columns): relocate(p -> q) and swap(p, q) for pairs of different routes,
// straight-line like eval_inter_kernel.  Everything a tile needs is fetched up front in one round of independent
// loads — the rows' records and unit vectors into shared memory, the column's into registers — and the cells are
// rebuilt from those vectors, so the row loop itself never waits on memory.  A thread owns column q and flushes its
// best key into inter[ra][rb] whenever the row's route changes.
template <bool COMPUTED>
__global__ void __launch_bounds__(kBandThreads) cache_inter_kernel(const CacheArgs a)
{
    __shared__ int4 sH[kTallRows], sA[kTallRows], sB[kTallRows], sC[kTallRows], sD[kTallRows];
    __shared__ double sPX[3][kTallRows + 2];
    __shared__ int s_tile;

    const int b = blockIdx.y;
    const int P = a.npos[b];
    const size_t po = static_cast<size_t>(b) * a.pstride;
    const int *rstart = a.rstart + static_cast<size_t>(b) * (a.rstride + 1);
    unsigned long long *inter = a.inter + static_cast<size_t>(b) * a.cr * a.cr;
    const int *tiles = a.tiles + static_cast<size_t>(b) * a.tile_cap;
    const int ntiles = a.ntiles[b];
    const int d0 = a.dirty[2 * b], d1 = a.dirty[2 * b + 1];
    const int tid = threadIdx.x;
    const double *px = COMPUTED ? a.PX + static_cast<size_t>(b) * 3 * a.pstride : nullptr;
    for (;;) {
        __syncthreads();  // the previous tile's shared data is no longer needed
        if (tid == 0) s_tile = atomicAdd(a.cursor + b, 1);
        __syncthreads();
        const int ti = s_tile;
        if (ti >= ntiles) break;
        const int packed = tiles[ti];
        const int trows = (packed & kTallFlag) ? kTallRows : kTileRows;
        const int p0 = ((packed & ~kTallFlag) >> 16) * trows, q = (packed & 0xFFFF) * kTileCols + tid;
        // one round of loads: rows -> shared memory, column -> registers
        if (tid < trows) {
            const int p = min(p0 + tid, P - 1);
            sH[tid] = a.H[po + p];
            sA[tid] = a.RA[po + p];
            sB[tid] = a.RB[po + p];
            sC[tid] = a.RC[po + p];
            sD[tid] = a.RD[po + p];
        }
        if (COMPUTED) {
            for (int k = tid; k < 3 * (trows + 2); k += kBandThreads) {
                const int comp = k / (trows + 2), i = k - comp * (trows + 2);
                sPX[comp][i] = px[static_cast<size_t>(comp) * a.pstride + min(max(p0 - 1 + i, 0), P - 1)];
            }
        }
        const bool q_in = q < P;
        const int qq = min(q, P - 1);
        const int4 Hq = a.H[po + qq], qA = a.RA[po + qq], qB = a.RB[po + qq], qC = a.RC[po + qq], qD = a.RD[po + qq];
        double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0}, cz[3] = {0, 0, 0};  // unit vectors of q-1, q, q+1
        if (COMPUTED) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int c = min(max(qq - 1 + k, 0), P - 1);
                cx[k] = px[c];
                cy[k] = px[a.pstride + c];
                cz[k] = px[2 * static_cast<size_t>(a.pstride) + c];
            }
        }
        __syncthreads();
        if (!q_in) continue;
        auto cell = [&](int i, int k, int pa, int pb) -> int2 {  // row slot i (0 = p0-1), column slot k (0 = q-1)
            if (COMPUTED) {
                const double ax = sPX[0][i], ay = sPX[1][i], az = sPX[2][i];
                const double dx = ax - cx[k], dy = ay - cy[k], dz = az - cz[k];
                const double s2 = fma(dx, dx, fma(dy, dy, dz * dz));
                return __double2hiint(s2) < a.lim_hi ? cell_i32_fast(s2, a.fd, a.ft)
                                                     : cell_i32(ax, ay, az, cx[k], cy[k], cz[k], a.fd, a.ft);
            }
            return pos_cell(a, b, pa, pb);
        };
        const int rb = Hq.w & 0x7FFF;
        const bool q_cust = (Hq.w >> 30) & 1;
        const bool q_dirty = is_dirty(rb, d0, d1);
        const bool rel_q = (a.families & 2) && q <= P - 2, swap_q = (a.families & 4) && q_cust;
        const int qs = rstart[min(rb, P - a.n)];
        int cur_ra = -1;
        unsigned long long best = kNoMove;
        const int rows = min(trows, P - p0);
        for (int rr = 0; rr < rows; ++rr) {
            const int p = p0 + rr;
            const int4 Hp = sH[rr];
            if (!((Hp.w >> 30) & 1)) continue;  // separators are never moved
            const int ra = Hp.w & 0x7FFF;
            if (ra == rb || !(q_dirty || is_dirty(ra, d0, d1))) continue;
            if (ra != cur_ra) {
                if (best != kNoMove) atomicMin(inter + static_cast<size_t>(cur_ra) * a.cr + rb, best);
                best = kNoMove;
                cur_ra = ra;
            }
            const int ps = rstart[ra];
            const int4 pA = sA[rr], pB = sB[rr], pC = sC[rr], pD = sD[rr];
            const int2 c_cur = cell(rr + 1, 1, p, q), c_rt = cell(rr + 1, 2, p, min(q + 1, P - 1));
            if (rel_q) {
                const int ins = c_cur.x + c_rt.x - Hq.z;
                const int tu = max(pB.x, qA.x + c_cur.y);
                const bool ok = (pC.w & 2) && pB.w <= qD.x && ins <= qD.y && tu <= pB.y && tu + pB.z + c_rt.y <= qA.y;
                if (ok) {
                    const unsigned long long k = rel_key(Hp.x + ins, 1, p - ps, q - qs);
                    best = k < best ? k : best;
                }
            }
            if (swap_q && p < q) {
                const int2 c_up = cell(rr, 1, p - 1, q), c_dn = cell(rr + 2, 1, p + 1, q), c_lf = cell(rr + 1, 0, p, q - 1);
                const int dA = c_up.x + c_dn.x - Hp.y, dB = c_lf.x + c_rt.x - Hq.y;
                const int tv = max(qB.x, pC.x + c_up.y), tw = max(pB.x, qC.x + c_lf.y);
                const bool ok = qB.w - pB.w <= pD.x && pB.w - qB.w <= qD.x && dA <= pD.y && dB <= qD.y && tv <= qB.y &&
                                tv + qB.z + c_dn.y <= pA.y && tw <= pB.y && tw + pB.z + c_rt.y <= qA.y;
                if (ok) {
                    const unsigned long long k = rel_key(dA + dB, 2, p - ps, q - qs);
                    best = k < best ? k : best;
                }
            }
        }
        if (best != kNoMove) atomicMin(inter + static_cast<size_t>(cur_ra) * a.cr + rb, best);
    }
}

struct CSeg {
    int a, b, L;
    bool ok;
};
__device__ __forceinline__ CSeg cseg_single(int ready, int due) { return CSeg{0, ready, due, true}; }
__device__ __forceinline__ CSeg cseg_concat(const CSeg &s1, int w, const CSeg &s2)
{
    CSeg s;
    s.a = s1.a + w + s2.a;
    s.b = max(s1.b + w + s2.a, s2.b);
    s.L = min(s1.L, s2.L - s1.a - w);
    s.ok = s1.ok && s2.ok && (s1.b + w <= s2.L);
    return s;
}
__device__ __forceinline__ bool cseg_fits(const CSeg &m, int t, int tail, int limit)
{
    return m.ok && t <= m.L && max(t + m.a, m.b) + tail <= limit;
}

// Re-sweeps touched routes (the sweep of eval_intra_kernel, symmetric matrices).  One block per route; a touched
// route's records and unit vectors are staged in shared memory first (when it has at most kStage positions), so
// the sequential sweeps of its anchors run without a memory round trip per step.
constexpr int kStage = 384;  // positions of one route held in shared memory (88 B each)
constexpr int kTable = 64;   // routes up to this many positions also get their cell block in shared memory (32 KB)

__global__ void __launch_bounds__(128) cache_intra_kernel(const CacheArgs a)
{
    __shared__ int4 sA[kStage], sB[kStage], sC[kStage], sD[kStage];
    __shared__ double sX[kStage], sY[kStage], sZ[kStage];
    __shared__ int sN[kStage];

    const int b = blockIdx.y, r = blockIdx.x;
    const int P = a.npos[b];
    if (r >= P - a.n) return;
    const int d0 = a.dirty[2 * b], d1 = a.dirty[2 * b + 1];
    if (!is_dirty(r, d0, d1)) return;
    const size_t po = static_cast<size_t>(b) * a.pstride;
    const int *rstart = a.rstart + static_cast<size_t>(b) * (a.rstride + 1);
    const int rs = rstart[r], re = rstart[r + 1];  // both separators
    const int lo = max(rs - 1, 0), cnt = re - lo + 1;
    const bool staged = cnt <= kStage;
    const double *gx = a.PX ? a.PX + static_cast<size_t>(b) * 3 * a.pstride : nullptr;
    if (staged) {
        for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
            const size_t o = po + lo + i;
            sA[i] = a.RA[o];
            sB[i] = a.RB[o];
            sC[i] = a.RC[o];
            sD[i] = a.RD[o];
            sN[i] = a.node[o];
            if (gx) {
                sX[i] = gx[lo + i];
                sY[i] = gx[a.pstride + lo + i];
                sZ[i] = gx[2 * static_cast<size_t>(a.pstride) + lo + i];
            }
        }
        __syncthreads();
    }
    // one indexing scheme for both cases: arrays addressed by absolute position
    const int4 *A = staged ? sA - lo : a.RA + po, *B = staged ? sB - lo : a.RB + po;
    const int4 *C = staged ? sC - lo : a.RC + po, *D = staged ? sD - lo : a.RD + po;
    const int *Nd = staged ? sN - lo : a.node + po;
    const double *X = staged ? sX - lo : gx, *Y = staged ? sY - lo : (gx ? gx + a.pstride : nullptr),
                 *Z = staged ? sZ - lo : (gx ? gx + 2 * static_cast<size_t>(a.pstride) : nullptr);
    // A short route (at most kTable positions with its two neighbours) gets its whole block of cells computed once,
    // by all threads in parallel, into shared memory: the sweeps below are a single warp's dependent chain, and three
    // cell evaluations per step (FP64 chord, sqrt, series) were two thirds of that chain.
    extern __shared__ int2 s_tab[];
    const bool tabled = gx && staged && cnt <= kTable;
    if (tabled) {
        for (int i = threadIdx.x; i < cnt * cnt; i += blockDim.x) {
            const int ia = i / cnt, ib = i - ia * cnt;
            const double ax = sX[ia], ay = sY[ia], az = sZ[ia], bx = sX[ib], by = sY[ib], bz = sZ[ib];
            const double dx = ax - bx, dy = ay - by, dz = az - bz;
            const double s2 = fma(dx, dx, fma(dy, dy, dz * dz));
            s_tab[i] = __double2hiint(s2) < a.lim_hi ? cell_i32_fast(s2, a.fd, a.ft) : cell_i32(ax, ay, az, bx, by, bz, a.fd, a.ft);
        }
        __syncthreads();
    }
    auto cell = [&](int pa, int pb) -> int2 {
        if (tabled) return s_tab[(pa - lo) * cnt + (pb - lo)];
        if (gx) {
            const double ax = X[pa], ay = Y[pa], az = Z[pa], bx = X[pb], by = Y[pb], bz = Z[pb];
            const double dx = ax - bx, dy = ay - by, dz = az - bz;
            const double s2 = fma(dx, dx, fma(dy, dy, dz * dz));
            return __double2hiint(s2) < a.lim_hi ? cell_i32_fast(s2, a.fd, a.ft) : cell_i32(ax, ay, az, bx, by, bz, a.fd, a.ft);
        }
        return __ldg(&a.cells[static_cast<size_t>(b) * a.n * a.pitch + static_cast<size_t>(Nd[pa]) * a.pitch + Nd[pb]]);
    };
    const bool f2 = a.families & 1, f_rel = a.families & 2, f_swap = a.families & 4;
    unsigned long long best = kNoMove;
    for (int t = rs + threadIdx.x; t < re - 1; t += blockDim.x) {  // anchors: the start separator and every stop but the last
        const bool t_cust = t != rs;
        const int4 tA = A[t], tB = B[t], tC = C[t], tD = D[t];
        const int dslack = tD.y;
        const int2 Et = make_int2(tA.z, tD.z);
        const int2 SKt = t_cust ? a.SK[po + t] : make_int2(0, 0);
        const int4 t1B = B[t + 1];
        const int dep_t = tA.x, dep_tm1 = tC.x;
        int2 cm = SKt, c0 = Et, cp = make_int2(0, 0), c0_prev = make_int2(0, 0);
        CSeg F = CSeg{0, 0, 0, true}, Rv = CSeg{0, 0, 0, true};
        int svc_xm1 = 0, Et_xm1 = 0;
        for (int x = t + 1; x < re; ++x) {
            const int4 xA = A[x], xB = B[x], xC = C[x], xD = D[x];
            const int2 Ex = make_int2(xA.z, xD.z);
            const int zn = xA.y;
            const int2 cm1 = t_cust ? cell(t - 1, x + 1) : make_int2(0, 0);
            const int2 c01 = cell(t, x + 1), cp1 = cell(t + 1, x + 1);
            const CSeg sx = cseg_single(xB.x, xB.y);
            Rv = (x == t + 1) ? sx : cseg_concat(sx, xB.z + Et_xm1, Rv);
            if (f2 && x >= t + 2) {
                const int d = c0.x + cp1.x - Et.x - Ex.x;
                if (d <= dslack && cseg_fits(Rv, dep_t + c0.y, t1B.z + cp1.y, zn)) {
                    const unsigned long long k = rel_key(d, 0, t - rs, x - rs);
                    best = k < best ? k : best;
                }
            }
            if (f_rel) {
                if (t_cust) {
                    CSeg m = (x == t + 1) ? sx : cseg_concat(F, svc_xm1 + Et_xm1, sx);
                    m = cseg_concat(m, xB.z + c0.y, cseg_single(tB.x, tB.y));
                    const int d = tC.y + c0.x + c01.x - Ex.x;
                    if (d <= dslack && cseg_fits(m, dep_tm1 + SKt.y, tB.z + c01.y, zn)) {
                        const unsigned long long k = rel_key(d, 1, t - rs, x - rs);
                        best = k < best ? k : best;
                    }
                }
                if (x >= t + 2) {
                    const CSeg m = cseg_concat(sx, xB.z + cp.y, F);
                    const int d = xC.y + c0.x + cp.x - Et.x;
                    if (d <= dslack && cseg_fits(m, dep_t + c0.y, svc_xm1 + xD.w, zn)) {
                        const unsigned long long k = rel_key(d, 1, x - rs, t - rs);
                        best = k < best ? k : best;
                    }
                }
            }
            if (f_swap && t_cust) {
                CSeg m;
                int d;
                if (x == t + 1) {
                    m = cseg_concat(sx, xB.z + Et.y, cseg_single(tB.x, tB.y));
                    d = cm.x + c01.x - (tC.z - Et.x) - Ex.x;
                } else {
                    m = cseg_concat(sx, xB.z + cp.y, F);
                    m = cseg_concat(m, svc_xm1 + c0_prev.y, cseg_single(tB.x, tB.y));
                    d = cm.x + cp.x - tC.z + c0_prev.x + c01.x - xC.z;
                }
                if (d <= dslack && cseg_fits(m, dep_tm1 + cm.y, tB.z + c01.y, zn)) {
                    const unsigned long long k = rel_key(d, 2, t - rs, x - rs);
                    best = k < best ? k : best;
                }
            }
            F = (x == t + 1) ? sx : cseg_concat(F, svc_xm1 + Et_xm1, sx);
            svc_xm1 = xB.z;
            Et_xm1 = Ex.y;
            c0_prev = c0;
            cm = cm1;
            c0 = c01;
            cp = cp1;
        }
    }
    if (best != kNoMove) atomicMin(a.intra + static_cast<size_t>(b) * a.cr + r, best);
}

// Minimum over the table with positions made absolute again -> best[b] (packed key of SPEC.md §7).  One block per
// source route ra (its row of the table plus its own intra entry); best[b] was reset by the state refresh.
__global__ void __launch_bounds__(128) cache_argmin_kernel(const CacheArgs a)
{
    __shared__ unsigned long long s_best[4];
    const int b = blockIdx.y, ra = blockIdx.x, tid = threadIdx.x;
    const int R = a.npos[b] - a.n;
    if (ra >= R) return;
    const int *rstart = a.rstart + static_cast<size_t>(b) * (a.rstride + 1);
    const unsigned long long *row = a.inter + (static_cast<size_t>(b) * a.cr + ra) * a.cr;
    const unsigned long long pa = static_cast<unsigned long long>(rstart[ra]);
    unsigned long long best = kNoMove;
    for (int rb = tid; rb <= R; rb += blockDim.x) {
        // entry R of the row stands for the route's own intra entry
        const bool self = rb == R;
        const unsigned long long k = self ? a.intra[static_cast<size_t>(b) * a.cr + ra] : row[rb];
        if (k == kNoMove) continue;
        const unsigned long long pr = (k >> 15) & 0x7FFF, qr = k & 0x7FFF;
        const unsigned long long qa = static_cast<unsigned long long>(rstart[self ? ra : rb]);
        const unsigned long long abs_k = (k & 0xFFFFFFFFC0000000ull) | ((pr + pa) << 15) | (qr + qa);
        best = abs_k < best ? abs_k : best;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, off);
        best = o < best ? o : best;
    }
    if ((tid & 31) == 0) s_best[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < 4; ++w) best = s_best[w] < best ? s_best[w] : best;
        if (best != kNoMove) atomicMin(a.best + b, best);
    }
}

__global__ void cache_all_dirty_kernel(int *dirty, int batch)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < batch) {
        dirty[2 * b] = kAllDirty;
        dirty[2 * b + 1] = -1;
    }
}

}  // namespace

int cache_capacity_routes(const rt_problem *p) { return std::max(1, p->grows - p->n); }

// The table fits (and the incremental path is used) up to this many routes per instance.
constexpr int kMaxCachedRoutes = 1024;

bool cache_usable(const rt_problem *p) { return p->symmetric && cache_capacity_routes(p) <= kMaxCachedRoutes; }

int cache_ensure(rt_problem *p)
{
    rt_engine *e = p->e;
    const int cr = cache_capacity_routes(p);
    const int nrt = (p->grows + kTileRows - 1) / kTileRows, nct = (p->grows + kTileCols - 1) / kTileCols;
    const int tile_cap = 4 * nrt * nct + 16;  // two touched routes, each listed as a row band and a column band
    const size_t B = p->batch;
    RT_TRY(reserve(e, p->cache_inter, B * cr * cr * 8));
    RT_TRY(reserve(e, p->cache_intra, B * cr * 8));
    RT_TRY(reserve(e, p->cache_tiles, B * tile_cap * 4));
    RT_TRY(reserve(e, p->cache_misc, B * 4 * 4));  // dirty[2], ntiles, cursor per instance
    p->cache_r = cr;
    p->cache_tile_cap = tile_cap;
    // the sweep kernel's cell table is dynamic shared memory on top of 35 KB of static arrays: past the 48 KB default
    RT_CUDA(e, cudaFuncSetAttribute(cache_intra_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    static_cast<int>(kTable * kTable * sizeof(int2))));
    return RT_OK;
}

static CacheArgs cache_args(rt_problem *p, int families)
{
    CacheArgs a{};
    a.node = p->node;
    a.npos = p->npos;
    a.route = p->route;
    a.rstart = p->rstart;
    a.RA = p->RA;
    a.RB = p->RB;
    a.RC = p->RC;
    a.RD = p->RD;
    a.H = p->H;
    a.E = p->E;
    a.SK = p->SK;
    a.PX = p->cells_from_coords ? p->PX : nullptr;
    a.cells = p->cells;
    a.fd = static_cast<float>(RT_EARTH_RADIUS_M * p->dist_scale);
    a.ft = static_cast<float>(RT_EARTH_RADIUS_M * p->time_scale);
    a.lim_hi = p->cells_from_coords ? fast_path_limit_hi(a.fd, a.ft) : 0;
    a.n = p->n;
    a.pitch = p->pitch;
    a.pstride = p->pstride;
    a.rstride = p->rstride;
    a.families = families;
    a.inter = static_cast<unsigned long long *>(p->cache_inter.p);
    a.intra = static_cast<unsigned long long *>(p->cache_intra.p);
    a.cr = p->cache_r;
    int *misc = static_cast<int *>(p->cache_misc.p);
    a.dirty = misc;
    a.ntiles = misc + 2 * p->batch;
    a.cursor = misc + 3 * p->batch;
    a.tiles = static_cast<int *>(p->cache_tiles.p);
    a.tile_cap = p->cache_tile_cap;
    a.best = p->best;
    return a;
}

// Route state without the route-order matrix: position vectors + tables.
static int cache_state_refresh(rt_problem *p, cudaStream_t st)
{
    return state_tables_refresh(p, st, /*with_vectors=*/true);
}

int cache_mark_all_dirty(rt_problem *p, cudaStream_t st)
{
    int *misc = static_cast<int *>(p->cache_misc.p);
    cache_all_dirty_kernel<<<(p->batch + 255) / 256, 256, 0, st>>>(misc, p->batch);
    RT_CUDA(p->e, cudaGetLastError());
    return RT_OK;
}

// One iteration: (state is fresh, dirty[] says what changed) -> best keys in p->best.
int cache_evaluate(rt_problem *p, int families, cudaStream_t st)
{
    rt_engine *e = p->e;
    const CacheArgs a = cache_args(p, families);
    if (!p->best_clean) {  // the arg-min accumulates into best[]: it must start from "no move"
        RT_CUDA(e, cudaMemsetAsync(p->best, 0xFF, static_cast<size_t>(p->batch) * 8, st));
    }
    cache_prepare_kernel<<<p->batch, 256, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    if (families & 6) {
        dim3 grid(kNumSMs * 4 / std::max(1, std::min(p->batch, kNumSMs * 4)) + 1, p->batch);
        if (a.PX) cache_inter_kernel<true><<<grid, kBandThreads, 0, st>>>(a);
        else cache_inter_kernel<false><<<grid, kBandThreads, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
    }
    dim3 gi(p->cache_r, p->batch);
    cache_intra_kernel<<<gi, 128, kTable * kTable * sizeof(int2), st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    dim3 ga(p->cache_r, p->batch);
    cache_argmin_kernel<<<ga, 128, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    p->best_clean = false;
    return RT_OK;
}

int cache_apply(rt_problem *p, bool improving_only, cudaStream_t st)
{
    int *misc = static_cast<int *>(p->cache_misc.p);
    RT_TRY(apply_only_launch(p, p->best, false, improving_only, st, misc));  // also notes the touched routes
    return cache_state_refresh(p, st);
}

int cache_begin(rt_problem *p, cudaStream_t st)
{
    RT_TRY(cache_ensure(p));
    RT_TRY(cache_state_refresh(p, st));
    return cache_mark_all_dirty(p, st);
}

}  // namespace rt
