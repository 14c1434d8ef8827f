// state_device.cuh — the device side of the route state (SPEC.md §6, §8): route scan, gather, per-route tables, packed
// records and the move application, as device functions, with the kernels of state.cu around them (compiled only where
// RT_STATE_KERNELS is defined).  Shared with moves.cu, whose persistent search kernel runs the same functions between
// grid-wide barriers.  Everything sits in an unnamed namespace: each translation unit gets its own copy.
//
// No reference counterpart: the reference's only solver loop is the greedy constructor (utils.py:111-139).
#pragma once
#include <algorithm>
#include <climits>

#include "geometry.cuh"
#include "problem.cuh"

namespace rt {

namespace {

struct RoutesArgs {
    const int *node, *npos;
    int *nroutes;
    int pstride, rstride;
    int *route, *rstart;
    long long *cost;
    int *feasible;
    unsigned long long *best;
    unsigned long long *ext_best;  // optional: the caller's key buffer (signed-order form), reset alongside
};


// The same scan for a 256-thread block that is the extra last block of a vector kernel (large instances): the route
// table then costs no launch of its own.  Warp w walks a contiguous slice of the positions 32 at a time (coalesced),
// the separators of a chunk are one ballot; pass 1 counts, pass 2 — after the eight warp totals have been combined —
// writes rstart[].  route[] is left to the per-route kernel.  Robust against a malformed node array (the plan entry
// points launch this before the host has finished validating it): nothing is written beyond rstride + 1 entries.
constexpr int kPrepThreads = 1024;  // the vector kernels are launched with this many threads when the scan rides along

template <int THREADS = kPrepThreads>
__device__ __forceinline__ void routes_scan_small_block(const RoutesArgs &a, int b)
{
    __shared__ int s_tot[THREADS / 32];
    __shared__ unsigned s_mask[kMaxPositions / 32];  // one ballot per 32 positions
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int P = min(a.npos[b], a.pstride);
    const int *node = a.node + static_cast<size_t>(b) * a.pstride;
    int *rstart = a.rstart + static_cast<size_t>(b) * (a.rstride + 1);
    const int chunks = (P + 31) / 32, per_warp = (chunks + THREADS / 32 - 1) / (THREADS / 32);
    const int c0 = min(chunks, wrp * per_warp), c1 = min(chunks, c0 + per_warp);
    // pass 1: eight independent loads in flight per lane (the ballots would otherwise serialise one memory round trip
    // per chunk), masks kept in shared memory for pass 2
    int cnt = 0;
    for (int c = c0; c < c1; c += 8) {
        int v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int p = (c + k) * 32 + lane;
            v[k] = (c + k < c1 && p < P) ? node[p] : 1;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const unsigned m = __ballot_sync(0xffffffffu, v[k] == 0);
            if (c + k < c1) {
                if (lane == 0) s_mask[c + k] = m;
                cnt += __popc(m);
            }
        }
    }
    if (lane == 0) s_tot[wrp] = cnt;
    __syncthreads();
    int before = 0, total = 0;
    for (int i = 0; i < THREADS / 32; ++i) {
        before += i < wrp ? s_tot[i] : 0;
        total += s_tot[i];
    }
    // pass 2: lane l takes the chunks c0 + l, c0 + l + 32, ... of its warp; the rank of a chunk's first separator is a
    // running sum over the warp's masks
    int r = before;
    for (int c = c0; c < c1; c += 32) {
        const unsigned m = c + lane < c1 ? s_mask[c + lane] : 0u;
        int pc = __popc(m), incl = pc;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += t;
        }
        int k = r + incl - pc;  // separators before this lane's chunk
        unsigned mm = m;
        while (mm) {
            const int bit = __ffs(mm) - 1;
            mm &= mm - 1;
            if (k <= a.rstride) rstart[k] = (c + lane) * 32 + bit;
            ++k;
        }
        r += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (tid == 0) {
        a.nroutes[b] = total - 1;
        a.cost[b] = 0;
        a.feasible[b] = 1;
        a.best[b] = kNoMove;
        if (a.ext_best) a.ext_best[b] = kNoMove ^ 0x8000000000000000ull;
    }
}

// PX[c][p] = xyz[c][node p]; positions beyond P repeat the depot so that G needs no bounds checks.
#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(kPrepThreads) position_vectors_kernel(const double *__restrict__ xyz,
                                                               const int *__restrict__ node,
                                                               const int *__restrict__ npos, int n, int pstride,
                                                               double *__restrict__ PX, const RoutesArgs ra, int with_scan)
{
    const int b = blockIdx.y;
    if (with_scan && blockIdx.x == gridDim.x - 1) {  // the extra block: route scan of this instance
        routes_scan_small_block(ra, b);
        return;
    }
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pstride) return;
    const int cur = p < npos[b] ? node[static_cast<size_t>(b) * pstride + p] : 0;
    const double *v = xyz + static_cast<size_t>(b) * 3 * n;
    double *px = PX + static_cast<size_t>(b) * 3 * pstride;
    px[p] = v[cur];
    px[pstride + p] = v[n + cur];
    px[2 * pstride + p] = v[2 * n + cur];
}
#endif

// Unit vectors straight into route order (SPEC.md §1 per position): PX[c][p] = unit vector of node[p], and, as every
// node occupies at least one position, the node-order array xyz on the side.  One launch instead of two on the plan
// path.  A node id outside [0, n) reads the depot instead (the host rejects such input, but only after this launch);
// a coordinate out of range is neutralised and reported through bad_flag = bad_value.
#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(kPrepThreads) position_unit_vectors_kernel(const double *__restrict__ lat,
                                                                    const double *__restrict__ lon,
                                                                    const int *__restrict__ node,
                                                                    const int *__restrict__ npos, int n, int pstride,
                                                                    double *__restrict__ xyz, double *__restrict__ PX,
                                                                    int *__restrict__ bad_flag, int bad_value,
                                                                    const RoutesArgs ra, int with_scan)
{
    const double DEG = 0.017453292519943295;
    const int b = blockIdx.y;
    if (with_scan && blockIdx.x == gridDim.x - 1) {  // the extra block: route scan of this instance
        routes_scan_small_block(ra, b);
        return;
    }
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pstride) return;
    int cur = p < npos[b] ? node[static_cast<size_t>(b) * pstride + p] : 0;
    if (static_cast<unsigned>(cur) >= static_cast<unsigned>(n)) cur = 0;
    double la = lat[static_cast<size_t>(b) * n + cur], lo = lon[static_cast<size_t>(b) * n + cur];
    if (bad_flag && (!(la >= -90.0 && la <= 90.0) || !(lo >= -180.0 && lo <= 180.0))) {
        *bad_flag = bad_value;
        la = 0.0;
        lo = 0.0;
    }
    double sp, cp, sl, cl;
    det_sincos(la * DEG, sp, cp);
    det_sincos(lo * DEG, sl, cl);
    const double x = cp * cl, y = cp * sl;
    double *px = PX + static_cast<size_t>(b) * 3 * pstride;
    px[p] = x;
    px[pstride + p] = y;
    px[2 * pstride + p] = sp;
    double *v = xyz + static_cast<size_t>(b) * 3 * n;  // duplicates (the depot) write identical values
    v[cur] = x;
    v[n + cur] = y;
    v[2 * n + cur] = sp;
}
#endif

// G[p][q] = M[node p][node q] for a stored (caller-supplied) matrix: coalesced writes, scattered reads within a row.
#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(256) permute_matrix_kernel(const int2 *__restrict__ cells,
                                                             const int *__restrict__ node,
                                                             const int *__restrict__ npos, int n, int pitch,
                                                             int pstride, int grows, int gpitch, int row0, int transpose,
                                                             int2 *__restrict__ G)
{
    const int b = blockIdx.z, p = row0 + blockIdx.y;
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= gpitch) return;
    const int P = npos[b];
    const int *nd = node + static_cast<size_t>(b) * pstride;
    const int a = p < P ? nd[p] : 0, c = q < P ? nd[q] : 0;
    const int from = transpose ? c : a, to = transpose ? a : c;
    G[(static_cast<size_t>(b) * grows + p) * gpitch + q] =
        __ldg(&cells[static_cast<size_t>(b) * n * pitch + static_cast<size_t>(from) * pitch + to]);
}
#endif

// K1: everything that depends only on node[] — one thread per position.  The two cells a position needs (edge to the
// successor, predecessor -> successor) are rebuilt from the position-ordered unit vectors (same cell function as the
// matrix kernel) or, for a caller-supplied matrix, gathered from it; either way not from G, so a replica that builds
// only its shard's rows of G still gets the state of every route.
// (The bodies of K1..K2c are device functions: the kernels below run them with the whole grid, state_fused_kernel runs
// all four for one instance in a single block.  No __restrict__ / read-only loads here: in the fused form the arrays a
// phase reads were written by the previous phase of the same kernel.)
struct GatherArgs {
    const double *PX;
    const int2 *cells;
    const int4 *stops;
    const int *node, *npos;
    int n, pitch, pstride;
    float fd, ft;
    int4 *RB;
    int2 *E, *SK, *ER;
    const double *xyz;
    double *PX_out;
};

__device__ __forceinline__ void gather_position(const GatherArgs &a, int b, int p)
{
    const double *PX = a.PX;
    const int2 *cells = a.cells;
    const int4 *stops = a.stops;
    const int n = a.n, pitch = a.pitch, pstride = a.pstride;
    const float fd = a.fd, ft = a.ft;
    int4 *RB = a.RB;
    int2 *E = a.E, *SK = a.SK, *ER = a.ER;
    const double *xyz = a.xyz;
    double *PX_out = a.PX_out;
    const int *node = a.node;
    const int P = a.npos[b];
    const int *nd = node + static_cast<size_t>(b) * pstride;
    if (xyz && p < pstride) {
        // fused position_vectors_kernel (the incremental search has no matrix kernel between the two): the vectors
        // of p-1, p, p+1 are read straight from the node-order array, p's own is also written out in route order
        const int cur = p < P ? nd[p] : 0;
        const double *v = xyz + static_cast<size_t>(b) * 3 * n;
        double *px = PX_out + static_cast<size_t>(b) * 3 * pstride;
        const double cx = v[cur], cy = v[n + cur], cz = v[2 * n + cur];
        px[p] = cx;
        px[pstride + p] = cy;
        px[2 * pstride + p] = cz;
        if (p >= P) return;
        const size_t o = static_cast<size_t>(b) * pstride + p;
        RB[o] = stops[static_cast<size_t>(b) * n + cur];
        int2 e = make_int2(0, 0), sk = make_int2(0, 0);
        if (p + 1 < P) {
            const int nx = nd[p + 1];
            const double nxx = v[nx], nxy = v[n + nx], nxz = v[2 * n + nx];
            e = cell_i32(cx, cy, cz, nxx, nxy, nxz, fd, ft);
            if (cur != 0) {
                const int pv = nd[p - 1];
                sk = cell_i32(v[pv], v[n + pv], v[2 * n + pv], nxx, nxy, nxz, fd, ft);
            }
        }
        E[o] = e;
        SK[o] = sk;
        return;
    }
    if (p >= P) return;
    const int cur = nd[p];
    const size_t o = static_cast<size_t>(b) * pstride + p;
    RB[o] = stops[static_cast<size_t>(b) * n + cur];
    int2 e = make_int2(0, 0), sk = make_int2(0, 0);
    if (PX) {
        const double *px = PX + static_cast<size_t>(b) * 3 * pstride, *py = px + pstride, *pz = py + pstride;
        if (p + 1 < P) {
            e = cell_i32(px[p], py[p], pz[p], px[p + 1], py[p + 1], pz[p + 1], fd, ft);
            if (cur != 0) sk = cell_i32(px[p - 1], py[p - 1], pz[p - 1], px[p + 1], py[p + 1], pz[p + 1], fd, ft);
        }
    } else {
        const int2 *M = cells + static_cast<size_t>(b) * n * pitch;
        if (p + 1 < P) {
            e = M[static_cast<size_t>(cur) * pitch + nd[p + 1]];
            if (cur != 0) sk = M[static_cast<size_t>(nd[p - 1]) * pitch + nd[p + 1]];
            if (ER) ER[o] = M[static_cast<size_t>(nd[p + 1]) * pitch + cur];
        } else if (ER) {
            ER[o] = make_int2(0, 0);
        }
    }
    E[o] = e;
    SK[o] = sk;
}

// K2a: one block per instance — route ids by a block scan over the separators; resets the per-instance results.
constexpr int kScanThreads = 1024;

// all kScanThreads threads of the block
__device__ __forceinline__ void routes_block(const RoutesArgs &a, int b)
{
    __shared__ int s_scan[kScanThreads];
    const int *node_all = a.node, *npos = a.npos;
    int *nroutes_out = a.nroutes;
    const int pstride = a.pstride, rstride = a.rstride;
    int *route_all = a.route, *rstart_all = a.rstart;
    long long *cost_all = a.cost;
    int *feasible_all = a.feasible;
    unsigned long long *best_all = a.best;
    const int tid = threadIdx.x;
    const int P = npos[b];
    const int *node = node_all + static_cast<size_t>(b) * pstride;
    int *route = route_all + static_cast<size_t>(b) * pstride;
    int *rstart = rstart_all + static_cast<size_t>(b) * (rstride + 1);
    // route[p] = #separators at positions <= p, minus 1
    const int chunk = (P + kScanThreads - 1) / kScanThreads;
    const int c0 = min(P, tid * chunk), c1 = min(P, c0 + chunk);
    int cnt = 0;
    for (int p = c0; p < c1; ++p) cnt += node[p] == 0;
    // inclusive scan of the per-thread counts: shuffles inside a warp, then the 32 warp totals by the first warp
    const int lane_ = tid & 31, warp_ = tid >> 5;
    int incl = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane_ >= off) incl += v;
    }
    __shared__ int s_wtot[kScanThreads / 32];
    if (lane_ == 31) s_wtot[warp_] = incl;
    __syncthreads();
    if (warp_ == 0) {
        int t = s_wtot[lane_];
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, t, off);
            if (lane_ >= off) t += v;
        }
        s_wtot[lane_] = t;  // inclusive over warps
    }
    __syncthreads();
    s_scan[tid] = incl + (warp_ > 0 ? s_wtot[warp_ - 1] : 0);
    __syncthreads();
    int r = s_scan[tid] - cnt - 1;  // separators before this chunk, minus 1
    for (int p = c0; p < c1; ++p) {
        if (node[p] == 0) {
            ++r;
            rstart[r] = p;
        }
        route[p] = r;
    }
    if (tid == 0) {
        nroutes_out[b] = s_scan[kScanThreads - 1] - 1;
        cost_all[b] = 0;
        feasible_all[b] = 1;
        best_all[b] = kNoMove;
        if (a.ext_best) a.ext_best[b] = kNoMove ^ 0x8000000000000000ull;
    }
}

#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(kScanThreads) state_routes_kernel(const RoutesArgs a) { routes_block(a, blockIdx.x); }
#endif

// K2b: one warp per route.  The recurrences for B (forwards) and Z (backwards) are sequential, but their inputs are
// not: each lane loads one position of a 32-position chunk (coalesced), then the warp replays the chunk in order
// from registers via shuffles — no dependent memory round trip per position.

struct WalkArgs {
    const int *npos;
    int n, pstride, rstride;
    const int4 *stops;
    const int *cap;
    const long long *maxd;
    const int4 *RB;
    const int2 *E;
    int *Bv, *Zv;
    const int *rstart;
    long long *rload, *rdist, *cost;
    int *feasible;
};

// one warp: route rr of instance b
__device__ __forceinline__ void walk_route(const WalkArgs &a, int b, int rr)
{
    const int *npos = a.npos;
    const int n = a.n, pstride = a.pstride, rstride = a.rstride;
    const int4 *stops_all = a.stops;
    const int *cap_all = a.cap;
    const long long *maxd_all = a.maxd;
    const int4 *RB_all = a.RB;
    const int2 *E_all = a.E;
    int *Bv_all = a.Bv, *Zv_all = a.Zv;
    const int *rstart_all = a.rstart;
    long long *rload_all = a.rload, *rdist_all = a.rdist, *cost_all = a.cost;
    int *feasible_all = a.feasible;
    const int lane = threadIdx.x & 31;
    const int P = npos[b];
    const int R = P - n;
    if (rr >= R) return;
    const size_t po = static_cast<size_t>(b) * pstride, ro = static_cast<size_t>(b) * rstride;
    const int4 *RB = RB_all + po;
    const int2 *E = E_all + po;
    int *Bv = Bv_all + po, *Zv = Zv_all + po;
    const int *rstart = rstart_all + ro + b;
    const int4 depot = stops_all[static_cast<size_t>(b) * n];
    const int s = rstart[rr], e = rstart[rr + 1];
    const unsigned full = 0xffffffffu;

    int Bcur = depot.x, svc_prev = depot.z;
    long long load = 0, dist = 0;
    bool feas = true;
    for (int base = s + 1; base < e; base += 32) {
        const int p = base + lane;
        const bool in = p < e;
        const int4 rb = in ? RB[p] : make_int4(0, 0, 0, 0);
        const int2 ed = in ? E[p - 1] : make_int2(0, 0);
        const int cntc = min(32, e - base);
        int myB = 0;
        for (int k = 0; k < cntc; ++k) {
            const int ready = __shfl_sync(full, rb.x, k), due = __shfl_sync(full, rb.y, k);
            const int svc = __shfl_sync(full, rb.z, k), dem = __shfl_sync(full, rb.w, k);
            const int ed_d = __shfl_sync(full, ed.x, k), ed_t = __shfl_sync(full, ed.y, k);
            Bcur = max(ready, Bcur + svc_prev + ed_t);
            feas = feas && Bcur <= due;
            if (k == lane) myB = Bcur;
            svc_prev = svc;
            load += dem;
            dist += ed_d;
        }
        if (in) Bv[p] = myB;
    }
    const int2 last = E[e - 1];
    feas = feas && (Bcur + svc_prev + last.y <= depot.y);
    dist += last.x;
    feas = feas && load <= cap_all[b] && dist <= maxd_all[b];

    int Zn = depot.y;
    for (int top = e - 1; top > s; top -= 32) {  // chunk [top-31, top], replayed from the highest position down
        const int p = top - lane;
        const bool in = p > s;
        const int4 rb = in ? RB[p] : make_int4(0, 0, 0, 0);
        const int2 ed = in ? E[p] : make_int2(0, 0);
        const int cntc = min(32, top - s);
        int myZ = 0;
        for (int k = 0; k < cntc; ++k) {
            const int due = __shfl_sync(full, rb.y, k), svc = __shfl_sync(full, rb.z, k);
            const int ed_t = __shfl_sync(full, ed.y, k);
            Zn = min(due, Zn - ed_t - svc);
            if (k == lane) myZ = Zn;
        }
        if (in) Zv[p] = myZ;
    }
    if (lane == 0) {
        Bv[s] = depot.x;
        Zv[s] = depot.y;
        if (rr == R - 1) {
            Bv[P - 1] = depot.x;
            Zv[P - 1] = depot.y;
        }
        rload_all[ro + rr] = load;
        rdist_all[ro + rr] = dist;
        atomicAdd(reinterpret_cast<unsigned long long *>(cost_all + b), static_cast<unsigned long long>(dist));
        if (!feas) feasible_all[b] = 0;
    }
}

// K2c: the packed per-position records, one thread per position.
struct RecordsArgs {
    const int *node, *npos;
    int n, pstride, rstride;
    const int4 *stops;
    const int *cap;
    const long long *maxd;
    const int4 *RB;
    const int2 *E, *SK;
    const int *Bv, *Zv, *route;
    const long long *rload, *rdist;
    int4 *RA, *RC, *RD, *H;
};

__device__ __forceinline__ void records_position(const RecordsArgs &a, int b, int p)
{
    const int *node_all = a.node, *npos = a.npos;
    const int n = a.n, pstride = a.pstride, rstride = a.rstride;
    const int4 *stops_all = a.stops;
    const int *cap_all = a.cap;
    const long long *maxd_all = a.maxd;
    const int4 *RB_all = a.RB;
    const int2 *E_all = a.E, *SK_all = a.SK;
    const int *Bv_all = a.Bv, *Zv_all = a.Zv, *route_all = a.route;
    const long long *rload_all = a.rload, *rdist_all = a.rdist;
    int4 *RA_all = a.RA, *RC_all = a.RC, *RD_all = a.RD, *H_all = a.H;
    const int P = npos[b];
    if (p >= P) return;
    const int R = P - n;
    const size_t po = static_cast<size_t>(b) * pstride, ro = static_cast<size_t>(b) * rstride, o = po + p;
    const int4 depot = stops_all[static_cast<size_t>(b) * n];
    const int rt_ = route_all[o];
    const int4 rb = RB_all[o];
    const int2 ed = E_all[o], sk = SK_all[o];
    const int dep = Bv_all[o] + rb.z;
    // Z of a separator is the depot's due time by definition: not read from memory, so a route's records depend on
    // nothing another route's warp writes
    const int zn = (p + 1 < P && node_all[o + 1] != 0) ? Zv_all[o + 1] : depot.y;
    RA_all[o] = make_int4(dep, zn, ed.x, rt_);
    // slacks are clamped to [-2^29, 2^30]: wide enough to compare against any demand / delta (both < 2^30 in
    // magnitude), narrow enough that "slack - value" never overflows in the move kernels
    constexpr long long kLo = -(1LL << 29), kHi = 1LL << 30;
    int capsl = -1, dsl = static_cast<int>(kLo);
    if (rt_ < R) {
        capsl = static_cast<int>(max(kLo, min(kHi, static_cast<long long>(cap_all[b]) - rload_all[ro + rt_])));
        dsl = static_cast<int>(max(kLo, min(kHi, maxd_all[b] - rdist_all[ro + rt_])));
    }
    RD_all[o] = make_int4(capsl, dsl, ed.y, sk.y);
    if (node_all[o] != 0) {
        const int2 ep = E_all[o - 1];
        const int dep_prev = Bv_all[o - 1] + RB_all[o - 1].z;
        const int rem_delta = sk.x - ep.x - ed.x;
        const bool rem_ok = dep_prev + sk.y <= zn && rem_delta <= dsl;
        RC_all[o] = make_int4(dep_prev, rem_delta, ep.x + ed.x, 1 | (rem_ok ? 2 : 0));
        H_all[o] = make_int4(rem_delta, ep.x + ed.x, ed.x, rt_ | (1 << 30));
    } else {
        RC_all[o] = make_int4(0, 0, 0, 0);
        H_all[o] = make_int4(0, 0, ed.x, rt_);
    }
}

// All four phases for one instance in one block (kScanThreads threads): what the four launches above do, minus three
// launch latencies — the tables of a small instance are a few microseconds of work behind ~20 µs of dependent launches.
struct StateArgs {
    GatherArgs g;
    RoutesArgs r;
    WalkArgs w;
    RecordsArgs c;
};

__device__ __forceinline__ void state_tables_block(const StateArgs &a, int b)
{
    const int tid = threadIdx.x;
    const int limit = a.g.xyz ? a.g.pstride : a.g.npos[b];
    for (int p = tid; p < limit; p += kScanThreads) gather_position(a.g, b, p);
    routes_block(a.r, b);  // reads node[] only; ends with every thread past its last barrier
    __syncthreads();
    const int P = a.g.npos[b], R = P - a.w.n;
    for (int rr = tid >> 5; rr < R; rr += kScanThreads / 32) walk_route(a.w, b, rr);
    __syncthreads();
    for (int p = tid; p < P; p += kScanThreads) records_position(a.c, b, p);
}

#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(kScanThreads) state_fused_kernel(const StateArgs a) { state_tables_block(a, blockIdx.x); }
#endif

// K2bc for large instances: ONE warp per route does everything the route needs — the gather of its positions (stop
// record, edge to the successor, skip edge), the forward/backward walk and the packed records — so that after the
// route scan a single launch finishes the tables (two dependent launches instead of four: the tables sit on the
// critical path of a sharded pass, where the matrix rows they used to hide under are 1/N of the work).  A route only
// reads what its own warp wrote (positions s .. e-1, s its opening separator), in program order behind __syncwarp().
constexpr int kRouteWarps = 4;

__device__ __forceinline__ void route_tables_warp(const StateArgs &a, int b, int rr)
{
    const WalkArgs &w = a.w;
    const RecordsArgs &c = a.c;
    const int lane = threadIdx.x & 31;
    const int P = w.npos[b];
    const int R = P - w.n;
    if (rr >= R) return;
    const size_t po = static_cast<size_t>(b) * w.pstride, ro = static_cast<size_t>(b) * w.rstride;
    const int *rstart = w.rstart + ro + b;
    const int s = rstart[rr], e = rstart[rr + 1];
    const bool last = rr == R - 1;
    const unsigned full = 0xffffffffu;
    int *route = a.r.route + po;
    // gather: this route's positions, plus (last route) the closing depot
    const int g_end = last ? P : e;
    for (int p = s + lane; p < g_end; p += 32) {
        gather_position(a.g, b, p);
        route[p] = p < e ? rr : R;
    }
    // fused-vectors form: the unused tail of the position arrays repeats the depot's vector; spread over all warps
    if (a.g.xyz)
        for (int p = P + rr * 32 + lane; p < a.g.pstride; p += 32 * R) gather_position(a.g, b, p);
    __syncwarp();
    const int4 *RB = w.RB + po;
    const int2 *E = w.E + po;
    int *Bv = w.Bv + po, *Zv = w.Zv + po;
    const int4 depot = w.stops[static_cast<size_t>(b) * w.n];
    // Forward: B[p] = max(ready[p], B[p-1] + service[p-1] + T[p-1][p]) is the composition of the maps
    // t -> max(t + w, ready); two of them compose to (w1 + w2, max(r1 + w2, r2)), so the B of 32 consecutive positions
    // is one warp scan (five shuffle steps) applied to the value carried in from the chunk below.  Z likewise,
    // downwards, with t -> min(t - c, due).  Exact integer max-plus algebra: the same values as the step-by-step walk.
    int Bcur = depot.x;
    long long load = 0, dist = 0;
    bool feas = true;
    int svc_last = depot.z;  // service time of the position just below the chunk
    for (int base = s + 1; base < e; base += 32) {
        const int p = base + lane;
        const bool in = p < e;
        const int4 rb = in ? RB[p] : make_int4(INT_MIN / 2, INT_MAX, 0, 0);
        const int2 ed = in ? E[p - 1] : make_int2(0, 0);
        int svc_prev = __shfl_up_sync(full, rb.z, 1);
        if (lane == 0) svc_prev = svc_last;
        int fa = in ? svc_prev + ed.y : 0, fb = in ? rb.x : INT_MIN / 2;   // identity map for lanes beyond the route
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int la = __shfl_up_sync(full, fa, off), lb = __shfl_up_sync(full, fb, off);
            if (lane >= off) {
                fb = max(lb + fa, fb);
                fa = la + fa;
            }
        }
        const int myB = max(Bcur + fa, fb);
        feas = feas && __all_sync(full, !in || myB <= rb.y);
        if (in) Bv[p] = myB;
        long long ld = in ? rb.w : 0, dd = in ? ed.x : 0;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            ld += __shfl_xor_sync(full, ld, off);
            dd += __shfl_xor_sync(full, dd, off);
        }
        load += ld;
        dist += dd;
        const int lastl = min(31, e - 1 - base);
        Bcur = __shfl_sync(full, myB, lastl);
        svc_last = __shfl_sync(full, rb.z, lastl);
    }
    const int2 lastE = E[e - 1];
    feas = feas && (Bcur + svc_last + lastE.y <= depot.y);
    dist += lastE.x;
    feas = feas && load <= w.cap[b] && dist <= w.maxd[b];
    int Zn = depot.y;
    for (int top = e - 1; top > s; top -= 32) {  // chunk [top-31, top], lane l holds position top - l
        const int p = top - lane;
        const bool in = p > s;
        const int4 rb = in ? RB[p] : make_int4(0, INT_MAX, 0, 0);
        const int2 ed = in ? E[p] : make_int2(0, 0);
        int gc = in ? ed.y + rb.z : 0, gd = rb.y;  // z -> min(z - gc, gd)
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int lc = __shfl_up_sync(full, gc, off), ldue = __shfl_up_sync(full, gd, off);
            if (lane >= off) {
                gd = min(ldue - gc, gd);
                gc = lc + gc;
            }
        }
        const int myZ = min(Zn - gc, gd);
        if (in) Zv[p] = myZ;
        Zn = __shfl_sync(full, myZ, min(31, top - s - 1));
    }
    if (lane == 0) {
        Bv[s] = depot.x;
        Zv[s] = depot.y;
        if (last) {
            Bv[P - 1] = depot.x;
            Zv[P - 1] = depot.y;
        }
        w.rload[ro + rr] = load;
        w.rdist[ro + rr] = dist;
        atomicAdd(reinterpret_cast<unsigned long long *>(w.cost + b), static_cast<unsigned long long>(dist));
        if (!feas) w.feasible[b] = 0;
    }
    __syncwarp();
    // records: position p reads B/Z/E of p-1, p, p+1 — all inside [s, e] (Z of the next separator is the depot's due
    // time, which records_position takes from the depot record when p + 1 is the last position; for an inner
    // separator e the next route's warp writes the same constant, so either order gives the same value)
    const int r_end = last ? P : e;
    for (int p = s + lane; p < r_end; p += 32) records_position(c, b, p);
}

#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(kRouteWarps * 32) state_route_tables_kernel(const StateArgs a)
{
    route_tables_warp(a, blockIdx.y, blockIdx.x * kRouteWarps + (threadIdx.x >> 5));
}
#endif


// Applies one move per instance (SPEC.md §8).  keys are the packed keys of SPEC.md §7, optionally in the
// order-preserving signed form (key ^ 2^63) that the _dev entry points exchange with NCCL.
struct ApplyArgs {
    const unsigned long long *keys;
    int signed_keys, improving_only;
    int *node_all, *tmp_all;
    const int *npos;
    int pstride;
    int *done, *moves;
    unsigned long long *hist;
    int hist_cap;
    const int *route_all;
    int *dirty;
};

// Every branch that leaves early is uniform over the block and comes before the one barrier.
__device__ __forceinline__ void apply_move_block(const ApplyArgs &ap, int b)
{
    const unsigned long long *__restrict__ keys = ap.keys;
    const int signed_keys = ap.signed_keys, improving_only = ap.improving_only, pstride = ap.pstride, hist_cap = ap.hist_cap;
    int *__restrict__ node_all = ap.node_all, *__restrict__ tmp_all = ap.tmp_all;
    const int *__restrict__ npos = ap.npos, *__restrict__ route_all = ap.route_all;
    int *__restrict__ done = ap.done, *__restrict__ moves = ap.moves, *__restrict__ dirty = ap.dirty;
    unsigned long long *__restrict__ hist = ap.hist;
    const int tid = threadIdx.x;
    unsigned long long key = keys[b];
    if (signed_keys) key ^= 0x8000000000000000ull;
    if (dirty && tid == 0) {  // incremental search: which routes this move touches (none if nothing is applied)
        dirty[2 * b] = -1;
        dirty[2 * b + 1] = -1;
    }
    if (done[b]) return;
    const int delta = static_cast<int>(static_cast<unsigned>(key >> 32) - 0x80000000u);
    if (key == kNoMove || (improving_only && delta >= 0)) {
        if (tid == 0) done[b] = 1;
        return;
    }
    const int family = static_cast<int>((key >> 30) & 3), p = static_cast<int>((key >> 15) & 0x7FFF),
              q = static_cast<int>(key & 0x7FFF);
    const int P = npos[b];
    int *node = node_all + static_cast<size_t>(b) * pstride;
    int *tmp = tmp_all + static_cast<size_t>(b) * pstride;
    // A key that is not a move of SPEC.md §7 is refused (done = 2) rather than applied: whatever the caller hands in,
    // the first and last positions stay the depot and the number of separators does not change, which is what every
    // other kernel's indexing relies on.
    bool good = family != 3 && p < P && q < P;
    if (good) {
        if (family == 2) good = node[p] != 0 && node[q] != 0;
        else if (family == 1) good = node[p] != 0 && q != p && q <= P - 2;
        else good = p < q && q <= P - 2;
    }
    if (!good) {
        if (tid == 0) done[b] = 2;
        return;
    }
    if (dirty && tid == 0) {  // route ids of the state BEFORE the move (the tables are refreshed afterwards)
        const int *route = route_all + static_cast<size_t>(b) * pstride;
        dirty[2 * b] = family == 0 ? route[p + 1] : route[p];
        dirty[2 * b + 1] = family == 0 ? -1 : route[q];
    }
    if (family == 2) {
        if (tid == 0) {
            int t = node[p];
            node[p] = node[q];
            node[q] = t;
        }
    } else {
        // affected span [lo, hi]; new[i] = old[src(i)]
        int lo, hi;
        if (family == 1) {
            lo = q > p ? p : q + 1;
            hi = q > p ? q : p;
        } else {
            lo = p + 1;
            hi = q;
        }
        for (int i = lo + tid; i <= hi; i += blockDim.x) tmp[i] = node[i];
        __syncthreads();
        for (int i = lo + tid; i <= hi; i += blockDim.x) {
            int src;
            if (family == 1) src = q > p ? (i == q ? p : i + 1) : (i == q + 1 ? p : i - 1);
            else src = lo + hi - i;
            node[i] = tmp[src];
        }
    }
    if (tid == 0) {
        const int m = moves[b];
        if (hist && m < hist_cap) hist[static_cast<size_t>(b) * hist_cap + m] = key;
        moves[b] = m + 1;
    }
}

#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(256) apply_moves_kernel(const ApplyArgs ap) { apply_move_block(ap, blockIdx.x); }
#endif

// The one-move search on small and medium instances is a chain of short launches, each waiting for the one before
// (about 3 us per link in a replayed graph, beside 4-5 us of dependent loads inside each).  This kernel is up to four
// links in one, one block per instance: the move, the route scan of the new order and — where every route gets a warp
// of its own (with_tables) — the per-route tables, whose gather also writes the position-ordered vectors.  The key
// is read before the first barrier, the scan's reset of best[] comes after it.
#ifdef RT_STATE_KERNELS
__global__ void __launch_bounds__(kPrepThreads) apply_state_kernel(const ApplyArgs ap, const StateArgs a, int with_tables)
{
    const int b = blockIdx.x;
    apply_move_block(ap, b);
    __syncthreads();
    routes_scan_small_block(a.r, b);
    if (!with_tables) return;
    __syncthreads();
    const int R = a.w.npos[b] - a.w.n;
    for (int rr = threadIdx.x >> 5; rr < R; rr += kPrepThreads / 32) route_tables_warp(a, b, rr);
}
#endif

// ---- the argument blocks of a problem (host) ----

static RoutesArgs routes_args(rt_problem *p)
{
    return RoutesArgs{p->node, p->npos, p->nroutes, p->pstride, p->rstride, p->route, p->rstart, p->cost, p->feasible, p->best,
                      p->ext_best};
}

static StateArgs state_args(rt_problem *p, bool fused_vectors)
{
    StateArgs a{};
    a.g = GatherArgs{p->cells_from_coords ? p->PX : nullptr, p->cells, p->stops, p->node, p->npos, p->n, p->pitch, p->pstride,
                     static_cast<float>(RT_EARTH_RADIUS_M * p->dist_scale), static_cast<float>(RT_EARTH_RADIUS_M * p->time_scale),
                     p->RB, p->E, p->SK, p->symmetric ? nullptr : p->ER, fused_vectors ? p->xyz : nullptr, p->PX};
    a.r = routes_args(p);
    a.w = WalkArgs{p->npos, p->n, p->pstride, p->rstride, p->stops, p->cap, p->max_dist, p->RB, p->E, p->Bv, p->Zv, p->rstart,
                   p->rload, p->rdist, p->cost, p->feasible};
    a.c = RecordsArgs{p->node, p->npos, p->n, p->pstride, p->rstride, p->stops, p->cap, p->max_dist, p->RB, p->E, p->SK, p->Bv,
                      p->Zv, p->route, p->rload, p->rdist, p->RA, p->RC, p->RD, p->H};
    return a;
}

static ApplyArgs apply_args(rt_problem *p, const unsigned long long *keys_d, bool signed_keys, bool improving_only,
                            int *dirty_d)
{
    return ApplyArgs{keys_d, signed_keys ? 1 : 0, improving_only ? 1 : 0, p->node, p->node_tmp, p->npos, p->pstride,
                     p->done,  p->moves, p->hist, p->hist_cap, p->route, dirty_d};
}

}  // namespace

}  // namespace rt
