// moves.cu — cost delta + capacity/distance/time-window feasibility of local-search moves and the best-move
// argmin (SPEC.md §7).  No reference counterpart (the reference has no improvement phase, utils.py:111-139).
//
// Where the matrix is SPEC.md §2 of the problem's own coordinates (the engine's normal case) NO matrix is read: every
// kernel rebuilds the cells it needs from the position-ordered unit vectors with the device functions the matrix
// kernel itself uses (geometry.cuh), so a pass moves O(n) bytes.  A caller-supplied matrix (road distances, possibly
// directional) is read in ROUTE ORDER instead, G[p][q] = M[node p][node q] (state.cu): a thread that owns column q
// and walks down the rows p touches G[p-1..p+1][q-1..q+1], each warp-wide load one contiguous 256-byte run.
//   * eval_inter_fused_kernel / eval_inter_kernel — candidates between two different routes (relocate, swap).  The
//     fused form scores the triangle p < q only: relocate(p -> q), relocate(q -> p) and swap(p, q) come from one
//     window of cells, so each cell is rebuilt once per unordered pair.  Delta
//     AND the full feasibility test (capacity, distance limit, time windows) for every pair, straight line — in a
//     realistic state a large share of the candidates beats the best feasible delta while being infeasible, so pruning
//     on the delta leaves almost every warp with some lane in the slow path.  What IS skipped is the best-key update:
//     one warp vote against the smallest best delta of the warp.
//   * eval_intra_kernel — candidates inside one route (2-opt, relocate, swap): one thread per anchor position sweeps
//     its route keeping two segment summaries (a, b, L) — forward and reversed — so every candidate is O(1) and exact
//     for time windows.  eval_intra_warp_kernel: the same summaries as two warp scans, one warp per anchor.
//   * search_persistent_kernel — the whole one-move search (SPEC.md §8) of small instances as ONE cooperative launch:
//     the device functions above and those of state_device.cuh in a loop, grid-wide barriers between the phases.
// Each thread keeps the minimum packed key of its feasible candidates; warp shuffles and one shared-memory step
// reduce it to one 64-bit atomicMin per block.  The minimum of packed keys is order independent, so the chosen
// move does not depend on the thread, block or GPU partition.
//
// ASYM: a caller-supplied matrix may be directional (road networks are).  The kernels then read a second array,
// GT[p][q] = M[node q][node p], wherever a cell is needed in the "towards p" orientation, so every access stays
// row-contiguous; for a symmetric matrix GT is G and the extra loads disappear at compile time.
#include <algorithm>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include <cooperative_groups.h>

#include "geometry.cuh"
#include "problem.cuh"
#include "state_device.cuh"

namespace rt {

int g_fused_eval = 1;  // see problem.cuh
int g_persistent_search = 1;  // rt_debug_set("persistent_search", v): 0 = never, 1 = where it is the faster form,
                              // 2 = wherever it can run (see search_persistent_launch)
int g_persist_blocks = 0;     // rt_debug_set("persist_blocks", k): at most k blocks per SM in that kernel (0 = all that fit)

namespace {

constexpr int kInterThreads = 128;  // one column q per thread
constexpr int kStripRows = 32;     // default rows p per block
constexpr int kMaxStripRows = 64;  // the launcher picks the height per launch (pick_strip_rows)
constexpr int kPersistMaxPositions = 4608;  // the one-launch search (search_persistent_kernel) can run up to here ...
constexpr int kPersistPositions = 256;      // ... and is the faster form up to here (see search_persistent_launch)

// a + b on the FMA pipe (IMAD with multiplier 1): the move kernels are bound by the ALU pipe (compares, min/max,
// selects, three-input adds), while the FMA pipe idles — plain sums go there.
__constant__ int kOne = 1;  // a multiplier the assembler cannot see through: with a literal 1 it folds pairs of these
                            // sums back into three-input IADD3s on the ALU pipe

__device__ __forceinline__ int fadd(int a, int b)
{
    int r;
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(kOne), "r"(b));
    return r;
}

__device__ __forceinline__ unsigned long long pack_key(int delta, int family, int p, int q)
{
    return (static_cast<unsigned long long>(static_cast<unsigned>(delta) + 0x80000000u) << 32) |
           (static_cast<unsigned long long>(family) << 30) | (static_cast<unsigned long long>(p) << 15) |
           static_cast<unsigned long long>(q);
}

__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        unsigned long long o = __shfl_xor_sync(0xffffffffu, v, off);
        v = o < v ? o : v;
    }
    return v;
}

// signed_dst: the destination holds keys in order-preserving signed form (key ^ 2^63, "none" = INT64_MAX) — what the
// replicas exchange — so the minimum is taken with the signed atomic and no conversion pass is needed afterwards.
template <int THREADS>
__device__ __forceinline__ void block_min_to_global(unsigned long long best, unsigned long long *dst, bool signed_dst = false)
{
    __shared__ unsigned long long s_best[THREADS / 32];
    best = warp_min_u64(best);
    if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned long long v = threadIdx.x < THREADS / 32 ? s_best[threadIdx.x] : kNoMove;
        v = warp_min_u64(v);
        if (threadIdx.x == 0 && v != kNoMove) {
            if (signed_dst) atomicMin(reinterpret_cast<long long *>(dst), static_cast<long long>(v ^ 0x8000000000000000ull));
            else atomicMin(dst, v);
        }
    }
}

// minimum into a table entry that holds either packed keys or their signed-order form (see block_min_to_global)
__device__ __forceinline__ void table_min(unsigned long long *dst, unsigned long long key, int signed_dst)
{
    if (signed_dst) atomicMin(reinterpret_cast<long long *>(dst), static_cast<long long>(key ^ 0x8000000000000000ull));
    else atomicMin(dst, key);
}

struct EvalArgs {
    const int2 *G;         // [batch][grows][gpitch] route-order matrix, G[p][q] = M[node p][node q]
    const int2 *GT;        // its transpose GT[p][q] = M[node q][node p] (asymmetric matrices only)
    const int2 *ER;        // ER[p] = M[node p+1][node p], the edge after p driven backwards (asymmetric only)
    int grows, gpitch;
    const int *node;
    const int *npos;
    const int4 *RA, *RB, *RC, *RD, *H;
    const int2 *E, *SK;
    const double *PX;        // [batch][3][pstride] position-ordered unit vectors (the no-G kernels rebuild cells from them)
    float fd, ft;            // R * dist_scale, R * time_scale
    int lim_hi;              // fast_path_limit_hi(fd, ft)
    const int *rstart;       // [batch][rstride + 1] separator positions (route r owns rstart[r] .. rstart[r+1]-1)
    int rstride;
    int n, pitch, pstride;
    int families;            // bit 0 = 2-opt, 1 = relocate, 2 = swap
    int shard_index, shard_count;
    double beta_inter, beta_intra;  // work model of the shard boundaries (problem.cuh: shard_boundary)
    unsigned long long *best;  // [batch]
    int signed_dst;            // best[] is in signed-order form (see block_min_to_global)
    int strip_rows;            // rows per block of the tile kernel (<= its ROWS template bound)
    // table mode (batched iterations, batch.cu): best key per ordered route pair / per route, absolute positions
    unsigned long long *inter, *intra;  // [batch][cr][cr], [batch][cr]
    int cr;
    // dump mode (one instance, one family)
    int dump_instance, dump_family;
    int *dump_delta;
    unsigned char *dump_feas;
};

__device__ __forceinline__ void shard_range(int P, int idx, int cnt, double beta, int &lo, int &hi)
{
    shard_bounds(P, idx, cnt, beta, lo, hi);
}

enum { kBest = 0, kDump = 1, kTable = 2 };

// Everything a thread of the tile kernels knows about its column q.
struct ColQ {
    int4 Hq, qB;
    int dep_q, zn_q, depprev_q, capsl_q, distsl_q;
    int q, q_route;
    bool rel_q, swap_q;
    // validity folded into data (score_pair_tri, non-dump modes): a candidate that is not one fails a comparison it
    // makes anyway instead of costing flag tests and predicate logic on the ALU pipe
    int capsl_rel;   // capacity slack of q's route for relocate(p -> q); below any demand where q takes no insertion
    int qdem_mir;    // demand of q for relocate(q -> p); above any slack where q cannot be moved
    int qdue_swap;   // due time of q for swap; -1 where q cannot be swapped
    int ncapsl_q;    // -capsl_q: swap's second capacity test as dd >= -capsl_q, no negation per pair
};

// relocate(p -> between q and q+1) and swap(p, q) for one customer row p against column q: delta AND feasibility,
// straight line.  Cells: c_up = (p-1, q), c_rt = (p, q+1); a_cur = (p, q), a_lf = (p, q-1), a_dn = (p+1, q) in the
// orientation that leads towards p's side (the same cells when the matrix is symmetric).
// thr: the delta a candidate has to reach to matter.  In best-key mode it is the smallest best delta any lane of the
// warp holds (the block's minimum is all that leaves the kernel, so a candidate above another lane's best can never be
// it): new records then become rare — a handful per warp and strip instead of a few per lane — and one warp vote
// skips the whole update; without it the compiler predicates the update, a dozen issue slots per row spent on nothing.
// In table mode every lane feeds its own (route, route) entry and thr is simply its own best delta.
template <int MODE>
__device__ __forceinline__ void score_pair(const EvalArgs &a, int P, int p, const int4 &Hp, const int4 &p1, const int4 &p2,
                                           const ColQ &c, bool swap_tile, int2 c_up, int2 c_rt, int2 a_cur, int2 a_lf,
                                           int2 a_dn, int &best_d, unsigned &best_i, int &thr)
{
    constexpr bool DUMP = MODE == kDump, TABLE = MODE == kTable;
    const int q = c.q;
    const bool other = (Hp.w & 0x7FFF) != c.q_route;
    // relocate(p -> between q and q+1): a -> u -> b replaces a -> b in route[q]
    const int ins = fadd(fadd(a_cur.x, c_rt.x), -c.Hq.z);
    const int d_rel = fadd(Hp.x, ins);
    const int tu = max(p1.x, fadd(c.dep_q, a_cur.y));
    const bool v_rel = c.rel_q && other;
    const bool ok_rel = v_rel && ((Hp.w >> 29) & 1) && p1.w <= c.capsl_q && ins <= c.distsl_q && tu <= p1.y &&
                        fadd(fadd(tu, p1.z), c_rt.y) <= c.zn_q;
    bool hit_rel = ok_rel && d_rel <= thr, hit_swap = false;
    int d_swap = 0;
    if (DUMP) {
        if (v_rel && a.dump_family == 1) {
            a.dump_delta[static_cast<size_t>(p) * P + q] = d_rel;
            a.dump_feas[static_cast<size_t>(p) * P + q] = ok_rel ? 1 : 0;
        }
    }
    if (swap_tile) {
        // swap(p, q), p < q: pu -> v -> nu in route[p], pw -> u -> nw in route[q]
        const int d1 = fadd(fadd(c_up.x, a_dn.x), -Hp.y);
        const int d2 = fadd(fadd(a_lf.x, c_rt.x), -c.Hq.y);
        d_swap = fadd(d1, d2);
        const int tv = max(c.qB.x, fadd(p2.y, c_up.y));
        const int tw = max(p1.x, fadd(c.depprev_q, a_lf.y));
        const int dd = fadd(c.qB.w, -p1.w);
        const bool v_swap = c.swap_q && other && p < q;
        const bool ok_swap = v_swap && dd <= p2.z && -dd <= c.capsl_q && d1 <= p2.w && d2 <= c.distsl_q &&
                             tv <= c.qB.y && fadd(fadd(tv, c.qB.z), a_dn.y) <= p2.x && tw <= p1.y &&
                             fadd(fadd(tw, p1.z), c_rt.y) <= c.zn_q;
        hit_swap = ok_swap && d_swap <= thr;
        if (DUMP) {
            if (v_swap && a.dump_family == 2) {
                a.dump_delta[static_cast<size_t>(p) * P + q] = d_swap;
                a.dump_feas[static_cast<size_t>(p) * P + q] = ok_swap ? 1 : 0;
            }
        }
    }
    if (DUMP) return;
    if (__any_sync(0xffffffffu, hit_rel || hit_swap)) {  // rare
        if (hit_rel) {
            const unsigned idx = (1u << 30) | (static_cast<unsigned>(p) << 15) | static_cast<unsigned>(q);
            if (d_rel < best_d || (d_rel == best_d && idx < best_i)) {
                best_d = d_rel;
                best_i = idx;
            }
        }
        if (hit_swap) {
            const unsigned idx = (2u << 30) | (static_cast<unsigned>(p) << 15) | static_cast<unsigned>(q);
            if (d_swap < best_d || (d_swap == best_d && idx < best_i)) {
                best_d = d_swap;
                best_i = idx;
            }
        }
        thr = TABLE ? best_d : __reduce_min_sync(0xffffffffu, best_d);
    }
}

// grid (column strips of 128, row strips of kStripRows within the shard, instances).
// Straight-line evaluation: in a realistic state a large share of the candidates has a delta below the best
// feasible one while being infeasible (capacity, windows), so pruning on the delta leaves almost every warp with
// some lane in the slow path; computing delta AND feasibility for every candidate without a branch is cheaper.
// Per row three broadcast LDS.128 bring everything about p; everything about q lives in registers.
// MODE: kBest = one key per instance, kDump = every candidate of one family (tests), kTable = one key per ordered
// route pair (the thread's column fixes the second route; its best is flushed whenever the row's route changes).
template <int MODE, bool ASYM>
__global__ void __launch_bounds__(kInterThreads, ASYM ? 6 : 9) eval_inter_kernel(const EvalArgs a)
{
    constexpr bool DUMP = MODE == kDump, TABLE = MODE == kTable;
    __shared__ int4 sH[kMaxStripRows];   // {rem_delta, oe, -, route | customer << 30 | rem_ok << 29}
    __shared__ int4 sP1[kMaxStripRows];  // {ready, due, service, demand}
    __shared__ int4 sP2[kMaxStripRows];  // {Z[p+1], B[p-1]+service[p-1], capacity slack, distance slack} of p's route
    const int kStripRows = a.strip_rows;  // this launch's strip height (block-uniform)

    const int b = DUMP ? a.dump_instance : blockIdx.z;
    const int P = a.npos[b];
    int p_lo, p_hi;
    shard_range(P, a.shard_index, a.shard_count, a.beta_inter, p_lo, p_hi);
    const int p0 = p_lo + blockIdx.y * kStripRows, q0 = blockIdx.x * kInterThreads;
    if (p0 >= p_hi || q0 >= P) return;

    const size_t po = static_cast<size_t>(b) * a.pstride;
    const int2 *G = a.G + static_cast<size_t>(b) * a.grows * a.gpitch;
    const int tid = threadIdx.x;
    const int q = q0 + tid;
    const bool q_in = q < P;
    if (tid < kStripRows) {
        int4 h = make_int4(0, 0, 0, 0), p1 = h, p2 = h;
        if (p0 + tid < P) {
            const size_t o = po + p0 + tid;
            const int4 rA = a.RA[o], rC = a.RC[o], rD = a.RD[o];
            h = a.H[o];
            h.w |= (rC.w & 2) << 28;  // rem_ok -> bit 29
            p1 = a.RB[o];
            p2 = make_int4(rA.y, rC.x, rD.x, rD.y);
        }
        sH[tid] = h;
        sP1[tid] = p1;
        sP2[tid] = p2;
    }
    int4 Hq = make_int4(0, 0, 0, -1), qB = make_int4(0, 0, 0, 0);
    int dep_q = 0, zn_q = 0, depprev_q = 0, capsl_q = 0, distsl_q = 0;
    if (q_in) {
        const int4 rA = a.RA[po + q], rC = a.RC[po + q], rD = a.RD[po + q];
        Hq = a.H[po + q];
        qB = a.RB[po + q];
        dep_q = rA.x;
        zn_q = rA.y;
        depprev_q = rC.x;
        capsl_q = rD.x;
        distsl_q = rD.y;
    }
    __syncthreads();

    const int qc = min(q, a.gpitch - 1);  // lanes beyond the last position re-read a pad column
    const int q_route = Hq.w & 0x7FFF;
    const bool q_cust = q_in && ((Hq.w >> 30) & 1);
    const bool rel_q = q_in && (a.families & 2) && q <= P - 2;
    const bool swap_q = q_cust && (a.families & 4);
    const ColQ col{Hq, qB, dep_q, zn_q, depprev_q, capsl_q, distsl_q, q, q_route, rel_q, swap_q, 0, 0, 0, 0};
    const bool swap_tile = (a.families & 4) && p0 < q0 + kInterThreads - 1;  // some pair of the tile has p < q

    int best_d = INT_MAX, thr = INT_MAX;
    unsigned best_i = 0xFFFFFFFFu;  // family << 30 | p << 15 | q
    int cur_ra = -1;                // table mode: the route the running best belongs to
    unsigned long long *trow = TABLE ? a.inter + static_cast<size_t>(b) * a.cr * a.cr + q_route : nullptr;
    auto flush = [&]() {
        if (best_i != 0xFFFFFFFFu)
            table_min(trow + static_cast<size_t>(cur_ra) * a.cr,
                      (static_cast<unsigned long long>(static_cast<unsigned>(best_d) + 0x80000000u) << 32) | best_i, a.signed_dst);
        best_d = INT_MAX;
        thr = INT_MAX;
        best_i = 0xFFFFFFFFu;
    };
    const int rows = min(kStripRows, p_hi - p0);
    // Register window over three rows of G, loaded two rows ahead of use.  G carries a pad cell in front and a dozen
    // pad rows behind (ensure_position_matrix), so q-1 / q+1 / p+2 need no clamping: whatever an out-of-range
    // neighbour holds is only ever combined into candidates that are not valid.
    const int2 *g = G + static_cast<size_t>(max(p0 - 1, 0)) * a.gpitch + qc;
    int2 c_up = __ldg(g);
    g = G + static_cast<size_t>(p0) * a.gpitch + qc;
    int2 c_lf = __ldg(g - 1), c_cur = __ldg(g), c_rt = __ldg(g + 1);
    g += a.gpitch;
    int2 n_lf = __ldg(g - 1), n_cur = __ldg(g), n_rt = __ldg(g + 1);
    // the same window over GT (cells towards p): only (row, q-1) and (row, q) are ever needed
    const int2 *gt = ASYM ? a.GT + static_cast<size_t>(b) * a.grows * a.gpitch + static_cast<size_t>(p0) * a.gpitch + qc : g;
    int2 t_lf = c_lf, t_cur = c_cur, tn_lf = n_lf, tn_cur = n_cur;
    if (ASYM) {
        t_lf = __ldg(gt - 1);
        t_cur = __ldg(gt);
        gt += a.gpitch;
        tn_lf = __ldg(gt - 1);
        tn_cur = __ldg(gt);
    }
#pragma unroll 2
    for (int rr = 0; rr < rows; ++rr) {
        const int p = p0 + rr;
        g += a.gpitch;
        const int2 m_lf = __ldg(g - 1), m_cur = __ldg(g), m_rt = __ldg(g + 1);
        // the register window covers two rows; pull the row after next-but-eight towards L2 already (one lane per
        // 128-byte line is enough)
        if ((threadIdx.x & 15) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(g + 8 * static_cast<size_t>(a.gpitch)));
        int2 tm_lf = m_lf, tm_cur = m_cur;
        if (ASYM) {
            gt += a.gpitch;
            tm_lf = __ldg(gt - 1);
            tm_cur = __ldg(gt);
        }
        const int2 c_dn = n_cur;
        // orientation: a_* are the cells that lead towards p's side (a -> u, v -> nu, pw -> u)
        const int2 a_cur = ASYM ? t_cur : c_cur, a_lf = ASYM ? t_lf : c_lf, a_dn = ASYM ? tn_cur : c_dn;
        const int4 Hp = sH[rr];
        if ((Hp.w >> 30) & 1) {  // p is a customer (warp-uniform)
            const int4 p1 = sP1[rr], p2 = sP2[rr];
            if (TABLE && (Hp.w & 0x7FFF) != cur_ra) {  // warp-uniform: p is the same for the whole block
                flush();
                cur_ra = Hp.w & 0x7FFF;
            }
            score_pair<MODE>(a, P, p, Hp, p1, p2, col, swap_tile, c_up, c_rt, a_cur, a_lf, a_dn, best_d, best_i, thr);
        }
        c_up = c_cur;
        c_lf = n_lf;
        c_cur = n_cur;
        c_rt = n_rt;
        n_lf = m_lf;
        n_cur = m_cur;
        n_rt = m_rt;
        t_lf = tn_lf;
        t_cur = tn_cur;
        tn_lf = tm_lf;
        tn_cur = tm_cur;
    }
    if (TABLE) {
        flush();
    } else if (!DUMP) {
        const unsigned long long best =
            best_i == 0xFFFFFFFFu
                ? kNoMove
                : (static_cast<unsigned long long>(static_cast<unsigned>(best_d) + 0x80000000u) << 32) | best_i;
        block_min_to_global<kInterThreads>(best, a.best + b, a.signed_dst != 0);
    }
}


// ---- the tile kernel without a matrix --------------------------------------------------------------------------
// For matrices that are SPEC.md §2 of the problem's own coordinates the route-order matrix need not exist: a cell
// is ~27 instructions from two unit vectors (6 FP64 ops for the chord, then the FP32 tail of geometry.cuh — the very
// functions matrix_kernel runs, hence the same bits), mostly on the FP64/FMA/XU pipes, which the ALU-bound scoring
// leaves idle.  A thread owns column q and keeps q's unit vector in registers; the strip's row vectors sit in shared
// memory (broadcast LDS); per row it computes ONE new cell, (p+1, q), rotates (p, q) and (p-1, q) down its register
// window, and gets (p, q-1) and (p, q+1) from its lane neighbours with four shuffles.  Lanes 0 and 31 of a warp only
// feed their neighbours (a warp scores 30 columns).  No G write, no G read: a pass touches O(n) bytes of HBM.
// (Two adjacent columns per thread — shared-memory loads and loop overhead paid once for two pairs, 2 of 64 columns halo
// instead of 2 of 32 — was built and measured: 94 registers instead of 56, 314 vs 313 us on one GPU, 64 vs 61 us on a
// 1/8 shard.  Not kept.)
constexpr int kFusedCols = 120;  // columns scored per block: 4 warps x 30

__device__ __forceinline__ int2 cell_from_vectors(double ax, double ay, double az, double bx, double by, double bz, float fd,
                                                  float ft, int lim_hi)
{
    const double dx = ax - bx, dy = ay - by, dz = az - bz;
    const double s2 = fma(dx, dx, fma(dy, dy, dz * dz));
    return __double2hiint(s2) < lim_hi ? cell_i32_fast(s2, fd, ft) : cell_i32(ax, ay, az, bx, by, bz, fd, ft);
}

// The pair (p, q), p < q, of two different routes, scored for everything it can mean: relocate(p -> after q),
// relocate(q -> after p) and swap(p, q).  The SPEC 2 matrix is symmetric, so the cells of the mirrored relocate are the
// ones already in the window — (p, q) and (p+1, q) — and the tiles below the diagonal never run: every cell is rebuilt
// once per unordered pair instead of twice.  best / thr: the direct candidates; mbest / mthr: in table mode the
// mirrored relocate belongs to the entry (route q, route p) and keeps its own running best (otherwise they alias the
// same variables).
constexpr int kNever = INT_MAX;  // a demand no capacity slack reaches (slacks are clamped to 2^30; only ever compared)

// NOLIM: no instance of the launch has a distance limit (every distance slack sits at its clamp 2^30, above any sum
// of three cells), so the four tests against the slacks are compiled out.
template <int MODE, bool DIAG, bool NOLIM>
__device__ __forceinline__ void score_pair_tri(const EvalArgs &a, int P, int p, const int4 &Hp, const int4 &p1, const int4 &p2,
                                               int dep_p, int reldem_p, const ColQ &c, bool q_rem_ok, bool q_cust, int2 c_up,
                                               int2 c_rt, int2 c_cur, int2 c_lf, int2 c_dn, int &best_d, unsigned &best_i,
                                               int &thr, int &mbest_d, unsigned &mbest_i, int &mthr)
{
    constexpr bool DUMP = MODE == kDump, TABLE = MODE == kTable;
    const int q = c.q;
    // DIAG: the tile lies near the diagonal — some of its columns may belong to the route of one of its rows, or not
    // lie to the right of them — so "another route" and p < q are tested per pair; everywhere else both hold for the
    // whole tile (routes are runs of consecutive positions).
    // (outside the dump the slot's H.w is the bare route id: the flags beside it moved into the data, see below)
    const bool pair = !DIAG || ((DUMP ? (Hp.w & 0x7FFF) : Hp.w) != c.q_route && p < q);
    // relocate(p -> between q and q+1)
    const int ins = fadd(fadd(c_cur.x, c_rt.x), -c.Hq.z);
    const int d_rel = fadd(Hp.x, ins);
    const int tu = __viaddmax_s32(c.dep_q, c_cur.y, p1.x);       // max(dep_q + t, ready_p): one VIADDMNMX
    // relocate(q -> between p and p+1): the same with the roles exchanged, cells (p, q) and (p+1, q)
    const int mins = fadd(fadd(c_cur.x, c_dn.x), -Hp.z);
    const int d_mir = fadd(c.Hq.x, mins);
    const int mtu = __viaddmax_s32(dep_p, c_cur.y, c.qB.x);
    // swap(p, q): pu -> v -> nu in route[p], pw -> u -> nw in route[q]
    const int d1 = fadd(fadd(c_up.x, c_dn.x), -Hp.y);
    const int d2 = fadd(fadd(c_lf.x, c_rt.x), -c.Hq.y);
    const int d_swap = fadd(d1, d2);
    const int tv = __viaddmax_s32(p2.y, c_up.y, c.qB.x);
    const int tw = __viaddmax_s32(c.depprev_q, c_lf.y, p1.x);
    const int dd = fadd(c.qB.w, -p1.w);
    bool ok_rel, ok_mir, ok_swap;
    if (DUMP) {   // explicit validity: the dump reports which pairs are candidates at all (the slots hold plain data)
        const bool p_cust = (Hp.w >> 30) & 1;
        const bool v_rel = c.rel_q && pair && p_cust;
        const bool v_mir = q_cust && (a.families & 2) && pair && p <= P - 2;
        const bool v_swap = c.swap_q && pair && p_cust;
        ok_rel = v_rel && ((Hp.w >> 29) & 1) && p1.w <= c.capsl_q && ins <= c.distsl_q && tu <= p1.y &&
                 fadd(fadd(tu, p1.z), c_rt.y) <= c.zn_q;
        ok_mir = v_mir && q_rem_ok && c.qB.w <= p2.z && mins <= p2.w && mtu <= c.qB.y && fadd(fadd(mtu, c.qB.z), c_dn.y) <= p2.x;
        ok_swap = v_swap && dd <= p2.z && -dd <= c.capsl_q && d1 <= p2.w && d2 <= c.distsl_q && tv <= c.qB.y &&
                  fadd(fadd(tv, c.qB.z), c_dn.y) <= p2.x && tw <= p1.y && fadd(fadd(tw, p1.z), c_rt.y) <= c.zn_q;
        if (a.dump_family == 1) {
            if (v_rel) {
                a.dump_delta[static_cast<size_t>(p) * P + q] = d_rel;
                a.dump_feas[static_cast<size_t>(p) * P + q] = ok_rel ? 1 : 0;
            }
            if (v_mir) {
                a.dump_delta[static_cast<size_t>(q) * P + p] = d_mir;
                a.dump_feas[static_cast<size_t>(q) * P + p] = ok_mir ? 1 : 0;
            }
        } else if (a.dump_family == 2 && v_swap) {
            a.dump_delta[static_cast<size_t>(p) * P + q] = d_swap;
            a.dump_feas[static_cast<size_t>(p) * P + q] = ok_swap ? 1 : 0;
        }
        return;
    }
    // validity rides on the data: a row that is no customer carries due = -1 (no arrival is that early) and, like a stop
    // that cannot be removed, the demand kNever for relocate; a column that takes no insertion / cannot be moved /
    // cannot be swapped carries capsl_rel / qdem_mir / qdue_swap to the same effect; the closing depot's row has
    // capacity slack -1.  So every test below is one the candidate needs anyway.
    ok_rel = pair && reldem_p <= c.capsl_rel && (NOLIM || ins <= c.distsl_q) && tu <= p1.y &&
             fadd(fadd(tu, p1.z), c_rt.y) <= c.zn_q;
    ok_mir = pair && c.qdem_mir <= p2.z && (NOLIM || mins <= p2.w) && mtu <= c.qB.y &&
             fadd(fadd(mtu, c.qB.z), c_dn.y) <= p2.x;
    // (bitwise ands: with && the compiler turns the second half of this chain into a divergent branch, which a warp
    // almost never takes as a whole)
    ok_swap = pair & (dd <= p2.z) & (dd >= c.ncapsl_q) & (NOLIM || ((d1 <= p2.w) & (d2 <= c.distsl_q))) &
              (tv <= c.qdue_swap) & (fadd(fadd(tv, c.qB.z), c_dn.y) <= p2.x) & (tw <= p1.y) &
              (fadd(fadd(tw, p1.z), c_rt.y) <= c.zn_q);
    const bool hit_rel = ok_rel && d_rel <= thr, hit_swap = ok_swap && d_swap <= thr, hit_mir = ok_mir && d_mir <= mthr;
    if (__any_sync(0xffffffffu, hit_rel || hit_swap || hit_mir)) {  // rare
        auto take = [](int d, unsigned idx, int &bd, unsigned &bi) {
            if (d < bd || (d == bd && idx < bi)) {
                bd = d;
                bi = idx;
            }
        };
        if (hit_rel) take(d_rel, (1u << 30) | (static_cast<unsigned>(p) << 15) | static_cast<unsigned>(q), best_d, best_i);
        if (hit_swap) take(d_swap, (2u << 30) | (static_cast<unsigned>(p) << 15) | static_cast<unsigned>(q), best_d, best_i);
        if (hit_mir) take(d_mir, (1u << 30) | (static_cast<unsigned>(q) << 15) | static_cast<unsigned>(p), mbest_d, mbest_i);
        if (TABLE) {
            thr = best_d;
            mthr = mbest_d;
        } else {
            thr = __reduce_min_sync(0xffffffffu, best_d);   // best / mbest are the same variables here, and so are thr / mthr
            mthr = thr;
        }
    }
}

__device__ __forceinline__ int wrap_add(int a, unsigned b) { return static_cast<int>(static_cast<unsigned>(a) + b); }

// One shared-memory slot per row of the strip (+2): slot i holds the unit vector of position p0 - 1 + i and, for
// i < rows, the three records of strip row i.  One array and explicit 32-bit shared addresses: with three separate
// arrays indexed by the (uniform) loop counter the compiler rebuilt every array's shared-window address in the uniform
// datapath on every iteration — a quarter of the loop's instructions.
struct __align__(16) RowSlot {
    int4 H, P1, P2;
    double vx, vy, vz;
    // the last 16 bytes of slot i: vz of its vector, and dep / reldem of the row TWO slots back — the row loop reads the
    // vector two slots ahead of the records, so one LDS.128 fetches both
    int dep;        // B[p] + service[p]: what an insertion after p starts from
    int reldem;     // demand of p for relocate(p -> q), kNever where p cannot be moved (see score_pair_tri)
};
static_assert(sizeof(RowSlot) == 80, "row slot layout");

__device__ __forceinline__ int4 lds128(unsigned addr)
{
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// FAST: the launcher has proved from the coordinates' bounding box that no pair of the batch can leave the
// straight-line range of the cell function (any metro-sized instance), so the per-cell range test is compiled out.
// (The body is a device function of the block's place in the grid: the persistent search kernel runs the same tiles
// from a loop.)
// FORM: 0 = any coordinates, 1 = FAST, 2 = FAST and no instance has a distance limit (score_pair_tri's NOLIM).
template <int MODE, int FORM>
__device__ __forceinline__ void eval_inter_fused_tile(const EvalArgs &a, int grid_x, int grid_y, int grid_z)
{
    constexpr bool DUMP = MODE == kDump, TABLE = MODE == kTable, FAST = FORM >= 1, NOLIM = FORM == 2 && !DUMP;
    __shared__ RowSlot slots[kMaxStripRows + 2];
    const int kStripRows = a.strip_rows;

    const int b = DUMP ? a.dump_instance : grid_z;
    const int P = a.npos[b];
    int p_lo, p_hi;
    shard_range(P, a.shard_index, a.shard_count, a.beta_inter, p_lo, p_hi);
    // Only pairs p < q are scored (score_pair_tri), so a strip's tiles start at the diagonal: strip y has
    // n(y) = nxb - (p0(y) + 1) / C of them, fewer and fewer towards the bottom.  A rectangular grid over that triangle
    // would be half empty blocks (each still costs a launch slot and a load of npos: 6 us on the last rank of eight),
    // so strip y shares its grid row with strip S-1-y: n(y) + n(S-1-y) is nearly constant.
    const int S = (p_hi - p_lo + kStripRows - 1) / kStripRows, nxb = (P + kFusedCols - 1) / kFusedCols;
    int y = grid_y, tile = grid_x;
    if (2 * y >= S) return;
    {
        const int n1 = max(0, nxb - (p_lo + y * kStripRows + 1) / kFusedCols);
        if (tile >= n1) {
            tile -= n1;
            if (S - 1 - y == y) return;
            y = S - 1 - y;
        }
    }
    const int p0 = p_lo + y * kStripRows;
    const int q0 = ((p0 + 1) / kFusedCols + tile) * kFusedCols;
    if (p0 >= p_hi || q0 >= P) return;

    const size_t po = static_cast<size_t>(b) * a.pstride;
    const double *__restrict__ PX = a.PX + static_cast<size_t>(b) * 3 * a.pstride;
    const int tid = threadIdx.x, lane = tid & 31;
    const int q = q0 + (tid >> 5) * 30 + lane - 1;            // lanes 0 and 31 are the warp's halo columns
    const bool owner = lane >= 1 && lane <= 30;
    const bool q_in = owner && q < P;
    const int rows = min(kStripRows, p_hi - p0);
    // first position after the route of the strip's last row (what decides the form of the row loop at the bottom;
    // two dependent loads, issued here so that they are back by then)
    int route_end;
    {
        const int p_last = min(p0 + rows, P) - 1;
        const int r_last = a.H[po + p_last].w & 0x7FFF;
        route_end = r_last < P - a.n ? a.rstart[static_cast<size_t>(b) * (a.rstride + 1) + r_last + 1] : P;
        route_end = max(route_end, p0 + rows);   // (never below the strip itself, whatever the arrays hold)
    }
    if (tid < rows + 2) {
        RowSlot sl;
        sl.H = sl.P1 = sl.P2 = make_int4(0, 0, 0, 0);
        sl.dep = 0;
        sl.reldem = kNever;
        sl.P1.y = -1;   // beyond the instance: no customer
        sl.P2.z = -1;   //   ... and no room for an insertion
        if (tid < rows && p0 + tid < P) {
            const size_t o = po + p0 + tid;
            const int4 rA = a.RA[o], rC = a.RC[o], rD = a.RD[o];
            sl.H = a.H[o];
            sl.H.w |= (rC.w & 2) << 28;  // rem_ok -> bit 29
            sl.P1 = a.RB[o];
            sl.P2 = make_int4(rA.y, rC.x, rD.x, rD.y);
            sl.dep = rA.x;
            if (!DUMP) {  // validity folded into the data (score_pair_tri)
                const bool cust = rC.w & 1;
                sl.reldem = (cust && (rC.w & 2) && (a.families & 2)) ? sl.P1.w : kNever;
                if (!cust) sl.P1.y = -1;
                sl.H.w &= 0x7FFF;
                // the cells of the row loop carry kCellBias (cell_i32_fast_biased): every sum they enter has a constant
                // of the row or of the column in it, which takes the bias out again (wrap-around arithmetic, exact)
                sl.H.y = wrap_add(sl.H.y, 2u * kCellBias);    // d1   = c_up.x + c_dn.x - H.y
                sl.H.z = wrap_add(sl.H.z, 2u * kCellBias);    // mins = c_cur.x + c_dn.x - H.z
                sl.P1.z = wrap_add(sl.P1.z, 0u - kCellBias);  // arrival = t + service + c_rt.y
                sl.P2.y = wrap_add(sl.P2.y, 0u - kCellBias);  // tv   = max(P2.y + c_up.y, ready)
                sl.dep = wrap_add(sl.dep, 0u - kCellBias);    // mtu  = max(dep + c_cur.y, ready)
            }
        }
        const int pv = min(max(p0 - 1 + tid, 0), a.pstride - 1);
        sl.vx = PX[pv];
        sl.vy = PX[a.pstride + pv];
        sl.vz = PX[2 * static_cast<size_t>(a.pstride) + pv];
        slots[tid].H = sl.H;
        slots[tid].P1 = sl.P1;
        slots[tid].P2 = sl.P2;
        slots[tid].vx = sl.vx;
        slots[tid].vy = sl.vy;
        slots[tid].vz = sl.vz;
        if (tid < rows) {   // (slots 0 and 1 keep no row's dep / reldem: nothing reads them)
            slots[tid + 2].dep = sl.dep;
            slots[tid + 2].reldem = sl.reldem;
        }
    }
    int4 Hq = make_int4(0, 0, 0, -1), qB = make_int4(0, 0, 0, 0);
    int dep_q = 0, zn_q = 0, depprev_q = 0, capsl_q = 0, distsl_q = 0;
    bool q_rem_ok = false;
    if (q_in) {
        const int4 rA = a.RA[po + q], rC = a.RC[po + q], rD = a.RD[po + q];
        Hq = a.H[po + q];
        qB = a.RB[po + q];
        dep_q = rA.x;
        zn_q = rA.y;
        depprev_q = rC.x;
        capsl_q = rD.x;
        distsl_q = rD.y;
        q_rem_ok = rC.w & 2;
    }
    const int qv = min(max(q, 0), a.pstride - 1);   // positions beyond P repeat the depot's vector
    const double bx = PX[qv], by = PX[a.pstride + qv], bz = PX[2 * static_cast<size_t>(a.pstride) + qv];
    __syncthreads();

    const int q_route = Hq.w & 0x7FFF;
    const bool q_cust = q_in && ((Hq.w >> 30) & 1);
    if (!DUMP) {  // the column's share of taking kCellBias out of the sums (see the slots)
        Hq.y = wrap_add(Hq.y, 2u * kCellBias);        // d2  = c_lf.x + c_rt.x - Hq.y
        Hq.z = wrap_add(Hq.z, 2u * kCellBias);        // ins = c_cur.x + c_rt.x - Hq.z
        dep_q = wrap_add(dep_q, 0u - kCellBias);      // tu  = max(dep_q + c_cur.y, ready)
        depprev_q = wrap_add(depprev_q, 0u - kCellBias);
        qB.z = wrap_add(qB.z, 0u - kCellBias);        // arrival = t + service + c_dn.y
    }
    const bool rel_q = q_in && (a.families & 2) && q <= P - 2;
    const bool swap_q = q_cust && (a.families & 4);
    const ColQ col{Hq, qB, dep_q, zn_q, depprev_q, capsl_q, distsl_q, q, q_route, rel_q, swap_q,
                   rel_q ? capsl_q : -1, (q_cust && q_rem_ok && (a.families & 2)) ? qB.w : kNever, swap_q ? qB.y : -1,
                   -capsl_q};

    int best_d = INT_MAX, thr = INT_MAX, tb_d = INT_MAX, tb_thr = INT_MAX;   // tb_*: table mode's mirrored-relocate entry
    unsigned best_i = 0xFFFFFFFFu, tb_i = 0xFFFFFFFFu;
    int cur_ra = -1;
    auto flush = [&]() {
        unsigned long long *tab = a.inter + static_cast<size_t>(b) * a.cr * a.cr;
        if (best_i != 0xFFFFFFFFu)
            table_min(tab + static_cast<size_t>(cur_ra) * a.cr + q_route,
                      (static_cast<unsigned long long>(static_cast<unsigned>(best_d) + 0x80000000u) << 32) | best_i, a.signed_dst);
        if (tb_i != 0xFFFFFFFFu)
            table_min(tab + static_cast<size_t>(q_route) * a.cr + cur_ra,
                      (static_cast<unsigned long long>(static_cast<unsigned>(tb_d) + 0x80000000u) << 32) | tb_i, a.signed_dst);
        best_d = thr = tb_d = tb_thr = INT_MAX;
        best_i = tb_i = 0xFFFFFFFFu;
    };
    // slot address in a per-thread register (the shift is zero for every thread of a 128-thread block; it only keeps
    // the compiler from proving the address uniform and moving it back into the uniform datapath)
    unsigned sa = static_cast<unsigned>(__cvta_generic_to_shared(slots)) + (threadIdx.x >> 10);
    const float fd = a.fd, ft = a.ft;
    const int lim_hi = a.lim_hi;
    auto cell_slot = [&](unsigned addr, int &dep_back2, int &reldem_back2) {  // cell(position of that slot, q)
        double ax, ay, az;
        asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(ax), "=d"(ay) : "r"(addr + 48));
        asm volatile("{ .reg .b64 t; ld.shared.v2.b64 {%0, t}, [%3]; mov.b64 {%1, %2}, t; }"
                     : "=d"(az), "=r"(dep_back2), "=r"(reldem_back2)
                     : "r"(addr + 64));
        if (FAST) {
            const double dx = ax - bx, dy = ay - by, dz = az - bz;
            const double s2 = fma(dx, dx, fma(dy, dy, dz * dz));
            return DUMP ? cell_i32_fast(s2, fd, ft) : cell_i32_fast_biased(s2, fd, ft);
        }
        const int2 c = cell_from_vectors(ax, ay, az, bx, by, bz, fd, ft, lim_hi);
        return DUMP ? c : make_int2(wrap_add(c.x, kCellBias), wrap_add(c.y, kCellBias));
    };
    int unused0, unused1;
    int2 c_up = cell_slot(sa, unused0, unused1), c_cur = cell_slot(sa + 80, unused0, unused1);
    const unsigned full = 0xffffffffu;
    auto run_rows = [&](auto diag_tag) {
        constexpr bool DIAG = decltype(diag_tag)::value;
#pragma unroll(TABLE ? 3 : 6)  // the register window rotates with period 3; the table form is out of registers at 6
        for (int rr = 0; rr < rows; ++rr, sa += 80) {
            const int p = p0 + rr;
            int dep_p, reldem_p;
            const int2 c_dn = cell_slot(sa + 160, dep_p, reldem_p);
            const int4 Hp = lds128(sa);
            // every row: a separator is never moved but is an insertion point of the mirrored relocate
            const int2 c_lf = make_int2(__shfl_up_sync(full, c_cur.x, 1), __shfl_up_sync(full, c_cur.y, 1));
            const int2 c_rt = make_int2(__shfl_down_sync(full, c_cur.x, 1), __shfl_down_sync(full, c_cur.y, 1));
            const int4 p1 = lds128(sa + 16), p2 = lds128(sa + 32);
            if (TABLE && (Hp.w & 0x7FFF) != cur_ra) {
                flush();
                cur_ra = Hp.w & 0x7FFF;
            }
            if (TABLE)
                score_pair_tri<MODE, DIAG, NOLIM>(a, P, p, Hp, p1, p2, dep_p, reldem_p, col, q_rem_ok, q_cust, c_up, c_rt, c_cur, c_lf,
                                           c_dn, best_d, best_i, thr, tb_d, tb_i, tb_thr);
            else
                score_pair_tri<MODE, DIAG, NOLIM>(a, P, p, Hp, p1, p2, dep_p, reldem_p, col, q_rem_ok, q_cust, c_up, c_rt, c_cur, c_lf,
                                           c_dn, best_d, best_i, thr, best_d, best_i, thr);
            c_up = c_cur;
            c_cur = c_dn;
        }
    };
    // Every pair of a tile whose first column lies beyond the ROUTE of the strip's last row has p < q and two different
    // routes: route_end is the first position after that route (loaded at the top of the kernel)
    if (q0 >= route_end) run_rows(std::false_type{});
    else run_rows(std::true_type{});
    if (TABLE) {
        flush();
    } else if (!DUMP) {
        const unsigned long long best =
            best_i == 0xFFFFFFFFu
                ? kNoMove
                : (static_cast<unsigned long long>(static_cast<unsigned>(best_d) + 0x80000000u) << 32) | best_i;
        block_min_to_global<kInterThreads>(best, a.best + b, a.signed_dst != 0);
    }
}

template <int MODE, int FORM>
__global__ void __launch_bounds__(kInterThreads, 8) eval_inter_fused_kernel(const EvalArgs a)
{
    eval_inter_fused_tile<MODE, FORM>(a, blockIdx.x, blockIdx.y, blockIdx.z);
}

// ---- same-route candidates ------------------------------------------------------------------------------------

struct Seg {  // service start at the last node = max(t + a, b) for arrival t at the first node; admissible iff ok && t <= L
    int a, b, L;
    bool ok;
};

__device__ __forceinline__ Seg seg_single(int ready, int due) { return Seg{0, ready, due, true}; }

// s1 then s2, w = service[last of s1] + T[last of s1][first of s2]
__device__ __forceinline__ Seg seg_concat(const Seg &s1, int w, const Seg &s2)
{
    Seg s;
    s.a = s1.a + w + s2.a;
    s.b = max(s1.b + w + s2.a, s2.b);
    s.L = min(s1.L, s2.L - s1.a - w);
    s.ok = s1.ok && s2.ok && (s1.b + w <= s2.L);
    return s;
}

// prefix arriving at the segment's first node at time t; afterwards tail = service[last] + T[last][next], limit = Z[next]
__device__ __forceinline__ bool seg_fits(const Seg &m, int t, int tail, int limit)
{
    return m.ok && t <= m.L && max(t + m.a, m.b) + tail <= limit;
}

constexpr int kIntraThreads = 128;

template <int MODE, bool ASYM, bool FUSED = false>
__global__ void __launch_bounds__(kIntraThreads) eval_intra_kernel(const EvalArgs a)
{
    static_assert(!(ASYM && FUSED), "cells are rebuilt from unit vectors only for the symmetric SPEC 2 matrix");
    constexpr bool DUMP = MODE == kDump, TABLE = MODE == kTable;
    const int b = DUMP ? a.dump_instance : blockIdx.y;
    const int P = a.npos[b];
    int t_lo, t_hi;
    shard_range(P, a.shard_index, a.shard_count, a.beta_intra, t_lo, t_hi);  // same-route work is sharded by anchor
    const int t = t_lo + blockIdx.x * kIntraThreads + threadIdx.x;
    unsigned long long best = kNoMove;
    int t_route = 0;

    if (t < t_hi && t < P - 1) {
        const size_t po = static_cast<size_t>(b) * a.pstride;
        const int *node = a.node + po;
        const int2 *G = a.G + static_cast<size_t>(b) * a.grows * a.gpitch;
        const int4 *RA = a.RA + po, *RB = a.RB + po, *RC = a.RC + po, *RD = a.RD + po;
        const int2 *E = a.E + po, *SK = a.SK + po;

        const int nt = node[t];
        const bool t_cust = nt != 0;
        const int4 tA = RA[t], tB = RB[t], tC = RC[t], tD = RD[t];
        t_route = tA.w;                    // every candidate of this sweep lies inside the anchor's route
        const int dslack = tD.y;           // max_dist - rdist of this route (clamped)
        const int2 Et = E[t];              // cell(node t, node t+1)
        const int4 t1B = RB[t + 1];        // stop data of node t+1 (first node of F)
        const int dep_t = tA.x;            // B[t] + service[node t]
        const int dep_tm1 = tC.x;          // B[t-1] + service[node t-1]   (customers only)
        const int2 SKt = SK[t];            // cell(node t-1, node t+1)
        // rows t-1, t, t+1 of the route-order matrix; the sweep reads them at consecutive columns x+1 — or, without a
        // matrix, rebuilds those cells from the unit vectors of the three rows (registers) and of the column
        const int2 *row_m = G + static_cast<size_t>(max(t - 1, 0)) * a.gpitch;
        const int2 *row_0 = G + static_cast<size_t>(t) * a.gpitch;
        const int2 *row_p = G + static_cast<size_t>(t + 1) * a.gpitch;
        const double *PXb = a.PX + static_cast<size_t>(b) * 3 * a.pstride;
        double vm[3] = {0, 0, 0}, v0[3] = {0, 0, 0}, vp[3] = {0, 0, 0};
        if (FUSED) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                vm[k] = PXb[static_cast<size_t>(k) * a.pstride + max(t - 1, 0)];
                v0[k] = PXb[static_cast<size_t>(k) * a.pstride + t];
                vp[k] = PXb[static_cast<size_t>(k) * a.pstride + t + 1];
            }
        }
        auto cells_at = [&](int col, int2 &o_m, int2 &o_0, int2 &o_p) {
            if (FUSED) {
                const double cx = PXb[col], cy = PXb[a.pstride + col], cz = PXb[2 * static_cast<size_t>(a.pstride) + col];
                o_m = cell_from_vectors(vm[0], vm[1], vm[2], cx, cy, cz, a.fd, a.ft, a.lim_hi);
                o_0 = cell_from_vectors(v0[0], v0[1], v0[2], cx, cy, cz, a.fd, a.ft, a.lim_hi);
                o_p = cell_from_vectors(vp[0], vp[1], vp[2], cx, cy, cz, a.fd, a.ft, a.lim_hi);
            } else {
                o_m = __ldg(row_m + col);
                o_0 = __ldg(row_0 + col);
                o_p = __ldg(row_p + col);
            }
        };
        // asymmetric matrices: the cells that lead from x back to t / t+1 come from the transposed array
        const int2 *GTb = ASYM ? a.GT + static_cast<size_t>(b) * a.grows * a.gpitch : G;
        const int2 *trow_0 = GTb + static_cast<size_t>(t) * a.gpitch;
        const int2 *trow_p = GTb + static_cast<size_t>(t + 1) * a.gpitch;
        const int2 *ER = ASYM ? a.ER + po : E;
        const bool f2 = a.families & 1, f_rel = a.families & 2, f_swap = a.families & 4;

        // cells of rows node[t-1], node[t], node[t+1] at column node[x]: start at x = t+1
        int2 cm = SKt, c0 = Et, cp = make_int2(0, 0);
        // r0 = cell(node x, node t), rp = cell(node x, node t+1): equal to c0 / cp when the matrix is symmetric
        int2 r0 = ASYM ? ER[t] : Et, rp = make_int2(0, 0), r0_prev = make_int2(0, 0);
        Seg F = Seg{0, 0, 0, true}, Rv = Seg{0, 0, 0, true};
        int svc_xm1 = 0, Et_xm1 = 0;  // service[node x-1], T[node x-1][node x]
        int Er_xm1 = 0;               // T[node x][node x-1]
        int asym = 0;                 // sum over the inner edges of (reversed - forward) distance, for 2-opt

        // The sweep ends at the route's closing separator, known from the route table — no data-dependent exit, so the
        // records and cells of step x+1 are fetched while step x is being scored (the sweep is a chain of dependent
        // steps: one memory round trip per step instead of two).
        const int x_end = a.rstart[static_cast<size_t>(b) * (a.rstride + 1) + t_route + 1];
        int4 nA = RA[t + 1], nB = t1B, nC = RC[t + 1], nD = RD[t + 1];
        int2 ncm, nc0, ncp;
        cells_at(t + 2, ncm, nc0, ncp);
        for (int x = t + 1; x < x_end; ++x) {
            const int4 xA = nA, xB = nB, xC = nC, xD = nD;
            const int2 Ex = make_int2(xA.z, xD.z);  // cell(node x, node x+1)
            const int zn = xA.y;                    // Z[x+1]
            // next column's cells (rows fixed per thread): the only scattered loads of the step
            const int2 cm1 = ncm, c01 = nc0, cp1 = ncp;
            // step x+1 (x+1 <= x_end <= P-1 and column x+2 <= P stay inside the padded arrays)
            nA = RA[x + 1];
            nB = RB[x + 1];
            nC = RC[x + 1];
            nD = RD[x + 1];
            cells_at(x + 2, ncm, nc0, ncp);
            const int2 r01 = ASYM ? __ldg(trow_0 + x + 1) : c01;
            const int2 rp1 = ASYM ? __ldg(trow_p + x + 1) : cp1;
            const Seg sx = seg_single(xB.x, xB.y);

            // reversed segment node[x..t+1]: prepend x (it is entered over the edge x -> x-1)
            Rv = (x == t + 1) ? sx : seg_concat(sx, xB.z + (ASYM ? Er_xm1 : Et_xm1), Rv);

            // (A) 2-opt(t, x)
            if (f2 && x >= t + 2) {
                const int d = c0.x + cp1.x - Et.x - Ex.x + asym;
                const bool ok = d <= dslack && seg_fits(Rv, dep_t + c0.y, t1B.z + cp1.y, zn);
                if (DUMP) {
                    if (a.dump_family == 0) {
                        a.dump_delta[static_cast<size_t>(t) * P + x] = d;
                        a.dump_feas[static_cast<size_t>(t) * P + x] = ok ? 1 : 0;
                    }
                } else if (ok) {
                    const unsigned long long k = pack_key(d, 0, t, x);
                    best = k < best ? k : best;
                }
            }
            if (f_rel) {
                // (B) relocate(t -> after x): prefix(..t-1) + F + [x] + [t] + suffix(x+1..)
                if (t_cust) {
                    Seg m = (x == t + 1) ? sx : seg_concat(F, svc_xm1 + Et_xm1, sx);
                    m = seg_concat(m, xB.z + r0.y, seg_single(tB.x, tB.y));
                    const int d = tC.y + r0.x + c01.x - Ex.x;
                    const bool ok = d <= dslack && seg_fits(m, dep_tm1 + SKt.y, tB.z + c01.y, zn);
                    if (DUMP) {
                        if (a.dump_family == 1) {
                            a.dump_delta[static_cast<size_t>(t) * P + x] = d;
                            a.dump_feas[static_cast<size_t>(t) * P + x] = ok ? 1 : 0;
                        }
                    } else if (ok) {
                        const unsigned long long k = pack_key(d, 1, t, x);
                        best = k < best ? k : best;
                    }
                }
                // (C) relocate(x -> after t): prefix(..t) + [x] + F + suffix(x+1..)
                if (x >= t + 2) {
                    const Seg m = seg_concat(sx, xB.z + rp.y, F);
                    const int d = xC.y + c0.x + rp.x - Et.x;
                    const bool ok = d <= dslack && seg_fits(m, dep_t + c0.y, svc_xm1 + xD.w, zn);
                    if (DUMP) {
                        if (a.dump_family == 1) {
                            a.dump_delta[static_cast<size_t>(x) * P + t] = d;
                            a.dump_feas[static_cast<size_t>(x) * P + t] = ok ? 1 : 0;
                        }
                    } else if (ok) {
                        const unsigned long long k = pack_key(d, 1, x, t);
                        best = k < best ? k : best;
                    }
                }
            }
            // (D) swap(t, x)
            if (f_swap && t_cust) {
                Seg m;
                int d;
                if (x == t + 1) {
                    m = seg_concat(sx, xB.z + r0.y, seg_single(tB.x, tB.y));
                    d = cm.x + r0.x + c01.x - tC.z - Ex.x;  // pv->x + x->t + t->nxt - (pv->t + t->x) - x->nxt
                } else {
                    m = seg_concat(sx, xB.z + rp.y, F);
                    m = seg_concat(m, svc_xm1 + r0_prev.y, seg_single(tB.x, tB.y));
                    d = cm.x + rp.x - tC.z + r0_prev.x + c01.x - xC.z;
                }
                const bool ok = d <= dslack && seg_fits(m, dep_tm1 + cm.y, tB.z + c01.y, zn);
                if (DUMP) {
                    if (a.dump_family == 2) {
                        a.dump_delta[static_cast<size_t>(t) * P + x] = d;
                        a.dump_feas[static_cast<size_t>(t) * P + x] = ok ? 1 : 0;
                    }
                } else if (ok) {
                    const unsigned long long k = pack_key(d, 2, t, x);
                    best = k < best ? k : best;
                }
            }
            // forward segment node[t+1..x]: append x
            F = (x == t + 1) ? sx : seg_concat(F, svc_xm1 + Et_xm1, sx);
            svc_xm1 = xB.z;
            Et_xm1 = Ex.y;
            if (ASYM) {
                const int2 er = ER[x];  // cell(node x+1, node x)
                Er_xm1 = er.y;
                asym += er.x - Ex.x;
            }
            r0_prev = r0;
            cm = cm1;
            c0 = c01;
            cp = cp1;
            r0 = r01;
            rp = rp1;
        }
    }
    if (TABLE) {
        if (best != kNoMove) table_min(a.intra + static_cast<size_t>(b) * a.cr + t_route, best, a.signed_dst);
    } else if (!DUMP) {
        block_min_to_global<kIntraThreads>(best, a.best + b, a.signed_dst != 0);
    }
}


// ---- same-route candidates, one WARP per anchor ------------------------------------------------------------------
// The sweep above is a chain of dependent steps: ~0.4 us per position of the route whatever the GPU has left to do,
// i.e. 35-50 us for routes of a hundred stops — invisible underneath a 270 us tile kernel, but the floor of a pass once
// the neighbourhood is sharded over several GPUs (a 1/8 shard of the tile kernel takes less).  Segment concatenation is
// associative, so the forward segment node[t+1..x] and the reversed one node[x..t+1] of EVERY x are two warp scans:
// lane j owns x = base + j, five shuffle steps build both segments, and all candidates of 32 consecutive x are scored
// at once; routes longer than 32 positions carry the last lane's segments into the next chunk.  Same candidates, same
// keys as the sweep (the composition is exact integer max-plus algebra, so the association order cannot change a
// value).  Symmetric matrices only; the sweep stays for directional ones and for launches with enough anchors to hide
// its latency (it needs half the instructions).
struct SegW {  // a segment plus the link weight the next concatenation on its open side needs
    int a, b, L, ok, w;
};

__device__ __forceinline__ SegW segw_shfl_up(const SegW &s, int off)
{
    SegW r;
    r.a = __shfl_up_sync(0xffffffffu, s.a, off);
    r.b = __shfl_up_sync(0xffffffffu, s.b, off);
    r.L = __shfl_up_sync(0xffffffffu, s.L, off);
    r.ok = __shfl_up_sync(0xffffffffu, s.ok, off);
    r.w = __shfl_up_sync(0xffffffffu, s.w, off);
    return r;
}

__device__ __forceinline__ Seg seg_of(const SegW &s) { return Seg{s.a, s.b, s.L, s.ok != 0}; }

constexpr int kIntraWarps = 4;

template <int MODE, bool FUSED>
__device__ __forceinline__ void eval_intra_warp_block(const EvalArgs &a, int grid_x, int grid_y)
{
    constexpr bool DUMP = MODE == kDump, TABLE = MODE == kTable;
    const int b = DUMP ? a.dump_instance : grid_y;
    const int P = a.npos[b];
    int t_lo, t_hi;
    shard_range(P, a.shard_index, a.shard_count, a.beta_intra, t_lo, t_hi);  // same-route work is sharded by anchor
    const int lane = threadIdx.x & 31;
    const int t = t_lo + grid_x * kIntraWarps + (threadIdx.x >> 5);
    unsigned long long best = kNoMove;
    int t_route = 0;

    if (t < t_hi && t < P - 1) {  // warp-uniform
        const size_t po = static_cast<size_t>(b) * a.pstride;
        const int2 *G = a.G + static_cast<size_t>(b) * a.grows * a.gpitch;
        const int4 *RA = a.RA + po, *RB = a.RB + po, *RC = a.RC + po, *RD = a.RD + po;
        const int4 tA = RA[t], tB = RB[t], tC = RC[t], tD = RD[t];
        t_route = tA.w;
        const int x_end = a.rstart[static_cast<size_t>(b) * (a.rstride + 1) + t_route + 1];
        const bool t_cust = tC.w & 1;
        const int dslack = tD.y;
        const int2 Et = a.E[po + t], SKt = a.SK[po + t];
        const int svc_t1 = RB[t + 1].z;           // service time of node t+1
        const int dep_t = tA.x, dep_tm1 = tC.x;
        const int2 *row_m = G + static_cast<size_t>(max(t - 1, 0)) * a.gpitch;
        const int2 *row_0 = G + static_cast<size_t>(t) * a.gpitch;
        const int2 *row_p = G + static_cast<size_t>(t + 1) * a.gpitch;
        const bool f2 = a.families & 1, f_rel = a.families & 2, f_swap = a.families & 4;
        const Seg st_single = seg_single(tB.x, tB.y);
        const double *PXb = a.PX + static_cast<size_t>(b) * 3 * a.pstride;
        double vm[3] = {0, 0, 0}, v0[3] = {0, 0, 0}, vp[3] = {0, 0, 0};  // unit vectors of t-1, t, t+1 (no-G form)
        if (FUSED) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                vm[k] = PXb[static_cast<size_t>(k) * a.pstride + max(t - 1, 0)];
                v0[k] = PXb[static_cast<size_t>(k) * a.pstride + t];
                vp[k] = PXb[static_cast<size_t>(k) * a.pstride + t + 1];
            }
        }

        // carries from the previous chunk (positions below `base`): forward segment t+1..base-1, reversed base-1..t+1,
        // service time and outgoing edge of base-1, cell (t, base-1)
        Seg Fc = Seg{0, 0, 0, true}, Rc = Fc;
        int svc_c = 0, et_c = 0;
        int2 c0_c = make_int2(0, 0);
        for (int base = t + 1; base < x_end; base += 32) {
            const int x = base + lane;
            const bool in = x < x_end;
            const int xx = in ? x : x_end - 1;     // lanes beyond the route re-read its last position; never reported
            const int4 xA = RA[xx], xB = RB[xx], xC = RC[xx], xD = RD[xx];
            int2 cm, c0, c01, cp, cp1;
            if (FUSED) {
                const double ax = PXb[xx], ay = PXb[a.pstride + xx], az = PXb[2 * static_cast<size_t>(a.pstride) + xx];
                const double ux = PXb[xx + 1], uy = PXb[a.pstride + xx + 1], uz = PXb[2 * static_cast<size_t>(a.pstride) + xx + 1];
                cm = cell_from_vectors(vm[0], vm[1], vm[2], ax, ay, az, a.fd, a.ft, a.lim_hi);
                c0 = cell_from_vectors(v0[0], v0[1], v0[2], ax, ay, az, a.fd, a.ft, a.lim_hi);
                cp = cell_from_vectors(vp[0], vp[1], vp[2], ax, ay, az, a.fd, a.ft, a.lim_hi);
                c01 = cell_from_vectors(v0[0], v0[1], v0[2], ux, uy, uz, a.fd, a.ft, a.lim_hi);
                cp1 = cell_from_vectors(vp[0], vp[1], vp[2], ux, uy, uz, a.fd, a.ft, a.lim_hi);
            } else {
                cm = __ldg(row_m + xx);
                c0 = __ldg(row_0 + xx);
                c01 = __ldg(row_0 + xx + 1);
                cp = __ldg(row_p + xx);
                cp1 = __ldg(row_p + xx + 1);
            }
            const int2 Ex = make_int2(xA.z, xD.z);  // cell(node x, node x+1)
            const int zn = xA.y;                    // Z[x+1]
            int svc_xm1 = __shfl_up_sync(0xffffffffu, xB.z, 1);
            int et_xm1 = __shfl_up_sync(0xffffffffu, Ex.y, 1);
            int2 r0_prev = make_int2(__shfl_up_sync(0xffffffffu, c0.x, 1), __shfl_up_sync(0xffffffffu, c0.y, 1));
            if (lane == 0) {
                svc_xm1 = svc_c;
                et_xm1 = et_c;
                r0_prev = c0_c;
            }
            const Seg sx = seg_single(xB.x, xB.y);
            // forward: F = node[base..x], open on its low side, link weight service[x-1] + T[x-1][x]
            // reversed: R = node[x..base], open on its low side too, link weight service[x] + T[x][x-1]
            SegW F{sx.a, sx.b, sx.L, 1, svc_xm1 + et_xm1};
            SegW R{sx.a, sx.b, sx.L, 1, xB.z + et_xm1};
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const SegW Fl = segw_shfl_up(F, off), Rl = segw_shfl_up(R, off);
                if (lane >= off) {
                    const Seg f = seg_concat(seg_of(Fl), F.w, seg_of(F));
                    F = SegW{f.a, f.b, f.L, f.ok ? 1 : 0, Fl.w};
                    const Seg r = seg_concat(seg_of(R), R.w, seg_of(Rl));
                    R = SegW{r.a, r.b, r.L, r.ok ? 1 : 0, Rl.w};
                }
            }
            Seg Fin = seg_of(F), Rv = seg_of(R);   // node[t+1..x] and node[x..t+1] once the carries are attached
            if (base > t + 1) {
                Fin = seg_concat(Fc, F.w, Fin);
                Rv = seg_concat(Rv, R.w, Rc);
            }
            // node[t+1..x-1]: the left neighbour's inclusive segment
            SegW Fw{Fin.a, Fin.b, Fin.L, Fin.ok ? 1 : 0, 0};
            const SegW Fp_ = segw_shfl_up(Fw, 1);
            const Seg Fprev = lane == 0 ? Fc : seg_of(Fp_);

            if (in) {
                // (A) 2-opt(t, x)
                if (f2 && x >= t + 2) {
                    const int d = c0.x + cp1.x - Et.x - Ex.x;
                    const bool ok = d <= dslack && seg_fits(Rv, dep_t + c0.y, svc_t1 + cp1.y, zn);
                    if (DUMP) {
                        if (a.dump_family == 0) {
                            a.dump_delta[static_cast<size_t>(t) * P + x] = d;
                            a.dump_feas[static_cast<size_t>(t) * P + x] = ok ? 1 : 0;
                        }
                    } else if (ok) {
                        const unsigned long long k = pack_key(d, 0, t, x);
                        best = k < best ? k : best;
                    }
                }
                if (f_rel) {
                    // (B) relocate(t -> after x): prefix(..t-1) + node[t+1..x] + [t] + suffix(x+1..)
                    if (t_cust) {
                        const Seg m = seg_concat(Fin, xB.z + c0.y, st_single);
                        const int d = tC.y + c0.x + c01.x - Ex.x;
                        const bool ok = d <= dslack && seg_fits(m, dep_tm1 + SKt.y, tB.z + c01.y, zn);
                        if (DUMP) {
                            if (a.dump_family == 1) {
                                a.dump_delta[static_cast<size_t>(t) * P + x] = d;
                                a.dump_feas[static_cast<size_t>(t) * P + x] = ok ? 1 : 0;
                            }
                        } else if (ok) {
                            const unsigned long long k = pack_key(d, 1, t, x);
                            best = k < best ? k : best;
                        }
                    }
                    // (C) relocate(x -> after t): prefix(..t) + [x] + node[t+1..x-1] + suffix(x+1..)
                    if (x >= t + 2) {
                        const Seg m = seg_concat(sx, xB.z + cp.y, Fprev);
                        const int d = xC.y + c0.x + cp.x - Et.x;
                        const bool ok = d <= dslack && seg_fits(m, dep_t + c0.y, svc_xm1 + xD.w, zn);
                        if (DUMP) {
                            if (a.dump_family == 1) {
                                a.dump_delta[static_cast<size_t>(x) * P + t] = d;
                                a.dump_feas[static_cast<size_t>(x) * P + t] = ok ? 1 : 0;
                            }
                        } else if (ok) {
                            const unsigned long long k = pack_key(d, 1, x, t);
                            best = k < best ? k : best;
                        }
                    }
                }
                // (D) swap(t, x)
                if (f_swap && t_cust) {
                    Seg m;
                    int d;
                    if (x == t + 1) {
                        m = seg_concat(sx, xB.z + c0.y, st_single);
                        d = cm.x + c0.x + c01.x - tC.z - Ex.x;
                    } else {
                        m = seg_concat(sx, xB.z + cp.y, Fprev);
                        m = seg_concat(m, svc_xm1 + r0_prev.y, st_single);
                        d = cm.x + cp.x - tC.z + r0_prev.x + c01.x - xC.z;
                    }
                    const bool ok = d <= dslack && seg_fits(m, dep_tm1 + cm.y, tB.z + c01.y, zn);
                    if (DUMP) {
                        if (a.dump_family == 2) {
                            a.dump_delta[static_cast<size_t>(t) * P + x] = d;
                            a.dump_feas[static_cast<size_t>(t) * P + x] = ok ? 1 : 0;
                        }
                    } else if (ok) {
                        const unsigned long long k = pack_key(d, 2, t, x);
                        best = k < best ? k : best;
                    }
                }
            }
            // carries for the next chunk: everything of lane 31 (x = base + 31 < x_end whenever there is a next chunk)
            Fc = Seg{__shfl_sync(0xffffffffu, Fin.a, 31), __shfl_sync(0xffffffffu, Fin.b, 31),
                     __shfl_sync(0xffffffffu, Fin.L, 31), __shfl_sync(0xffffffffu, Fin.ok ? 1 : 0, 31) != 0};
            Rc = Seg{__shfl_sync(0xffffffffu, Rv.a, 31), __shfl_sync(0xffffffffu, Rv.b, 31),
                     __shfl_sync(0xffffffffu, Rv.L, 31), __shfl_sync(0xffffffffu, Rv.ok ? 1 : 0, 31) != 0};
            svc_c = __shfl_sync(0xffffffffu, xB.z, 31);
            et_c = __shfl_sync(0xffffffffu, Ex.y, 31);
            c0_c = make_int2(__shfl_sync(0xffffffffu, c0.x, 31), __shfl_sync(0xffffffffu, c0.y, 31));
        }
    }
    if (TABLE) {
        best = warp_min_u64(best);
        if (lane == 0 && best != kNoMove) table_min(a.intra + static_cast<size_t>(b) * a.cr + t_route, best, a.signed_dst);
    } else if (!DUMP) {
        block_min_to_global<kIntraWarps * 32>(best, a.best + b, a.signed_dst != 0);
    }
}

template <int MODE, bool FUSED = false>
__global__ void __launch_bounds__(kIntraWarps * 32) eval_intra_warp_kernel(const EvalArgs a)
{
    eval_intra_warp_block<MODE, FUSED>(a, blockIdx.x, blockIdx.y);
}

// Strip height of the tile kernel.  Taller strips amortise the per-block prologue (the column's records, the window's
// first rows, the block reduction: worth about four rows — 0.249 -> 0.236 ms on the 64 x 1,000 batch going from 32 to
// 64), shorter ones shorten the tail in which the last blocks leave SMs idle.  64 rows where at least four waves of
// blocks remain, 32 otherwise.  (Measured on a 1/8 shard of the 10,000-stop instance, 2.5 waves at 32 rows: 44 us at
// 32 rows, 47 us at 41 rows = exactly two full waves — filling whole waves does not pay, the dispatcher back-fills
// and a half-empty last wave simply runs faster.)  ROUTEST_STRIP_ROWS overrides, for experiments.
int g_strip_rows = 0;    // rt_debug_set("strip_rows", h): 0 = pick, else the height to use
int g_small_strips = 1;  // rt_debug_set("small_strips", 0/1): short strips where the whole grid fits one wave
int g_intra_kernel = 0;  // rt_debug_set("intra_kernel", k): 0 = pick, 1 = one thread per anchor, 2 = one warp per anchor

int pick_strip_rows(int pmax, int rows, int batch, bool asym, int cols_per_block)
{
    static const int env = getenv("ROUTEST_STRIP_ROWS") ? atoi(getenv("ROUTEST_STRIP_ROWS")) : 0;
    const int forced = g_strip_rows ? g_strip_rows : env;
    if (forced >= 8 && forced <= kMaxStripRows) return forced;
    const long long col_blocks = (pmax + cols_per_block - 1) / cols_per_block;
    // (the kernel without a matrix runs only the tiles above the diagonal: about half of the grid)
    const int tri = cols_per_block == kFusedCols ? 2 : 1;
    const long long blocks64 = col_blocks * ((rows + 63) / 64) * batch / tri;
    if (!asym && blocks64 >= 4LL * kNumSMs * 9) return 64;
    // A small grid is bound by the latency of one block (a dependent chain of ~0.15 us per row behind a ~3 us
    // prologue), not by throughput: while every block is resident at once, shorter strips mean a shorter kernel
    // (1,010 positions: 11.3 us at 32 rows, 144 blocks).
    if (g_small_strips && tri == 2)
        for (int h = 8; h < kStripRows; h *= 2)
            if (col_blocks * ((rows + h - 1) / h) * batch / tri <= 8LL * kNumSMs) return h;
    return kStripRows;
}

// One warp per anchor while the launch has too few anchors to hide the sweep's latency behind other work (a single
// instance, a shard of one); one thread per anchor for big batches, where the sweep's lower instruction count wins.
bool intra_by_warp(const rt_problem *p, int shard_rows)
{
    if (!p->symmetric) return false;
    if (g_intra_kernel) return g_intra_kernel == 2;
    return static_cast<long long>(shard_rows) * p->batch <= 16384;
}

template <int MODE>
int launch_intra(rt_problem *p, const EvalArgs &a, int shard_rows, int batch, cudaStream_t st)
{
    const bool fused = eval_without_matrix(p);
    if (intra_by_warp(p, shard_rows)) {
        dim3 grid((shard_rows + kIntraWarps - 1) / kIntraWarps, batch);
        if (fused) eval_intra_warp_kernel<MODE, true><<<grid, kIntraWarps * 32, 0, st>>>(a);
        else eval_intra_warp_kernel<MODE, false><<<grid, kIntraWarps * 32, 0, st>>>(a);
    } else {
        dim3 grid((shard_rows + kIntraThreads - 1) / kIntraThreads, batch);
        if (fused) eval_intra_kernel<MODE, false, true><<<grid, kIntraThreads, 0, st>>>(a);
        else if (p->symmetric) eval_intra_kernel<MODE, false><<<grid, kIntraThreads, 0, st>>>(a);
        else eval_intra_kernel<MODE, true><<<grid, kIntraThreads, 0, st>>>(a);
    }
    RT_CUDA(p->e, cudaGetLastError());
    return RT_OK;
}

// chord^2 <= angle^2 <= dlat^2 + dlon^2 over the bounding box of every coordinate of the batch (p->angle2_bound, kept
// by the entry points that receive coordinates): below the fast path's limit, with a margin, no cell can need the
// guarded form.
bool all_pairs_in_fast_range(const rt_problem *p, const EvalArgs &a)
{
    long long bits = static_cast<long long>(a.lim_hi) << 32;  // the smallest double with that high word
    double lim;
    memcpy(&lim, &bits, sizeof(lim));
    return p->angle2_bound >= 0.0 && p->angle2_bound * 1.001 < lim;
}

// Grid of the tile kernel without a matrix: the triangle above the diagonal, two strips per grid row (see the kernel);
// sized for the largest instance of the batch from this shard's first row on, with slack for instances whose
// boundaries differ by a row or two.  a.strip_rows must be set.
dim3 fused_tile_grid(const EvalArgs &a, int pmax, int batch)
{
    const int cols = kFusedCols;
    int lo, hi;
    shard_bounds(pmax, a.shard_index, a.shard_count, a.beta_inter, lo, hi);
    const int h = a.strip_rows, S = (hi - lo + 2 + h - 1) / h, nxb = (pmax + cols - 1) / cols;
    auto n_of = [&](int yy) { return std::max(0, nxb - (std::max(lo - 2, 0) + yy * h + 1) / cols); };
    int width = 1;
    for (int yy = 0; 2 * yy < S; ++yy) width = std::max(width, n_of(yy) + (S - 1 - yy != yy ? n_of(S - 1 - yy) : 0));
    return dim3(width + 2, (S + 1) / 2 + 1, batch);
}

// The tile kernel of this launch: without a matrix where the cells are SPEC 2 of the problem's coordinates, on G
// (and GT for a directional matrix) otherwise.
template <int MODE>
int launch_inter(rt_problem *p, EvalArgs &a, int pmax, int shard_rows, int batch, int strip_rows, cudaStream_t st)
{
    const bool fused = eval_without_matrix(p);
    const int cols = fused ? kFusedCols : kInterThreads;
    a.strip_rows = strip_rows > 0 ? strip_rows : pick_strip_rows(pmax, shard_rows, batch, !p->symmetric, cols);
    dim3 grid((pmax + cols - 1) / cols, (shard_rows + a.strip_rows - 1) / a.strip_rows, batch);
    if (fused) grid = fused_tile_grid(a, pmax, batch);
    const bool fast = fused && all_pairs_in_fast_range(p, a);
    if (fast && p->dist_unlimited && MODE != kDump) eval_inter_fused_kernel<MODE, 2><<<grid, kInterThreads, 0, st>>>(a);
    else if (fast) eval_inter_fused_kernel<MODE, 1><<<grid, kInterThreads, 0, st>>>(a);
    else if (fused) eval_inter_fused_kernel<MODE, 0><<<grid, kInterThreads, 0, st>>>(a);
    else if (p->symmetric) eval_inter_kernel<MODE, false><<<grid, kInterThreads, 0, st>>>(a);
    else eval_inter_kernel<MODE, true><<<grid, kInterThreads, 0, st>>>(a);
    RT_CUDA(p->e, cudaGetLastError());
    return RT_OK;
}

// Rows the grids must cover for shard idx of cnt: instances of a batch differ in P (pmin..pmax), and so do their
// boundaries; the widest shard of the largest instance plus one row of slack bounds them all.
int max_shard_rows(int pmax, int idx, int cnt, double beta)
{
    int lo, hi;
    shard_bounds(pmax, idx, cnt, beta, lo, hi);
    return hi - lo + 2;
}

EvalArgs make_args(rt_problem *p, int families, int shard_index, int shard_count, unsigned long long *best_d)
{
    EvalArgs a{};
    a.G = position_matrix(p);
    a.GT = position_matrix_t(p);
    a.ER = p->ER;
    a.grows = p->grows;
    a.gpitch = p->gpitch;
    a.node = p->node;
    a.npos = p->npos;
    a.RA = p->RA;
    a.RB = p->RB;
    a.RC = p->RC;
    a.RD = p->RD;
    a.H = p->H;
    a.E = p->E;
    a.SK = p->SK;
    a.PX = p->PX;
    a.fd = static_cast<float>(RT_EARTH_RADIUS_M * p->dist_scale);
    a.ft = static_cast<float>(RT_EARTH_RADIUS_M * p->time_scale);
    a.lim_hi = fast_path_limit_hi(a.fd, a.ft);
    a.rstart = p->rstart;
    a.rstride = p->rstride;
    a.n = p->n;
    a.pitch = p->pitch;
    a.pstride = p->pstride;
    a.families = families;
    a.shard_index = shard_index;
    a.shard_count = shard_count;
    // without a matrix the tile kernel scores the triangle p < q and nothing ties the same-route anchors to its rows;
    // on a stored matrix both kernels read the rows of G this replica keeps fresh, so they share one range
    a.beta_inter = eval_without_matrix(p) ? kBetaTriangle : kBetaStoredMatrix;
    a.beta_intra = eval_without_matrix(p) ? kBetaUniform : kBetaStoredMatrix;
    a.best = best_d;
    a.strip_rows = kStripRows;
    return a;
}

// ---- one-move search as ONE launch (SPEC.md §8) ----
// Up to a few thousand positions an iteration of the best-improvement search is three short phases — score every
// candidate, apply the best one and rescan the routes, rebuild the per-route tables — each a chain of dependent loads
// a few microseconds long.  As separate launches (even replayed from a graph) every phase also pays ~3 us of launch
// dependency; here the phases are loops over the same device functions inside one cooperative kernel, separated by
// grid-wide barriers, and the host is out of the loop until the search is over (or pa.iters have run).  Used for
// small instances, where it is the faster form (kPersistPositions; the measurements are in search_persistent_launch).
//   E  every block takes tiles of the triangle and same-route anchors (eval_inter_fused_tile, eval_intra_warp_block)
//      of every instance that is not done; minima go to best[] by atomicMin as in the separate kernels
//   A  the block that leaves E last applies best[b] to node[] (apply_move_block: sets done[b] where nothing
//      improves) and, if a move was applied, rescans the routes (resets best[b], cost[b]); with at most four routes
//      it rebuilds the tables as well and the iteration has a single barrier
//   T  otherwise one warp per route, over the whole grid, rebuilds vectors, tables and records of the instances
//      still live (route_tables_warp)
// done[] only changes in A, so after the barrier behind it every block takes the same decision to leave.  None of the
// loads below may use the read-only path: what E reads is what T of the same launch wrote (no __restrict__ on them).
struct PersistArgs {
    int iters;               // iterations this launch may run
    int batch;
    int inter_gx, inter_gy;  // the tile kernel's grid per instance
    int intra_gx;            // the same-route kernel's blocks per instance
    int rmax;                // most routes of any instance
    int tables_in_block;     // rmax <= warps of a block: the block that applies the move rebuilds the tables too
    unsigned *ticket;        // arrival counter, zero at launch
};

constexpr int kPersistThreads = 128;
static_assert(kPersistThreads == kInterThreads && kPersistThreads == kIntraWarps * 32, "the phases share one block shape");

template <bool FAST>
__global__ void __launch_bounds__(kPersistThreads, 4) search_persistent_kernel(const EvalArgs ea, const StateArgs sa,
                                                                               const ApplyArgs ap, const PersistArgs pa)
{
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    __shared__ int s_last;
    const int n_inter = pa.inter_gx * pa.inter_gy, per_instance = pa.intra_gx + n_inter;
    const int warps = gridDim.x * (kPersistThreads / 32);
    for (int it = 0; it < pa.iters; ++it) {
        // E (the same-route anchors first: theirs is the longer chain)
        for (int v = blockIdx.x; v < per_instance * pa.batch; v += gridDim.x) {
            const int b = v / per_instance, w = v - b * per_instance;
            if (!ap.done[b]) {  // uniform over the block
                if (w < pa.intra_gx) eval_intra_warp_block<kBest, true>(ea, w, b);
                else eval_inter_fused_tile<kBest, FAST ? 1 : 0>(ea, (w - pa.intra_gx) % pa.inter_gx, (w - pa.intra_gx) / pa.inter_gx, b);
            }
            __syncthreads();  // the shared memory of a tile is reused by the next
        }
        // A: whichever block finishes E last goes straight on (its atomicMins are ordered before its ticket by the
        // fence, everybody else's before theirs): one grid-wide barrier less per iteration
        if (threadIdx.x == 0) {
            __threadfence();
            s_last = atomicAdd(pa.ticket, 1u) == static_cast<unsigned>(gridDim.x) * (it + 1) - 1u;
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            for (int b = 0; b < pa.batch; ++b) {
                apply_move_block(ap, b);
                __syncthreads();
                if (!ap.done[b]) {
                    routes_scan_small_block<kPersistThreads>(sa.r, b);
                    if (pa.tables_in_block) {
                        __syncthreads();
                        route_tables_warp(sa, b, threadIdx.x >> 5);  // returns at once beyond the instance's routes
                    }
                }
                __syncthreads();
            }
        }
        grid.sync();
        // done[] only changed in A: every block sees the same values here
        int live = 0;
        for (int b = threadIdx.x; b < pa.batch; b += kPersistThreads) live |= ap.done[b] == 0;
        if (!__syncthreads_or(live)) break;
        if (pa.tables_in_block) continue;
        // T
        for (int v = blockIdx.x * (kPersistThreads / 32) + (threadIdx.x >> 5); v < pa.rmax * pa.batch; v += warps) {
            const int b = v / pa.rmax;
            if (!ap.done[b]) route_tables_warp(sa, b, v - b * pa.rmax);  // returns at once beyond the instance's routes
        }
        grid.sync();
    }
}

__global__ void finish_keys_kernel(const unsigned long long *__restrict__ src, unsigned long long *__restrict__ dst,
                                   int batch, int to_signed)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < batch) dst[i] = to_signed ? src[i] ^ 0x8000000000000000ull : src[i];
}

__global__ void fill_keys_kernel(unsigned long long *dst, int batch)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < batch) dst[i] = kNoMove;
}

__global__ void fill_signed_none_kernel(unsigned long long *keys, size_t count)
{
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < count) keys[i] = 0x7FFFFFFFFFFFFFFFull;
}

}  // namespace

// Evaluates this shard's candidates of every instance and leaves one key per instance in best_d (device),
// optionally in signed-order form.  The state must be fresh (state_refresh after the last change).
int eval_moves_launch(rt_problem *p, int families, int shard_index, int shard_count, unsigned long long *best_d,
                      bool signed_keys, cudaStream_t st)
{
    rt_engine *e = p->e;
    // Where the minimum is accumulated.  Normally p->best (reset by every state refresh; a second evaluation of the same
    // state clears it itself) followed by a small conversion/copy kernel.  When the state refresh that has just run
    // was told about the caller's buffer (p->ext_best, the plan entry points), that buffer already holds "none" in
    // signed-order form and the kernels reduce straight into it: no fill, no conversion launch.
    const bool direct = signed_keys && best_d == p->ext_best && p->ext_best_clean;
    p->ext_best_clean = false;
    if (!direct) {
        if (!p->best_clean) fill_keys_kernel<<<(p->batch + 255) / 256, 256, 0, st>>>(p->best, p->batch);
        p->best_clean = false;
    }
    EvalArgs a = make_args(p, families, shard_index, shard_count, direct ? best_d : p->best);
    a.signed_dst = direct ? 1 : 0;
    const int pmax = p->pmax > 0 ? p->pmax : p->pstride;  // largest P of the batch
    const int shard_rows = max_shard_rows(pmax, shard_index, shard_count, a.beta_inter);
    const int anchor_rows = max_shard_rows(pmax, shard_index, shard_count, a.beta_intra);
    // The sweep kernel is latency bound (a sequential walk per anchor, a few warps per SM) and the tile kernel is ALU
    // bound: they only share the best[] atomics, so the sweep runs on the side stream underneath the tile kernel.
    const bool overlap = (families & 6) && p->side_stream;
    cudaStream_t st_intra = st;
    if (overlap) {
        RT_CUDA(e, cudaEventRecord(p->ev_fork, st));
        RT_CUDA(e, cudaStreamWaitEvent(p->side_stream, p->ev_fork, 0));
        st_intra = p->side_stream;
    }
    RT_TRY(launch_intra<kBest>(p, a, anchor_rows, p->batch, st_intra));
    if (families & 6) RT_TRY(launch_inter<kBest>(p, a, pmax, shard_rows, p->batch, 0, st));
    if (overlap) {
        RT_CUDA(e, cudaEventRecord(p->ev_join, p->side_stream));
        RT_CUDA(e, cudaStreamWaitEvent(st, p->ev_join, 0));
    }
    if (!direct && (best_d != p->best || signed_keys))
        finish_keys_kernel<<<(p->batch + 255) / 256, 256, 0, st>>>(p->best, best_d, p->batch, signed_keys ? 1 : 0);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

// Runs up to `iters` iterations of the one-move search in one cooperative launch (search_persistent_kernel); the state
// must be fresh.  Returns RT_OK after enqueueing, -1 where this form does not apply (the caller then launches the
// iterations kernel by kernel): a stored matrix, a sharded replica, an instance too large for one block to rescan,
// a device without cooperative launch.
int search_persistent_launch(rt_problem *p, int families, int iters, bool improving_only, cudaStream_t st)
{
    rt_engine *e = p->e;
    const int pmax = p->pmax > 0 ? p->pmax : p->pstride;
    // Measured per applied move, one launch vs the graph-replayed chain of three (B200, us): 13.9 vs 16.1 at 100 stops,
    // 20.1 vs 19.2 at 400, 27.9 vs 23.7 at 1,000, 43.1 vs 31.9 at 2,000.  The barriers cost less than the launch
    // dependencies they replace, but this kernel carries the registers of all three phases (121: four blocks per SM)
    // and runs the two evaluation kernels one after the other inside a block instead of side by side — a price that
    // grows with the instance.  So: the one launch for small instances, the chain above.
    const int limit = g_persistent_search >= 2 ? kPersistMaxPositions : kPersistPositions;
    if (!g_persistent_search || !eval_without_matrix(p) || !p->xyz || p->shard_count != 1 || pmax > limit ||
        !intra_by_warp(p, pmax))
        return -1;
    static int coop = -1, blocks_per_sm[2] = {0, 0};
    if (coop < 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) != cudaSuccess) coop = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm[0], search_persistent_kernel<false>,
                                                          kPersistThreads, 0) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm[1], search_persistent_kernel<true>,
                                                          kPersistThreads, 0) != cudaSuccess)
            coop = 0;
        (void)cudaGetLastError();
    }
    if (!coop) return -1;
    if (!p->best_clean) {
        fill_keys_kernel<<<(p->batch + 255) / 256, 256, 0, st>>>(p->best, p->batch);
        RT_CUDA(e, cudaGetLastError());
    }
    EvalArgs ea = make_args(p, families, 0, 1, p->best);
    ea.strip_rows = pick_strip_rows(pmax, pmax, p->batch, false, kFusedCols);
    const dim3 tg = fused_tile_grid(ea, pmax, p->batch);
    const bool fast = all_pairs_in_fast_range(p, ea);
    PersistArgs pa{};
    pa.iters = iters;
    pa.batch = p->batch;
    pa.inter_gx = (families & 6) ? static_cast<int>(tg.x) : 0;
    pa.inter_gy = (families & 6) ? static_cast<int>(tg.y) : 0;
    pa.intra_gx = (pmax + 2 + kIntraWarps - 1) / kIntraWarps;
    pa.rmax = std::max(1, pmax - p->n);
    pa.tables_in_block = pa.rmax <= kPersistThreads / 32 ? 1 : 0;
    pa.ticket = p->ticket;
    RT_CUDA(e, cudaMemsetAsync(p->ticket, 0, sizeof(unsigned), st));
    StateArgs sa = state_args(p, /*fused_vectors=*/true);
    ApplyArgs ap = apply_args(p, p->best, false, improving_only, nullptr);
    const long long want = std::max<long long>({static_cast<long long>(pa.inter_gx * pa.inter_gy + pa.intra_gx) * p->batch,
                                                 p->batch, (static_cast<long long>(pa.rmax) * p->batch + 3) / 4});
    const int per_sm = g_persist_blocks > 0 ? std::min(g_persist_blocks, blocks_per_sm[fast ? 1 : 0]) : blocks_per_sm[fast ? 1 : 0];
    const int grid = static_cast<int>(std::min<long long>(want, static_cast<long long>(per_sm) * kNumSMs));
    if (grid < 1) return -1;
    void *args[] = {&ea, &sa, &ap, &pa};
    const void *fn = fast ? reinterpret_cast<const void *>(search_persistent_kernel<true>)
                          : reinterpret_cast<const void *>(search_persistent_kernel<false>);
    const cudaError_t ce = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kPersistThreads), args, 0, st);
    if (ce == cudaErrorCooperativeLaunchTooLarge || ce == cudaErrorNotSupported) {
        // fewer SMs than the occupancy query promised (a partitioned GPU): the chain of launches serves from now on
        (void)cudaGetLastError();
        coop = 0;
        return -1;
    }
    RT_CUDA(e, ce);
    p->best_clean = false;  // a search that ended at its optimum leaves the last (non-improving) keys behind
    p->ext_best_clean = false;
    p->scan_done = false;
    return RT_OK;
}

// Full evaluation of this shard's candidates into a per-route-pair table (batch.cu): inter[ra][rb] for every ordered
// pair of routes followed by intra[r], [batch][cr][cr] + [batch][cr] keys with absolute positions.  The state must
// be fresh.  signed_keys: leave the table in order-preserving signed form (key ^ 2^63) so that replicas can combine
// their shards with an int64 MIN all-reduce.
int eval_table_launch(rt_problem *p, int families, int shard_index, int shard_count, unsigned long long *table_d,
                      bool signed_keys, cudaStream_t st)
{
    rt_engine *e = p->e;
    EvalArgs a = make_args(p, families, shard_index, shard_count, p->best);
    const size_t B = p->batch, cr = p->cache_r;
    a.inter = table_d;
    a.intra = table_d + B * cr * cr;
    a.cr = p->cache_r;
    const size_t count = B * (cr * cr + cr);
    if (signed_keys) {  // "none" in signed-order form is INT64_MAX: the kernels reduce with the signed atomic, no conversion pass
        fill_signed_none_kernel<<<static_cast<unsigned>((count + 255) / 256), 256, 0, st>>>(table_d, count);
        RT_CUDA(e, cudaGetLastError());
        a.signed_dst = 1;
    } else {
        RT_CUDA(e, cudaMemsetAsync(table_d, 0xFF, count * 8, st));
    }
    const int pmax = p->pmax > 0 ? p->pmax : p->pstride;
    const int shard_rows = max_shard_rows(pmax, shard_index, shard_count, a.beta_inter);
    const int anchor_rows = max_shard_rows(pmax, shard_index, shard_count, a.beta_intra);
    const bool overlap = (families & 6) && p->side_stream;
    cudaStream_t st_intra = st;
    if (overlap) {
        RT_CUDA(e, cudaEventRecord(p->ev_fork, st));
        RT_CUDA(e, cudaStreamWaitEvent(p->side_stream, p->ev_fork, 0));
        st_intra = p->side_stream;
    }
    RT_TRY(launch_intra<kTable>(p, a, anchor_rows, p->batch, st_intra));
    // table mode flushes per route change: short strips keep the flush count low
    if (families & 6) RT_TRY(launch_inter<kTable>(p, a, pmax, shard_rows, p->batch, kStripRows, st));
    if (overlap) {
        RT_CUDA(e, cudaEventRecord(p->ev_join, p->side_stream));
        RT_CUDA(e, cudaStreamWaitEvent(st, p->ev_join, 0));
    }
    return RT_OK;
}

int eval_dump_launch(rt_problem *p, int instance, int family, int *delta_d, unsigned char *feas_d, cudaStream_t st)
{
    EvalArgs a = make_args(p, 1 << family, 0, 1, p->best);
    a.dump_instance = instance;
    a.dump_family = family;
    a.dump_delta = delta_d;
    a.dump_feas = feas_d;
    const int pmax = p->pmax > 0 ? p->pmax : p->pstride;
    if (family != 0) RT_TRY(launch_inter<kDump>(p, a, pmax, pmax, 1, kStripRows, st));
    return launch_intra<kDump>(p, a, pmax, 1, st);
}

}  // namespace rt

extern "C" int rt_debug_set(const char *name, int value)
{
    if (!name) return RT_ERR_INVALID_ARG;
    if (!strcmp(name, "strip_rows")) rt::g_strip_rows = value;
    else if (!strcmp(name, "state_fused_max")) rt::g_state_fused_max = value;
    else if (!strcmp(name, "apply_prep")) rt::g_apply_prep = value;
    else if (!strcmp(name, "small_strips")) rt::g_small_strips = value;
    else if (!strcmp(name, "intra_kernel")) rt::g_intra_kernel = value;
    else if (!strcmp(name, "fused_eval")) rt::g_fused_eval = value;
    else if (!strcmp(name, "persistent_search")) rt::g_persistent_search = value;
    else if (!strcmp(name, "persist_blocks")) rt::g_persist_blocks = value;
    else return RT_ERR_INVALID_ARG;
    return RT_OK;
}
