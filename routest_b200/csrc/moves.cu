// moves.cu — cost delta + capacity/distance/time-window feasibility of local-search moves and the best-move
// argmin (SPEC.md §7).  No reference counterpart (the reference has no improvement phase, utils.py:111-139).
//
// Two kernels share the work:
//   * eval_inter_kernel — candidates between two different routes (relocate, swap): O(1) each.  A block owns a
//     64 x 64 tile of (p, q) position pairs; the 66 x 66 haloed tile of matrix cells M[node p][node q] is gathered
//     ONCE into shared memory and then serves every candidate of the tile (each cell is read by up to five
//     (p,q) sites and two families); the per-position records of the q side live in registers, those of the p
//     side are broadcast from shared memory.
//   * eval_intra_kernel — candidates inside one route (2-opt, relocate, swap): one thread per anchor position
//     sweeps its route keeping two segment summaries (a, b, L) — forward and reversed — so every candidate is
//     again O(1) and exact for time windows.
// Each thread keeps the minimum packed key of its feasible candidates; warp shuffles and one shared-memory step
// reduce it to one 64-bit atomicMin per block.  The minimum of packed keys is order independent, so the chosen
// move does not depend on the thread, block or GPU partition.
//
// The matrix is assumed symmetric here (true for SPEC.md §2 matrices): M[a][b] is read where M[b][a] is meant.
#include <climits>

#include "geometry.cuh"
#include "problem.cuh"

namespace rt {

namespace {

constexpr int kTP = 64, kTQ = 64;  // tile of (p, q) pairs per block
constexpr int kInterThreads = 256;

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

template <int THREADS>
__device__ __forceinline__ void block_min_to_global(unsigned long long best, unsigned long long *dst)
{
    __shared__ unsigned long long s_best[THREADS / 32];
    best = warp_min_u64(best);
    if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned long long v = threadIdx.x < THREADS / 32 ? s_best[threadIdx.x] : kNoMove;
        v = warp_min_u64(v);
        if (threadIdx.x == 0 && v != kNoMove) atomicMin(dst, v);
    }
}

struct EvalArgs {
    const int2 *cells;
    const int *node;
    const int *npos;
    const int4 *RA, *RB, *RC, *RD, *H;
    const int2 *E, *SK;
    const double *PX;      // unit vectors in position order [3][pstride] per instance
    float fd, ft;          // cell scales and fast-path limit (geometry.cuh), for rebuilt cells
    int lim_hi;
    int n, pitch, pstride;
    int families;            // bit 0 = 2-opt, 1 = relocate, 2 = swap
    int shard_index, shard_count;
    unsigned long long *best;  // [batch]
    // dump mode (one instance, one family)
    int dump_instance, dump_family;
    int *dump_delta;
    unsigned char *dump_feas;
};

__device__ __forceinline__ void shard_range(int P, int idx, int cnt, int &lo, int &hi)
{
    lo = static_cast<int>(static_cast<long long>(P) * idx / cnt);
    hi = static_cast<int>(static_cast<long long>(P) * (idx + 1) / cnt);
}

// Full evaluation of one inter-route candidate from the cold records (global memory): only reached by the few
// candidates whose delta beats the current bound, or by every candidate in dump mode.
__device__ __noinline__ bool check_relocate(const EvalArgs &a, size_t po, int p, int q, int2 c_pq, int2 c_pq1)
{
    const int4 pB = a.RB[po + p], pC = a.RC[po + p];
    const int4 qA = a.RA[po + q], qD = a.RD[po + q];
    const int ins = c_pq.x + c_pq1.x - qA.z;
    const int tu = max(pB.x, qA.x + c_pq.y);
    return (pC.w & 2) && pB.w <= qD.x && ins <= qD.y && tu <= pB.y && tu + pB.z + c_pq1.y <= qA.y;
}

__device__ __noinline__ bool check_swap(const EvalArgs &a, size_t po, int p, int q, int2 c_up, int2 c_dn, int2 c_lf,
                                        int2 c_rt)
{
    const int4 pA = a.RA[po + p], pB = a.RB[po + p], pC = a.RC[po + p], pD = a.RD[po + p];
    const int4 qA = a.RA[po + q], qB = a.RB[po + q], qC = a.RC[po + q], qD = a.RD[po + q];
    const int d1 = c_up.x + c_dn.x - pC.z;  // route of p: pu -> v -> nu
    const int d2 = c_lf.x + c_rt.x - qC.z;  // route of q: pw -> u -> nw
    const int tv = max(qB.x, pC.x + c_up.y);
    const int tu = max(pB.x, qC.x + c_lf.y);
    return qB.w - pB.w <= pD.x && pB.w - qB.w <= qD.x && d1 <= pD.y && d2 <= qD.y && tv <= qB.y &&
           tv + qB.z + c_dn.y <= pA.y && tu <= pB.y && tu + pB.z + c_rt.y <= qA.y;
}

// COMPUTED: the tile's cells are rebuilt from the unit vectors of the 66 + 66 positions (SPEC.md §2, the same
// function as the matrix kernel, so the same bits) instead of being gathered from the stored matrix: ~25 ALU
// instructions per cell replace a scattered 32-byte sector fetch per cell.  Only valid while the stored matrix is
// the great-circle matrix of the problem's coordinates; a caller-supplied matrix takes the gather path.
template <bool DUMP, bool COMPUTED>
__global__ void __launch_bounds__(kInterThreads) eval_inter_kernel(const EvalArgs a)
{
    __shared__ int2 S[kTP + 2][kTQ + 2];
    __shared__ int4 sH[kTP];
    __shared__ int s_nodeP[kTP + 2], s_nodeQ[kTQ + 2];
    __shared__ double sPX[3][kTP + 2], sQX[3][kTQ + 2];

    const int b = DUMP ? a.dump_instance : blockIdx.z;
    const int P = a.npos[b];
    int p_lo, p_hi;
    shard_range(P, a.shard_index, a.shard_count, p_lo, p_hi);
    const int p0 = p_lo + blockIdx.y * kTP, q0 = blockIdx.x * kTQ;
    if (p0 >= p_hi || q0 >= P) return;

    const size_t po = static_cast<size_t>(b) * a.pstride;
    const int tid = threadIdx.x;

    for (int k = tid; k < kTP; k += kInterThreads) sH[k] = p0 + k < P ? a.H[po + p0 + k] : make_int4(0, 0, 0, 0);
    if (COMPUTED) {
        const double *px = a.PX + static_cast<size_t>(b) * 3 * a.pstride;
        for (int k = tid; k < 3 * (kTP + 2); k += kInterThreads) {
            const int comp = k / (kTP + 2), i = k - comp * (kTP + 2), p = min(max(p0 - 1 + i, 0), P - 1);
            sPX[comp][i] = px[static_cast<size_t>(comp) * a.pstride + p];
        }
        for (int k = tid; k < 3 * (kTQ + 2); k += kInterThreads) {
            const int comp = k / (kTQ + 2), j = k - comp * (kTQ + 2), q = min(max(q0 - 1 + j, 0), P - 1);
            sQX[comp][j] = px[static_cast<size_t>(comp) * a.pstride + q];
        }
        __syncthreads();
        for (int k = tid; k < (kTP + 2) * (kTQ + 2); k += kInterThreads) {
            const int i = k / (kTQ + 2), j = k - i * (kTQ + 2);
            const double ax = sPX[0][i], ay = sPX[1][i], az = sPX[2][i];
            const double bx = sQX[0][j], by = sQX[1][j], bz = sQX[2][j];
            const double dx = ax - bx, dy = ay - by, dz = az - bz;
            const double s2 = fma(dx, dx, fma(dy, dy, dz * dz));
            S[i][j] = __double2hiint(s2) < a.lim_hi ? cell_i32_fast(s2, a.fd, a.ft)
                                                    : cell_i32(ax, ay, az, bx, by, bz, a.fd, a.ft);
        }
    } else {
        const int *node = a.node + po;
        const int2 *M = a.cells + static_cast<size_t>(b) * a.n * a.pitch;
        for (int k = tid; k < kTP + 2; k += kInterThreads) {
            const int p = p0 - 1 + k;
            s_nodeP[k] = (p >= 0 && p < P) ? node[p] : -1;
        }
        for (int k = tid; k < kTQ + 2; k += kInterThreads) {
            const int q = q0 - 1 + k;
            s_nodeQ[k] = (q >= 0 && q < P) ? node[q] : -1;
        }
        __syncthreads();
        // the one gather of the tile: 66 x 66 cells, scattered within matrix rows
        for (int k = tid; k < (kTP + 2) * (kTQ + 2); k += kInterThreads) {
            const int i = k / (kTQ + 2), j = k - i * (kTQ + 2);
            const int na = s_nodeP[i], nb = s_nodeQ[j];
            S[i][j] = (na >= 0 && nb >= 0) ? __ldg(&M[static_cast<size_t>(na) * a.pitch + nb]) : make_int2(0, 0);
        }
    }
    __syncthreads();

    const int c = tid & (kTQ - 1), g = tid / kTQ;  // column of the tile, group of rows
    const int q = q0 + c, j = c + 1;
    const bool q_in = q < P;
    const int4 Hq = q_in ? a.H[po + q] : make_int4(0, 0, 0, -1);
    const int q_route = Hq.w & 0x3FFFFFFF;
    const bool q_cust = q_in && ((Hq.w >> 30) & 1);
    const bool rel_q = q_in && (a.families & 2) && q <= P - 2;
    const bool swap_q = q_cust && (a.families & 4);

    // Only the minimum key matters, so a candidate is fully evaluated (feasibility, cold records) only if its delta
    // can still beat the best feasible key seen so far by this warp; the bound starts from whatever other blocks
    // have already published for this instance.  Exact: a pruned candidate has a larger key than an existing one.
    unsigned long long best = kNoMove;
    int bound = INT_MAX;
    if (!DUMP) {
        const unsigned long long seen = *reinterpret_cast<const volatile unsigned long long *>(a.best + b);
        if (seen != kNoMove) bound = static_cast<int>(static_cast<unsigned>(seen >> 32) - 0x80000000u);
    }
    constexpr int kRows = kTP / (kInterThreads / kTQ);
    const int pi0 = g * kRows;
    int2 c_up = S[pi0][j], c_cur = S[pi0 + 1][j];
    for (int rr = 0; rr < kRows; ++rr) {
        const int pi = pi0 + rr, p = p0 + pi, i = pi + 1;
        if (p >= p_hi) break;  // warp-uniform
        const int2 c_dn = S[i + 1][j], c_lf = S[i][j - 1], c_rt = S[i][j + 1];
        const int4 Hp = sH[pi];
        if ((Hp.w >> 30) & 1) {  // p is a customer (warp-uniform)
            const bool other = (Hp.w & 0x3FFFFFFF) != q_route;
            const int d_rel = Hp.x + c_cur.x + c_rt.x - Hq.z;
            const int d_swap = c_up.x + c_dn.x - Hp.y + c_lf.x + c_rt.x - Hq.y;
            const bool v_rel = rel_q && other, v_swap = swap_q && other && p < q;
            if (DUMP) {
                if (v_rel && a.dump_family == 1) {
                    a.dump_delta[static_cast<size_t>(p) * P + q] = d_rel;
                    a.dump_feas[static_cast<size_t>(p) * P + q] = check_relocate(a, po, p, q, c_cur, c_rt) ? 1 : 0;
                }
                if (v_swap && a.dump_family == 2) {
                    a.dump_delta[static_cast<size_t>(p) * P + q] = d_swap;
                    a.dump_feas[static_cast<size_t>(p) * P + q] = check_swap(a, po, p, q, c_up, c_dn, c_lf, c_rt) ? 1 : 0;
                }
            } else {
                if (v_rel && d_rel <= bound && check_relocate(a, po, p, q, c_cur, c_rt)) {
                    const unsigned long long k = pack_key(d_rel, 1, p, q);
                    if (k < best) {
                        best = k;
                        bound = d_rel;
                    }
                }
                if (v_swap && d_swap <= bound && check_swap(a, po, p, q, c_up, c_dn, c_lf, c_rt)) {
                    const unsigned long long k = pack_key(d_swap, 2, p, q);
                    if (k < best) {
                        best = k;
                        bound = d_swap;
                    }
                }
                bound = __reduce_min_sync(0xffffffffu, bound);  // share the bound across the warp (one REDUX)
            }
        }
        c_up = c_cur;
        c_cur = c_dn;
    }
    if (!DUMP) block_min_to_global<kInterThreads>(best, a.best + b);
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

template <bool DUMP>
__global__ void __launch_bounds__(kIntraThreads) eval_intra_kernel(const EvalArgs a)
{
    const int b = DUMP ? a.dump_instance : blockIdx.y;
    const int P = a.npos[b];
    int t_lo, t_hi;
    shard_range(P, a.shard_index, a.shard_count, t_lo, t_hi);  // same-route work is sharded by anchor
    const int t = t_lo + blockIdx.x * kIntraThreads + threadIdx.x;
    unsigned long long best = kNoMove;

    if (t < t_hi && t < P - 1) {
        const size_t po = static_cast<size_t>(b) * a.pstride;
        const int *node = a.node + po;
        const int2 *M = a.cells + static_cast<size_t>(b) * a.n * a.pitch;
        const int4 *RA = a.RA + po, *RB = a.RB + po, *RC = a.RC + po, *RD = a.RD + po;
        const int2 *E = a.E + po, *SK = a.SK + po;

        const int nt = node[t];
        const bool t_cust = nt != 0;
        const int4 tA = RA[t], tB = RB[t], tC = RC[t], tD = RD[t];
        const int dslack = tD.y;           // max_dist - rdist of this route (clamped)
        const int2 Et = E[t];              // cell(node t, node t+1)
        const int4 t1B = RB[t + 1];        // stop data of node t+1 (first node of F)
        const int dep_t = tA.x;            // B[t] + service[node t]
        const int dep_tm1 = tC.x;          // B[t-1] + service[node t-1]   (customers only)
        const int2 SKt = SK[t];            // cell(node t-1, node t+1)
        const int2 *row_m = t_cust ? M + static_cast<size_t>(node[t - 1]) * a.pitch : M;
        const int2 *row_0 = M + static_cast<size_t>(nt) * a.pitch;
        const int2 *row_p = M + static_cast<size_t>(node[t + 1]) * a.pitch;
        const bool f2 = a.families & 1, f_rel = a.families & 2, f_swap = a.families & 4;

        // cells of rows node[t-1], node[t], node[t+1] at column node[x]: start at x = t+1
        int2 cm = SKt, c0 = Et, cp = make_int2(0, 0);
        int2 c0_prev = make_int2(0, 0);  // c0 at x-1
        Seg F = Seg{0, 0, 0, true}, Rv = Seg{0, 0, 0, true};
        int svc_xm1 = 0, Et_xm1 = 0;  // service[node x-1], T[node x-1][node x]

        for (int x = t + 1; x < P; ++x) {
            const int nx = node[x];
            if (nx == 0) break;
            const int nxn = node[x + 1];
            const int4 xA = RA[x], xB = RB[x], xC = RC[x], xD = RD[x];
            const int2 Ex = make_int2(xA.z, xD.z);  // cell(node x, node x+1)
            const int zn = xA.y;                    // Z[x+1]
            // next column's cells (rows fixed per thread): the only scattered loads of the step
            const int2 cm1 = t_cust ? __ldg(row_m + nxn) : make_int2(0, 0);
            const int2 c01 = __ldg(row_0 + nxn);
            const int2 cp1 = __ldg(row_p + nxn);
            const Seg sx = seg_single(xB.x, xB.y);

            // reversed segment node[x..t+1]: prepend x
            Rv = (x == t + 1) ? sx : seg_concat(sx, xB.z + Et_xm1, Rv);

            // (A) 2-opt(t, x)
            if (f2 && x >= t + 2) {
                const int d = c0.x + cp1.x - Et.x - Ex.x;
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
                    m = seg_concat(m, xB.z + c0.y, seg_single(tB.x, tB.y));
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
                // (C) relocate(x -> after t): prefix(..t) + [x] + F + suffix(x+1..)
                if (x >= t + 2) {
                    const Seg m = seg_concat(sx, xB.z + cp.y, F);
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
                    m = seg_concat(sx, xB.z + Et.y, seg_single(tB.x, tB.y));
                    d = cm.x + c01.x - (tC.z - Et.x) - Ex.x;  // pv->x + x->t + t->nxt - pv->t - t->x - x->nxt
                } else {
                    m = seg_concat(sx, xB.z + cp.y, F);
                    m = seg_concat(m, svc_xm1 + c0_prev.y, seg_single(tB.x, tB.y));
                    d = cm.x + cp.x - tC.z + c0_prev.x + c01.x - xC.z;
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
            c0_prev = c0;
            cm = cm1;
            c0 = c01;
            cp = cp1;
        }
    }
    if (!DUMP) block_min_to_global<kIntraThreads>(best, a.best + b);
}

EvalArgs make_args(rt_problem *p, int families, int shard_index, int shard_count, unsigned long long *best_d)
{
    EvalArgs a{};
    a.cells = p->cells;
    a.node = p->node;
    a.npos = p->npos;
    a.RA = p->RA;
    a.RB = p->RB;
    a.RC = p->RC;
    a.RD = p->RD;
    a.H = p->H;
    a.PX = p->PX;
    a.fd = static_cast<float>(RT_EARTH_RADIUS_M * p->dist_scale);
    a.ft = static_cast<float>(RT_EARTH_RADIUS_M * p->time_scale);
    a.lim_hi = fast_path_limit_hi(a.fd, a.ft);
    a.E = p->E;
    a.SK = p->SK;
    a.n = p->n;
    a.pitch = p->pitch;
    a.pstride = p->pstride;
    a.families = families;
    a.shard_index = shard_index;
    a.shard_count = shard_count;
    a.best = best_d;
    return a;
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

}  // namespace

// Evaluates this shard's candidates of every instance and leaves one key per instance in best_d (device),
// optionally in signed-order form.  The state must be fresh (state_refresh after the last change).
int eval_moves_launch(rt_problem *p, int families, int shard_index, int shard_count, unsigned long long *best_d,
                      bool signed_keys, cudaStream_t st)
{
    rt_engine *e = p->e;
    fill_keys_kernel<<<(p->batch + 255) / 256, 256, 0, st>>>(p->best, p->batch);
    EvalArgs a = make_args(p, families, shard_index, shard_count, p->best);
    const int pmax = p->pmax > 0 ? p->pmax : p->pstride;  // largest P of the batch
    const int shard_rows = (pmax + shard_count - 1) / shard_count + 1;
    if (families & 6) {
        dim3 grid((pmax + kTQ - 1) / kTQ, (shard_rows + kTP - 1) / kTP, p->batch);
        if (p->cells_from_coords) eval_inter_kernel<false, true><<<grid, kInterThreads, 0, st>>>(a);
        else eval_inter_kernel<false, false><<<grid, kInterThreads, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
    }
    {
        dim3 grid((shard_rows + kIntraThreads - 1) / kIntraThreads, p->batch);
        eval_intra_kernel<false><<<grid, kIntraThreads, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
    }
    finish_keys_kernel<<<(p->batch + 255) / 256, 256, 0, st>>>(p->best, best_d, p->batch, signed_keys ? 1 : 0);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

int eval_dump_launch(rt_problem *p, int instance, int family, int *delta_d, unsigned char *feas_d, cudaStream_t st)
{
    rt_engine *e = p->e;
    EvalArgs a = make_args(p, 1 << family, 0, 1, p->best);
    a.dump_instance = instance;
    a.dump_family = family;
    a.dump_delta = delta_d;
    a.dump_feas = feas_d;
    const int pmax = p->pmax > 0 ? p->pmax : p->pstride;
    if (family != 0) {
        dim3 grid((pmax + kTQ - 1) / kTQ, (pmax + kTP - 1) / kTP, 1);
        if (p->cells_from_coords) eval_inter_kernel<true, true><<<grid, kInterThreads, 0, st>>>(a);
        else eval_inter_kernel<true, false><<<grid, kInterThreads, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
    }
    dim3 grid((pmax + kIntraThreads - 1) / kIntraThreads, 1);
    eval_intra_kernel<true><<<grid, kIntraThreads, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

}  // namespace rt
