// moves.cu — cost delta + capacity/distance/time-window feasibility of local-search moves and the best-move
// argmin (SPEC.md §7).  No reference counterpart (the reference has no improvement phase, utils.py:111-139).
//
// Both kernels read the stop matrix in ROUTE ORDER, G[p][q] = M[node p][node q] (state.cu), so a thread that owns
// insertion point / partner q and walks down the moved positions p touches G[p-1..p+1][q-1..q+1]: each warp-wide
// load is one contiguous 256-byte run of a row, every cell comes from DRAM/L2 once per strip and its four
// re-uses (left/right/up/down neighbours) are L1 hits or register rotations.
//   * eval_inter_kernel — candidates between two different routes (relocate, swap): O(1) each from the 16-byte
//     "hot" record of p and of q (removal delta, old-edge sum, edge after, route id) plus five cells.  Only the
//     minimum key matters, so the full feasibility test (capacity, distance limit, time windows; cold records)
//     runs only for candidates whose delta can still beat the best feasible key seen so far.
//   * eval_intra_kernel — candidates inside one route (2-opt, relocate, swap): one thread per anchor position
//     sweeps its route keeping two segment summaries (a, b, L) — forward and reversed — so every candidate is
//     again O(1) and exact for time windows.
// Each thread keeps the minimum packed key of its feasible candidates; warp shuffles and one shared-memory step
// reduce it to one 64-bit atomicMin per block.  The minimum of packed keys is order independent, so the chosen
// move does not depend on the thread, block or GPU partition.
//
// The matrix is assumed symmetric here (true for SPEC.md §2 matrices): G[a][b] is read where G[b][a] is meant.
#include <climits>

#include "problem.cuh"

namespace rt {

namespace {

constexpr int kInterThreads = 128;  // one column q per thread
constexpr int kStripRows = 32;     // rows p per block

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
    const int2 *G;         // [batch][grows][gpitch] route-order matrix
    int grows, gpitch;
    const int *node;
    const int *npos;
    const int4 *RA, *RB, *RC, *RD, *H;
    const int2 *E, *SK;
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

// Full feasibility of one inter-route candidate from the cold records (SPEC.md §7): capacity, distance limit and
// time windows of both routes.  Only reached by candidates whose delta can still win, or by all in dump mode.
__device__ __forceinline__ bool check_relocate(int4 pB, int4 pC, int4 qA, int4 qD, int2 c_pq, int2 c_pq1)
{
    const int ins = c_pq.x + c_pq1.x - qA.z;
    const int tu = max(pB.x, qA.x + c_pq.y);
    return (pC.w & 2) && pB.w <= qD.x && ins <= qD.y && tu <= pB.y && tu + pB.z + c_pq1.y <= qA.y;
}

__device__ __forceinline__ bool check_swap(int4 pA, int4 pB, int4 pC, int4 pD, int4 qA, int4 qB, int4 qC, int4 qD,
                                           int2 c_up, int2 c_dn, int2 c_lf, int2 c_rt)
{
    const int d1 = c_up.x + c_dn.x - pC.z;  // route of p: pu -> v -> nu
    const int d2 = c_lf.x + c_rt.x - qC.z;  // route of q: pw -> u -> nw
    const int tv = max(qB.x, pC.x + c_up.y);
    const int tu = max(pB.x, qC.x + c_lf.y);
    return qB.w - pB.w <= pD.x && pB.w - qB.w <= qD.x && d1 <= pD.y && d2 <= qD.y && tv <= qB.y &&
           tv + qB.z + c_dn.y <= pA.y && tu <= pB.y && tu + pB.z + c_rt.y <= qA.y;
}

// grid (column strips of 128, row strips of kStripRows within the shard, instances)
template <bool DUMP>
__global__ void __launch_bounds__(kInterThreads, 8) eval_inter_kernel(const EvalArgs a)
{
    // hot and cold records of the strip's rows (p side) and of the block's columns (q side): the slow path never
    // waits on global memory
    __shared__ int4 sH[kStripRows], sPA[kStripRows], sPB[kStripRows], sPC[kStripRows], sPD[kStripRows];
    __shared__ int4 sQA[kInterThreads], sQB[kInterThreads], sQC[kInterThreads], sQD[kInterThreads];

    const int b = DUMP ? a.dump_instance : blockIdx.z;
    const int P = a.npos[b];
    int p_lo, p_hi;
    shard_range(P, a.shard_index, a.shard_count, p_lo, p_hi);
    const int p0 = p_lo + blockIdx.y * kStripRows, q0 = blockIdx.x * kInterThreads;
    if (p0 >= p_hi || q0 >= P) return;

    const size_t po = static_cast<size_t>(b) * a.pstride;
    const int2 *G = a.G + static_cast<size_t>(b) * a.grows * a.gpitch;
    const int tid = threadIdx.x;
    const int q = q0 + tid;
    const bool q_in = q < P;
    if (tid < kStripRows) {
        const bool in = p0 + tid < P;
        const size_t o = po + p0 + tid;
        sH[tid] = in ? a.H[o] : make_int4(0, 0, 0, 0);
        sPA[tid] = in ? a.RA[o] : make_int4(0, 0, 0, 0);
        sPB[tid] = in ? a.RB[o] : make_int4(0, 0, 0, 0);
        sPC[tid] = in ? a.RC[o] : make_int4(0, 0, 0, 0);
        sPD[tid] = in ? a.RD[o] : make_int4(0, 0, 0, 0);
    }
    const int4 Hq = q_in ? a.H[po + q] : make_int4(0, 0, 0, -1);
    sQA[tid] = q_in ? a.RA[po + q] : make_int4(0, 0, 0, 0);
    sQB[tid] = q_in ? a.RB[po + q] : make_int4(0, 0, 0, 0);
    sQC[tid] = q_in ? a.RC[po + q] : make_int4(0, 0, 0, 0);
    sQD[tid] = q_in ? a.RD[po + q] : make_int4(0, 0, 0, 0);
    __syncthreads();

    // columns are clamped so that every lane can load unconditionally (pad columns of G are valid depot cells)
    const int qc = min(q, a.gpitch - 1), ql = max(qc - 1, 0), qr = min(qc + 1, a.gpitch - 1);
    const int q_route = Hq.w & 0x3FFFFFFF;
    const bool q_cust = q_in && ((Hq.w >> 30) & 1);
    const bool rel_q = q_in && (a.families & 2) && q <= P - 2;
    const bool swap_q = q_cust && (a.families & 4);

    // Only the minimum key matters, so a candidate is fully evaluated only if its delta can still beat the best
    // feasible key seen so far by this warp; the bound starts from whatever other blocks have already published
    // for this instance.  Exact: a pruned candidate has a larger key than an existing feasible one.
    unsigned long long best = kNoMove;
    int bound = INT_MAX;
    if (!DUMP) {
        const unsigned long long seen = *reinterpret_cast<const volatile unsigned long long *>(a.best + b);
        if (seen != kNoMove) bound = static_cast<int>(static_cast<unsigned>(seen >> 32) - 0x80000000u);
    }
    const int rows = min(kStripRows, p_hi - p0);
    const int last_row = a.grows - 1;
    auto row_ptr = [&](int r) { return G + static_cast<size_t>(min(max(r, 0), last_row)) * a.gpitch; };
    // register window over three rows of G, loaded two rows ahead of use
    int2 c_up = __ldg(row_ptr(p0 - 1) + qc);
    const int2 *g0 = row_ptr(p0), *g1 = row_ptr(p0 + 1);
    int2 c_lf = __ldg(g0 + ql), c_cur = __ldg(g0 + qc), c_rt = __ldg(g0 + qr);
    int2 n_lf = __ldg(g1 + ql), n_cur = __ldg(g1 + qc), n_rt = __ldg(g1 + qr);
#pragma unroll 2
    for (int rr = 0; rr < rows; ++rr) {
        const int p = p0 + rr;
        const int2 *g2 = row_ptr(p + 2);
        const int2 m_lf = __ldg(g2 + ql), m_cur = __ldg(g2 + qc), m_rt = __ldg(g2 + qr);
        const int2 c_dn = n_cur;
        const int4 Hp = sH[rr];
        if ((Hp.w >> 30) & 1) {  // p is a customer (warp-uniform)
            const bool other = (Hp.w & 0x3FFFFFFF) != q_route;
            const int d_rel = Hp.x + c_cur.x + c_rt.x - Hq.z;
            const int d_swap = c_up.x + c_dn.x - Hp.y + c_lf.x + c_rt.x - Hq.y;
            const bool v_rel = rel_q && other, v_swap = swap_q && other && p < q;
            if (DUMP) {
                if (v_rel && a.dump_family == 1) {
                    a.dump_delta[static_cast<size_t>(p) * P + q] = d_rel;
                    a.dump_feas[static_cast<size_t>(p) * P + q] =
                        check_relocate(sPB[rr], sPC[rr], sQA[tid], sQD[tid], c_cur, c_rt) ? 1 : 0;
                }
                if (v_swap && a.dump_family == 2) {
                    a.dump_delta[static_cast<size_t>(p) * P + q] = d_swap;
                    a.dump_feas[static_cast<size_t>(p) * P + q] =
                        check_swap(sPA[rr], sPB[rr], sPC[rr], sPD[rr], sQA[tid], sQB[tid], sQC[tid], sQD[tid], c_up, c_dn,
                                   c_lf, c_rt) ? 1 : 0;
                }
            } else {
                if (v_rel && d_rel <= bound && check_relocate(sPB[rr], sPC[rr], sQA[tid], sQD[tid], c_cur, c_rt)) {
                    const unsigned long long k = pack_key(d_rel, 1, p, q);
                    if (k < best) {
                        best = k;
                        bound = d_rel;
                    }
                }
                if (v_swap && d_swap <= bound &&
                    check_swap(sPA[rr], sPB[rr], sPC[rr], sPD[rr], sQA[tid], sQB[tid], sQC[tid], sQD[tid], c_up, c_dn, c_lf,
                               c_rt)) {
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
        c_lf = n_lf;
        c_cur = n_cur;
        c_rt = n_rt;
        n_lf = m_lf;
        n_cur = m_cur;
        n_rt = m_rt;
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
        const int2 *G = a.G + static_cast<size_t>(b) * a.grows * a.gpitch;
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
        // rows t-1, t, t+1 of the route-order matrix; the sweep reads them at consecutive columns x+1
        const int2 *row_m = G + static_cast<size_t>(max(t - 1, 0)) * a.gpitch;
        const int2 *row_0 = G + static_cast<size_t>(t) * a.gpitch;
        const int2 *row_p = G + static_cast<size_t>(t + 1) * a.gpitch;
        const bool f2 = a.families & 1, f_rel = a.families & 2, f_swap = a.families & 4;

        // cells of rows node[t-1], node[t], node[t+1] at column node[x]: start at x = t+1
        int2 cm = SKt, c0 = Et, cp = make_int2(0, 0);
        int2 c0_prev = make_int2(0, 0);  // c0 at x-1
        Seg F = Seg{0, 0, 0, true}, Rv = Seg{0, 0, 0, true};
        int svc_xm1 = 0, Et_xm1 = 0;  // service[node x-1], T[node x-1][node x]

        for (int x = t + 1; x < P; ++x) {
            const int nx = node[x];
            if (nx == 0) break;
            const int4 xA = RA[x], xB = RB[x], xC = RC[x], xD = RD[x];
            const int2 Ex = make_int2(xA.z, xD.z);  // cell(node x, node x+1)
            const int zn = xA.y;                    // Z[x+1]
            // next column's cells (rows fixed per thread): the only scattered loads of the step
            const int2 cm1 = __ldg(row_m + x + 1);
            const int2 c01 = __ldg(row_0 + x + 1);
            const int2 cp1 = __ldg(row_p + x + 1);
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
    a.G = static_cast<const int2 *>(p->Gbuf.p);
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
        dim3 grid((pmax + kInterThreads - 1) / kInterThreads, (shard_rows + kStripRows - 1) / kStripRows, p->batch);
        eval_inter_kernel<false><<<grid, kInterThreads, 0, st>>>(a);
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
        dim3 grid((pmax + kInterThreads - 1) / kInterThreads, (pmax + kStripRows - 1) / kStripRows, 1);
        eval_inter_kernel<true><<<grid, kInterThreads, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
    }
    dim3 grid((pmax + kIntraThreads - 1) / kIntraThreads, 1);
    eval_intra_kernel<true><<<grid, kIntraThreads, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

}  // namespace rt
