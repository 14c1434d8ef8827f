/*
 * routest_moves.c — CPU oracle for route state, move evaluation and local search (SPEC.md §6-§8).
 * TEST INFRASTRUCTURE ONLY (see routest_oracle.c for the rules on who may load this).
 *
 * The reference has no time windows and no improvement phase (utils.py:111-139 is its only solver
 * loop), so nothing here restates reference code: parity for these parts is UNPINNED IN THE REFERENCE.  What pins
 * them instead is internal: every candidate is evaluated twice, by brute force (apply the move to a copy, recompute
 * §6 from scratch) and by the O(1)/sweep algebra the engine uses, and the two must agree (tests/test_moves_oracle.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int n;                 /* nodes incl. depot */
    const int32_t *D, *T;  /* dense n x n */
    const int32_t *stops;  /* n x {ready,due,service,demand} */
    int32_t cap;
    int64_t max_dist;
} or_inst;

#define RDY(I, x) ((I)->stops[4 * (x)])
#define DUE(I, x) ((I)->stops[4 * (x) + 1])
#define SVC(I, x) ((I)->stops[4 * (x) + 2])
#define DEM(I, x) ((I)->stops[4 * (x) + 3])
#define DD(I, a, b) ((I)->D[(size_t)(a) * (I)->n + (b)])
#define TT(I, a, b) ((I)->T[(size_t)(a) * (I)->n + (b)])

#define OR_NO_MOVE 0xFFFFFFFFFFFFFFFFull

static inline uint64_t or_key(int32_t delta, int family, int p, int q)
{
    return ((uint64_t)(uint32_t)(delta + (int32_t)0x80000000) << 32) | ((uint64_t)family << 30) | ((uint64_t)p << 15) |
           (uint64_t)q;
}

/* ------------------------------------------------------------------ §6 route state */

/* order/trip_off as produced by the greedy: customers in visiting order, trip k = order[trip_off[k]..trip_off[k+1]) */
int or_build_positions(const int32_t *order, const int32_t *trip_off, int n_trips, int32_t *node)
{
    int P = 0;
    node[P++] = 0;
    for (int k = 0; k < n_trips; ++k) {
        for (int i = trip_off[k]; i < trip_off[k + 1]; ++i) node[P++] = order[i];
        node[P++] = 0;
    }
    return P;
}

/*
 * Recomputes §6 from scratch.  route/B/Z have length P; rload/rdist have length R (= P - n).
 * Returns 1 if the state is feasible, 0 otherwise.  *cost_out = sum of rdist.
 */
int or_route_state(int n, const int32_t *D, const int32_t *T, const int32_t *stops, int32_t cap, int64_t max_dist,
                   const int32_t *node, int P, int32_t *route, int32_t *B, int32_t *Z, int64_t *rload, int64_t *rdist,
                   int64_t *cost_out)
{
    or_inst I = {n, D, T, stops, cap, max_dist};
    int feas = 1, r = -1;
    int64_t cost = 0;
    for (int p = 0; p < P; ++p) {
        if (node[p] == 0) {
            if (p > 0) { /* close route r */
                int64_t arr = (int64_t)B[p - 1] + SVC(&I, node[p - 1]) + TT(&I, node[p - 1], 0);
                if (arr > DUE(&I, 0)) feas = 0;
                rdist[r] += DD(&I, node[p - 1], 0);
                if (rload[r] > cap || rdist[r] > max_dist) feas = 0;
                cost += rdist[r];
            }
            ++r;
            route[p] = r;
            B[p] = RDY(&I, 0);
            if (p < P - 1) { rload[r] = 0; rdist[r] = 0; }
        } else {
            int x = node[p], pv = node[p - 1];
            route[p] = r;
            int64_t arr = (int64_t)B[p - 1] + SVC(&I, pv) + TT(&I, pv, x);
            B[p] = (int32_t)(arr > RDY(&I, x) ? arr : RDY(&I, x));
            if (B[p] > DUE(&I, x)) feas = 0;
            rload[r] += DEM(&I, x);
            rdist[r] += DD(&I, pv, x);
        }
    }
    for (int p = P - 1; p >= 0; --p) {
        if (node[p] == 0) Z[p] = DUE(&I, 0);
        else {
            int x = node[p];
            int64_t z = (int64_t)Z[p + 1] - TT(&I, x, node[p + 1]) - SVC(&I, x);
            Z[p] = (int32_t)(z < DUE(&I, x) ? z : DUE(&I, x));
        }
    }
    if (cost_out) *cost_out = cost;
    return feas;
}

/* ------------------------------------------------------------------ §7 validity and application */

static int or_valid(const int32_t *node, const int32_t *route, int P, int family, int p, int q)
{
    if (p < 0 || q < 0 || p >= P || q >= P) return 0;
    if (family == 1) return node[p] != 0 && q <= P - 2 && q != p && q != p - 1;
    if (family == 2) return p < q && node[p] != 0 && node[q] != 0;
    /* 2-opt */
    if (q < p + 2 || node[p + 1] == 0 || node[q] == 0) return 0;
    return route[p + 1] == route[q];
}

void or_apply_move(int32_t *node, int family, int p, int q)
{
    if (family == 1) {
        int32_t u = node[p];
        if (q > p) { memmove(node + p, node + p + 1, sizeof(int32_t) * (size_t)(q - p)); node[q] = u; }
        else { memmove(node + q + 2, node + q + 1, sizeof(int32_t) * (size_t)(p - q - 1)); node[q + 1] = u; }
    } else if (family == 2) {
        int32_t t = node[p]; node[p] = node[q]; node[q] = t;
    } else {
        for (int i = p + 1, j = q; i < j; ++i, --j) { int32_t t = node[i]; node[i] = node[j]; node[j] = t; }
    }
}

/* ------------------------------------------------------------------ §7 brute force: apply, recompute, compare */

typedef struct { int32_t *node, *route, *B, *Z; int64_t *rload, *rdist; } or_scratch;

static or_scratch or_scratch_new(int P)
{
    or_scratch s;
    s.node = (int32_t *)malloc(sizeof(int32_t) * (size_t)P);
    s.route = (int32_t *)malloc(sizeof(int32_t) * (size_t)P);
    s.B = (int32_t *)malloc(sizeof(int32_t) * (size_t)P);
    s.Z = (int32_t *)malloc(sizeof(int32_t) * (size_t)P);
    s.rload = (int64_t *)malloc(sizeof(int64_t) * (size_t)P);
    s.rdist = (int64_t *)malloc(sizeof(int64_t) * (size_t)P);
    return s;
}
static void or_scratch_free(or_scratch *s)
{
    free(s->node); free(s->route); free(s->B); free(s->Z); free(s->rload); free(s->rdist);
}

/*
 * For one family: delta[p*P+q] and feas[p*P+q] (0 infeasible, 1 feasible, 2 not a valid candidate) for every (p,q).
 */
void or_eval_bruteforce(int n, const int32_t *D, const int32_t *T, const int32_t *stops, int32_t cap, int64_t max_dist,
                        const int32_t *node, int P, int family, int32_t *delta, uint8_t *feas)
{
    or_scratch base = or_scratch_new(P);
    int64_t cost0;
    or_route_state(n, D, T, stops, cap, max_dist, node, P, base.route, base.B, base.Z, base.rload, base.rdist, &cost0);
#pragma omp parallel
    {
        or_scratch s = or_scratch_new(P);
#pragma omp for schedule(dynamic, 4)
        for (int p = 0; p < P; ++p)
            for (int q = 0; q < P; ++q) {
                size_t o = (size_t)p * P + q;
                if (!or_valid(node, base.route, P, family, p, q)) { delta[o] = 0; feas[o] = 2; continue; }
                memcpy(s.node, node, sizeof(int32_t) * (size_t)P);
                or_apply_move(s.node, family, p, q);
                int64_t cost1;
                int ok = or_route_state(n, D, T, stops, cap, max_dist, s.node, P, s.route, s.B, s.Z, s.rload, s.rdist, &cost1);
                delta[o] = (int32_t)(cost1 - cost0);
                feas[o] = (uint8_t)(ok ? 1 : 0);
            }
        or_scratch_free(&s);
    }
    or_scratch_free(&base);
}

/* ------------------------------------------------------------------ §7 the O(1)/sweep algebra (what the engine does) */

typedef struct { int64_t a, b, L; int ok; } or_seg; /* service start at last node = max(t + a, b); admissible iff ok && t <= L */

static inline or_seg seg_single(const or_inst *I, int x)
{
    or_seg s = {0, RDY(I, x), DUE(I, x), 1};
    return s;
}
/* S1 (last node l1) followed by S2 (first node f2) */
static inline or_seg seg_concat(const or_inst *I, or_seg s1, int l1, or_seg s2, int f2)
{
    int64_t w = (int64_t)SVC(I, l1) + TT(I, l1, f2);
    or_seg s;
    s.a = s1.a + w + s2.a;
    s.b = s1.b + w + s2.a > s2.b ? s1.b + w + s2.a : s2.b;
    s.L = s1.L < s2.L - s1.a - w ? s1.L : s2.L - s1.a - w;
    s.ok = s1.ok && s2.ok && (s1.b + w <= s2.L);
    return s;
}
/* prefix ending at position i, then segment M (first node f, last node l), then suffix starting at position j */
static inline int seg_fits(const or_inst *I, const int32_t *node, const int32_t *B, const int32_t *Z, int i, or_seg M,
                           int f, int l, int j)
{
    int64_t t = (int64_t)B[i] + SVC(I, node[i]) + TT(I, node[i], f);
    if (!M.ok || t > M.L) return 0;
    int64_t out = t + M.a > M.b ? t + M.a : M.b;
    return out + SVC(I, l) + TT(I, l, node[j]) <= Z[j];
}

typedef struct {
    const or_inst *I;
    const int32_t *node, *route, *B, *Z;
    const int64_t *rload, *rdist;
    int P;
} or_view;

/* moves between two different routes: O(1) */
static int eval_relocate_inter(const or_view *V, int p, int q, int32_t *delta)
{
    const or_inst *I = V->I;
    const int32_t *nd = V->node;
    int u = nd[p], pv = nd[p - 1], nx = nd[p + 1], a = nd[q], b = nd[q + 1];
    int r1 = V->route[p], r2 = V->route[q];
    int32_t d_rem = DD(I, pv, nx) - DD(I, pv, u) - DD(I, u, nx);
    int32_t d_ins = DD(I, a, u) + DD(I, u, b) - DD(I, a, b);
    *delta = d_rem + d_ins;
    if (V->rload[r2] + DEM(I, u) > I->cap) return 0;
    if (V->rdist[r2] + d_ins > I->max_dist || V->rdist[r1] + d_rem > I->max_dist) return 0;
    int64_t tu = (int64_t)V->B[q] + SVC(I, a) + TT(I, a, u);
    if (tu < RDY(I, u)) tu = RDY(I, u);
    if (tu > DUE(I, u)) return 0;
    if (tu + SVC(I, u) + TT(I, u, b) > V->Z[q + 1]) return 0;
    if ((int64_t)V->B[p - 1] + SVC(I, pv) + TT(I, pv, nx) > V->Z[p + 1]) return 0;
    return 1;
}

static int eval_swap_inter(const or_view *V, int p, int q, int32_t *delta)
{
    const or_inst *I = V->I;
    const int32_t *nd = V->node;
    int u = nd[p], v = nd[q], pu = nd[p - 1], nu = nd[p + 1], pw = nd[q - 1], nw = nd[q + 1];
    int r1 = V->route[p], r2 = V->route[q];
    int32_t d1 = DD(I, pu, v) + DD(I, v, nu) - DD(I, pu, u) - DD(I, u, nu);
    int32_t d2 = DD(I, pw, u) + DD(I, u, nw) - DD(I, pw, v) - DD(I, v, nw);
    *delta = d1 + d2;
    if (V->rload[r1] - DEM(I, u) + DEM(I, v) > I->cap || V->rload[r2] - DEM(I, v) + DEM(I, u) > I->cap) return 0;
    if (V->rdist[r1] + d1 > I->max_dist || V->rdist[r2] + d2 > I->max_dist) return 0;
    int64_t tv = (int64_t)V->B[p - 1] + SVC(I, pu) + TT(I, pu, v);
    if (tv < RDY(I, v)) tv = RDY(I, v);
    if (tv > DUE(I, v) || tv + SVC(I, v) + TT(I, v, nu) > V->Z[p + 1]) return 0;
    int64_t tu = (int64_t)V->B[q - 1] + SVC(I, pw) + TT(I, pw, u);
    if (tu < RDY(I, u)) tu = RDY(I, u);
    if (tu > DUE(I, u) || tu + SVC(I, u) + TT(I, u, nw) > V->Z[q + 1]) return 0;
    return 1;
}

/*
 * Moves inside one route, anchored at position t: sweeps x = t+1.. to the end of t's route keeping
 *   F  = forward segment  node[t+1..x-1]   (grows by appending)
 *   Rv = reversed segment node[x..t+1]     (grows by prepending)
 * and reports, per x, up to four candidates through `emit`:
 *   2-opt(t,x), relocate(t -> after x), relocate(x -> after t), swap(t,x).
 */
typedef void (*or_emit)(void *ctx, int family, int p, int q, int32_t delta, int feas);

static void sweep_anchor(const or_view *V, int t, int families, or_emit emit, void *ctx)
{
    const or_inst *I = V->I;
    const int32_t *nd = V->node;
    const int P = V->P;
    if (t >= P - 1) return;
    const int r = V->route[t];
    const int nt = nd[t];
    const int t_is_cust = nt != 0;
    or_seg F = {0, 0, 0, 0}, Rv = {0, 0, 0, 0};
    int have_F = 0;
    int64_t asym = 0; /* sum over inner edges of (reverse - forward) distance, for 2-opt on asymmetric matrices */
    for (int x = t + 1; x < P && nd[x] != 0; ++x) {
        const int nx_ = nd[x], nxn = nd[x + 1];
        /* grow the reversed segment: prepend node x */
        if (x == t + 1) Rv = seg_single(I, nx_);
        else {
            Rv = seg_concat(I, seg_single(I, nx_), nx_, Rv, nd[x - 1]);
            asym += (int64_t)DD(I, nx_, nd[x - 1]) - DD(I, nd[x - 1], nx_);
        }
        const int64_t base = V->rdist[r];
        /* (A) 2-opt(t, x) */
        if ((families & 1) && x >= t + 2) {
            int32_t d = (int32_t)(DD(I, nt, nx_) + DD(I, nd[t + 1], nxn) - DD(I, nt, nd[t + 1]) - DD(I, nx_, nxn) + asym);
            int ok = base + d <= I->max_dist && seg_fits(I, nd, V->B, V->Z, t, Rv, nx_, nd[t + 1], x + 1);
            emit(ctx, 0, t, x, d, ok);
        }
        if (families & 2) {
            /* (B) relocate(t -> after x): prefix(..t-1) + F + [x] + [t] + suffix(x+1..) */
            if (t_is_cust) {
                or_seg M = have_F ? seg_concat(I, F, nd[x - 1], seg_single(I, nx_), nx_) : seg_single(I, nx_);
                M = seg_concat(I, M, nx_, seg_single(I, nt), nt);
                int32_t d = DD(I, nd[t - 1], nd[t + 1]) - DD(I, nd[t - 1], nt) - DD(I, nt, nd[t + 1]) + DD(I, nx_, nt) +
                            DD(I, nt, nxn) - DD(I, nx_, nxn);
                int ok = base + d <= I->max_dist && seg_fits(I, nd, V->B, V->Z, t - 1, M, nd[t + 1], nt, x + 1);
                emit(ctx, 1, t, x, d, ok);
            }
            /* (C) relocate(x -> after t): prefix(..t) + [x] + F + suffix(x+1..) */
            if (x >= t + 2) {
                or_seg M = seg_concat(I, seg_single(I, nx_), nx_, F, nd[t + 1]);
                int32_t d = DD(I, nd[x - 1], nxn) - DD(I, nd[x - 1], nx_) - DD(I, nx_, nxn) + DD(I, nt, nx_) +
                            DD(I, nx_, nd[t + 1]) - DD(I, nt, nd[t + 1]);
                int ok = base + d <= I->max_dist && seg_fits(I, nd, V->B, V->Z, t, M, nx_, nd[x - 1], x + 1);
                emit(ctx, 1, x, t, d, ok);
            }
        }
        /* (D) swap(t, x) */
        if ((families & 4) && t_is_cust) {
            or_seg M;
            int32_t d;
            if (x == t + 1) {
                M = seg_concat(I, seg_single(I, nx_), nx_, seg_single(I, nt), nt);
                d = DD(I, nd[t - 1], nx_) + DD(I, nx_, nt) + DD(I, nt, nxn) - DD(I, nd[t - 1], nt) - DD(I, nt, nx_) -
                    DD(I, nx_, nxn);
            } else {
                M = seg_concat(I, seg_single(I, nx_), nx_, F, nd[t + 1]);
                M = seg_concat(I, M, nd[x - 1], seg_single(I, nt), nt);
                d = DD(I, nd[t - 1], nx_) + DD(I, nx_, nd[t + 1]) - DD(I, nd[t - 1], nt) - DD(I, nt, nd[t + 1]) +
                    DD(I, nd[x - 1], nt) + DD(I, nt, nxn) - DD(I, nd[x - 1], nx_) - DD(I, nx_, nxn);
            }
            int ok = base + d <= I->max_dist && seg_fits(I, nd, V->B, V->Z, t - 1, M, nx_, nt, x + 1);
            emit(ctx, 2, t, x, d, ok);
        }
        /* grow the forward segment: append node x */
        if (have_F) F = seg_concat(I, F, nd[x - 1], seg_single(I, nx_), nx_);
        else { F = seg_single(I, nx_); have_F = 1; }
    }
}

/* ---- dump through the algebra (for the cross-check against brute force) */

typedef struct { int P; int family; int32_t *delta; uint8_t *feas; } dump_ctx;

static void emit_dump(void *c, int family, int p, int q, int32_t delta, int feas)
{
    dump_ctx *d = (dump_ctx *)c;
    if (family != d->family) return;
    size_t o = (size_t)p * d->P + q;
    d->delta[o] = delta;
    d->feas[o] = (uint8_t)(feas ? 1 : 0);
}

void or_eval_fast(int n, const int32_t *D, const int32_t *T, const int32_t *stops, int32_t cap, int64_t max_dist,
                  const int32_t *node, int P, int family, int32_t *delta, uint8_t *feas)
{
    or_inst I = {n, D, T, stops, cap, max_dist};
    or_scratch s = or_scratch_new(P);
    or_route_state(n, D, T, stops, cap, max_dist, node, P, s.route, s.B, s.Z, s.rload, s.rdist, NULL);
    or_view V = {&I, node, s.route, s.B, s.Z, s.rload, s.rdist, P};
    for (size_t o = 0; o < (size_t)P * P; ++o) { delta[o] = 0; feas[o] = 2; }
    dump_ctx ctx = {P, family, delta, feas};
    for (int t = 0; t < P; ++t) sweep_anchor(&V, t, 1 << family, emit_dump, &ctx);
    if (family != 0)
        for (int p = 0; p < P; ++p)
            for (int q = 0; q < P; ++q) {
                if (!or_valid(node, s.route, P, family, p, q)) continue;
                if (s.route[p] == s.route[q]) continue; /* same route: done by the sweep */
                int32_t d;
                int ok = family == 1 ? eval_relocate_inter(&V, p, q, &d) : eval_swap_inter(&V, p, q, &d);
                delta[(size_t)p * P + q] = d;
                feas[(size_t)p * P + q] = (uint8_t)ok;
            }
    or_scratch_free(&s);
}

/* ---- best move through the algebra, positions p in [p_begin, p_end) (the shard a rank owns) */

typedef struct { uint64_t best; int p_begin, p_end; } best_ctx;

static void emit_best(void *c, int family, int p, int q, int32_t delta, int feas)
{
    best_ctx *b = (best_ctx *)c;
    if (!feas || p < b->p_begin || p >= b->p_end) return;
    uint64_t k = or_key(delta, family, p, q);
    if (k < b->best) b->best = k;
}

/*
 * families: bit 0 = 2-opt, bit 1 = relocate, bit 2 = swap.  A candidate (family,p,q) belongs to the shard that
 * contains p.  evals_out (may be NULL) receives the number of valid candidates scored.
 */
uint64_t or_best_move(int n, const int32_t *D, const int32_t *T, const int32_t *stops, int32_t cap, int64_t max_dist,
                      const int32_t *node, int P, int families, int p_begin, int p_end, int threads, int64_t *evals_out)
{
    or_inst I = {n, D, T, stops, cap, max_dist};
    or_scratch s = or_scratch_new(P);
    or_route_state(n, D, T, stops, cap, max_dist, node, P, s.route, s.B, s.Z, s.rload, s.rdist, NULL);
    or_view V = {&I, node, s.route, s.B, s.Z, s.rload, s.rdist, P};
    uint64_t best = OR_NO_MOVE;
    int64_t evals = 0;
    (void)threads;
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
    {
        best_ctx ctx = {OR_NO_MOVE, p_begin, p_end};
        int64_t my = 0;
        /* same-route candidates: the sweep emits (p,q) with p on either side of the anchor, so every anchor is swept
           and emit_best filters on p */
#pragma omp for schedule(dynamic, 16) nowait
        for (int t = 0; t < P; ++t) sweep_anchor(&V, t, families, emit_best, &ctx);
#pragma omp for schedule(dynamic, 16) nowait
        for (int p = p_begin; p < p_end; ++p) {
            if (node[p] == 0) continue;
            for (int q = 0; q < P; ++q) {
                if ((families & 2) && s.route[p] != s.route[q] && or_valid(node, s.route, P, 1, p, q)) {
                    int32_t d;
                    ++my;
                    if (eval_relocate_inter(&V, p, q, &d)) { uint64_t k = or_key(d, 1, p, q); if (k < ctx.best) ctx.best = k; }
                }
                if ((families & 4) && q > p && node[q] != 0 && s.route[p] != s.route[q]) {
                    int32_t d;
                    ++my;
                    if (eval_swap_inter(&V, p, q, &d)) { uint64_t k = or_key(d, 2, p, q); if (k < ctx.best) ctx.best = k; }
                }
            }
        }
#pragma omp critical
        {
            if (ctx.best < best) best = ctx.best;
            evals += my;
        }
    }
    if (evals_out) *evals_out = evals;
    or_scratch_free(&s);
    return best;
}

/* Number of valid candidates per family for a state (the unit of the move-evals/s metric). */
void or_count_candidates(const int32_t *node, int P, int64_t counts[3])
{
    int32_t *route = (int32_t *)malloc(sizeof(int32_t) * (size_t)P);
    int r = -1;
    for (int p = 0; p < P; ++p) { if (node[p] == 0) ++r; route[p] = r; }
    counts[0] = counts[1] = counts[2] = 0;
    int64_t ncust = 0;
    for (int p = 0; p < P; ++p) ncust += node[p] != 0;
    /* relocate: every customer x every q in [0,P-2] except p and p-1 */
    counts[1] = ncust * (int64_t)(P - 3);
    counts[2] = ncust * (ncust - 1) / 2;
    for (int p = 0; p < P; ++p) {
        if (node[p] != 0) continue;
        int64_t L = 0;
        for (int x = p + 1; x < P && node[x] != 0; ++x) ++L;
        /* 2-opt(p,q): p in [start, start+L-2], q in [p+2, start+L]  =>  sum_{k=2..L} (L-k+1) = L(L-1)/2 */
        counts[0] += L * (L - 1) / 2;
    }
    free(route);
}

/* ------------------------------------------------------------------ §8 local search */

/*
 * Applies best moves until none improves or max_moves is reached.  keys_out (may be NULL) receives the applied keys.
 * Returns the number of moves applied; node[] holds the final state; *cost_out the final cost.
 */
int or_local_search(int n, const int32_t *D, const int32_t *T, const int32_t *stops, int32_t cap, int64_t max_dist,
                    int32_t *node, int P, int families, int max_moves, int threads, uint64_t *keys_out, int64_t *cost_out)
{
    int moves = 0;
    while (moves < max_moves) {
        uint64_t k = or_best_move(n, D, T, stops, cap, max_dist, node, P, families, 0, P, threads, NULL);
        if (k == OR_NO_MOVE) break;
        int32_t delta = (int32_t)((uint32_t)(k >> 32) - 0x80000000u);
        if (delta >= 0) break;
        or_apply_move(node, (int)((k >> 30) & 3), (int)((k >> 15) & 0x7FFF), (int)(k & 0x7FFF));
        if (keys_out) keys_out[moves] = k;
        ++moves;
    }
    if (cost_out) {
        or_scratch s = or_scratch_new(P);
        or_route_state(n, D, T, stops, cap, max_dist, node, P, s.route, s.B, s.Z, s.rload, s.rdist, cost_out);
        or_scratch_free(&s);
    }
    return moves;
}

/* ------------------------------------------------------------------ §8.1 batched iterations */

typedef struct { uint64_t *m; const int32_t *route; const uint8_t *taken; } rmin_ctx;

static inline void rmin_note(rmin_ctx *c, uint64_t k, int p, int q)
{
    const int ra = c->route[p], rb = c->route[q];
    if (c->taken[ra] || c->taken[rb]) return; /* a route already claimed by an earlier round of this batch */
    if (k < c->m[ra]) c->m[ra] = k;
    if (k < c->m[rb]) c->m[rb] = k;
}

static void emit_rmin(void *c, int family, int p, int q, int32_t delta, int feas)
{
    if (feas) rmin_note((rmin_ctx *)c, or_key(delta, family, p, q), p, q);
}

/*
 * One batch of SPEC §8.1.  Round 1: m[r] (r < R = P - n) = minimum key over the valid feasible candidates that touch
 * route r; every candidate with delta < 0 that is the minimum of all its routes is selected and its routes are
 * claimed.  Rounds 2..rounds repeat that over the candidates whose routes are all still unclaimed.  keys_out receives
 * the batch round after round, ascending route[p] within a round (at most R keys); m_out (may be NULL) the round-1
 * minima.  Returns the batch size.
 */
int or_select_batch(int n, const int32_t *D, const int32_t *T, const int32_t *stops, int32_t cap, int64_t max_dist,
                    const int32_t *node, int P, int families, int rounds, int threads, uint64_t *m_out, uint64_t *keys_out)
{
    or_inst I = {n, D, T, stops, cap, max_dist};
    or_scratch s = or_scratch_new(P);
    or_route_state(n, D, T, stops, cap, max_dist, node, P, s.route, s.B, s.Z, s.rload, s.rdist, NULL);
    or_view V = {&I, node, s.route, s.B, s.Z, s.rload, s.rdist, P};
    const int R = P - n;
    uint64_t *m = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(R + 1)); /* +1: the closing separator's pseudo route */
    uint8_t *taken = (uint8_t *)calloc((size_t)(R + 1), 1);
    int count = 0;
    (void)threads;
    for (int round = 0; round < rounds; ++round) {
        for (int r = 0; r <= R; ++r) m[r] = OR_NO_MOVE;
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
        {
            uint64_t *mine = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(R + 1));
            for (int r = 0; r <= R; ++r) mine[r] = OR_NO_MOVE;
            rmin_ctx ctx = {mine, s.route, taken};
#pragma omp for schedule(dynamic, 16) nowait
            for (int t = 0; t < P; ++t) sweep_anchor(&V, t, families, emit_rmin, &ctx);
#pragma omp for schedule(dynamic, 16) nowait
            for (int p = 0; p < P; ++p) {
                if (node[p] == 0) continue;
                for (int q = 0; q < P; ++q) {
                    int32_t d;
                    if ((families & 2) && s.route[p] != s.route[q] && or_valid(node, s.route, P, 1, p, q) &&
                        eval_relocate_inter(&V, p, q, &d))
                        rmin_note(&ctx, or_key(d, 1, p, q), p, q);
                    if ((families & 4) && q > p && node[q] != 0 && s.route[p] != s.route[q] && eval_swap_inter(&V, p, q, &d))
                        rmin_note(&ctx, or_key(d, 2, p, q), p, q);
                }
            }
#pragma omp critical
            for (int r = 0; r <= R; ++r)
                if (mine[r] < m[r]) m[r] = mine[r];
            free(mine);
        }
        if (round == 0 && m_out && R > 0) memcpy(m_out, m, sizeof(uint64_t) * (size_t)R);
        const int before = count;
        for (int r = 0; r < R; ++r) {
            const uint64_t k = m[r];
            if (k == OR_NO_MOVE || (int32_t)((uint32_t)(k >> 32) - 0x80000000u) >= 0) continue;
            const int p = (int)((k >> 15) & 0x7FFF), q = (int)(k & 0x7FFF);
            if (s.route[p] != r) continue; /* reported once, at the route of p */
            if (m[s.route[q]] != k) continue;
            keys_out[count++] = k;
        }
        if (count == before) break;
        for (int i = before; i < count; ++i) {
            taken[s.route[(keys_out[i] >> 15) & 0x7FFF]] = 1;
            taken[s.route[keys_out[i] & 0x7FFF]] = 1;
        }
    }
    free(m);
    free(taken);
    or_scratch_free(&s);
    return count;
}

/* Applies a batch selected on the state node[] currently holds (disjoint routes), in the order given. */
void or_apply_batch(int n, int32_t *node, int P, const uint64_t *keys, int count)
{
    const int R = P - n;
    int32_t *route = (int32_t *)malloc(sizeof(int32_t) * (size_t)P);
    int32_t *start = (int32_t *)malloc(sizeof(int32_t) * (size_t)(R + 2));
    int32_t *rel = (int32_t *)malloc(sizeof(int32_t) * 4 * (size_t)(count > 0 ? count : 1));
    int r = -1;
    for (int p = 0; p < P; ++p) { if (node[p] == 0) { ++r; start[r] = p; } route[p] = r; }
    for (int i = 0; i < count; ++i) { /* route-relative form on the state the batch was selected on */
        const int p = (int)((keys[i] >> 15) & 0x7FFF), q = (int)(keys[i] & 0x7FFF);
        rel[4 * i] = route[p]; rel[4 * i + 1] = p - start[route[p]];
        rel[4 * i + 2] = route[q]; rel[4 * i + 3] = q - start[route[q]];
    }
    for (int i = 0; i < count; ++i) {
        r = -1;
        for (int p = 0; p < P; ++p) if (node[p] == 0) start[++r] = p;
        or_apply_move(node, (int)((keys[i] >> 30) & 3), start[rel[4 * i]] + rel[4 * i + 1], start[rel[4 * i + 2]] + rel[4 * i + 3]);
    }
    free(route); free(start); free(rel);
}

/*
 * §8.1 to a local optimum or max_iters batches.  keys_out (capacity keys_cap, may be NULL) receives the applied keys
 * batch after batch; batch_sizes_out (capacity max_iters or until done, may be NULL) the size of every batch.
 * Returns the number of moves applied; *iters_out the number of non-empty batches.
 */
int or_local_search_batched(int n, const int32_t *D, const int32_t *T, const int32_t *stops, int32_t cap, int64_t max_dist,
                            int32_t *node, int P, int families, int rounds, int max_iters, int threads, uint64_t *keys_out,
                            int keys_cap, int32_t *batch_sizes_out, int sizes_cap, int32_t *iters_out, int64_t *cost_out)
{
    const int R = P - n;
    uint64_t *batch = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(R > 0 ? R : 1));
    int moves = 0, iters = 0;
    while (iters < max_iters) {
        const int c = or_select_batch(n, D, T, stops, cap, max_dist, node, P, families, rounds, threads, NULL, batch);
        if (c == 0) break;
        or_apply_batch(n, node, P, batch, c);
        for (int i = 0; i < c; ++i)
            if (keys_out && moves + i < keys_cap) keys_out[moves + i] = batch[i];
        if (batch_sizes_out && iters < sizes_cap) batch_sizes_out[iters] = c;
        moves += c;
        ++iters;
    }
    free(batch);
    if (iters_out) *iters_out = iters;
    if (cost_out) {
        or_scratch s = or_scratch_new(P);
        or_route_state(n, D, T, stops, cap, max_dist, node, P, s.route, s.B, s.Z, s.rload, s.rdist, cost_out);
        or_scratch_free(&s);
    }
    return moves;
}
