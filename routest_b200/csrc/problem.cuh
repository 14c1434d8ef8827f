// problem.cuh — a batch of device-resident VRPTW instances: coordinates, stop data, the interleaved {dist,time}
// matrix, the current solution in position space (SPEC.md §6) and everything derived from it.
#pragma once

#include "common.cuh"

namespace rt {

constexpr unsigned long long kNoMove = 0xFFFFFFFFFFFFFFFFull;
constexpr int kMaxPositions = 32768;  // key layout: 15 bits per position (SPEC.md §7)

// Per-position records written by the state refresh and read by the move kernels (all int4 => one LDG.128 each).
//   RA = {dep = B[p]+service[node p], zn = Z[p+1], E.d = D[node p][node p+1], route[p]}      "edge after p"
//   RB = {ready, due, service, demand} of node p
//   RC = {dep[p-1], rem_delta[p], oe[p] = D[node p-1][node p]+D[node p][node p+1], flags}    flags: 1 customer, 2 rem_ok
//   RD = {cap - rload[route p], min(max_dist - rdist[route p], INT_MAX), E.t, SK.t}
//   H  = {rem_delta[p], oe[p], E.d, route[p] | customer << 30}   the "hot" record: all a cost delta needs
struct StateView {
    int n, pitch, pstride, rstride;
    const int2 *cells;   // [n][pitch]
    const int4 *stops;   // [n]
    int *node;           // [pstride]
    int4 *RA, *RB, *RC, *RD;
    int2 *E, *SK;        // E[p] = cell(node p, node p+1), SK[p] = cell(node p-1, node p+1)
    int *Bv, *Zv, *route;  // [pstride] plain copies for the getters
    int *rstart;         // [rstride+1] separator positions
    long long *rload, *rdist;  // [rstride]
};

}  // namespace rt

struct rt_problem {
    rt_engine *e = nullptr;
    int n = 0, batch = 0, pitch = 0, pstride = 0, rstride = 0;
    bool has_coords = false, has_stops = false, has_matrix = false, has_routes = false;
    double dist_scale = 0, time_scale = 0;
    // device arrays, all [batch][...]
    double *lat = nullptr, *lon = nullptr, *xyz = nullptr;
    int2 *cells = nullptr;
    int4 *stops = nullptr;
    int *cap = nullptr;
    long long *max_dist = nullptr;
    int *npos = nullptr, *nroutes = nullptr;
    int *node = nullptr, *node_tmp = nullptr;
    int4 *RA = nullptr, *RB = nullptr, *RC = nullptr, *RD = nullptr, *H = nullptr;
    int2 *E = nullptr, *SK = nullptr;
    double *PX = nullptr;                  // [batch][3][pstride] unit vectors in position order
    rt::DevBuf GTbuf;                      // its transpose (only for asymmetric caller-supplied matrices)
    int2 *ER = nullptr;                    // [batch][pstride] cell(node p+1, node p) (asymmetric only)
    int supplied = 0;                      // matrices installed with rt_problem_set_matrix since the last build
    bool symmetric = true;                 // the stored matrix is symmetric (always true for SPEC §2 matrices)
    rt::DevBuf Gbuf;                       // position-space matrix G[batch][grows][gpitch], G[p][q] = M[node p][node q]
    int grows = 0, gpitch = 0;             // grows = pmax at the time G was sized
    int pmax = 0, pmin = 0;                // max / min npos over the batch (host copies; size grids and shards)
    int shard_index = 0, shard_count = 1;  // which part of the candidate space this replica evaluates (multi-GPU)
    bool best_clean = false;               // p->best holds "no move" for every instance (set by the state refresh)
    unsigned long long *ext_best = nullptr;  // caller's key buffer of the plan call in progress: the refresh resets it too
    bool ext_best_clean = false;           //   ... and it holds "none" in signed-order form right now
    bool scan_done = false;                // the vector kernel just enqueued also scanned the routes (large instances)
    bool cells_from_coords = false;        // the matrix is SPEC §2 of this problem's coordinates
    double angle2_bound = -1.0;            // dlat^2 + dlon^2 (radians) over the bounding box of all coordinates; < 0 = unknown
    int *Bv = nullptr, *Zv = nullptr, *route = nullptr, *rstart = nullptr;
    long long *rload = nullptr, *rdist = nullptr;
    long long *cost = nullptr;             // [batch]
    int *feasible = nullptr;               // [batch]
    unsigned long long *best = nullptr;    // [batch] packed keys (SPEC §7)
    int *done = nullptr, *moves = nullptr; // [batch] local-search bookkeeping
    bool dist_unlimited = false;           // every instance's max_dist as last uploaded from the host is >= 2^61
    unsigned *ticket = nullptr;            // arrival counter of the one-launch search (search_persistent_kernel)
    int *coord_flag = nullptr;             // [1] set to plan_seq by the vector kernel on a bad coordinate (rt_problem_plan)
    int plan_seq = 0;                      // rt_problem_plan calls so far: the flag needs no reset between calls
    int plan_flag_seq = 0;                 // the number the pass in flight writes into the flag
    bool plan_pending = false;             // a pass has been enqueued and not collected (rt_problem_plan_begin/_end)
    cudaEvent_t ev_in[2] = {nullptr, nullptr};  // inputs of the pass in flight are on the device (matrix part / the rest)
    cudaEvent_t ev_done = nullptr;         // its keys are in the staging buffer
    double plan_t[4] = {0, 0, 0, 0};       // host timestamps of the pass (ROUTEST_DEBUG_PLAN)
    unsigned long long *hist = nullptr;    // [batch][hist_cap] applied keys
    int hist_cap = 0;
    // greedy scratch
    unsigned long long *gkeys = nullptr;   // [batch][npow2]
    int *gorder = nullptr;                 // [batch][n]
    double *gdemand = nullptr;             // [batch][n]
    double *gcap = nullptr, *gmaxd = nullptr;  // [batch]
    int *g_order_out = nullptr, *g_trip_off = nullptr, *g_ntrips = nullptr, *g_status = nullptr;
    void *host_stage = nullptr;            // pinned staging for small D2H results
    size_t host_stage_bytes = 0;
    // incremental local search (cache.cu)
    rt::DevBuf cache_inter, cache_intra, cache_tiles, cache_misc;
    int cache_r = 0, cache_tile_cap = 0;
    rt::DevBuf batch_buf;                  // batched iterations (batch.cu): route minima, selection, claims
    cudaStream_t side_stream = nullptr;    // state tables run here underneath the matrix kernel (state.cu)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    // peer exchange (exchange.cu)
    unsigned long long *xbuf = nullptr;        // this replica's slot buffer (IPC-exported)
    unsigned long long **xpeers = nullptr;     // device table of every replica's slot buffer
    void *xpeer_mapped[64] = {nullptr};
    int n_xpeer_mapped = 0, xrank = 0, xworld = 0, xiter = 0;
    // peer exchange of the batched search's table (exchange.cu)
    unsigned long long *xtab = nullptr;        // this replica's buffer: control words + [2][world][entries] keys
    size_t xtab_words = 0;
    unsigned long long **xtab_peers = nullptr; // device table of every replica's buffer (+ a block counter behind it)
    void *xtab_mapped[64] = {nullptr};
    int n_xtab_mapped = 0, xtab_rank = 0, xtab_world = 0, xtab_iter = 0;
    void *all_allocs[64] = {nullptr};
    int n_allocs = 0;
};

namespace rt {

// Row range [lo, hi) of shard idx of cnt over P positions.  Rows are not equally expensive, and a pass is as slow as
// its slowest rank, so the boundaries split the cumulative WORK evenly, not the rows.  With cost per row 1 - beta x
// (x = p / P) the cumulative work is x - (beta / 2) x^2:
//   beta = 1    the tile kernel without a matrix scores only pairs p < q: row p costs P - p columns (a triangle);
//   beta = 0.3  the tile kernel on a stored matrix scores a relocate against all P columns and a swap against the
//               columns to the right of p: 82 against 58 instructions per pair (rank 0 of eight took 68 us where rank 4
//               took 61 with equal row counts);
//   beta = 0    equal row counts (same-route candidates: sharded by anchor).
// Plain IEEE double arithmetic with the correctly rounded sqrt, one expression on host and device: every rank computes
// the same boundaries.
__host__ __device__ inline int shard_boundary(int P, int k, int cnt, double beta)
{
    if (k <= 0) return 0;
    if (k >= cnt) return P;
    const double t = static_cast<double>(k) / static_cast<double>(cnt);
    const double x = beta > 0.0 ? (1.0 - sqrt(1.0 - 2.0 * beta * t * (1.0 - 0.5 * beta))) / beta : t;
    int b = static_cast<int>(x * static_cast<double>(P));
    return b < 0 ? 0 : (b > P ? P : b);
}
__host__ __device__ inline void shard_bounds(int P, int idx, int cnt, double beta, int &lo, int &hi)
{
    lo = shard_boundary(P, idx, cnt, beta);
    hi = shard_boundary(P, idx + 1, cnt, beta);
}
constexpr double kBetaTriangle = 1.0, kBetaStoredMatrix = 0.3, kBetaUniform = 0.0;

// rt_debug_set("fused_eval", v): 1 (default) = the move kernels rebuild matrix cells from the position-ordered unit
// vectors wherever the matrix is SPEC 2 of the problem's own coordinates, so no route-order matrix is written or read;
// 0 = always materialise G and read it (the write-then-read form, kept for measurement and for supplied matrices).
extern int g_fused_eval, g_state_fused_max, g_apply_prep;
inline bool eval_without_matrix(const rt_problem *p) { return g_fused_eval != 0 && p->cells_from_coords && p->symmetric; }

// first cell of G (the buffer starts with 16 pad cells)
inline int2 *position_matrix(const rt_problem *p) { return p->Gbuf.p ? static_cast<int2 *>(p->Gbuf.p) + 16 : nullptr; }
inline int2 *position_matrix_t(const rt_problem *p) { return p->GTbuf.p ? static_cast<int2 *>(p->GTbuf.p) + 16 : nullptr; }

int state_refresh(rt_problem *p, cudaStream_t st, cudaEvent_t tables_wait_for = nullptr, bool vectors_ready = false);
// unit vectors of the current coordinates straight into route order (and node order): replaces the unit-vector kernel
// plus the position gather on the plan path
int position_unit_vectors_launch(rt_problem *p, cudaStream_t st, int *bad_flag = nullptr, int bad_value = 1);
int position_matrix_refresh(rt_problem *p, cudaStream_t st, bool vectors_ready = false);
int state_tables_refresh(rt_problem *p, cudaStream_t st, bool with_vectors = false);
// up to `iters` iterations of the one-move search in one cooperative launch; -1 = not applicable here (moves.cu)
int search_persistent_launch(rt_problem *p, int families, int iters, bool improving_only, cudaStream_t st);
int eval_moves_launch(rt_problem *p, int families, int shard_index, int shard_count, unsigned long long *best_d,
                      bool signed_keys, cudaStream_t st);
int eval_dump_launch(rt_problem *p, int instance, int family, int *delta_d, unsigned char *feas_d, cudaStream_t st);
int apply_moves_launch(rt_problem *p, const unsigned long long *keys_d, bool signed_keys, bool improving_only,
                       cudaStream_t st);
int greedy_launch(rt_problem *p, bool use_tw, cudaStream_t st);
int position_vectors_refresh(rt_problem *p, cudaStream_t st);
int apply_only_launch(rt_problem *p, const unsigned long long *keys_d, bool signed_keys, bool improving_only,
                      cudaStream_t st, int *dirty_d = nullptr);
// incremental local search (cache.cu)
bool cache_usable(const rt_problem *p);
int cache_begin(rt_problem *p, cudaStream_t st);
int cache_evaluate(rt_problem *p, int families, cudaStream_t st);
int cache_apply(rt_problem *p, bool improving_only, cudaStream_t st);
int cache_ensure(rt_problem *p);
// batched iterations (batch.cu, SPEC.md §8.1)
bool batch_usable(const rt_problem *p);
int batch_begin(rt_problem *p, cudaStream_t st);
int batch_iteration(rt_problem *p, int families, cudaStream_t st);
int batch_ensure(rt_problem *p);
size_t batch_table_entries(const rt_problem *p);
int batch_select_apply(rt_problem *p, const unsigned long long *table, bool signed_keys, cudaStream_t st);
int eval_table_launch(rt_problem *p, int families, int shard_index, int shard_count, unsigned long long *table_d,
                      bool signed_keys, cudaStream_t st);
int ensure_position_matrix(rt_problem *p, bool need_buffer = false);  // sizes G for p->pmax (may allocate; not stream-ordered)

}  // namespace rt
