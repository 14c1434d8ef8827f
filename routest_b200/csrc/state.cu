// state.cu — route state in position space (SPEC.md §6): the refresh that derives everything the move kernels read,
// and the kernel that applies a chosen move (SPEC.md §8).
//
// No reference counterpart: the reference's only solver loop is the greedy constructor (utils.py:111-139).
//
// The refresh first lays the stop matrix out in ROUTE ORDER: G[p][q] = M[node p][node q] for positions p, q of
// the current solution.  When the matrix is the great-circle matrix of the problem's coordinates, G is produced by
// the matrix kernel itself running on the position-ordered unit vectors (same cell function, so the same bits, at
// the HBM write roofline); a caller-supplied matrix is permuted instead.  Every later access of the move kernels
// — a row of G per moved stop, consecutive columns per insertion point — is then contiguous.  (Where the move kernels
// rebuild their cells from the unit vectors — eval_without_matrix — no G is written at all.)
//
// The device code lives in state_device.cuh (shared with the one-launch search of moves.cu); this file holds the
// launches: state_refresh / state_tables_refresh, apply_moves_launch.
#define RT_STATE_KERNELS
#include "state_device.cuh"

namespace rt {

constexpr int kFusedStatePositions = 0;  // (the one-block state kernel: measured slower than scan + per-route warps from 400 stops up, equal below)
int g_state_fused_max = -1;  // rt_debug_set("state_fused_max", positions): -1 = the default below
int g_apply_prep = 1;        // rt_debug_set("apply_prep", 0/1): the merged apply + vectors + scan launch
static void launch_position_vectors(rt_problem *p, cudaStream_t st);

int ensure_position_matrix(rt_problem *p, bool need_buffer)
{
    const int rows = p->pmax > 0 ? p->pmax : p->pstride;
    const int pitch = (rows + 15) & ~15;
    p->grows = rows;
    p->gpitch = pitch;
    // Where the move kernels rebuild their cells from the unit vectors no buffer is needed until somebody asks for the
    // matrix itself (rt_problem_get_position_matrix, rt_problem_refresh_dev bit 2): 843 MB at 10,000 stops.
    if (!need_buffer && eval_without_matrix(p)) return RT_OK;
    // 16 pad cells in front and pad rows behind: the move kernels read q-1, q+1 and p+2 without clamping and
    // prefetch up to ten rows ahead
    if (!p->symmetric)
        RT_TRY(reserve(p->e, p->GTbuf, (static_cast<size_t>(p->batch) * rows * pitch + 16 + 12 * static_cast<size_t>(pitch)) * sizeof(int2)));
    return reserve(p->e, p->Gbuf, (static_cast<size_t>(p->batch) * rows * pitch + 16 + 12 * static_cast<size_t>(pitch)) * sizeof(int2));
}

// node[] -> G (route-order matrix), rows of this replica's shard only (plus the halo rows the move kernels read).
// Stream-ordered, no allocation (G is sized by ensure_position_matrix when routes are installed).
int position_matrix_refresh(rt_problem *p, cudaStream_t st, bool vectors_ready)
{
    rt_engine *e = p->e;
    if (p->grows <= 0) return fail(e, RT_ERR_STATE, "route state refresh before routes were installed");
    RT_TRY(ensure_position_matrix(p, /*need_buffer=*/true));  // a no-op once the buffer has its size
    int2 *G = position_matrix(p);
    // rows any instance of the batch can need: shard [P*i/c, P*(i+1)/c) over P in [pmin, pmax], one halo row each side
    int r0 = 0, r1 = p->grows;
    if (p->shard_count > 1) {
        const long long lo = static_cast<long long>(shard_boundary(p->pmin, p->shard_index, p->shard_count, kBetaStoredMatrix)) - 1;
        const long long hi = static_cast<long long>(shard_boundary(p->pmax, p->shard_index + 1, p->shard_count, kBetaStoredMatrix)) + 2;
        r0 = static_cast<int>(std::max<long long>(lo, 0));
        r1 = static_cast<int>(std::min<long long>(hi, p->grows));
    }
    if (p->cells_from_coords) {
        if (!vectors_ready) {
            launch_position_vectors(p, st);
            p->scan_done = false;  // the tables are not refreshed here: whoever does it next scans for itself
            RT_CUDA(e, cudaGetLastError());
        }
        RT_TRY(launch_matrix_cells_ld(e, p->PX, p->grows, p->pstride, p->batch, r0, r1, p->dist_scale, p->time_scale, G,
                                      p->gpitch, st));
    } else {
        dim3 gp((p->gpitch + 255) / 256, r1 - r0, p->batch);
        permute_matrix_kernel<<<gp, 256, 0, st>>>(p->cells, p->node, p->npos, p->n, p->pitch, p->pstride, p->grows,
                                                  p->gpitch, r0, 0, G);
        RT_CUDA(e, cudaGetLastError());
        if (!p->symmetric) {
            permute_matrix_kernel<<<gp, 256, 0, st>>>(p->cells, p->node, p->npos, p->n, p->pitch, p->pstride, p->grows,
                                                      p->gpitch, r0, 1, position_matrix_t(p));
            RT_CUDA(e, cudaGetLastError());
        }
    }
    return RT_OK;
}

// E/SK/RB, B/Z/load/dist, the packed records; resets the best keys.  Up to kFusedStatePositions positions per instance
// one block per instance does it all in a single launch.


// large instances: the vector kernel that precedes the tables carries the route scan as an extra block
static int state_fused_max() { return g_state_fused_max >= 0 ? g_state_fused_max : kFusedStatePositions; }
static bool scan_rides_along(const rt_problem *p) { return p->grows > state_fused_max(); }

static void launch_position_vectors(rt_problem *p, cudaStream_t st)
{
    const int ws = scan_rides_along(p) ? 1 : 0, th = ws ? kPrepThreads : 256;
    dim3 g0((p->pstride + th - 1) / th + ws, p->batch);
    position_vectors_kernel<<<g0, th, 0, st>>>(p->xyz, p->node, p->npos, p->n, p->pstride, p->PX, routes_args(p), ws);
    p->scan_done = ws != 0;
    p->best_clean = p->best_clean || ws != 0;
}


int state_tables_refresh(rt_problem *p, cudaStream_t st, bool with_vectors)
{
    rt_engine *e = p->e;
    const bool fused = with_vectors && p->cells_from_coords;  // also produce the position-ordered vectors
    const StateArgs a = state_args(p, fused);
    p->best_clean = true;  // the route scan resets the best keys
    p->ext_best_clean = p->ext_best != nullptr;
    if (!p->scan_done && p->grows <= state_fused_max()) {
        state_fused_kernel<<<p->batch, kScanThreads, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
        return RT_OK;
    }
    // route scan (unless the vector kernel in front carried it), then one warp per route: gather + walk + records
    if (!p->scan_done) {
        state_routes_kernel<<<p->batch, kScanThreads, 0, st>>>(a.r);
        RT_CUDA(e, cudaGetLastError());
    }
    p->scan_done = false;
    const int rmax = std::max(1, p->grows - p->n);  // most routes of any instance
    dim3 gw((rmax + kRouteWarps - 1) / kRouteWarps, p->batch);
    state_route_tables_kernel<<<gw, kRouteWarps * 32, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

// Both halves.  The state tables do not read G (the gather step rebuilds the cells it needs), so their four small
// latency-bound kernels run on a side stream underneath the matrix kernel and join before anything evaluates moves.
int position_unit_vectors_launch(rt_problem *p, cudaStream_t st, int *bad_flag, int bad_value)
{
    const int ws = scan_rides_along(p) ? 1 : 0, th = ws ? kPrepThreads : 256;
    dim3 g0((p->pstride + th - 1) / th + ws, p->batch);
    position_unit_vectors_kernel<<<g0, th, 0, st>>>(p->lat, p->lon, p->node, p->npos, p->n, p->pstride, p->xyz, p->PX,
                                                     bad_flag, bad_value, routes_args(p), ws);
    p->scan_done = ws != 0;
    RT_CUDA(p->e, cudaGetLastError());
    return RT_OK;
}

int state_refresh(rt_problem *p, cudaStream_t st, cudaEvent_t tables_wait_for, bool vectors_ready)
{
    rt_engine *e = p->e;
    if (!p->side_stream) {
        RT_CUDA(e, cudaStreamCreateWithFlags(&p->side_stream, cudaStreamNonBlocking));
        RT_CUDA(e, cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
        RT_CUDA(e, cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
    }
    if (p->cells_from_coords && !vectors_ready) {
        // the position-ordered vectors feed both halves: produce them first, then fork
        launch_position_vectors(p, st);
        RT_CUDA(e, cudaGetLastError());
    }
    if (eval_without_matrix(p)) {
        // no route-order matrix to rebuild: the kernels that follow read the unit vectors and the tables only
        if (tables_wait_for) RT_CUDA(e, cudaStreamWaitEvent(st, tables_wait_for, 0));
        return state_tables_refresh(p, st);
    }
    RT_CUDA(e, cudaEventRecord(p->ev_fork, st));
    RT_CUDA(e, cudaStreamWaitEvent(p->side_stream, p->ev_fork, 0));
    if (tables_wait_for) RT_CUDA(e, cudaStreamWaitEvent(p->side_stream, tables_wait_for, 0));  // e.g. stop records still in flight
    RT_TRY(position_matrix_refresh(p, st, /*vectors_ready=*/true));
    RT_TRY(state_tables_refresh(p, p->side_stream));
    RT_CUDA(e, cudaEventRecord(p->ev_join, p->side_stream));
    RT_CUDA(e, cudaStreamWaitEvent(st, p->ev_join, 0));
    return RT_OK;
}

int position_vectors_refresh(rt_problem *p, cudaStream_t st)
{
    launch_position_vectors(p, st);
    p->scan_done = false;
    RT_CUDA(p->e, cudaGetLastError());
    return RT_OK;
}


int apply_only_launch(rt_problem *p, const unsigned long long *keys_d, bool signed_keys, bool improving_only,
                      cudaStream_t st, int *dirty_d)
{
    apply_moves_kernel<<<p->batch, 256, 0, st>>>(apply_args(p, keys_d, signed_keys, improving_only, dirty_d));
    RT_CUDA(p->e, cudaGetLastError());
    return RT_OK;
}

// Up to this many positions one block per instance applies the move and scans the routes (apply_state_kernel); with
// at most one route per warp of that block it fills the tables too, otherwise the per-route kernel follows.
constexpr int kApplyStatePositions = 4096;

int apply_moves_launch(rt_problem *p, const unsigned long long *keys_d, bool signed_keys, bool improving_only,
                       cudaStream_t st)
{
    if (g_apply_prep && p->cells_from_coords && eval_without_matrix(p) && p->xyz && p->grows <= kApplyStatePositions) {
        const int rmax = std::max(1, (p->pmax > 0 ? p->pmax : p->grows) - p->n);  // most routes of any instance
        const bool with_tables = rmax <= kPrepThreads / 32;
        apply_state_kernel<<<p->batch, kPrepThreads, 0, st>>>(apply_args(p, keys_d, signed_keys, improving_only, nullptr),
                                                              state_args(p, /*fused_vectors=*/true), with_tables ? 1 : 0);
        RT_CUDA(p->e, cudaGetLastError());
        p->best_clean = true;  // the route scan resets the best keys
        p->ext_best_clean = p->ext_best != nullptr;
        p->scan_done = !with_tables;
        return with_tables ? RT_OK : state_tables_refresh(p, st, /*with_vectors=*/true);
    }
    RT_TRY(apply_only_launch(p, keys_d, signed_keys, improving_only, st));
    return state_refresh(p, st);
}

}  // namespace rt
