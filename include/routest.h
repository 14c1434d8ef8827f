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
 * lazily loaded model global, ml.py:4,11-21, whose unlocked initialisation this handle makes explicit).
 * Problems (rt_problem_create) point into their engine: destroy them before rt_destroy. */
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

/*
 * The same device-resident matrix as rt_unit_vectors_dev + rt_matrix_cells_dev (one instance, all rows), for callers
 * that cannot allocate: processes that gave all GPU memory to a pool at start-up, or that must keep cudaMalloc's
 * device-wide synchronisation off their request path, have no buffer to offer for the per-stop unit vectors.  Here
 * that scratch lives in a static device array inside the library (384 KiB per GPU, held until exit), so the call
 * allocates nothing; the caller passes only its inputs (lat_d, lon_d: n doubles each) and its output (cells_d:
 * n x rt_matrix_pitch(n) cells), all device pointers.
 *   - n may not exceed rt_matrix_static_max_nodes() (16384); larger sizes are refused with RT_ERR_INVALID_ARG;
 *   - the first call on a GPU looks the array's device address up once (cudaGetSymbolAddress): make that call
 *     before starting any stream capture;
 *   - every call on a GPU shares the one array, which keeps the previous call's contents (nothing needs clearing:
 *     each call rewrites the 3n entries it reads); calls that can overlap on the GPU race — keep them on one stream
 *     or order them with events.
 * This entry point has no counterpart in the reference (which neither builds a matrix nor manages device memory).
 */
int rt_matrix_static_max_nodes(void);
int rt_matrix_cells_static_dev(rt_engine *e, const double *lat_d, const double *lon_d, int n, double dist_scale,
                               double time_scale, void *cells_d, void *stream);

/* ------------------------------------------------------------------------------------------------------------
 * Greedy trip construction on a caller-supplied matrix — takes over utils.py:111-139 decision for decision (sort
 * by distance from the depot with ties by index, capacity/max-distance acceptance in FP64, a rejected stop does not
 * end the trip).  D is the n x n FP64 matrix in host memory, exactly what the reference holds after utils.py:103
 * (floats, possibly asymmetric); demand[0] is ignored; cap / max_dist as at utils.py:112-113.
 *   order_out    int32[n-1]  stops (matrix indices 1..n-1) in visiting order
 *   trip_off_out int32[n]    trip k is order_out[trip_off_out[k] .. trip_off_out[k+1])
 * Returns RT_ERR_INFEASIBLE_STOP (with the trips built so far) where the reference would never return.
 */
int rt_greedy_trips(rt_engine *e, const double *D, int n, const double *demand, double cap, double max_dist,
                    int32_t *order_out, int32_t *trip_off_out, int32_t *n_trips_out);

/* ------------------------------------------------------------------------------------------------------------
 * rt_problem: `batch` independent instances of n nodes each (node 0 = depot) kept on the GPU — coordinates, stop
 * data, the {dist,time} matrix, the current routes and everything derived from them (SPEC.md §6).  None of this
 * has a counterpart in the reference, whose only solver state is local to utils.py:_multi_stop; the calls are
 * ordered the way that function is: matrix (utils.py:94-109) -> trips (utils.py:111-139) -> [improvement, new].
 * All array arguments are host memory, instance-major: [batch][...].
 */
typedef struct rt_problem rt_problem;

int rt_problem_create(rt_engine *e, int n, int batch, rt_problem **out);
void rt_problem_destroy(rt_problem *p);
/* Length of the per-instance position arrays returned by rt_problem_get_state. */
int rt_problem_position_stride(const rt_problem *p);

/* lat/lon [batch][n] in degrees; computes the unit vectors (SPEC.md §1). */
int rt_problem_set_coords(rt_problem *p, const double *lat, const double *lon);
/* stops [batch][n][4] = {ready, due, service, demand} (int32, time units of the matrix); cap [batch]; max_dist
 * [batch] in distance units of the matrix.  The reference's request has payload only (routes.py:32-42); windows and
 * service times are the optional extension of SPEC.md §4/§6. */
int rt_problem_set_stops(rt_problem *p, const int32_t *stops, const int32_t *cap, const int64_t *max_dist);
/* Builds every instance's matrix on the device (same arithmetic as rt_matrix_build). */
int rt_problem_build_matrix(rt_problem *p, double dist_scale, double time_scale);
/* Installs a caller-supplied matrix for one instance (dense n x n int32 planes, entries in [0, 2^29)) — what the
 * reference holds after utils.py:103, integerised by the caller.  It may be directional (road networks are): the
 * greedy reads D[cur][idx] and D[idx][0] in that orientation (utils.py:128) and the move kernels then keep a
 * transposed route-order copy next to the forward one. */
int rt_problem_set_matrix(rt_problem *p, int instance, const int32_t *dist, const int32_t *time);
int rt_problem_get_matrix(rt_problem *p, int instance, int32_t *dist, int32_t *time);

/* Greedy construction for every instance on its device matrix (utils.py:111-139); demand [batch][n] FP64, cap and
 * max_dist [batch] FP64.  use_tw != 0 adds the time-window acceptance test of SPEC.md §4 (needs set_stops).
 * order_out [batch][n-1], trip_off_out [batch][n], n_trips_out [batch], status_out [batch] (RT_OK or
 * RT_ERR_INFEASIBLE_STOP per instance). */
int rt_problem_greedy(rt_problem *p, const double *demand, const double *cap, const double *max_dist, int use_tw,
                      int32_t *order_out, int32_t *trip_off_out, int32_t *n_trips_out, int32_t *status_out);

/* Installs routes (same layout as the greedy's output) as the current solution and derives the route state. */
int rt_problem_set_routes(rt_problem *p, const int32_t *order, const int32_t *trip_off, const int32_t *n_trips);
int rt_problem_get_routes(rt_problem *p, int32_t *order_out, int32_t *trip_off_out, int32_t *n_trips_out);
/* Route state of one instance (SPEC.md §6); position arrays have rt_problem_position_stride() entries, route
 * arrays n entries; any output may be NULL except npos_out. */
int rt_problem_get_state(rt_problem *p, int instance, int32_t *npos_out, int32_t *node, int32_t *route, int32_t *B,
                         int32_t *Z, int64_t *rload, int64_t *rdist, int64_t *cost, int32_t *feasible);

/* The stop matrix in ROUTE ORDER that the move kernels read: G[p][q] = M[node p][node q] for the P positions of the
 * current solution (SPEC.md §6), dense int32 [P][P] planes.  It is rebuilt by every state refresh. */
int rt_problem_get_position_matrix(rt_problem *p, int instance, int32_t *dist, int32_t *time);

/* Scores this shard's local-search candidates of every instance (families: bit 0 = 2-opt, 1 = relocate, 2 = swap;
 * delta + capacity + distance-limit + time-window feasibility each) and returns the minimum packed key per instance
 * (SPEC.md §7; 2^64-1 = none).  The candidate space is split into shard_count parts; the minimum over all shards
 * is the best move. */
int rt_problem_eval_moves(rt_problem *p, int families, int shard_index, int shard_count, uint64_t *best_keys_out);
/* Every candidate of one family for one instance: delta_out int32[P*P], feas_out uint8[P*P] indexed [p*P+q];
 * feas 0 = infeasible, 1 = feasible, 2 = not a candidate.  For tests and diagnostics. */
int rt_problem_eval_dump(rt_problem *p, int instance, int family, int32_t *delta_out, uint8_t *feas_out);
/* Applies one move per instance (keys [batch]; 2^64-1 = leave unchanged) and refreshes the state.  A key that is
 * not a move of the instance's current routes (wrong family bits, positions outside the routes, a depot where a
 * customer is needed) leaves that instance unchanged and makes the call return RT_ERR_INVALID_ARG after the other
 * instances' moves have been applied. */
int rt_problem_apply_moves(rt_problem *p, const uint64_t *keys);
/* Best-improvement local search to a local optimum or max_moves applied moves per instance (SPEC.md §8).
 * moves_out [batch], cost_out [batch]; keys_out [batch][keys_capacity] receives the applied keys (may be NULL).
 * Above ~4,600 positions per instance the search keeps a per-route-pair table and re-scores only what a move touched;
 * below that it re-scores everything each iteration, which is faster there (an iteration is then two dependent
 * launches replayed from a CUDA graph; up to 256 positions the whole search is ONE cooperative launch with grid-wide
 * barriers between its phases).  Same keys, same routes either way. */
int rt_problem_local_search(rt_problem *p, int families, int max_moves, int32_t *moves_out, int64_t *cost_out,
                            uint64_t *keys_out, int keys_capacity);
/* The same search with every iteration re-scoring the whole neighbourhood on the route-order matrix, i.e. without
 * the per-route-pair table rt_problem_local_search keeps (it re-scores only what the last move touched; same keys,
 * same routes).  For cross-checking and for asymmetric matrices / more than 1,024 routes, where the table is not used. */
int rt_problem_local_search_full(rt_problem *p, int families, int max_moves, int32_t *moves_out, int64_t *cost_out,
                                 uint64_t *keys_out, int keys_capacity);
/* Batched iterations (SPEC.md §8.1): every iteration applies a whole set of improving moves that touch pairwise
 * disjoint routes (selected in three rounds from the per-route-pair table of a full evaluation; the best move is
 * always among them), so a local optimum of the same neighbourhood — in general a different one — is reached in
 * about a tenth of the iterations.  max_iterations bounds the iterations, not the moves; moves_out counts moves;
 * keys_out receives them batch after batch (positions are those of the state the batch was selected on).  At most
 * 1,024 routes per instance.  No reference counterpart (utils.py:111-139 has no improvement phase). */
int rt_problem_local_search_batched(rt_problem *p, int families, int max_iterations, int32_t *moves_out,
                                    int64_t *cost_out, uint64_t *keys_out, int keys_capacity);

/* One pass of the hot path from host inputs to best moves, for every instance: uploads coordinates, stop data
 * and the current solution, builds the matrices (utils.py:94-109) — in route order, G[p][q] = M[node p][node q],
 * which is the layout the move kernels stream — derives the route state and scores the full neighbourhood; only
 * the best keys [batch] come back (the matrices stay on the GPU; with a shard set by rt_problem_set_shard, the best
 * keys of that shard).  node is the position array
 * of SPEC.md §6, [batch][rt_problem_position_stride()], npos [batch] its used length. */
int rt_problem_plan(rt_problem *p, const double *lat, const double *lon, const int32_t *stops, const int32_t *cap,
                    const int64_t *max_dist, const int32_t *node, const int32_t *npos, double dist_scale,
                    double time_scale, int families, uint64_t *best_keys_out);

/* The same pass in two halves, for a caller that keeps requests in flight (a service between two of its requests):
 * rt_problem_plan_begin enqueues the uploads and the kernels and returns; rt_problem_plan_end waits for that pass and
 * hands out its keys.  All inputs travel on the engine's copy stream, so with two problems of one engine used in
 * turn, the upload of one request overlaps the kernels of the other and the GPU never waits for PCIe.  The input
 * buffers must stay unchanged until rt_problem_plan_end returns; one pass per problem may be in flight;
 * rt_problem_plan_end does not take the engine's lock (other threads can go on enqueueing).  No reference
 * counterpart (the reference handles one request per thread, synchronously: routes.py:89-127). */
int rt_problem_plan_begin(rt_problem *p, const double *lat, const double *lon, const int32_t *stops,
                          const int32_t *cap, const int64_t *max_dist, const int32_t *node, const int32_t *npos,
                          double dist_scale, double time_scale, int families);
int rt_problem_plan_end(rt_problem *p, uint64_t *best_keys_out);

/* The same pass for a replica of a multi-GPU search (one process per GPU): host inputs in, but every copy and kernel
 * goes to the CALLER's stream, the replica scores only its own shard (rt_problem_set_shard; default everything), and
 * the keys stay on the device — best_keys_d is device int64[batch] in signed-order form, ready for the MIN
 * all-reduce that combines the shards (NCCL through torch.distributed, or the peer exchange below).  Nothing is
 * synchronised: the host buffers must stay unchanged until the stream has passed the copies (pinned memory makes them
 * asynchronous).  Coordinates and the position arrays are validated on the host while the copies are in flight; a
 * refused call waits for them before it returns.
 * No reference counterpart (the reference handles one request on one thread: routes.py:89-127). */
int rt_problem_plan_upload_dev(rt_problem *p, const double *lat, const double *lon, const int32_t *stops,
                               const int32_t *cap, const int64_t *max_dist, const int32_t *node, const int32_t *npos,
                               double dist_scale, double time_scale, int families, void *best_keys_d, void *stream);

/* The same pass for inputs that are already on the device (coordinates, stops and routes installed earlier), on
 * the caller's stream, without synchronisation: unit vectors -> route-order matrix -> route state -> best keys
 * (device int64[batch] in signed-order form, see rt_problem_eval_moves_dev).  A replica whose shard was set with
 * rt_problem_set_shard rebuilds only its rows of the matrix and scores only its shard: the MIN over the replicas'
 * keys is the best move. */
int rt_problem_plan_dev(rt_problem *p, double dist_scale, double time_scale, int families, void *best_keys_d,
                        void *stream);
/* Its pieces, stream-ordered: the node-order matrix of rt_problem_build_matrix; the refresh (what: bit 0 = rebuild
 * what the move kernels read of the route-order matrix from the current routes, bit 1 = re-derive the route state
 * tables; either bit works on its own after a route change, and with both set the tables run on a side stream
 * underneath the matrix kernel.  Where the matrix is SPEC.md §2 of the problem's own coordinates the move kernels
 * rebuild their cells from the unit vectors and read no matrix, so bit 0 has nothing to do; bit 2 materialises the
 * route-order matrix regardless (rt_problem_get_position_matrix does that by itself)). */
int rt_problem_build_matrix_dev(rt_problem *p, double dist_scale, double time_scale, void *stream);
int rt_problem_refresh_dev(rt_problem *p, int what, void *stream);

/* Declares which part of the candidate space this replica evaluates (multi-GPU search: rank i of c): a contiguous range
 * of rows p of the (p, q) plane, cut so that the c parts cost the same (early rows also score swaps against every later
 * column, so they are about 1.4x as expensive as late ones and get fewer rows).  From then on every refresh rebuilds
 * only what that shard needs, and rt_problem_eval_moves[_dev] accepts only that shard.  Default 0 of 1 = everything. */
int rt_problem_set_shard(rt_problem *p, int shard_index, int shard_count);

/* The same two steps on a caller's stream, for a multi-GPU search: every rank holds a replica of the problem,
 * evaluates its shard, all-reduces the keys with MIN (e.g. NCCL through torch.distributed) and applies the winner.
 * Keys are exchanged as int64 in order-preserving signed form (packed key XOR 2^63, so INT64_MAX = none).
 * best_keys_d / keys_d: device int64[batch].  apply marks an instance done when there is no move or, with
 * improving_only, when the best delta is >= 0. */
int rt_problem_eval_moves_dev(rt_problem *p, int families, int shard_index, int shard_count, void *best_keys_d,
                              void *stream);
int rt_problem_apply_moves_dev(rt_problem *p, const void *keys_d, int improving_only, void *stream);
/* Batched iterations (SPEC.md §8.1) over replicas: every rank scores its shard into a table of best keys per ordered
 * route pair (followed by one key per route for the same-route families), rt_problem_table_entries int64 values in
 * order-preserving signed form; the ranks combine their tables with ONE int64 MIN all-reduce; every rank then selects
 * the same batch from the combined table, applies it and refreshes its state.  table_d: device int64
 * [rt_problem_table_entries] owned by the caller (e.g. a torch tensor, so the collective can run on it directly).
 * rt_problem_table_entries returns 0 when no routes are installed or an instance has more than 1,024 routes; call
 * it once before the loop (it sizes the internal selection buffers).  An instance whose batch comes out empty is
 * marked done (rt_problem_search_status). */
int64_t rt_problem_table_entries(rt_problem *p);
int rt_problem_eval_table_dev(rt_problem *p, int families, int shard_index, int shard_count, void *table_d,
                              void *stream);
int rt_problem_apply_batch_dev(rt_problem *p, const void *table_d, void *stream);
/* The same exchange without a collective library: every replica (one process per GPU) exports a small slot buffer
 * through CUDA IPC, maps its peers', and from then on the evaluation step MIN-reduces its best keys straight into
 * every peer's buffer with system-scope atomics over NVLink, and the apply step waits until all replicas have
 * published before applying the winner.  The wait is bounded (ROUTEST_XCHG_TIMEOUT_MS, default 10 s); a replica that
 * gives up raises an abort word in every replica's buffer, so ALL replicas stop (done = 3 in
 * rt_problem_search_status) instead of drifting apart.  8 bytes per instance, no NCCL call in the
 * loop.  Order of calls: rt_problem_exchange_export on every rank -> all-gather the handles (any transport) ->
 * rt_problem_exchange_connect -> per iteration rt_problem_eval_moves_xchg_dev, rt_problem_apply_moves_xchg_dev.
 * The replica's shard (rt_problem_set_shard) must be (rank, world) or unset.  rt_problem_exchange_connect may be called
 * again (it closes the earlier mappings, re-arms the slots and clears the abort word). */
int rt_problem_exchange_handle_bytes(void);
int rt_problem_exchange_export(rt_problem *p, void *handle_out);
int rt_problem_exchange_connect(rt_problem *p, const void *handles, int rank, int world);
int rt_problem_eval_moves_xchg_dev(rt_problem *p, int families, void *stream);
int rt_problem_apply_moves_xchg_dev(rt_problem *p, int improving_only, void *stream);
/* The wait alone: the reduced keys (packed, unsigned form) land in keys_d [batch] and nothing is applied. */
int rt_problem_collect_keys_xchg_dev(rt_problem *p, void *keys_d, void *stream);

/* The same for the batched iterations' table: the replicas all-gather their shard tables into each other's memory
 * (plain coalesced stores over NVLink, one copy per rank and slot) and every replica takes the minimum locally, so the
 * batched search needs no collective call in its loop either.  Order of calls: rt_problem_table_exchange_export(world)
 * on every rank (after the routes are installed: it sizes the buffer for the table) -> all-gather the handles ->
 * rt_problem_table_exchange_connect -> per iteration rt_problem_eval_table_xchg_dev (scores this replica's shard into
 * table_d and publishes it), rt_problem_apply_batch_xchg_dev (waits, reduces into table_d, selects and applies the
 * batch, refreshes the state).  table_d: device int64 [rt_problem_table_entries + 1] owned by the caller.  Same bounded
 * wait and collective abort as above.  No reference counterpart. */
int rt_problem_table_exchange_export(rt_problem *p, int world, void *handle_out);
int rt_problem_table_exchange_connect(rt_problem *p, const void *handles, int rank, int world);
int rt_problem_eval_table_xchg_dev(rt_problem *p, int families, void *table_d, void *stream);
int rt_problem_apply_batch_xchg_dev(rt_problem *p, void *table_d, void *stream);

/* Synchronises the device and reads the search bookkeeping: done [batch], moves [batch], cost [batch]. */
int rt_problem_search_status(rt_problem *p, int32_t *done_out, int32_t *moves_out, int64_t *cost_out);

/* ------------------------------------------------------------------------------------------------------------
 * ETA scoring (SPEC.md §9) — takes over Flaskr/ml.py:35-53: the 12-feature row (ml.py:35-48) and model.predict on it
 * (ml.py:51-53), batched.  Inputs are SoA over n stop pairs: distance_m FP64 (the value routes.py:100-103 reads from
 * properties.summary.distance), weather u8 {0 Cloudy,1 Stormy,2 Sunny,3 Windy, other = none}, traffic u8 {0 High,
 * 1 Jam, 2 Low, 3 Medium, other = none}, weekday u8 (0 = Monday, ml.py:44), hour u8 (ml.py:45), driver_age f32
 * (0 means 30, ml.py:47).  The model replaces the pickle of ml.py:6-21: an ensemble in the engine's tree format
 * (SPEC.md §9.2), copied to the GPU by rt_eta_load_model.  Without a model the scoring calls return RT_ERR_STATE
 * (the reference returns (None, None) there, ml.py:25-26).
 */
int rt_eta_load_model(rt_engine *e, const void *blob, size_t bytes);
int rt_eta_model_info(rt_engine *e, int32_t *n_trees, int32_t *n_nodes); /* -1 when no model is loaded */
int rt_eta_score(rt_engine *e, const double *distance_m, const uint8_t *weather, const uint8_t *traffic,
                 const uint8_t *weekday, const uint8_t *hour, const float *driver_age, int64_t n, float *minutes_out);
/* The feature rows alone, float32 [n][12] in the column order of ml.py:35-48 (what model.predict is handed). */
int rt_eta_features(rt_engine *e, const double *distance_m, const uint8_t *weather, const uint8_t *traffic,
                    const uint8_t *weekday, const uint8_t *hour, const float *driver_age, int64_t n, float *rows_out);
int rt_eta_score_dev(rt_engine *e, const double *distance_m_d, const uint8_t *weather_d, const uint8_t *traffic_d,
                     const uint8_t *weekday_d, const uint8_t *hour_d, const float *driver_age_d, int64_t n,
                     float *minutes_d, void *stream);

/* Self-test: counts floats with bit pattern in [lo_bits, hi_bits) for which the matrix kernel's straight-line
 * square root differs from IEEE sqrtf (must be 0; SPEC.md §2 relies on it).  No reference counterpart. */
int rt_selftest_fast_sqrt(rt_engine *e, uint32_t lo_bits, uint32_t hi_bits, uint64_t *mismatches);

/* On-box instruction-rate peak, the roofline denominator of the integer move-evaluation kernels (SURVEY.md §8(d) asks
 * for ALU peaks measured on the box): a kernel of independent register chains, timed with CUDA events on the engine's
 * stream (the call synchronises).  mode 0 = ALU pipe only (LOP3/IMNMX), 1 = FMA pipe only (IMAD), 2 = ALU + IMAD
 * interleaved, 3 = ALU + immediate-form FFMA interleaved, 4 = ALU + IMAD + FFMA (2:1:1, three pipes: the closest to the
 * issue limit of one warp instruction per scheduler per clock).  iters = loop trips per thread (32 instructions each); the result is warp instructions per second
 * over the whole GPU.  No reference counterpart. */
int rt_selftest_pipe_peak(rt_engine *e, int mode, int iters, double *warp_inst_per_s, double *elapsed_ms);

/* Test and measurement knobs (process-wide, not for production use): "strip_rows" = rows per block of the tile kernel
 * (0 = let the launcher pick, 8..64), "intra_kernel" = which same-route kernel runs (0 = pick, 1 = one thread per
 * anchor, 2 = one warp per anchor), "fused_eval" = 1 (default): move kernels rebuild matrix cells from unit vectors
 * where the matrix is SPEC.md §2 of the problem's coordinates, 0: always write the route-order matrix and read it back;
 * "small_strips" = 1 (default): strips of 8 or 16 rows where the whole tile grid is resident at once; how the one-move
 * search cuts an iteration into launches: "persistent_search" = 0 never / 1 (default) the whole search as one
 * cooperative launch where that is the faster form (small instances) / 2 wherever it can run, "persist_blocks" = most
 * blocks per SM of that launch (0 = all that fit), "apply_prep" = 1 (default): move + route scan (+ per-route tables up
 * to 32 routes) in one launch, "state_fused_max" = positions up to which one block per instance builds the whole route
 * state (-1 = default: never, one warp per route instead).
 * Every choice gives the same keys; the tests run them against each other and the oracle.  No reference counterpart. */
int rt_debug_set(const char *name, int value);

#ifdef __cplusplus
}
#endif
#endif /* ROUTEST_H */
