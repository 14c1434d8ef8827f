// batch.cu — batched iterations of the local search (SPEC.md §8.1): every iteration applies a whole set of
// improving moves that touch pairwise disjoint routes instead of the single best one.
//
// A candidate's delta and feasibility depend on its (at most two) routes only, so such a set is independent.  One
// iteration is
//   full evaluation into the per-route-pair table (moves.cu, table mode: inter[ra][rb], intra[r], absolute keys)
//   -> three selection rounds: m[r] = minimum key over the table entries that touch r and whose routes are all
//      unclaimed; an entry with delta < 0 that equals m[] of all its routes is selected, its routes are claimed
//   -> one kernel rebuilds node[] with all selected moves applied (route by route: every route is changed by at most
//      one move, so its new content is a simple index map of the old array)
//   -> state refresh.
// The best move of SPEC.md §7 is always in the batch, so the search ends in a local optimum of the same neighbourhood
// — after about a tenth of the iterations at 1,000 stops.  The batches are those of the CPU oracle
// (or_local_search_batched), key for key.  No reference counterpart (utils.py:111-139 has no improvement phase).
#include <algorithm>

#include "problem.cuh"

namespace rt {

namespace {

constexpr int kSelectThreads = 1024;  // one route per thread in the selection and apply kernels
constexpr int kRounds = 3;            // SPEC.md §8.1

struct BatchArgs {
    const unsigned long long *inter, *intra;  // [batch][cr][cr], [batch][cr]
    unsigned long long xmask;                 // 2^63 when the table is in signed-order form (multi-GPU), else 0
    int cr;
    const int *npos, *route, *rstart;
    int n, pstride, rstride;
    unsigned long long *m, *rowbest, *sel;  // [batch][cr] each
    int *taken, *selcount;                  // [batch][cr], [batch]
    int *node, *tmp, *done, *moves;
    unsigned long long *hist;
    int hist_cap;
};

__device__ __forceinline__ int key_delta(unsigned long long k)
{
    return static_cast<int>(static_cast<unsigned>(k >> 32) - 0x80000000u);
}

__global__ void __launch_bounds__(256) batch_reset_kernel(const BatchArgs a)
{
    const int b = blockIdx.x;
    for (int r = threadIdx.x; r < a.cr; r += blockDim.x) {
        a.m[static_cast<size_t>(b) * a.cr + r] = kNoMove;
        a.taken[static_cast<size_t>(b) * a.cr + r] = 0;
    }
    if (threadIdx.x == 0) a.selcount[b] = 0;
}

// One block per (route ra, instance): the minimum over ra's row of the table (its own intra entry included) among
// unclaimed routes -> rowbest[ra]; every entry also lowers m[] of the two routes it touches.
__global__ void __launch_bounds__(128) batch_rowmin_kernel(const BatchArgs a)
{
    __shared__ unsigned long long s_best[4];
    const int b = blockIdx.y, ra = blockIdx.x, tid = threadIdx.x;
    const int R = a.npos[b] - a.n;
    if (ra >= R || a.done[b]) return;
    const size_t ro = static_cast<size_t>(b) * a.cr;
    const int *taken = a.taken + ro;
    unsigned long long *m = a.m + ro;
    if (taken[ra]) {
        if (tid == 0) a.rowbest[ro + ra] = kNoMove;
        return;
    }
    const unsigned long long *row = a.inter + (ro + ra) * a.cr;
    unsigned long long best = kNoMove;
    for (int rb = tid; rb < R; rb += blockDim.x) {
        if (taken[rb]) continue;
        const unsigned long long k = row[rb] ^ a.xmask;
        if (k == kNoMove) continue;
        best = k < best ? k : best;
        if (k < m[rb]) atomicMin(m + rb, k);  // m[] only ever decreases: a stale read costs one redundant atomic
    }
    if (tid == 0) {
        const unsigned long long k = a.intra[ro + ra] ^ a.xmask;
        best = k < best ? k : best;
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
        a.rowbest[ro + ra] = best;
        if (best != kNoMove) atomicMin(m + ra, best);
    }
}

// One block per instance, one thread per route: the row minimum of route r is selected when it improves and is also
// the minimum of every route it touches.  Selected keys are appended in ascending r; their routes are claimed and m[]
// is cleared for the next round.
__global__ void __launch_bounds__(kSelectThreads) batch_select_kernel(const BatchArgs a)
{
    __shared__ int s_warp[kSelectThreads / 32];
    const int b = blockIdx.x, r = threadIdx.x, lane = r & 31, w = r >> 5;
    const int R = a.npos[b] - a.n;
    if (a.done[b]) return;
    const size_t ro = static_cast<size_t>(b) * a.cr;
    unsigned long long *m = a.m + ro;
    unsigned long long K = kNoMove;
    int rb = 0;
    bool pick = false;
    if (r < R) {
        K = a.rowbest[ro + r];
        if (K != kNoMove && key_delta(K) < 0 && m[r] == K) {
            rb = a.route[static_cast<size_t>(b) * a.pstride + static_cast<int>(K & 0x7FFF)];
            pick = m[rb] == K;
        }
    }
    const unsigned ball = __ballot_sync(0xffffffffu, pick);
    if (lane == 0) s_warp[w] = __popc(ball);
    __syncthreads();
    int before = 0, total = 0;
    for (int i = 0; i < kSelectThreads / 32; ++i) {
        const int c = s_warp[i];
        before += i < w ? c : 0;
        total += c;
    }
    const int base = a.selcount[b];
    if (pick) {
        a.sel[ro + base + before + __popc(ball & ((1u << lane) - 1))] = K;
        a.taken[ro + r] = 1;
        a.taken[ro + rb] = 1;
    }
    __syncthreads();  // every thread has read m[] and selcount
    if (r < R) m[r] = kNoMove;
    if (r == 0) a.selcount[b] = base + total;
}

enum { kKeep = 0, kReverse, kMoveWithin, kSwapWithin, kRemove, kInsert, kReplace };

// node[] of instance b with the whole batch sel[0..cnt) applied (all threads of the block).  Route r owns positions
// rstart[r] .. rstart[r+1]-1 (its opening separator and its stops); at most one move changes it, described by
// (type, x, y, src).
__device__ __forceinline__ void apply_batch(const BatchArgs &a, int b, const unsigned long long *sel, int cnt)
{
    __shared__ int s_type[kSelectThreads], s_x[kSelectThreads], s_y[kSelectThreads], s_src[kSelectThreads];
    __shared__ int s_len[kSelectThreads], s_scan[kSelectThreads];
    const int tid = threadIdx.x;
    const int P = a.npos[b], R = P - a.n;
    if (cnt == 0) {  // local optimum: nothing improves any more
        if (tid == 0) a.done[b] = 1;
        return;
    }
    const size_t po = static_cast<size_t>(b) * a.pstride;
    const int *rstart = a.rstart + static_cast<size_t>(b) * (a.rstride + 1);
    const int *route = a.route + po;
    int *__restrict__ node = a.node + po;
    int *__restrict__ tmp = a.tmp + po;   // distinct arrays: lets the copy loops below keep several loads in flight
    s_type[tid] = kKeep;
    s_len[tid] = tid < R ? rstart[tid + 1] - rstart[tid] : 0;
    __syncthreads();
    if (tid < cnt) {
        const unsigned long long K = sel[tid];
        const int family = static_cast<int>((K >> 30) & 3), p = static_cast<int>((K >> 15) & 0x7FFF),
                  q = static_cast<int>(K & 0x7FFF);
        const int ra = route[p], rb = route[q];
        const int pr = p - rstart[ra], qr = q - rstart[rb];
        if (family == 0) {
            s_type[ra] = kReverse, s_x[ra] = pr, s_y[ra] = qr;
        } else if (ra == rb) {
            s_type[ra] = family == 1 ? kMoveWithin : kSwapWithin, s_x[ra] = pr, s_y[ra] = qr;
        } else if (family == 1) {
            s_type[ra] = kRemove, s_x[ra] = pr, s_len[ra] -= 1;
            s_type[rb] = kInsert, s_x[rb] = qr, s_src[rb] = p, s_len[rb] += 1;
        } else {
            s_type[ra] = kReplace, s_x[ra] = pr, s_src[ra] = q;
            s_type[rb] = kReplace, s_x[rb] = qr, s_src[rb] = p;
        }
        const int m0 = a.moves[b];
        if (a.hist && m0 + tid < a.hist_cap) a.hist[static_cast<size_t>(b) * a.hist_cap + m0 + tid] = K;
    }
    __syncthreads();
    // new route starts: exclusive scan of the new lengths
    const int len = s_len[tid];
    s_scan[tid] = len;
    __syncthreads();
    for (int off = 1; off < kSelectThreads; off <<= 1) {
        const int v = tid >= off ? s_scan[tid - off] : 0;
        __syncthreads();
        s_scan[tid] += v;
        __syncthreads();
    }
    if (tid == 0) a.moves[b] += cnt;
    // a warp per route: tmp[new start + j] = node[source of j]
    const int lane = tid & 31;
    for (int r = tid >> 5; r < R; r += kSelectThreads / 32) {
        const int os = rstart[r], ns = s_scan[r] - s_len[r], L = s_len[r];
        const int type = s_type[r], x = s_x[r], y = s_y[r], src = s_src[r];
#pragma unroll 4
        for (int j = lane; j < L; j += 32) {
            int from = os + j;
            switch (type) {
            case kReverse: if (j > x && j <= y) from = os + x + 1 + y - j; break;
            case kMoveWithin:
                if (y > x) from = j < x || j > y ? os + j : (j == y ? os + x : os + j + 1);
                else from = j <= y || j > x ? os + j : (j == y + 1 ? os + x : os + j - 1);
                break;
            case kSwapWithin: from = j == x ? os + y : (j == y ? os + x : os + j); break;
            case kRemove: from = j < x ? os + j : os + j + 1; break;
            case kInsert: from = j <= x ? os + j : (j == x + 1 ? src : os + j - 1); break;
            case kReplace: from = j == x ? src : os + j; break;
            default: break;
            }
            tmp[ns + j] = node[from];
        }
    }
    if (tid == 0) tmp[P - 1] = 0;
    __syncthreads();
#pragma unroll 8
    for (int i = tid; i < P; i += kSelectThreads) node[i] = tmp[i];
}

__global__ void __launch_bounds__(kSelectThreads) batch_apply_kernel(const BatchArgs a)
{
    const int b = blockIdx.x;
    if (a.done[b]) return;
    apply_batch(a, b, a.sel + static_cast<size_t>(b) * a.cr, a.selcount[b]);
}

// Up to kFusedRoutes routes per instance the whole selection is small enough for one block: the three rounds and the
// apply step run in ONE kernel (row minima by a warp per route, m[] / claims / the batch in shared memory) instead of
// eight dependent launches.  The R x R table is copied into (dynamic) shared memory once and the three rounds read it
// from there (row minima by a warp per row, column minima by a warp per column, no atomics): 8 R^2 bytes next to the ~30 KB of the selection and apply arrays caps R at 152 (a 5,000-stop / 50-vehicle
// instance has 140 routes).  Same selection, same order.  Beyond that the separate launches are used (one block reading
// a 262 x 262 table three times from global memory was slower than them: 0.32 s instead of 0.28 s for the 10,000-stop
// solve).
constexpr int kFusedRoutes = 152;

__global__ void __launch_bounds__(kSelectThreads) batch_fused_kernel(const BatchArgs a)
{
    __shared__ unsigned long long s_m[kFusedRoutes], s_rowbest[kFusedRoutes], s_sel[kFusedRoutes];
    __shared__ int s_taken[kFusedRoutes], s_warp[kSelectThreads / 32], s_count;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (a.done[b]) return;
    const int R = a.npos[b] - a.n;
    const size_t ro = static_cast<size_t>(b) * a.cr;
    const int *route = a.route + static_cast<size_t>(b) * a.pstride;
    if (tid < kFusedRoutes) s_taken[tid] = 0;
    if (tid == 0) s_count = 0;
    // the table, once, into shared memory: [R][RS] packed keys (unsigned form), odd row stride so that a column walk
    // touches different banks
    extern __shared__ unsigned long long s_tab[];
    const int RS = R | 1;
    __shared__ unsigned long long s_intra[kFusedRoutes];
    constexpr int kColChunks = (kFusedRoutes + 31) / 32;
    for (int ra = w; ra < R; ra += kSelectThreads / 32) {  // a warp per row, the row's loads all in flight at once
        const unsigned long long *row = a.inter + (ro + ra) * a.cr;
        unsigned long long v[kColChunks];
#pragma unroll
        for (int k = 0; k < kColChunks; ++k) v[k] = lane + 32 * k < R ? row[lane + 32 * k] : 0ull;
#pragma unroll
        for (int k = 0; k < kColChunks; ++k)
            if (lane + 32 * k < R) s_tab[ra * RS + lane + 32 * k] = v[k] ^ a.xmask;
    }
    if (tid < R) s_intra[tid] = a.intra[ro + tid] ^ a.xmask;
    __syncthreads();
    for (int round = 0; round < kRounds; ++round) {
        // m[r] = minimum over the entries that touch r with both routes unclaimed = min(row r, column r, intra[r]):
        // one warp per row, then one warp per column — no atomics (64-bit minima on shared memory are CAS loops)
        for (int ra = w; ra < R; ra += kSelectThreads / 32) {
            unsigned long long best = kNoMove;
            if (!s_taken[ra]) {
                const unsigned long long *row = s_tab + ra * RS;
                for (int rb = lane; rb < R; rb += 32) {
                    const unsigned long long k = s_taken[rb] ? kNoMove : row[rb];
                    best = k < best ? k : best;
                }
                if (lane == 0) best = s_intra[ra] < best ? s_intra[ra] : best;
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, off);
                    best = o < best ? o : best;
                }
            }
            if (lane == 0) s_rowbest[ra] = best;
        }
        for (int rb = w; rb < R; rb += kSelectThreads / 32) {
            unsigned long long best = kNoMove;
            if (!s_taken[rb]) {
                for (int ra = lane; ra < R; ra += 32) {
                    const unsigned long long k = s_taken[ra] ? kNoMove : s_tab[ra * RS + rb];
                    best = k < best ? k : best;
                }
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, off);
                    best = o < best ? o : best;
                }
            }
            if (lane == 0) s_m[rb] = best;   // column part; the row part joins below
        }
        __syncthreads();
        if (tid < R) s_m[tid] = s_rowbest[tid] < s_m[tid] ? s_rowbest[tid] : s_m[tid];
        __syncthreads();
        unsigned long long K = kNoMove;
        int rb = 0;
        bool pick = false;
        if (tid < R) {
            K = s_rowbest[tid];
            if (K != kNoMove && key_delta(K) < 0 && s_m[tid] == K) {
                rb = route[static_cast<int>(K & 0x7FFF)];
                pick = s_m[rb] == K;
            }
        }
        const unsigned ball = __ballot_sync(0xffffffffu, pick);
        if (lane == 0) s_warp[w] = __popc(ball);
        __syncthreads();
        int before = 0, total = 0;
        for (int i = 0; i < (kFusedRoutes + 31) / 32; ++i) {  // routes live in the first kFusedRoutes threads
            const int c = s_warp[i];
            before += i < w ? c : 0;
            total += c;
        }
        const int base = s_count;
        if (pick) {
            s_sel[base + before + __popc(ball & ((1u << lane) - 1))] = K;
            s_taken[tid] = 1;
            s_taken[rb] = 1;
        }
        __syncthreads();
        if (tid == 0) s_count = base + total;
        if (total == 0) break;  // block-uniform
    }
    __syncthreads();
    apply_batch(a, b, s_sel, s_count);
}

}  // namespace

bool batch_usable(const rt_problem *p) { return std::max(1, p->grows - p->n) <= kSelectThreads; }

// the search's own table: inter and intra back to back in the incremental search's buffer
static unsigned long long *own_table(rt_problem *p) { return static_cast<unsigned long long *>(p->cache_inter.p); }

size_t batch_table_entries(const rt_problem *p)
{
    const size_t B = p->batch, cr = std::max(1, p->grows - p->n);
    return B * (cr * cr + cr);
}

int batch_ensure(rt_problem *p)
{
    RT_TRY(cache_ensure(p));  // sizes the table (cache_r) and, for the single-GPU driver, provides its memory
    const size_t B = p->batch, cr = p->cache_r;
    RT_TRY(reserve(p->e, p->cache_inter, batch_table_entries(p) * 8));
    return reserve(p->e, p->batch_buf, B * cr * (3 * 8 + 4) + B * 4 + 64);
}

static BatchArgs batch_args(rt_problem *p, const unsigned long long *table, bool signed_keys)
{
    BatchArgs a{};
    const size_t B = p->batch, cr = p->cache_r;
    a.inter = table;
    a.intra = table + B * cr * cr;
    a.xmask = signed_keys ? 0x8000000000000000ull : 0ull;
    a.cr = p->cache_r;
    a.npos = p->npos;
    a.route = p->route;
    a.rstart = p->rstart;
    a.n = p->n;
    a.pstride = p->pstride;
    a.rstride = p->rstride;
    unsigned long long *u = static_cast<unsigned long long *>(p->batch_buf.p);
    a.m = u;
    a.rowbest = u + B * cr;
    a.sel = u + 2 * B * cr;
    a.taken = reinterpret_cast<int *>(u + 3 * B * cr);
    a.selcount = a.taken + B * cr;
    a.node = p->node;
    a.tmp = p->node_tmp;
    a.done = p->done;
    a.moves = p->moves;
    a.hist = p->hist;
    a.hist_cap = p->hist_cap;
    return a;
}

int batch_begin(rt_problem *p, cudaStream_t st)
{
    RT_TRY(batch_ensure(p));
    return state_refresh(p, st);
}

// Table (complete: every shard's minima combined) -> selection rounds -> apply -> state refresh.
int batch_select_apply(rt_problem *p, const unsigned long long *table, bool signed_keys, cudaStream_t st)
{
    rt_engine *e = p->e;
    const BatchArgs a = batch_args(p, table, signed_keys);
    if (p->cache_r <= kFusedRoutes) {
        const size_t tab_bytes = static_cast<size_t>(p->cache_r) * (p->cache_r | 1) * 8;
        static bool attr_set = false;  // once per process: the kernel may use up to 8 * kFusedRoutes^2 bytes on top of its statics
        if (!attr_set) {
            RT_CUDA(e, cudaFuncSetAttribute(batch_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            kFusedRoutes * (kFusedRoutes | 1) * 8));
            attr_set = true;
        }
        batch_fused_kernel<<<p->batch, kSelectThreads, tab_bytes, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
        return state_refresh(p, st);
    }
    batch_reset_kernel<<<p->batch, 256, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    for (int round = 0; round < kRounds; ++round) {
        dim3 grid(p->cache_r, p->batch);
        batch_rowmin_kernel<<<grid, 128, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
        batch_select_kernel<<<p->batch, kSelectThreads, 0, st>>>(a);
        RT_CUDA(e, cudaGetLastError());
    }
    batch_apply_kernel<<<p->batch, kSelectThreads, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    return state_refresh(p, st);
}

// One batched iteration on a fresh state; leaves the state fresh again.
int batch_iteration(rt_problem *p, int families, cudaStream_t st)
{
    RT_TRY(eval_table_launch(p, families, 0, 1, own_table(p), false, st));
    return batch_select_apply(p, own_table(p), false, st);
}

}  // namespace rt
