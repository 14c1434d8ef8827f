// greedy.cu — capacity/distance-aware greedy trip construction on the device-resident matrix (SPEC.md §4).
//
// Takes over Flaskr/utils.py:111-139 of the reference and reproduces it decision for decision:
//   * candidates are ordered ONCE by distance from the depot, ties by index (sorted() evaluates its key with
//     current == 0 and is stable over an ascending list: utils.py:116,122-125);
//   * a stop is accepted iff load+demand <= cap and trip_dist + (D[cur][idx] + D[idx][0]) <= max_dist, in FP64 and
//     with this association (utils.py:126-129); a rejected stop does not end the trip;
//   * accepted: load += demand, trip_dist += D[cur][idx], cur = idx (utils.py:130-133).
// The chain is sequential, so one warp per instance walks it.  A pass fetches the next 128 candidates (four per lane,
// all loads in flight together), then tests them 32 at a time against the current (load, trip_dist, cur): the first
// lane that fits is accepted, lanes before it are rejected for this trip (exactly what the sequential loop would
// have decided), lanes after it — and the later chunks of the pass — re-read D[cur][.] and are tested again.
// Deviation, documented in SPEC.md §4: where the reference would loop forever (a stop that fits no trip on its
// own) the kernel stops with RT_ERR_INFEASIBLE_STOP.
#include "problem.cuh"

namespace rt {

namespace {

__device__ __forceinline__ unsigned long long ordered_bits(double d)
{
    unsigned long long u = static_cast<unsigned long long>(__double_as_longlong(d));
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

// keys/idx for stops 1..n-1, padded to n2 (a power of two) with +inf keys
template <bool F64>
__global__ void greedy_keys_kernel(const int2 *__restrict__ cells, const double *__restrict__ Df, int n, int pitch,
                                   int n2, unsigned long long *__restrict__ keys, int *__restrict__ order)
{
    const int b = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n2) return;
    const size_t o = static_cast<size_t>(b) * n2 + i;
    const int stop = i + 1;
    if (stop < n) {
        keys[o] = F64 ? ordered_bits(Df[static_cast<size_t>(b) * n * n + stop])
                      : static_cast<unsigned long long>(
                            static_cast<unsigned>(cells[static_cast<size_t>(b) * n * pitch + stop].x) ^ 0x80000000u);
        order[o] = stop;
    } else {
        keys[o] = ~0ull;
        order[o] = 0x7fffffff;
    }
}

// one block per instance: bitonic sort of (key, idx) pairs in global memory
__global__ void __launch_bounds__(1024) greedy_sort_kernel(unsigned long long *__restrict__ keys_all,
                                                           int *__restrict__ order_all, int n2)
{
    unsigned long long *keys = keys_all + static_cast<size_t>(blockIdx.x) * n2;
    int *order = order_all + static_cast<size_t>(blockIdx.x) * n2;
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n2; i += blockDim.x) {
                const int l = i ^ j;
                if (l > i) {
                    const unsigned long long ki = keys[i], kl = keys[l];
                    const int oi = order[i], ol = order[l];
                    const bool gt = ki > kl || (ki == kl && oi > ol);
                    if (gt == ((i & k) == 0)) {
                        keys[i] = kl;
                        keys[l] = ki;
                        order[i] = ol;
                        order[l] = oi;
                    }
                }
            }
            __syncthreads();
        }
    }
}

struct GreedyArgs {
    const int2 *cells;
    const double *Df;
    const int4 *stops;
    const double *demand, *cap, *maxd;
    int *order;  // sorted candidate list, consumed
    int n, pitch, n2, use_tw;
    int *order_out, *trip_off, *ntrips, *status;
};

template <bool F64>
__global__ void __launch_bounds__(32) greedy_trips_kernel(const GreedyArgs a)
{
    const int b = blockIdx.x, lane = threadIdx.x, n = a.n;
    const int2 *M = a.cells + static_cast<size_t>(b) * n * a.pitch;
    const double *Df = a.Df + static_cast<size_t>(b) * n * n;
    const double *demand = a.demand + static_cast<size_t>(b) * n;
    const int4 *stops = a.stops + static_cast<size_t>(b) * n;
    int *order = a.order + static_cast<size_t>(b) * a.n2;
    int *order_out = a.order_out + static_cast<size_t>(b) * n;
    int *trip_off = a.trip_off + static_cast<size_t>(b) * (n + 1);
    const double cap = a.cap[b], maxd = a.maxd[b];
    const bool tw = a.use_tw;
    const int4 depot = tw ? stops[0] : make_int4(0, 0, 0, 0);

    auto dist = [&](int from, int to) -> double {
        return F64 ? Df[static_cast<size_t>(from) * n + to] : static_cast<double>(M[static_cast<size_t>(from) * a.pitch + to].x);
    };

    int m = n - 1, done = 0, trips = 0, status = RT_OK;
    if (lane == 0) trip_off[0] = 0;
    constexpr int kChunks = 4;  // 128 candidates per pass: their loads are issued together, then tested chunk by chunk
    while (m > 0) {
        double load = 0.0, trip_dist = 0.0;
        int cur = 0, accepted = 0;
        long long clock = static_cast<long long>(depot.x) + depot.z;
        for (int pos = 0; pos < m; pos += 32 * kChunks) {
            int idx[kChunks], tcur[kChunks], tback[kChunks];
            double dem[kChunks], dback[kChunks], dcur[kChunks];
            int4 st[kChunks];
            const int cur_at_load = cur;
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                const int k = pos + 32 * c + lane;
                idx[c] = k < m ? order[k] : -1;
                dem[c] = dback[c] = dcur[c] = 0.0;
                tcur[c] = tback[c] = 0;
                st[c] = make_int4(0, 0, 0, 0);
                if (idx[c] > 0) {
                    dem[c] = demand[idx[c]];
                    dback[c] = dist(idx[c], 0);
                    dcur[c] = dist(cur, idx[c]);
                    if (tw) {
                        st[c] = stops[idx[c]];
                        tcur[c] = M[static_cast<size_t>(cur) * a.pitch + idx[c]].y;
                        tback[c] = M[static_cast<size_t>(idx[c]) * a.pitch].y;
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                const int k = pos + 32 * c + lane;
                if (pos + 32 * c >= m) break;
                if (cur != cur_at_load && idx[c] > 0) {  // an earlier chunk of this pass accepted a stop
                    dcur[c] = dist(cur, idx[c]);
                    if (tw) tcur[c] = M[static_cast<size_t>(cur) * a.pitch + idx[c]].y;
                }
                unsigned alive = __ballot_sync(0xffffffffu, idx[c] > 0);
                while (alive) {
                    bool ok = false;
                    long long arr = 0;
                    if ((alive >> lane) & 1u) {
                        ok = (load + dem[c]) <= cap && (trip_dist + (dcur[c] + dback[c])) <= maxd;
                        if (ok && tw) {
                            arr = clock + tcur[c];
                            if (arr < st[c].x) arr = st[c].x;
                            ok = arr <= st[c].y && arr + st[c].z + tback[c] <= depot.y;
                        }
                    }
                    const unsigned mask = __ballot_sync(0xffffffffu, ok);
                    if (!mask) break;
                    const int f = __ffs(mask) - 1;
                    const int idx_f = __shfl_sync(0xffffffffu, idx[c], f);
                    const double dem_f = __shfl_sync(0xffffffffu, dem[c], f), dcur_f = __shfl_sync(0xffffffffu, dcur[c], f);
                    const long long arr_f = __shfl_sync(0xffffffffu, arr, f);
                    const int svc_f = __shfl_sync(0xffffffffu, st[c].z, f);
                    if (lane == 0) order_out[done] = idx_f;
                    ++done;
                    ++accepted;
                    load += dem_f;
                    trip_dist += dcur_f;
                    cur = idx_f;
                    clock = arr_f + svc_f;
                    if (lane == f) order[k] = -idx[c];  // visited
                    alive &= ~((2u << f) - 1u);          // lanes up to f are decided for this trip
                    if ((alive >> lane) & 1u) {
                        dcur[c] = dist(cur, idx[c]);
                        if (tw) tcur[c] = M[static_cast<size_t>(cur) * a.pitch + idx[c]].y;
                    }
                }
            }
        }
        if (!accepted) {
            status = RT_ERR_INFEASIBLE_STOP;
            break;
        }
        ++trips;
        if (lane == 0) trip_off[trips] = done;
        // drop the visited stops from the candidate list (utils.py:138-139)
        // (128 entries per step: the four loads are in flight together — one memory round trip per step instead of
        // four; writes land at or below the lowest index read, so reading ahead of them is safe)
        __syncwarp();
        int w = 0;
        for (int pos = 0; pos < m; pos += 32 * kChunks) {
            int v[kChunks];
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                const int k = pos + 32 * c + lane;
                v[c] = k < m ? order[k] : -1;
            }
            __syncwarp();
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                const unsigned keep = __ballot_sync(0xffffffffu, v[c] > 0);
                if (v[c] > 0) order[w + __popc(keep & ((1u << lane) - 1u))] = v[c];
                w += __popc(keep);
            }
            __syncwarp();
        }
        m = w;
    }
    if (lane == 0) {
        a.ntrips[b] = trips;
        a.status[b] = status;
    }
}

int next_pow2(int v)
{
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

}  // namespace

// Runs the construction for every instance of the problem on its device-resident matrix.  Inputs (demand, cap,
// max_dist) must already be in p->gdemand / p->gcap / p->gmaxd; outputs land in p->g_* (device).
int greedy_launch(rt_problem *p, bool use_tw, cudaStream_t st)
{
    rt_engine *e = p->e;
    const int n2 = next_pow2(p->n > 1 ? p->n - 1 : 1);
    dim3 gk((n2 + 255) / 256, p->batch);
    greedy_keys_kernel<false><<<gk, 256, 0, st>>>(p->cells, nullptr, p->n, p->pitch, n2, p->gkeys, p->gorder);
    RT_CUDA(e, cudaGetLastError());
    greedy_sort_kernel<<<p->batch, 1024, 0, st>>>(p->gkeys, p->gorder, n2);
    RT_CUDA(e, cudaGetLastError());
    GreedyArgs a{};
    a.cells = p->cells;
    a.Df = nullptr;
    a.stops = p->stops;
    a.demand = p->gdemand;
    a.cap = p->gcap;
    a.maxd = p->gmaxd;
    a.order = p->gorder;
    a.n = p->n;
    a.pitch = p->pitch;
    a.n2 = n2;
    a.use_tw = use_tw ? 1 : 0;
    a.order_out = p->g_order_out;
    a.trip_off = p->g_trip_off;
    a.ntrips = p->g_ntrips;
    a.status = p->g_status;
    greedy_trips_kernel<false><<<p->batch, 32, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

// Stand-alone construction on a caller-supplied FP64 matrix (any matrix the reference could have been handed by its
// matrix service, utils.py:97-103: floats, asymmetric).  All pointers are device pointers.
int greedy_f64_launch(rt_engine *e, const double *D_d, int n, const double *demand_d, const double *cap_d,
                      const double *maxd_d, unsigned long long *keys_d, int *order_d, int *order_out_d, int *trip_off_d,
                      int *ntrips_d, int *status_d, cudaStream_t st)
{
    const int n2 = next_pow2(n > 1 ? n - 1 : 1);
    dim3 gk((n2 + 255) / 256, 1);
    greedy_keys_kernel<true><<<gk, 256, 0, st>>>(nullptr, D_d, n, 0, n2, keys_d, order_d);
    RT_CUDA(e, cudaGetLastError());
    greedy_sort_kernel<<<1, 1024, 0, st>>>(keys_d, order_d, n2);
    RT_CUDA(e, cudaGetLastError());
    GreedyArgs a{};
    a.Df = D_d;
    a.demand = demand_d;
    a.cap = cap_d;
    a.maxd = maxd_d;
    a.order = order_d;
    a.n = n;
    a.n2 = n2;
    a.order_out = order_out_d;
    a.trip_off = trip_off_d;
    a.ntrips = ntrips_d;
    a.status = status_d;
    greedy_trips_kernel<true><<<1, 32, 0, st>>>(a);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

}  // namespace rt
