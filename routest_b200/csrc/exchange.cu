// exchange.cu — the one exchange step of a multi-GPU search, done over NVLink peer memory instead of a collective
// library call: 8 bytes per instance, latency bound.
//
// Every replica owns a small slot buffer in its own HBM that its peers map through CUDA IPC (one process per GPU):
//   keys  [2][batch]  packed move keys (SPEC.md §7), initialised to "no move"
//   count [2]         how many replicas have published into keys[slot]
// Iteration i uses slot i & 1.
//   publish (after the evaluation kernels, same stream): every replica MIN-reduces its shard's best key straight
//           into keys[slot][b] of EVERY replica with system-scope atomics over NVLink, fences, and bumps every
//           replica's count[slot] — the reduction happens in the remote L2 atomic units, no message is sent twice
//           and nobody waits for a ring to go round;
//   collect (before the apply kernel): waits until its own count[slot] has reached `world`, hands keys[slot] to the
//           apply kernel and re-arms the slot.  A slot is reused two iterations later; by then every peer has
//           passed its own collect of the iteration in between, which it can only do after this replica's publish
//           of that iteration, which is stream-ordered after this collect — so the re-arm cannot race a publish.
// The wait is bounded (a peer that died must not hang this GPU): on timeout every instance is marked done = 3.
// No reference counterpart (the reference has no solver state to exchange).
#include <vector>

#include "problem.cuh"

namespace rt {

namespace {

__global__ void xchg_publish_kernel(const unsigned long long *__restrict__ local_best,
                                    unsigned long long *const *__restrict__ peers, int world, int batch, int slot)
{
    for (int b = threadIdx.x; b < batch; b += blockDim.x) {
        const unsigned long long k = local_best[b];
        if (k != kNoMove)
            for (int r = 0; r < world; ++r) atomicMin_system(peers[r] + static_cast<size_t>(slot) * batch + b, k);
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world)
        atomicAdd_system(peers[threadIdx.x] + 2 * static_cast<size_t>(batch) + slot, 1ull);
}

__global__ void xchg_collect_kernel(unsigned long long *__restrict__ xbuf, int world, int batch, int slot,
                                    unsigned long long *__restrict__ best_out, int *__restrict__ done,
                                    long long timeout_cycles)
{
    __shared__ int s_ok;
    volatile unsigned long long *count = xbuf + 2 * static_cast<size_t>(batch) + slot;
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        int ok = 1;
        while (*count < static_cast<unsigned long long>(world)) {
            __nanosleep(200);
            if (clock64() - t0 > timeout_cycles) {
                ok = 0;
                break;
            }
        }
        s_ok = ok;
    }
    __syncthreads();
    __threadfence_system();
    const unsigned long long none = 0xFFFFFFFFFFFFFFFFull;
    for (int b = threadIdx.x; b < batch; b += blockDim.x) {
        unsigned long long *cell = xbuf + static_cast<size_t>(slot) * batch + b;
        const unsigned long long got = *reinterpret_cast<volatile unsigned long long *>(cell);
        best_out[b] = s_ok ? got : none;
        *cell = none;  // re-arm
        if (!s_ok) done[b] = 3;
    }
    __syncthreads();
    if (threadIdx.x == 0) *count = 0;
}

__global__ void xchg_init_kernel(unsigned long long *xbuf, int batch)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 2 * batch) xbuf[i] = kNoMove;
    else if (i < 2 * batch + 2) xbuf[i] = 0;
}

}  // namespace

}  // namespace rt

using namespace rt;

extern "C" {

int rt_problem_exchange_handle_bytes(void) { return static_cast<int>(sizeof(cudaIpcMemHandle_t)); }

int rt_problem_exchange_export(rt_problem *p, void *handle_out)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!handle_out) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_exchange_export: NULL argument");
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    if (!p->xbuf) {
        const size_t cnt = 2 * static_cast<size_t>(p->batch) + 2;
        void *raw = nullptr;
        RT_CUDA(e, cudaMalloc(&raw, cnt * 8));  // plain cudaMalloc: IPC handles cannot refer to pool memory
        p->xbuf = static_cast<unsigned long long *>(raw);
        xchg_init_kernel<<<static_cast<unsigned>((cnt + 255) / 256), 256, 0, e->stream>>>(p->xbuf, p->batch);
        RT_CUDA(e, cudaGetLastError());
        RT_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    cudaIpcMemHandle_t h;
    RT_CUDA(e, cudaIpcGetMemHandle(&h, p->xbuf));
    memcpy(handle_out, &h, sizeof(h));
    return RT_OK;
}

int rt_problem_exchange_connect(rt_problem *p, const void *handles, int rank, int world)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!handles || world < 1 || world > 64 || rank < 0 || rank >= world)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_exchange_connect: bad argument");
    if (!p->xbuf) return fail(e, RT_ERR_STATE, "rt_problem_exchange_connect: export this replica's handle first");
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    std::vector<unsigned long long *> ptrs(world);
    for (int r = 0; r < world; ++r) {
        if (r == rank) {
            ptrs[r] = p->xbuf;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char *>(handles) + static_cast<size_t>(r) * sizeof(h), sizeof(h));
        void *mapped = nullptr;
        RT_CUDA(e, cudaIpcOpenMemHandle(&mapped, h, cudaIpcMemLazyEnablePeerAccess));
        ptrs[r] = static_cast<unsigned long long *>(mapped);
        p->xpeer_mapped[p->n_xpeer_mapped++] = mapped;
    }
    if (!p->xpeers) {
        void *raw = nullptr;
        RT_CUDA(e, cudaMalloc(&raw, 64 * sizeof(void *)));
        p->xpeers = static_cast<unsigned long long **>(raw);
    }
    RT_CUDA(e, cudaMemcpy(p->xpeers, ptrs.data(), world * sizeof(void *), cudaMemcpyHostToDevice));
    p->xrank = rank;
    p->xworld = world;
    p->xiter = 0;
    return RT_OK;
}

int rt_problem_eval_moves_xchg_dev(rt_problem *p, int families, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!(families & 7)) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_eval_moves_xchg_dev: bad argument");
    if (!p->has_routes || p->xworld < 1) return fail(e, RT_ERR_STATE, "rt_problem_eval_moves_xchg_dev: routes and exchange first");
    if (!(p->shard_count == 1 || (p->shard_count == p->xworld && p->shard_index == p->xrank)))
        return fail(e, RT_ERR_STATE, "rt_problem_eval_moves_xchg_dev: shard does not match the exchange rank");
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    RT_TRY(eval_moves_launch(p, families, p->xrank, p->xworld, p->best, false, st));
    xchg_publish_kernel<<<1, 256, 0, st>>>(p->best, p->xpeers, p->xworld, p->batch, p->xiter & 1);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

int rt_problem_collect_keys_xchg_dev(rt_problem *p, void *keys_d, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!keys_d) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_collect_keys_xchg_dev: NULL argument");
    if (!p->has_routes || p->xworld < 1) return fail(e, RT_ERR_STATE, "rt_problem_collect_keys_xchg_dev: routes and exchange first");
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    xchg_collect_kernel<<<1, 256, 0, st>>>(p->xbuf, p->xworld, p->batch, p->xiter & 1,
                                           static_cast<unsigned long long *>(keys_d), p->done, 4000000000LL);
    RT_CUDA(e, cudaGetLastError());
    ++p->xiter;
    return RT_OK;
}

int rt_problem_apply_moves_xchg_dev(rt_problem *p, int improving_only, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!p->has_routes || p->xworld < 1) return fail(e, RT_ERR_STATE, "rt_problem_apply_moves_xchg_dev: routes and exchange first");
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    const long long timeout_cycles = 4000000000LL;  // ~2 s at 1.9 GHz: a dead peer must not hang this GPU
    xchg_collect_kernel<<<1, 256, 0, st>>>(p->xbuf, p->xworld, p->batch, p->xiter & 1, p->best, p->done, timeout_cycles);
    RT_CUDA(e, cudaGetLastError());
    ++p->xiter;
    return apply_moves_launch(p, p->best, false, improving_only != 0, st);
}

}  // extern "C"
