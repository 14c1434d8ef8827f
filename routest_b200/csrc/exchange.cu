// exchange.cu — the one exchange step of a multi-GPU search, done over NVLink peer memory instead of a collective
// library call: 8 bytes per instance, latency bound.
//
// Every replica owns a small slot buffer in its own HBM that its peers map through CUDA IPC (one process per GPU):
//   keys  [2][batch]  packed move keys (SPEC.md §7), initialised to "no move"
//   count [2]         how many replicas have published into keys[slot]
//   abort [1]         non-zero once ANY replica has given up waiting (written into every replica's buffer)
// Iteration i uses slot i & 1.
//   publish (after the evaluation kernels, same stream): every replica MIN-reduces its shard's best key straight
//           into keys[slot][b] of EVERY replica with system-scope atomics over NVLink, fences, and bumps every
//           replica's count[slot] — the reduction happens in the remote L2 atomic units, no message is sent twice
//           and nobody waits for a ring to go round;
//   collect (before the apply kernel): waits until its own count[slot] has reached `world`, hands keys[slot] to the
//           apply kernel and re-arms the slot.  A slot is reused two iterations later; by then every peer has
//           passed its own collect of the iteration in between, which it can only do after this replica's publish
//           of that iteration, which is stream-ordered after this collect — so the re-arm cannot race a publish.
// The wait is bounded (a peer that died must not hang this GPU; ROUTEST_XCHG_TIMEOUT_MS, default 10 s of wall time).
// Giving up is collective: the replica that times out raises the abort word in EVERY replica's buffer, stops
// publishing keys (it still bumps the counters so that nobody waits for it) and marks its instances done = 3; every
// other replica sees the word at its next collect and does the same, so no replica goes on applying keys that were
// computed from routes another one no longer follows.  DistributedSearch.run turns done = 3 into an exception.
// No reference counterpart (the reference has no solver state to exchange).
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "problem.cuh"

namespace rt {

namespace {

__device__ __forceinline__ unsigned long long wall_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// slot buffer layout: keys[2][batch] | count[2] | abort
__device__ __forceinline__ size_t count_at(int batch, int slot) { return 2 * static_cast<size_t>(batch) + slot; }
__device__ __forceinline__ size_t abort_at(int batch) { return 2 * static_cast<size_t>(batch) + 2; }

__global__ void xchg_publish_kernel(const unsigned long long *__restrict__ local_best,
                                    unsigned long long *const *__restrict__ peers, int rank, int world, int batch,
                                    int slot)
{
    // a replica that has aborted keeps the counters moving but contributes no keys (its routes are frozen)
    const bool aborted = *reinterpret_cast<volatile unsigned long long *>(peers[rank] + abort_at(batch)) != 0;
    if (!aborted)
        for (int b = threadIdx.x; b < batch; b += blockDim.x) {
            const unsigned long long k = local_best[b];
            if (k != kNoMove)
                for (int r = 0; r < world; ++r) atomicMin_system(peers[r] + static_cast<size_t>(slot) * batch + b, k);
        }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) atomicAdd_system(peers[threadIdx.x] + count_at(batch, slot), 1ull);
}

__global__ void xchg_collect_kernel(unsigned long long *__restrict__ xbuf,
                                    unsigned long long *const *__restrict__ peers, int world, int batch, int slot,
                                    unsigned long long *__restrict__ best_out, int *__restrict__ done,
                                    unsigned long long timeout_ns)
{
    __shared__ int s_ok;
    volatile unsigned long long *count = xbuf + count_at(batch, slot);
    volatile unsigned long long *abort_word = xbuf + abort_at(batch);
    if (threadIdx.x == 0) {
        const unsigned long long t0 = wall_ns();
        int ok = 1;
        while (*count < static_cast<unsigned long long>(world)) {
            __nanosleep(200);
            if (*abort_word != 0 || wall_ns() - t0 > timeout_ns) {
                ok = 0;
                break;
            }
        }
        if (*abort_word != 0) ok = 0;
        s_ok = ok;
    }
    __syncthreads();
    __threadfence_system();
    if (!s_ok && threadIdx.x < world)  // collective abort: every replica learns of it at its next collect at the latest
        atomicExch_system(peers[threadIdx.x] + abort_at(batch), 1ull);
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
    else if (i < 2 * batch + 3) xbuf[i] = 0;
}


// ---- the table of a batched iteration (SPEC.md §8.1) -----------------------------------------------------------
// R^2 + R keys per instance instead of one: too many for an atomic each, so the exchange is an all-gather into peer
// memory followed by a local minimum.  Every replica owns, per slot, one copy of the table PER RANK:
//   tab [2][world][entries]   rank r's shard table lands in tab[slot][r] of EVERY replica (plain 16-byte stores over
//                             NVLink, coalesced), then a fence, then count[slot] of every replica is bumped
//   count[2], uses[2], abort  count only grows: use number u of a slot is complete at world * (u + 1); uses[] lives on
//                             the device so that the kernels can be replayed from a CUDA graph
// collect waits for its own count, takes the minimum over the world copies (L1 bypassed: the lines were written by
// other GPUs) into the caller's table and bumps uses[slot].  Same slot discipline and same bounded, collective abort
// as the single-key exchange above.
constexpr int kTabCtlWords = 8;  // buffer head: count[2], uses[2], abort, blocks-done[2], pad; the tables follow

__global__ void __launch_bounds__(256) xtab_publish_kernel(const long long *__restrict__ table,
                                                           unsigned long long *const *__restrict__ peers, int rank,
                                                           int world, size_t entries, int slot)
{
    __shared__ int s_last;
    unsigned long long *own = peers[rank];
    const bool aborted = *reinterpret_cast<volatile unsigned long long *>(own + 4) != 0;
    if (!aborted) {
        const size_t pairs = entries / 2;  // entries is even (padded by the host)
        const longlong2 *src = reinterpret_cast<const longlong2 *>(table);
        for (int r = 0; r < world; ++r) {
            longlong2 *dst = reinterpret_cast<longlong2 *>(reinterpret_cast<long long *>(peers[r] + kTabCtlWords) +
                                                           (static_cast<size_t>(slot) * world + rank) * entries);
            for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < pairs;
                 i += static_cast<size_t>(gridDim.x) * blockDim.x)
                dst[i] = src[i];
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(own + 5 + slot, 1ull) == gridDim.x - 1;  // blocks of this launch done
    __syncthreads();
    if (s_last) {  // the last block: every block's stores are fenced, tell every replica
        if (threadIdx.x == 0) own[5 + slot] = 0;
        if (threadIdx.x < world) atomicAdd_system(peers[threadIdx.x] + slot, 1ull);
    }
}

__global__ void __launch_bounds__(256) xtab_collect_kernel(unsigned long long *__restrict__ own,
                                                           unsigned long long *const *__restrict__ peers, int world,
                                                           size_t entries, int slot, long long *__restrict__ table,
                                                           int *__restrict__ done, int batch,
                                                           unsigned long long timeout_ns, unsigned *__restrict__ finished)
{
    __shared__ int s_ok;
    volatile unsigned long long *count = own + slot;
    volatile unsigned long long *abort_word = own + 4;
    if (threadIdx.x == 0) {
        const unsigned long long want = static_cast<unsigned long long>(world) * (own[2 + slot] + 1);
        const unsigned long long t0 = wall_ns();
        int ok = 1;
        while (*count < want) {
            __nanosleep(200);
            if (*abort_word != 0 || wall_ns() - t0 > timeout_ns) {
                ok = 0;
                break;
            }
        }
        if (*abort_word != 0) ok = 0;
        s_ok = ok;
    }
    __syncthreads();
    __threadfence_system();
    if (!s_ok && blockIdx.x == 0 && threadIdx.x < world) atomicExch_system(peers[threadIdx.x] + 4, 1ull);
    const long long *tab = reinterpret_cast<const long long *>(own + kTabCtlWords) + static_cast<size_t>(slot) * world * entries;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < entries;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        long long m = 0x7FFFFFFFFFFFFFFFll;
        if (s_ok)
            for (int r = 0; r < world; ++r) m = min(m, __ldcg(tab + static_cast<size_t>(r) * entries + i));
        table[i] = m;
    }
    if (!s_ok && blockIdx.x == 0)
        for (int b = threadIdx.x; b < batch; b += blockDim.x) done[b] = 3;
    // the last block to finish advances the slot's use count (every block has read it by then)
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(finished, 1u) == gridDim.x - 1) {
        *finished = 0;
        own[2 + slot] += 1;
    }
}

__global__ void xtab_init_kernel(unsigned long long *base) { if (threadIdx.x < kTabCtlWords) base[threadIdx.x] = 0; }

unsigned long long xchg_timeout_ns()
{
    long long ms = 10000;
    if (const char *ev = getenv("ROUTEST_XCHG_TIMEOUT_MS")) ms = atoll(ev);
    if (ms < 1) ms = 1;
    return static_cast<unsigned long long>(ms) * 1000000ull;
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
        const size_t cnt = 2 * static_cast<size_t>(p->batch) + 3;
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
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    if (!p->xbuf) return fail(e, RT_ERR_STATE, "rt_problem_exchange_connect: export this replica's handle first");
    // a second connect (a search constructed again, a reconnect after an abort) starts from a clean slate: earlier
    // mappings are closed, the slots re-armed and the abort word cleared
    RT_CUDA(e, cudaDeviceSynchronize());
    for (int i = 0; i < p->n_xpeer_mapped; ++i) cudaIpcCloseMemHandle(p->xpeer_mapped[i]);
    p->n_xpeer_mapped = 0;
    p->xworld = 0;
    {
        const size_t cnt = 2 * static_cast<size_t>(p->batch) + 3;
        xchg_init_kernel<<<static_cast<unsigned>((cnt + 255) / 256), 256, 0, e->stream>>>(p->xbuf, p->batch);
        RT_CUDA(e, cudaGetLastError());
        RT_CUDA(e, cudaStreamSynchronize(e->stream));
    }
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
        if (p->n_xpeer_mapped >= 64) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_exchange_connect: too many peers");
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
    xchg_publish_kernel<<<1, 256, 0, st>>>(p->best, p->xpeers, p->xrank, p->xworld, p->batch, p->xiter & 1);
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
    xchg_collect_kernel<<<1, 256, 0, st>>>(p->xbuf, p->xpeers, p->xworld, p->batch, p->xiter & 1,
                                           static_cast<unsigned long long *>(keys_d), p->done, xchg_timeout_ns());
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
    // bounded wait: a dead peer must not hang this GPU
    xchg_collect_kernel<<<1, 256, 0, st>>>(p->xbuf, p->xpeers, p->xworld, p->batch, p->xiter & 1, p->best, p->done,
                                           xchg_timeout_ns());
    RT_CUDA(e, cudaGetLastError());
    ++p->xiter;
    return apply_moves_launch(p, p->best, false, improving_only != 0, st);
}


// ---- table exchange entry points (batched iterations over peer memory) -----------------------------------------

static size_t padded_entries(const rt_problem *p) { return (rt::batch_table_entries(p) + 1) & ~static_cast<size_t>(1); }

int rt_problem_table_exchange_export(rt_problem *p, int world, void *handle_out)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!handle_out || world < 1 || world > 64) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_table_exchange_export: bad argument");
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    if (!p->has_routes || !batch_usable(p) || batch_ensure(p) != RT_OK)
        return fail(e, RT_ERR_STATE, "rt_problem_table_exchange_export: routes first (at most 1,024 per instance)");
    const size_t words = kTabCtlWords + 2 * static_cast<size_t>(world) * padded_entries(p) + 2;
    if (p->xtab && p->xtab_words < words) {  // the table grew (more routes): a new buffer, the peers reconnect
        RT_CUDA(e, cudaDeviceSynchronize());
        RT_CUDA(e, cudaFree(p->xtab));
        p->xtab = nullptr;
    }
    if (!p->xtab) {
        void *raw = nullptr;
        RT_CUDA(e, cudaMalloc(&raw, words * 8));  // plain cudaMalloc: IPC handles cannot refer to pool memory
        p->xtab = static_cast<unsigned long long *>(raw);
        p->xtab_words = words;
    }
    xtab_init_kernel<<<1, 32, 0, e->stream>>>(p->xtab);
    RT_CUDA(e, cudaGetLastError());
    RT_CUDA(e, cudaStreamSynchronize(e->stream));
    cudaIpcMemHandle_t h;
    RT_CUDA(e, cudaIpcGetMemHandle(&h, p->xtab));
    memcpy(handle_out, &h, sizeof(h));
    return RT_OK;
}

int rt_problem_table_exchange_connect(rt_problem *p, const void *handles, int rank, int world)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!handles || world < 1 || world > 64 || rank < 0 || rank >= world)
        return fail(e, RT_ERR_INVALID_ARG, "rt_problem_table_exchange_connect: bad argument");
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    if (!p->xtab) return fail(e, RT_ERR_STATE, "rt_problem_table_exchange_connect: export this replica's handle first");
    RT_CUDA(e, cudaDeviceSynchronize());
    for (int i = 0; i < p->n_xtab_mapped; ++i) cudaIpcCloseMemHandle(p->xtab_mapped[i]);
    p->n_xtab_mapped = 0;
    p->xtab_world = 0;
    std::vector<unsigned long long *> ptrs(world);
    for (int r = 0; r < world; ++r) {
        if (r == rank) {
            ptrs[r] = p->xtab;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char *>(handles) + static_cast<size_t>(r) * sizeof(h), sizeof(h));
        void *mapped = nullptr;
        RT_CUDA(e, cudaIpcOpenMemHandle(&mapped, h, cudaIpcMemLazyEnablePeerAccess));
        ptrs[r] = static_cast<unsigned long long *>(mapped);
        p->xtab_mapped[p->n_xtab_mapped++] = mapped;
    }
    if (!p->xtab_peers) {
        void *raw = nullptr;
        RT_CUDA(e, cudaMalloc(&raw, 64 * sizeof(void *) + 16));
        p->xtab_peers = static_cast<unsigned long long **>(raw);
        RT_CUDA(e, cudaMemset(raw, 0, 64 * sizeof(void *) + 16));
    }
    RT_CUDA(e, cudaMemcpy(p->xtab_peers, ptrs.data(), world * sizeof(void *), cudaMemcpyHostToDevice));
    p->xtab_rank = rank;
    p->xtab_world = world;
    p->xtab_iter = 0;
    return RT_OK;
}

int rt_problem_eval_table_xchg_dev(rt_problem *p, int families, void *table_d, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!table_d || !(families & 7)) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_eval_table_xchg_dev: bad argument");
    if (!p->has_routes || p->xtab_world < 1) return fail(e, RT_ERR_STATE, "rt_problem_eval_table_xchg_dev: routes and table exchange first");
    if (!(p->shard_count == 1 || (p->shard_count == p->xtab_world && p->shard_index == p->xtab_rank)))
        return fail(e, RT_ERR_STATE, "rt_problem_eval_table_xchg_dev: shard does not match the exchange rank");
    const size_t entries = padded_entries(p);
    if (kTabCtlWords + 2 * static_cast<size_t>(p->xtab_world) * entries + 2 > p->xtab_words)
        return fail(e, RT_ERR_STATE, "rt_problem_eval_table_xchg_dev: the table outgrew the exchange buffer (export and connect again)");
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    RT_TRY(batch_ensure(p));
    RT_TRY(eval_table_launch(p, families, p->xtab_rank, p->xtab_world, static_cast<unsigned long long *>(table_d), true, st));
    const unsigned blocks = static_cast<unsigned>(std::min<size_t>(kNumSMs, (entries / 2 + 255) / 256));
    xtab_publish_kernel<<<blocks, 256, 0, st>>>(static_cast<const long long *>(table_d), p->xtab_peers, p->xtab_rank,
                                                p->xtab_world, entries, p->xtab_iter & 1);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

int rt_problem_apply_batch_xchg_dev(rt_problem *p, void *table_d, void *stream)
{
    if (!p) return RT_ERR_INVALID_ARG;
    rt_engine *e = p->e;
    if (!table_d) return fail(e, RT_ERR_INVALID_ARG, "rt_problem_apply_batch_xchg_dev: NULL argument");
    if (!p->has_routes || p->xtab_world < 1) return fail(e, RT_ERR_STATE, "rt_problem_apply_batch_xchg_dev: routes and table exchange first");
    DeviceGuard dg(e->device);
    cudaStream_t st = pick_stream(e, stream);
    const size_t entries = padded_entries(p);
    const unsigned blocks = static_cast<unsigned>(std::min<size_t>(kNumSMs, (entries + 255) / 256));
    unsigned *finished = reinterpret_cast<unsigned *>(p->xtab_peers + 64);
    xtab_collect_kernel<<<blocks, 256, 0, st>>>(p->xtab, p->xtab_peers, p->xtab_world, entries, p->xtab_iter & 1,
                                                static_cast<long long *>(table_d), p->done, p->batch, xchg_timeout_ns(),
                                                finished);
    RT_CUDA(e, cudaGetLastError());
    ++p->xtab_iter;
    RT_TRY(batch_ensure(p));
    return batch_select_apply(p, static_cast<const unsigned long long *>(table_d), true, st);
}

}  // extern "C"
