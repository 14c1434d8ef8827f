// eta.cu — batched ETA scoring: fused feature build + tree-ensemble evaluation (SPEC.md §9).
//
// Takes over Flaskr/ml.py:35-53 of the reference: the 12-feature row (ml.py:35-48) and model.predict on it
// (ml.py:51-53), which the reference does one row per call through a pandas DataFrame.  Here a thread owns a few
// stop pairs; their feature rows are built in shared memory ([feature][slot] layout, so the data-dependent feature
// index of a tree node is a conflict-free LDS), the ensemble streams through shared memory in chunks of whole trees,
// and every pair accumulates its leaves in tree order in FP32 (bit-reproducible against the scalar evaluator).
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace rt {

struct TreeNode {
    float value;
    unsigned short left;
    unsigned char feature;
    unsigned char flags;  // bit 0 leaf, bit 1 default-left
};
static_assert(sizeof(TreeNode) == 8, "node layout is part of the model format");

struct EtaModel {
    int n_trees = 0, n_features = 0;
    float base = 0.f;
    unsigned *offsets_d = nullptr;   // [n_trees + 1]
    TreeNode *nodes_d = nullptr;     // [offsets[n_trees]]
    int *chunk_first_d = nullptr;    // [n_chunks + 1] first tree of each shared-memory chunk
    int n_chunks = 0, total_nodes = 0;
    int max_depth = 0;               // edges on the longest root-to-leaf path over all trees
};

namespace {

constexpr int kEtaThreads = 256;
constexpr int kPairsPerThread = 4;
constexpr int kFeatures = 12;
constexpr int kFeatRows = kFeatures + 1;  // row 12 holds -inf: the feature a leaf "tests" so that it loops on itself
constexpr int kChunkNodes = 6144;   // 48 KB of nodes per chunk + 48 KB of feature rows => 2 blocks per SM
constexpr int kChunkTrees = 512;    // trees per chunk (their offsets sit in static shared memory)

__device__ __forceinline__ void build_features(double dist_m, unsigned char w, unsigned char t, unsigned char wd,
                                               unsigned char hr, float age, float (&f)[kFeatures])
{
#pragma unroll
    for (int k = 0; k < 4; ++k) f[k] = w == k ? 1.0f : 0.0f;       // ml.py:36-39
#pragma unroll
    for (int k = 0; k < 4; ++k) f[4 + k] = t == k ? 1.0f : 0.0f;   // ml.py:40-43
    f[8] = static_cast<float>(wd);                                  // ml.py:44
    f[9] = static_cast<float>(hr);                                  // ml.py:45
    f[10] = static_cast<float>(dist_m / 1000.0);                    // ml.py:46, then the model's float32 view
    f[11] = age == 0.0f ? 30.0f : age;                              // ml.py:47: float(driver_age or 30.0)
}

__global__ void __launch_bounds__(kEtaThreads) eta_features_kernel(const double *__restrict__ dist_m,
                                                                   const unsigned char *__restrict__ weather,
                                                                   const unsigned char *__restrict__ traffic,
                                                                   const unsigned char *__restrict__ weekday,
                                                                   const unsigned char *__restrict__ hour,
                                                                   const float *__restrict__ age, long long n,
                                                                   float *__restrict__ rows)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float f[kFeatures];
    build_features(dist_m[i], weather[i], traffic[i], weekday[i], hour[i], age[i], f);
#pragma unroll
    for (int k = 0; k < kFeatures; ++k) rows[i * kFeatures + k] = f[k];
}

// Device copies of the nodes are re-encoded at load time so that the traversal needs no leaf test: a leaf points to
// itself (left = own index) and tests feature row 12 (-inf), so "x < value" holds and it stays where it is.  Every
// pair then takes exactly max_depth steps per tree; the four pairs of a thread advance in lock step (four
// independent LDS chains in flight instead of one).  Rows with a NaN feature are rare and take the exact,
// branchy walk (default directions) instead.
__global__ void __launch_bounds__(kEtaThreads) eta_score_kernel(
    const double *__restrict__ dist_m, const unsigned char *__restrict__ weather,
    const unsigned char *__restrict__ traffic, const unsigned char *__restrict__ weekday,
    const unsigned char *__restrict__ hour, const float *__restrict__ age, long long n, float *__restrict__ out,
    const TreeNode *__restrict__ nodes, const unsigned *__restrict__ offsets, const int *__restrict__ chunk_first,
    int n_chunks, float base, int max_depth)
{
    extern __shared__ __align__(16) unsigned char smem[];
    float *sfeat = reinterpret_cast<float *>(smem);  // [kFeatRows][kPairsPerThread * kEtaThreads]
    constexpr int kSlots = kPairsPerThread * kEtaThreads;
    // the chunk's nodes as two 4-byte arrays (value | left, feature, flags): nodes of one tree level are adjacent, so up
    // to 32 of them sit in 32 different banks and a warp's data-dependent fetch is conflict-free whatever its lanes
    // point at (as 8-byte structs two nodes 16 apart shared a bank pair: +60 % shared-memory wavefronts measured)
    float *svals = reinterpret_cast<float *>(smem + sizeof(float) * kFeatRows * kSlots);
    unsigned *smeta = reinterpret_cast<unsigned *>(svals + kChunkNodes);
    __shared__ unsigned s_off[kChunkTrees + 1];  // node offset of every tree of the chunk

    const int tid = threadIdx.x;
    const long long base_i = static_cast<long long>(blockIdx.x) * kSlots;
    float acc[kPairsPerThread];
    bool any_nan = false;
#pragma unroll
    for (int j = 0; j < kPairsPerThread; ++j) {
        const int slot = j * kEtaThreads + tid;  // coalesced input loads
        const long long i = base_i + slot;
        float f[kFeatures];
        if (i < n) build_features(dist_m[i], weather[i], traffic[i], weekday[i], hour[i], age[i], f);
        else {
#pragma unroll
            for (int k = 0; k < kFeatures; ++k) f[k] = 0.f;
        }
#pragma unroll
        for (int k = 0; k < kFeatures; ++k) {
            sfeat[k * kSlots + slot] = f[k];
            any_nan = any_nan || (f[k] != f[k]);
        }
        sfeat[kFeatures * kSlots + slot] = -INFINITY;
        acc[j] = base;
    }

    for (int c = 0; c < n_chunks; ++c) {
        const int t0 = chunk_first[c], t1 = chunk_first[c + 1];
        const unsigned n0 = offsets[t0], n1 = offsets[t1];
        __syncthreads();  // previous chunk fully consumed (and the feature rows visible)
        {
            const uint2 *src = reinterpret_cast<const uint2 *>(nodes + n0);
            for (unsigned k = tid; k < n1 - n0; k += kEtaThreads) {
                const uint2 nd = src[k];
                svals[k] = __uint_as_float(nd.x);
                smeta[k] = nd.y;  // left | feature << 16 | flags << 24
            }
            for (int k = tid; k <= t1 - t0; k += kEtaThreads) s_off[k] = offsets[t0 + k] - n0;
        }
        __syncthreads();
        if (!any_nan) {
            for (int t = 0; t < t1 - t0; ++t) {
                const float *tv = svals + s_off[t];
                const unsigned *tm = smeta + s_off[t];
                float val[kPairsPerThread];
                unsigned meta[kPairsPerThread];
#pragma unroll
                for (int j = 0; j < kPairsPerThread; ++j) {
                    val[j] = tv[0];
                    meta[j] = tm[0];
                }
                for (int d = 0; d < max_depth; ++d) {
#pragma unroll
                    for (int j = 0; j < kPairsPerThread; ++j) {
                        const float v = sfeat[((meta[j] >> 16) & 0xFFu) * kSlots + j * kEtaThreads + tid];
                        const unsigned next = (meta[j] & 0xFFFFu) + (v < val[j] ? 0u : 1u);
                        val[j] = tv[next];
                        meta[j] = tm[next];
                    }
                }
#pragma unroll
                for (int j = 0; j < kPairsPerThread; ++j) acc[j] += val[j];
            }
        } else {
            for (int t = 0; t < t1 - t0; ++t) {
                const float *tv = svals + s_off[t];
                const unsigned *tm = smeta + s_off[t];
#pragma unroll
                for (int j = 0; j < kPairsPerThread; ++j) {
                    const int slot = j * kEtaThreads + tid;
                    float val = tv[0];
                    unsigned meta = tm[0];
                    while (!((meta >> 24) & 1u)) {
                        const float v = sfeat[((meta >> 16) & 0xFFu) * kSlots + slot];
                        const bool left = isnan(v) ? ((meta >> 25) & 1u) : (v < val);
                        const unsigned next = (meta & 0xFFFFu) + (left ? 0u : 1u);
                        val = tv[next];
                        meta = tm[next];
                    }
                    acc[j] += val;
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kPairsPerThread; ++j) {
        const long long i = base_i + j * kEtaThreads + tid;
        if (i < n) out[i] = acc[j];
    }
}

constexpr size_t kEtaSmem = sizeof(float) * kFeatRows * kPairsPerThread * kEtaThreads + sizeof(TreeNode) * kChunkNodes;

}  // namespace

int eta_model_free(rt_engine *e);

int eta_model_load(rt_engine *e, const unsigned char *blob, size_t bytes)
{
    if (bytes < 16 || memcmp(blob, "RTE1", 4) != 0) return fail(e, RT_ERR_INVALID_ARG, "model blob: bad magic (want RTE1)");
    unsigned n_trees, n_feat;
    float base;
    memcpy(&n_trees, blob + 4, 4);
    memcpy(&n_feat, blob + 8, 4);
    memcpy(&base, blob + 12, 4);
    if (n_feat != kFeatures) return fail(e, RT_ERR_INVALID_ARG, "model blob: %u features, this engine builds %d", n_feat, kFeatures);
    if (n_trees > (1u << 20) || bytes < 16 + 4 * static_cast<size_t>(n_trees + 1))
        return fail(e, RT_ERR_INVALID_ARG, "model blob: truncated header");
    std::vector<unsigned> off(n_trees + 1);
    memcpy(off.data(), blob + 16, 4 * static_cast<size_t>(n_trees + 1));
    const size_t total = off[n_trees], hdr = 16 + 4 * static_cast<size_t>(n_trees + 1);
    if (off[0] != 0 || bytes < hdr + 8 * total) return fail(e, RT_ERR_INVALID_ARG, "model blob: truncated node table");
    const TreeNode *nodes = reinterpret_cast<const TreeNode *>(blob + hdr);
    std::vector<TreeNode> dev_nodes(nodes, nodes + total);   // device encoding: leaves loop on themselves
    std::vector<int> depth(total, 0);
    int max_depth = 0;
    // validate every tree: offsets increasing, children inside the tree, features in range, every walk ends in a leaf
    std::vector<int> chunk_first{0};
    unsigned chunk_nodes = 0;
    for (unsigned t = 0; t < n_trees; ++t) {
        if (off[t + 1] <= off[t]) return fail(e, RT_ERR_INVALID_ARG, "model blob: tree %u is empty", t);
        const unsigned cnt = off[t + 1] - off[t];
        if (cnt > 65535 || cnt > kChunkNodes)
            return fail(e, RT_ERR_INVALID_ARG, "model blob: tree %u has %u nodes (limit %d)", t, cnt, kChunkNodes);
        for (unsigned k = 0; k < cnt; ++k) {
            const TreeNode &nd = nodes[off[t] + k];
            if (nd.flags & 1) {
                if (!std::isfinite(nd.value)) return fail(e, RT_ERR_INVALID_ARG, "model blob: tree %u leaf %u is not finite", t, k);
                TreeNode &dn = dev_nodes[off[t] + k];
                dn.left = static_cast<unsigned short>(k);   // self loop
                dn.feature = kFeatures;                     // the constant -inf row: "x < value" always holds
                max_depth = std::max(max_depth, depth[off[t] + k]);
                continue;
            }
            if (nd.feature >= kFeatures || nd.left <= k || static_cast<unsigned>(nd.left) + 1 >= cnt)
                return fail(e, RT_ERR_INVALID_ARG, "model blob: tree %u node %u is malformed", t, k);
            depth[off[t] + nd.left] = depth[off[t] + nd.left + 1] = depth[off[t] + k] + 1;  // children come after parents
        }
        if (chunk_nodes + cnt > kChunkNodes || static_cast<size_t>(t - chunk_first.back()) + 1 > kChunkTrees) {
            chunk_first.push_back(static_cast<int>(t));
            chunk_nodes = 0;
        }
        chunk_nodes += cnt;
    }
    chunk_first.push_back(static_cast<int>(n_trees));
    RT_TRY(eta_model_free(e));
    EtaModel *m = new EtaModel();
    m->n_trees = static_cast<int>(n_trees);
    m->n_features = static_cast<int>(n_feat);
    m->base = base;
    m->total_nodes = static_cast<int>(total);
    m->max_depth = max_depth;
    m->n_chunks = static_cast<int>(chunk_first.size()) - 1;
    if (n_trees == 0) m->n_chunks = 0;
    cudaError_t st = cudaMalloc(&m->offsets_d, 4 * (n_trees + 1));
    if (st == cudaSuccess) st = cudaMalloc(&m->nodes_d, 8 * std::max<size_t>(total, 1));
    if (st == cudaSuccess) st = cudaMalloc(&m->chunk_first_d, 4 * chunk_first.size());
    if (st == cudaSuccess) st = cudaMemcpy(m->offsets_d, off.data(), 4 * (n_trees + 1), cudaMemcpyHostToDevice);
    if (st == cudaSuccess && total) st = cudaMemcpy(m->nodes_d, dev_nodes.data(), 8 * total, cudaMemcpyHostToDevice);
    if (st == cudaSuccess) st = cudaMemcpy(m->chunk_first_d, chunk_first.data(), 4 * chunk_first.size(), cudaMemcpyHostToDevice);
    if (st == cudaSuccess)
        st = cudaFuncSetAttribute(eta_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kEtaSmem));
    if (st != cudaSuccess) {
        cudaFree(m->offsets_d);
        cudaFree(m->nodes_d);
        cudaFree(m->chunk_first_d);
        delete m;
        return fail(e, RT_ERR_CUDA, "loading the ETA model failed: %s", cudaGetErrorString(st));
    }
    e->eta_model = m;
    return RT_OK;
}

int eta_model_free(rt_engine *e)
{
    EtaModel *m = static_cast<EtaModel *>(e->eta_model);
    if (!m) return RT_OK;
    cudaStreamSynchronize(e->stream);
    cudaFree(m->offsets_d);
    cudaFree(m->nodes_d);
    cudaFree(m->chunk_first_d);
    delete m;
    e->eta_model = nullptr;
    return RT_OK;
}

int eta_score_launch(rt_engine *e, const double *dist_m, const unsigned char *weather, const unsigned char *traffic,
                     const unsigned char *weekday, const unsigned char *hour, const float *age, long long n, float *out,
                     cudaStream_t st)
{
    EtaModel *m = static_cast<EtaModel *>(e->eta_model);
    if (!m) return fail(e, RT_ERR_STATE, "no ETA model loaded (rt_eta_load_model)");
    if (n <= 0) return RT_OK;
    const long long per_block = static_cast<long long>(kPairsPerThread) * kEtaThreads;
    const long long blocks = (n + per_block - 1) / per_block;
    if (blocks > 0x7fffffffLL) return fail(e, RT_ERR_INVALID_ARG, "too many pairs for one launch");
    eta_score_kernel<<<static_cast<unsigned>(blocks), kEtaThreads, kEtaSmem, st>>>(
        dist_m, weather, traffic, weekday, hour, age, n, out, m->nodes_d, m->offsets_d, m->chunk_first_d, m->n_chunks,
        m->base, m->max_depth);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

int eta_features_launch(rt_engine *e, const double *dist_m, const unsigned char *weather, const unsigned char *traffic,
                        const unsigned char *weekday, const unsigned char *hour, const float *age, long long n,
                        float *rows, cudaStream_t st)
{
    if (n <= 0) return RT_OK;
    eta_features_kernel<<<static_cast<unsigned>((n + kEtaThreads - 1) / kEtaThreads), kEtaThreads, 0, st>>>(
        dist_m, weather, traffic, weekday, hour, age, n, rows);
    RT_CUDA(e, cudaGetLastError());
    return RT_OK;
}

}  // namespace rt

using namespace rt;

extern "C" {

int rt_eta_load_model(rt_engine *e, const void *blob, size_t bytes)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!blob) return fail(e, RT_ERR_INVALID_ARG, "rt_eta_load_model: NULL blob");
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard g(e->device);
    return eta_model_load(e, static_cast<const unsigned char *>(blob), bytes);
}

int rt_eta_model_info(rt_engine *e, int32_t *n_trees, int32_t *n_nodes)
{
    if (!e) return RT_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> lk(e->mu);
    EtaModel *m = static_cast<EtaModel *>(e->eta_model);
    if (n_trees) *n_trees = m ? m->n_trees : -1;
    if (n_nodes) *n_nodes = m ? m->total_nodes : -1;
    return RT_OK;
}

int rt_eta_score_dev(rt_engine *e, const double *distance_m_d, const uint8_t *weather_d, const uint8_t *traffic_d,
                     const uint8_t *weekday_d, const uint8_t *hour_d, const float *age_d, int64_t n, float *minutes_d,
                     void *stream)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (n < 0 || (n > 0 && (!distance_m_d || !weather_d || !traffic_d || !weekday_d || !hour_d || !age_d || !minutes_d)))
        return fail(e, RT_ERR_INVALID_ARG, "rt_eta_score_dev: bad argument");
    DeviceGuard g(e->device);
    return eta_score_launch(e, distance_m_d, weather_d, traffic_d, weekday_d, hour_d, age_d, n, minutes_d,
                            pick_stream(e, stream));
}

// host buffers: one staging block  dist | age | out/rows | 4 byte columns
static int eta_host(rt_engine *e, const double *distance_m, const uint8_t *weather, const uint8_t *traffic,
                    const uint8_t *weekday, const uint8_t *hour, const float *age, int64_t n, float *out, bool rows)
{
    if (n < 0 || (n > 0 && (!distance_m || !weather || !traffic || !weekday || !hour || !age || !out)))
        return fail(e, RT_ERR_INVALID_ARG, "rt_eta_*: bad argument");
    if (n == 0) return RT_OK;
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard g(e->device);
    const size_t N = static_cast<size_t>(n), out_f = rows ? 12 * N : N;
    const size_t oD = 0, oA = oD + 8 * N, oO = oA + 4 * ((N + 1) & ~size_t(1)), oB = oO + 4 * ((out_f + 3) & ~size_t(3));
    RT_TRY(reserve(e, e->plane_c[1], oB + 4 * N));
    char *base = static_cast<char *>(e->plane_c[1].p);
    unsigned char *bcols = reinterpret_cast<unsigned char *>(base + oB);
    cudaStream_t st = e->stream;
    RT_CUDA(e, cudaMemcpyAsync(base + oD, distance_m, 8 * N, cudaMemcpyHostToDevice, st));
    RT_CUDA(e, cudaMemcpyAsync(base + oA, age, 4 * N, cudaMemcpyHostToDevice, st));
    RT_CUDA(e, cudaMemcpyAsync(bcols, weather, N, cudaMemcpyHostToDevice, st));
    RT_CUDA(e, cudaMemcpyAsync(bcols + N, traffic, N, cudaMemcpyHostToDevice, st));
    RT_CUDA(e, cudaMemcpyAsync(bcols + 2 * N, weekday, N, cudaMemcpyHostToDevice, st));
    RT_CUDA(e, cudaMemcpyAsync(bcols + 3 * N, hour, N, cudaMemcpyHostToDevice, st));
    const double *dd = reinterpret_cast<const double *>(base + oD);
    const float *ad = reinterpret_cast<const float *>(base + oA);
    float *od = reinterpret_cast<float *>(base + oO);
    if (rows) RT_TRY(eta_features_launch(e, dd, bcols, bcols + N, bcols + 2 * N, bcols + 3 * N, ad, n, od, st));
    else RT_TRY(eta_score_launch(e, dd, bcols, bcols + N, bcols + 2 * N, bcols + 3 * N, ad, n, od, st));
    RT_CUDA(e, cudaMemcpyAsync(out, od, 4 * out_f, cudaMemcpyDeviceToHost, st));
    RT_CUDA(e, cudaStreamSynchronize(st));
    return RT_OK;
}

int rt_eta_score(rt_engine *e, const double *distance_m, const uint8_t *weather, const uint8_t *traffic,
                 const uint8_t *weekday, const uint8_t *hour, const float *age, int64_t n, float *minutes_out)
{
    if (!e) return RT_ERR_INVALID_ARG;
    return eta_host(e, distance_m, weather, traffic, weekday, hour, age, n, minutes_out, false);
}

int rt_eta_features(rt_engine *e, const double *distance_m, const uint8_t *weather, const uint8_t *traffic,
                    const uint8_t *weekday, const uint8_t *hour, const float *age, int64_t n, float *rows_out)
{
    if (!e) return RT_ERR_INVALID_ARG;
    return eta_host(e, distance_m, weather, traffic, weekday, hour, age, n, rows_out, true);
}

}  // extern "C"
