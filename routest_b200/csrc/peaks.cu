// peaks.cu — on-box instruction-rate peaks for the integer kernels' roofline ("FP32/FP64/INT32 peak
// must be measured on the box").  The move-evaluation kernels are bound by the ALU pipe (compares, min/max, selects,
// logic) with plain sums pushed to the FMA pipe as IMAD; their roofline denominator is therefore a warp-instruction
// rate, which this file measures with three straight-line kernels of independent register chains:
//   mode 0  ALU pipe only   (LOP3 / IMNMX alternating: neither fuses with the other)
//   mode 1  FMA pipe only   (IMAD)
//   mode 2  ALU + IMAD interleaved (two half-rate pipes)
//   mode 3  ALU + immediate-form FFMA interleaved
//   mode 4  ALU + IMAD + immediate-form FFMA (2:1:1) — three pipes, the closest of these to the issue limit of one warp
//           instruction per scheduler per clock
// No reference counterpart.
#include "common.cuh"

namespace rt {

namespace {

constexpr int kChains = 8;
constexpr int kOpsPerChainPerIter = 4;
constexpr int kPeakThreads = 256;
constexpr int kPeakBlocksPerSM = 8;

template <int MODE>
__global__ void __launch_bounds__(kPeakThreads) peak_kernel(int iters, int x, int y, int *sink)
{
    int a[kChains], g[kChains];
    float f[kChains];
#pragma unroll
    for (int k = 0; k < kChains; ++k) {
        a[k] = threadIdx.x * (k + 1) + x;
        g[k] = a[k] ^ y;
        f[k] = static_cast<float>(a[k]) * 1e-6f;
    }
    const int u = x ^ 0x5a5a5a5a, v = y | 1;
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < kChains; ++k) {
            if (MODE == 0) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[k]) : "r"(u), "r"(v));
                asm volatile("min.s32 %0, %0, %1;" : "+r"(a[k]) : "r"(y));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xe8;" : "+r"(a[k]) : "r"(v), "r"(u));
                asm volatile("max.s32 %0, %0, %1;" : "+r"(a[k]) : "r"(x));
            } else if (MODE == 1) {
                asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(v), "r"(u));
                asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(u), "r"(v));
                asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(v), "r"(x));
                asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(u), "r"(y));
            } else if (MODE == 2) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[k]) : "r"(u), "r"(v));
                asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(v), "r"(u));
                asm volatile("min.s32 %0, %0, %1;" : "+r"(a[k]) : "r"(y));
                asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[k]) : "r"(u), "r"(x));
            } else if (MODE == 3) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[k]) : "r"(u), "r"(v));
                asm volatile("fma.rn.f32 %0, %0, 0f3F7FFF00, 0f3A000000;" : "+f"(f[k]));
                asm volatile("min.s32 %0, %0, %1;" : "+r"(a[k]) : "r"(y));
                asm volatile("fma.rn.f32 %0, %0, 0f3F800100, 0fBA000000;" : "+f"(f[k]));
            } else {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[k]) : "r"(u), "r"(v));
                asm volatile("fma.rn.f32 %0, %0, 0f3F7FFF00, 0f3A000000;" : "+f"(f[k]));
                asm volatile("min.s32 %0, %0, %1;" : "+r"(a[k]) : "r"(y));
                asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(g[k]) : "r"(v), "r"(u));
            }
        }
    }
    int r = 0;
#pragma unroll
    for (int k = 0; k < kChains; ++k) r ^= a[k] ^ g[k] ^ __float_as_int(f[k]);
    if (r == 0x7fffffff) *sink = r;  // keeps the chains alive; practically never true
}

}  // namespace

}  // namespace rt

using namespace rt;

extern "C" int rt_selftest_pipe_peak(rt_engine *e, int mode, int iters, double *warp_inst_per_s, double *elapsed_ms)
{
    if (!e) return RT_ERR_INVALID_ARG;
    if (!warp_inst_per_s || mode < 0 || mode > 4 || iters < 1)
        return fail(e, RT_ERR_INVALID_ARG, "rt_selftest_pipe_peak: bad argument");
    std::lock_guard<std::mutex> lk(e->mu);
    DeviceGuard dg(e->device);
    RT_TRY(reserve(e, e->coords, 256));
    int *sink = static_cast<int *>(e->coords.p);
    const int blocks = kNumSMs * kPeakBlocksPerSM;
    auto launch = [&](int n) {
        if (mode == 0) peak_kernel<0><<<blocks, kPeakThreads, 0, e->stream>>>(n, 3, 5, sink);
        else if (mode == 1) peak_kernel<1><<<blocks, kPeakThreads, 0, e->stream>>>(n, 3, 5, sink);
        else if (mode == 2) peak_kernel<2><<<blocks, kPeakThreads, 0, e->stream>>>(n, 3, 5, sink);
        else if (mode == 3) peak_kernel<3><<<blocks, kPeakThreads, 0, e->stream>>>(n, 3, 5, sink);
        else peak_kernel<4><<<blocks, kPeakThreads, 0, e->stream>>>(n, 3, 5, sink);
    };
    launch(iters / 8 + 1);  // warm-up: module load, clocks
    RT_CUDA(e, cudaGetLastError());
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    RT_CUDA(e, cudaEventCreate(&t0));
    RT_CUDA(e, cudaEventCreate(&t1));
    cudaEventRecord(t0, e->stream);
    launch(iters);
    cudaEventRecord(t1, e->stream);
    cudaError_t st = cudaStreamSynchronize(e->stream);
    float ms = 0.f;
    if (st == cudaSuccess) st = cudaEventElapsedTime(&ms, t0, t1);
    cudaEventDestroy(t0);
    cudaEventDestroy(t1);
    if (st != cudaSuccess) return fail(e, RT_ERR_CUDA, "rt_selftest_pipe_peak: %s", cudaGetErrorString(st));
    const double warp_inst = static_cast<double>(blocks) * (kPeakThreads / 32) * iters * kChains * kOpsPerChainPerIter;
    *warp_inst_per_s = warp_inst / (ms * 1e-3);
    if (elapsed_ms) *elapsed_ms = ms;
    return RT_OK;
}
