// Pipe-rate probes: measured denominators for the Kernel SHAP rooflines (bench.py, tools/microbench).
//
// The Kernel SHAP evaluators are bound by instruction throughput, not by memory, and MEASURED_PEAKS.json holds no
// SFU or FP32-issue figure.  These kernels measure, on the device and at the clock it actually runs under load:
//   [0] ex2.approx.ftz.f32 results per second            (MUFU.EX2, the SFU line of the north star)
//   [1] rcp.approx.ftz.f32 results per second            (MUFU.RCP)
//   [2] independent FFMA results per second              (FP32 FMA pipe = warp-instruction issue rate x 32)
//   [3] sigmoids per second of K2f's clamp-free quad loop (E read as broadcast LDS.128) at 16 resident warps per SM -
//       the occupancy the evaluator's 16 compute warps give it
//   [4] the same loop at 64 resident warps per SM
//   [5] 0 (reserved)
// Every probe keeps 8 independent dependency chains per thread and 64 warps per SM, so the pipes, not latencies, set
// the rate.  No counterpart in the reference; a measurement aid only.

#include "fdb_internal.cuh"
#include "ptx_sm100.cuh"

namespace {

constexpr int kChains = 8;
constexpr int kProbeThreads = 1024;
constexpr int kCtasPerSm = 2;

__global__ void __launch_bounds__(kProbeThreads, kCtasPerSm) probe_ex2(float* sink, int iters) {
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = -0.01f * (float)(threadIdx.x + c);
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int c = 0; c < kChains; ++c) x[c] = ptx::ex2_approx(x[c]) - 1.0f;      // stays in (-1, 0]
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) *sink = s;
}

__global__ void __launch_bounds__(kProbeThreads, kCtasPerSm) probe_rcp(float* sink, int iters) {
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = 1.0f + 0.001f * (float)(threadIdx.x + c);
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int c = 0; c < kChains; ++c) x[c] = ptx::rcp_approx(x[c]) + 0.5f;      // stays in (0.5, 2.5)
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) *sink = s;
}

__global__ void __launch_bounds__(kProbeThreads, kCtasPerSm) probe_ffma(float* sink, int iters, float a, float b,
                                                                         long long* clocks) {
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = 0.001f * (float)(threadIdx.x + c);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int c = 0; c < kChains; ++c) x[c] = fmaf(x[c], a, b);
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) *sink = s;
    if (blockIdx.x == 0 && threadIdx.x == 0) *clocks = t1 - t0;
}

// K2f's clamp-free quad (kshap_eval_fx.cu, chunk_sum4<kFxBalanced>): 13 FP32-pipe instructions and one reciprocal per
// four sigmoids.  SMEM = false keeps E in registers (pure pipe rate), true reads it as LDS.128 like the kernel does.
template <bool SMEM>
__global__ void __launch_bounds__(kProbeThreads, kCtasPerSm) probe_quad(float* sink, int iters) {
    __shared__ float4 sE[4][32];
    if (threadIdx.x < 128) {
        const float v = 0.25f + 0.01f * (float)threadIdx.x;
        sE[threadIdx.x >> 5][threadIdx.x & 31] = make_float4(v, 1.5f * v, 2.f * v, 3.f * v);
    }
    __syncthreads();
    float a[4], b[4], sum[4];
#pragma unroll
    for (int m = 0; m < 4; ++m) {
        a[m] = 0.5f + 0.001f * (float)((threadIdx.x + m) & 63);
        b[m] = 1.0f;
        sum[m] = 0.f;
    }
    float4 er[4];
#pragma unroll
    for (int m = 0; m < 4; ++m) er[m] = sE[m][threadIdx.x & 31];
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int q = 0; q < 2; ++q)
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                const float4 e = SMEM ? sE[m][(i * 2 + q) & 31] : er[m];
                asm volatile("" : "+f"(a[m]));      // a[m] looks modified to the compiler: nothing is hoisted, no instruction is added
                const float d0 = fmaf(e.x, a[m], b[m]), d1 = fmaf(e.y, a[m], b[m]);
                const float d2 = fmaf(e.z, a[m], b[m]), d3 = fmaf(e.w, a[m], b[m]);
                const float n01 = d0 + d1, p01 = d0 * d1, n23 = d2 + d3, p23 = d2 * d3;
                sum[m] = fmaf(fmaf(n01, p23, n23 * p01), ptx::rcp_approx(p01 * p23), sum[m]);
            }
    }
    const float s = sum[0] + sum[1] + sum[2] + sum[3];
    if (s == 12345.678f) *sink = s;
}

}  // namespace

extern "C" int fdb_probe_pipes(fdb_ctx* ctx, double* out8) {
    FDB_REQUIRE(ctx != nullptr && out8 != nullptr, FDB_ERR_ARG, "NULL argument");
    int prev = -1;
    FDB_CUDA(cudaGetDevice(&prev));
    if (prev != ctx->device) FDB_CUDA(cudaSetDevice(ctx->device));
    float* sink = nullptr;
    long long* clocks = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = FDB_OK;
    do {
#define PTRY(x) if ((x) != cudaSuccess) { fdb_set_error("pipe probe failed: %s", cudaGetErrorString(cudaGetLastError())); rc = FDB_ERR_CUDA; break; }
        PTRY(cudaMalloc((void**)&sink, 256));
        clocks = reinterpret_cast<long long*>(sink + 8);
        PTRY(cudaEventCreate(&e0));
        PTRY(cudaEventCreate(&e1));
        const int grid = ctx->sm_count * kCtasPerSm;
        const double threads = (double)grid * kProbeThreads;
        auto timed = [&](auto launch, double per_thread, double* res) -> bool {
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep) {              // first repetition warms the clocks up
                if (cudaEventRecord(e0) != cudaSuccess) return false;
                launch();
                if (cudaEventRecord(e1) != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) return false;
                float ms = 0.f;
                if (cudaEventElapsedTime(&ms, e0, e1) != cudaSuccess) return false;
                if (rep > 0 && ms < best) best = ms;
            }
            *res = threads * per_thread / (best * 1e-3);
            return cudaGetLastError() == cudaSuccess;
        };
        const int it_mufu = 4096, it_fma = 32768, it_quad = 4096;
        bool ok = true;
        ok = ok && timed([&] { probe_ex2<<<grid, kProbeThreads>>>(sink, it_mufu); }, (double)kChains * it_mufu, &out8[0]);
        ok = ok && timed([&] { probe_rcp<<<grid, kProbeThreads>>>(sink, it_mufu); }, (double)kChains * it_mufu, &out8[1]);
        ok = ok && timed([&] { probe_ffma<<<grid, kProbeThreads>>>(sink, it_fma, 0.999f, 0.001f, clocks); }, (double)kChains * it_fma, &out8[2]);
        out8[5] = 0.0;
        {
            // the quad loop at the evaluator's own occupancy: one CTA of 512 threads per SM
            double r16 = 0.0;
            ok = ok && timed([&] { probe_quad<true><<<ctx->sm_count, 512>>>(sink, it_quad); }, 32.0 * it_quad, &r16);
            out8[3] = r16 * (512.0 * ctx->sm_count) / threads;          // `timed` multiplies by the full-grid thread count
        }
        ok = ok && timed([&] { probe_quad<true><<<grid, kProbeThreads>>>(sink, it_quad); }, 32.0 * it_quad, &out8[4]);
        out8[6] = (double)ctx->sm_count;
        out8[7] = 0.0;
        if (!ok) { fdb_set_error("pipe probe failed: %s", cudaGetErrorString(cudaGetLastError())); rc = FDB_ERR_CUDA; }
#undef PTRY
    } while (0);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(sink);
    if (prev >= 0 && prev != ctx->device) cudaSetDevice(prev);
    return rc;
}
