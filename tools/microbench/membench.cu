// HBM streaming micro-benchmarks used to pick K1's structure (run on the B200 box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/membench tools/microbench/membench.cu && /tmp/membench
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../fraud_detection_b200/csrc/ptx_sm100.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__global__ void copy_ldst(const float4* __restrict__ in, float4* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        float4 a = __ldcs(in + i), b = __ldcs(in + i + stride), c = __ldcs(in + i + 2 * stride), d = __ldcs(in + i + 3 * stride);
        __stcs(out + i, a); __stcs(out + i + stride, b); __stcs(out + i + 2 * stride, c); __stcs(out + i + 3 * stride, d);
    }
    for (; i < n; i += stride) __stcs(out + i, __ldcs(in + i));
}
__global__ void read_only(const float4* __restrict__ in, float* sink, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    float acc = 0.f;
    for (; i + 3 * stride < n; i += 4 * stride) {
        float4 a = __ldcs(in + i), b = __ldcs(in + i + stride), c = __ldcs(in + i + 2 * stride), d = __ldcs(in + i + 3 * stride);
        acc += a.x + b.y + c.z + d.w;
    }
    if (acc == 123.456f) *sink = acc;
}
__global__ void write_only(float4* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    float4 v = make_float4(1, 2, 3, 4);
    for (; i < n; i += stride) __stcs(out + i, v);
}

// TMA ring: loader thread + storer thread, optional consumer warps touching the tile in place
template <int STAGES, int TILE, int CONSUMERS>
__global__ void __launch_bounds__((CONSUMERS + 2) * 32) tma_copy(const unsigned char* __restrict__ in, unsigned char* __restrict__ out, size_t n_tiles, int touch) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)STAGES * TILE);
    uint64_t* empty = full + STAGES;
    uint64_t* computed = empty + STAGES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); ptx::mbar_init(&computed[s], CONSUMERS); }
        ptx::mbar_fence_init();
    }
    __syncthreads();
    if (warp == 0) {
        if (lane == 0) {
            int k = 0;
            for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
                int s = k % STAGES; uint32_t ph = (k / STAGES) & 1;
                ptx::mbar_wait(&empty[s], ph ^ 1);
                ptx::mbar_arrive_expect_tx(&full[s], TILE);
                ptx::bulk_g2s(smem + (size_t)s * TILE, in + t * TILE, TILE, &full[s]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            int k = 0;
            for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
                int s = k % STAGES; uint32_t ph = (k / STAGES) & 1;
                if (CONSUMERS > 0) ptx::mbar_wait(&computed[s], ph); else ptx::mbar_wait(&full[s], ph);
                ptx::bulk_s2g(out + t * TILE, smem + (size_t)s * TILE, TILE);
                ptx::bulk_commit();
                ptx::bulk_wait_read<1>();
                if (k >= 1) ptx::mbar_arrive(&empty[(k - 1) % STAGES]);
            }
            ptx::bulk_wait<0>();
        }
    } else {
        int cw = warp - 2, k = 0;
        for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
            int s = k % STAGES; uint32_t ph = (k / STAGES) & 1;
            ptx::mbar_wait(&full[s], ph);
            if (touch) {
                float4* p = reinterpret_cast<float4*>(smem + (size_t)s * TILE);
                for (int i = cw * 32 + lane; i < TILE / 16; i += CONSUMERS * 32) { float4 v = p[i]; v.x += 1.f; p[i] = v; }
                ptx::fence_proxy_async_smem();
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&computed[s]);
        }
    }
}

template <typename F> float time_ms(F f, int iters = 10) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); f(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int i = 0; i < iters; ++i) { cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
    return best;
}

template <int STAGES, int TILE, int CONSUMERS> void run_tma(const char* name, const unsigned char* in, unsigned char* out, size_t bytes, int ctas_per_sm, int touch, int sms) {
    size_t smem = (size_t)STAGES * TILE + 3 * STAGES * 8 + 64;
    CK(cudaFuncSetAttribute(tma_copy<STAGES, TILE, CONSUMERS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    size_t n_tiles = bytes / TILE;
    float ms = time_ms([&] { tma_copy<STAGES, TILE, CONSUMERS><<<sms * ctas_per_sm, (CONSUMERS + 2) * 32, smem>>>(in, out, n_tiles, touch); });
    CK(cudaGetLastError());
    printf("%-44s stages=%d tile=%6d cons=%d cta/sm=%d touch=%d : %8.3f ms  %8.1f GB/s (r+w)\n", name, STAGES, TILE, CONSUMERS, ctas_per_sm, touch, ms, 2.0 * n_tiles * TILE / ms / 1e6);
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    size_t bytes = (size_t)12 << 30;   // 12 GiB each way, like 1e8 rows
    unsigned char *in, *out; float* sink;
    CK(cudaMalloc(&in, bytes)); CK(cudaMalloc(&out, bytes)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(in, 1, bytes)); CK(cudaMemset(out, 0, bytes));
    size_t n4 = bytes / 16;
    printf("device %s, %d SMs\n", prop.name, sms);
    { float ms = time_ms([&] { cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice); }); printf("%-44s : %8.3f ms %8.1f GB/s (r+w)\n", "cudaMemcpy D2D", ms, 2.0 * bytes / ms / 1e6); }
    for (int bps : {4, 8}) for (int thr : {256, 512}) {
        float ms = time_ms([&] { copy_ldst<<<sms * bps, thr>>>((const float4*)in, (float4*)out, n4); });
        printf("copy_ldst blocks/sm=%d threads=%d               : %8.3f ms %8.1f GB/s (r+w)\n", bps, thr, ms, 2.0 * bytes / ms / 1e6);
    }
    { float ms = time_ms([&] { read_only<<<sms * 8, 256>>>((const float4*)in, sink, n4); }); printf("%-44s : %8.3f ms %8.1f GB/s (r)\n", "read_only", ms, 1.0 * bytes / ms / 1e6); }
    { float ms = time_ms([&] { write_only<<<sms * 8, 256>>>((float4*)out, n4); }); printf("%-44s : %8.3f ms %8.1f GB/s (w)\n", "write_only", ms, 1.0 * bytes / ms / 1e6); }
    run_tma<8, 15360, 0>("tma ring, no consumers", in, out, bytes, 1, 0, sms);
    run_tma<4, 15360, 0>("tma ring, no consumers", in, out, bytes, 2, 0, sms);
    run_tma<4, 15360, 0>("tma ring, no consumers", in, out, bytes, 3, 0, sms);
    run_tma<6, 32768, 0>("tma ring, no consumers", in, out, bytes, 1, 0, sms);
    run_tma<3, 65536, 0>("tma ring, no consumers", in, out, bytes, 1, 0, sms);
    run_tma<12, 15360, 0>("tma ring, no consumers", in, out, bytes, 1, 0, sms);
    run_tma<14, 7680, 0>("tma ring, no consumers", in, out, bytes, 1, 0, sms);
    run_tma<8, 15360, 8>("tma ring, consumers idle", in, out, bytes, 1, 0, sms);
    run_tma<8, 15360, 8>("tma ring, consumers touch", in, out, bytes, 1, 1, sms);
    run_tma<4, 15360, 8>("tma ring, consumers touch", in, out, bytes, 2, 1, sms);
    return 0;
}
