// K1 - fused linear pass over 30-feature rows (sm_100a).
//
// One streaming read of x produces every requested output of the reference's scoring and
// linear-SHAP calls (see include/fdb200.h, fdb_linear_pass_dev):
//     phi[r][k] = (x[r][k] - sub[k]) * mul[k]      (or (x - sub) / mul for StandardScaler.transform)
//     logit[r]  = bias + sum_k phi[r][k];  proba = sigmoid(logit);  cls = logit > 0
//     abs_sum[k] = sum_r |phi[r][k]|
//
// Layout.  A row is 30 elements = 120 B (f32) / 240 B (f64): not a multiple of 16 B per row
// in f32, so the matrix is read as a flat stream of 16-byte vectors.  A warp iteration moves
// one 480-byte "group" = 30 vectors (4 f32 rows or 2 f64 rows): each 16-lane half-warp takes
// 15 consecutive vectors (lane 15 of each half idles), so a half-warp owns exactly 2 f32 rows
// (60 floats) or 1 f64 row.  The feature index of every register element therefore depends
// only on the lane, so sub/mul live in registers, phi is pure element-wise on coalesced
// 16-byte loads/stores, and the row dot products are 4-step xor-shuffle reductions inside
// the half-warp (f32: the two rows of a half are reduced together by swapping partials in
// the first step).  The kernel is HBM-bound: 244 B/row (f32, proba+phi) against ~150 flops.
//
// |phi| column sums: per-lane accumulators (fixed feature per register), fixed-order block
// reduction, per-CTA partials in ctx scratch, last-arriving CTA folds them in CTA order ->
// bitwise deterministic for a given grid.

#include <stdlib.h>

#include "linear_common.cuh"

using namespace k1;

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

template <typename T, bool DIVIDE>
__global__ void __launch_bounds__(kThreads)
linear_pass_kernel(const T* __restrict__ x, int64_t n_rows, const KParams<T> prm, T* __restrict__ proba,
                   int proba_cols, T* __restrict__ logit, T* __restrict__ phi, uint8_t* __restrict__ cls,
                   double* __restrict__ abs_sum, double* __restrict__ partials, unsigned int* __restrict__ ticket) {
    using V = typename Lane<T>::V;
    constexpr int EPL = Lane<T>::EPL;
    constexpr int RPW = Lane<T>::RPW;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int j = lane & 15;
    const int half = lane >> 4;
    const bool active = j < 15;
    const bool want_rows = (proba != nullptr) | (logit != nullptr) | (cls != nullptr);
    const bool want_abs = abs_sum != nullptr;

    T sub[EPL], mul[EPL];
#pragma unroll
    for (int k = 0; k < EPL; ++k) {
        const int f = (EPL * j + k) % FDB_F;
        sub[k] = prm.sub[f];
        mul[k] = prm.mul[f];
    }

    const int64_t n_groups = n_rows / RPW;
    const int64_t total_warps = (int64_t)gridDim.x * kWarps;
    const int64_t warp_id = (int64_t)blockIdx.x * kWarps + warp;
    const V* __restrict__ xv = reinterpret_cast<const V*>(x);
    V* __restrict__ pv = reinterpret_cast<V*>(phi);
    const int lane_vec = half * 15 + j;

    float acc32[EPL];
    double acc64[EPL];
#pragma unroll
    for (int k = 0; k < EPL; ++k) { acc32[k] = 0.f; acc64[k] = 0.0; }
    int since_flush = 0;

    auto process = [&](const V& v, int64_t g) {
        T e[EPL];
        unpack(v, e);
#pragma unroll
        for (int k = 0; k < EPL; ++k) {
            T d = e[k] - sub[k];
            e[k] = DIVIDE ? d / mul[k] : d * mul[k];
        }
        if (phi != nullptr && active) __stcs(pv + g * 30 + lane_vec, pack(e));
        if (!DIVIDE) {
            if (want_abs && active) {
#pragma unroll
                for (int k = 0; k < EPL; ++k) {
                    if constexpr (sizeof(T) == 4) acc32[k] += fabsf((float)e[k]);   // flushed to f64 every 32 iterations
                    else acc64[k] += fabs((double)e[k]);
                }
            }
            if (want_rows) {
                T keep;
                if constexpr (EPL == 4) {
                    // half-warp = 2 rows: lanes 0..6 row A, lane 7 splits, lanes 8..14 row B
                    const T lo = e[0] + e[1], hi = e[2] + e[3];
                    T a = (j < 7) ? lo + hi : (j == 7 ? lo : T(0));
                    T b = (j > 7) ? lo + hi : (j == 7 ? hi : T(0));
                    if (!active) { a = T(0); b = T(0); }
                    const bool low = j < 8;
                    keep = low ? a : b;
                    const T send = low ? b : a;
                    keep += __shfl_xor_sync(0xffffffffu, send, 8);
                } else {
                    keep = active ? e[0] + e[1] : T(0);
                    keep += __shfl_xor_sync(0xffffffffu, keep, 8);
                }
                keep += __shfl_xor_sync(0xffffffffu, keep, 4);
                keep += __shfl_xor_sync(0xffffffffu, keep, 2);
                keep += __shfl_xor_sync(0xffffffffu, keep, 1);
                if constexpr (EPL == 4) {
                    if ((j & 7) == 0) emit_row<T>(g * RPW + half * 2 + (j >> 3), keep + prm.bias, proba, proba_cols, logit, cls);
                } else {
                    if (j == 0) emit_row<T>(g * RPW + half, keep + prm.bias, proba, proba_cols, logit, cls);
                }
            }
        }
    };

    const V zero = {};
    for (int64_t g = warp_id; g < n_groups; g += 2 * total_warps) {
        const int64_t g1 = g + total_warps;
        const bool has1 = g1 < n_groups;
        V v0 = zero, v1 = zero;
        if (active) {
            v0 = __ldcs(xv + g * 30 + lane_vec);
            if (has1) v1 = __ldcs(xv + g1 * 30 + lane_vec);
        }
        process(v0, g);
        if (has1) process(v1, g1);
        if (want_abs && ++since_flush == 32) {
#pragma unroll
            for (int k = 0; k < EPL; ++k) { acc64[k] += (double)acc32[k]; acc32[k] = 0.f; }
            since_flush = 0;
        }
    }

    // tail rows (n_rows % RPW of them): warp 0 of CTA 0, one feature per lane
    double tail_abs = 0.0;
    if (blockIdx.x == 0 && warp == 0) {
        for (int64_t r = n_groups * RPW; r < n_rows; ++r) {
            T e = T(0);
            if (lane < FDB_F) {
                const T d = x[r * FDB_F + lane] - prm.sub[lane];
                e = DIVIDE ? d / prm.mul[lane] : d * prm.mul[lane];
                if (phi != nullptr) phi[r * FDB_F + lane] = e;
                tail_abs += fabs((double)e);
            }
            if (!DIVIDE && want_rows) {
                T s = e;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                if (lane == 0) emit_row<T>(r, s + prm.bias, proba, proba_cols, logit, cls);
            }
        }
    }

    if constexpr (!DIVIDE) {
    if (!want_abs) return;

    // ---- |phi| column sums: fixed-order CTA reduction, then last-CTA fold ----------------
    __shared__ double s_acc[kWarps][32][EPL];
    __shared__ double s_tail[32];
    __shared__ bool s_last;
#pragma unroll
    for (int k = 0; k < EPL; ++k) s_acc[warp][lane][k] = active ? acc64[k] + (double)acc32[k] : 0.0;
    if (warp == 0) s_tail[lane] = tail_abs;
    __syncthreads();
    if (threadIdx.x < 32) {
        double tot = 0.0;
        if (threadIdx.x < FDB_F) {
            const int f = threadIdx.x;
            for (int w = 0; w < kWarps; ++w) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    // flat positions of feature f inside a half-warp's 15 vectors
                    for (int e = f; e < 15 * EPL; e += FDB_F) tot += s_acc[w][h * 16 + e / EPL][e % EPL];
                }
            }
            tot += s_tail[f];
        }
        partials[(size_t)blockIdx.x * 32 + threadIdx.x] = tot;
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        __threadfence();
        double tot = 0.0;
        if (threadIdx.x < FDB_F) {
            for (unsigned b = 0; b < gridDim.x; ++b) tot += __ldcg(partials + (size_t)b * 32 + threadIdx.x);
        } else if (threadIdx.x == FDB_F) {
            tot = (double)n_rows;
        }
        abs_sum[threadIdx.x] = tot;
        if (threadIdx.x == 0) *ticket = 0u;
    }
    }  // !DIVIDE
}

template <typename T>
int launch_typed(fdb_ctx* ctx, cudaStream_t st, int64_t n_rows, const void* x, const LinearParams& prm,
                 bool divide, void* proba, int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum,
                 double* partials, unsigned int* ticket) {
    KParams<T> kp;
    for (int k = 0; k < FDB_F; ++k) {
        kp.sub[k] = (T)prm.sub[k];
        kp.mul[k] = (T)prm.mul[k];
    }
    kp.bias = (T)prm.bias;
    const int64_t n_groups = n_rows / Lane<T>::RPW;
    int64_t want = (n_groups + 2 * kWarps - 1) / (2 * kWarps);   // 2 groups per warp minimum
    if (want < 1) want = 1;
    // persistent grid-stride launch: exactly the resident CTA count (multiple of the SM count)
    static int per_sm[FDB_MAX_DEVICES][2] = {};
    int& occ = per_sm[ctx->device][divide ? 1 : 0];
    if (occ == 0) {
        int v = 0;
        cudaError_t e = divide
            ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, linear_pass_kernel<T, true>, kThreads, 0)
            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, linear_pass_kernel<T, false>, kThreads, 0);
        FDB_CUDA(e);
        occ = v > 0 ? v : 1;
    }
    int64_t cap = (int64_t)ctx->sm_count * occ;
    if (cap > FDB_MAX_GRID) cap = FDB_MAX_GRID;
    const int grid = (int)(want < cap ? want : cap);
    if (divide) {
        linear_pass_kernel<T, true><<<grid, kThreads, 0, st>>>(
            (const T*)x, n_rows, kp, nullptr, 1, nullptr, (T*)phi, nullptr, nullptr, nullptr, nullptr);
    } else {
        linear_pass_kernel<T, false><<<grid, kThreads, 0, st>>>(
            (const T*)x, n_rows, kp, (T*)proba, proba_cols, (T*)logit, (T*)phi, cls, abs_sum, partials, ticket);
    }
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}

}  // namespace

int fdb_launch_linear_pass(fdb_ctx* ctx, cudaStream_t st, int dtype, int64_t n_rows, const void* x,
                           const LinearParams& prm, bool divide, void* proba, int proba_cols, void* logit,
                           void* phi, uint8_t* cls, double* abs_sum, double* partials, unsigned int* ticket) {
    if (n_rows == 0) {
        if (abs_sum) FDB_CUDA(cudaMemsetAsync(abs_sum, 0, 32 * sizeof(double), st));
        return FDB_OK;
    }
    if (n_rows >= fdb_linear_pass_tma_min_rows(ctx, dtype) && getenv("FDB_K1_NO_TMA") == nullptr)
        return fdb_launch_linear_pass_tma(ctx, st, dtype, n_rows, x, prm, divide, proba, proba_cols, logit, phi, cls,
                                          abs_sum, partials, ticket);
    if (dtype == FDB_F32)
        return launch_typed<float>(ctx, st, n_rows, x, prm, divide, proba, proba_cols, logit, phi, cls, abs_sum, partials, ticket);
    return launch_typed<double>(ctx, st, n_rows, x, prm, divide, proba, proba_cols, logit, phi, cls, abs_sum, partials, ticket);
}

// =======================================================================================
// fdb_linear_pass_noalloc - the fused linear pass for callers that must not allocate.
//
// Who this is for.  fdb_linear_pass_dev needs an fdb_ctx, and creating a context calls
// cudaMalloc for the |phi| column-sum scratch.  Some users of this library cannot do that:
//   * a scoring thread inside a serving process whose device memory is already owned in
//     full by another allocator (a framework's caching pool, a fixed MIG/MPS slice budget),
//     where a late cudaMalloc either fails or eats into memory promised to someone else;
//   * a latency-critical request path, where cudaMalloc/cudaFree synchronise the whole
//     device and stall every other stream for the duration of the call;
//   * code that is only allowed to enqueue work on a stream it was handed and has no
//     place in its lifecycle where allocating (and later freeing) would be legal.
// For them this one entry point runs the same kernel with its scratch in a fixed array that
// is part of the library image: nothing is allocated, by us or by any library we call.
//
// The price is a size limit.  The scratch holds the per-CTA partial sums for at most
// FDB_NOALLOC_MAX_CTAS CTAs (128 KiB + one 256-byte line for the arrival ticket), which
// covers FDB_NOALLOC_MAX_ROWS = 16384 rows per call in either dtype - the serving-size
// batches of /predict and the XAI task.  Larger batches are refused with FDB_ERR_ARG; use
// fdb_linear_pass_dev for those.  Every GPU that loads libfdb200.so carries the array until
// the process exits, whether or not this entry point is ever called.
//
// Sharing.  There is one array per device, shared by every call on that device.  It keeps
// whatever the previous call left in it; the only part that has to start at zero is the
// arrival ticket, which is cleared on the caller's stream before each launch.  Two calls
// whose kernels can overlap on the same GPU (different streams, or different host threads
// with their own streams) race on the array and give wrong column sums: order them on one
// stream or with events.  Calls that do not request abs_sum do not touch the array.
//
// Set-up.  The first call on each device looks up the array's device address
// (cudaGetSymbolAddress, with that device current) and remembers it.  That lookup is host
// work which may initialise the CUDA module, so make the first call on every device you
// use during start-up, before any stream capture begins; later calls only enqueue a 4-byte
// memset and the kernel.
// =======================================================================================
#include <mutex>

#define FDB_NOALLOC_MAX_CTAS 512
#define FDB_NOALLOC_MAX_DEVICES 64

// Per-CTA partial sums ([FDB_NOALLOC_MAX_CTAS][32] doubles) followed by the ticket line.
// In host code this name is only a handle for the runtime's symbol lookup - it is not a
// device address and must never be handed to a kernel as one.
static __device__ __align__(256) unsigned char g_noalloc_scratch[FDB_NOALLOC_MAX_CTAS * 32 * sizeof(double) + 256];

namespace {

std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[FDB_NOALLOC_MAX_DEVICES] = {};
int g_noalloc_sms[FDB_NOALLOC_MAX_DEVICES] = {};

// Device address of the scratch array on the CURRENT device (looked up once per device).
int noalloc_scratch(unsigned char** out, int* sm_count) {
    int dev = -1;
    FDB_CUDA(cudaGetDevice(&dev));
    FDB_REQUIRE(dev >= 0 && dev < FDB_NOALLOC_MAX_DEVICES, FDB_ERR_ARG, "device index %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {
        void* p = nullptr;
        FDB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        int sms = 0;
        FDB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        g_noalloc_sms[dev] = sms;
        g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[dev];
    *sm_count = g_noalloc_sms[dev];
    return FDB_OK;
}

template <typename T>
int launch_noalloc(cudaStream_t st, int64_t n_rows, const void* x, const LinearParams& prm, void* proba,
                   int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum) {
    KParams<T> kp;
    for (int k = 0; k < FDB_F; ++k) {
        kp.sub[k] = (T)prm.sub[k];
        kp.mul[k] = (T)prm.mul[k];
    }
    kp.bias = (T)prm.bias;
    // same CTA count as the allocating entry point picks for a batch this small (two groups
    // per warp), so the column sums fold in the same order and agree bit for bit
    const int64_t n_groups = n_rows / Lane<T>::RPW;
    int64_t want = (n_groups + 2 * kWarps - 1) / (2 * kWarps);
    if (want < 1) want = 1;
    FDB_REQUIRE(want <= FDB_NOALLOC_MAX_CTAS, FDB_ERR_ARG,
                "fdb_linear_pass_noalloc: %lld rows need %lld CTAs of scratch but the built-in array holds %d",
                (long long)n_rows, (long long)want, FDB_NOALLOC_MAX_CTAS);
    double* partials = nullptr;
    unsigned int* ticket = nullptr;
    if (abs_sum != nullptr) {
        unsigned char* base = nullptr;
        int sms = 0;
        int rc = noalloc_scratch(&base, &sms);
        if (rc != FDB_OK) return rc;
        partials = reinterpret_cast<double*>(base);
        ticket = reinterpret_cast<unsigned int*>(base + (size_t)FDB_NOALLOC_MAX_CTAS * 32 * sizeof(double));
        // the array keeps the last call's contents; only the ticket must start at zero
        FDB_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st));
    }
    linear_pass_kernel<T, false><<<(int)want, kThreads, 0, st>>>(
        (const T*)x, n_rows, kp, (T*)proba, proba_cols, (T*)logit, (T*)phi, cls, abs_sum, partials, ticket);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}

}  // namespace

extern "C" int fdb_linear_pass_noalloc(void* stream, int dtype, int64_t n_rows, const void* x, const double* sub,
                                       const double* mul, double bias, void* proba, int proba_cols, void* logit,
                                       void* phi, uint8_t* cls, double* abs_sum) {
    FDB_REQUIRE(dtype == FDB_F32 || dtype == FDB_F64, FDB_ERR_ARG, "unknown dtype %d", dtype);
    FDB_REQUIRE(n_rows >= 0, FDB_ERR_ARG, "n_rows is negative");
    FDB_REQUIRE(n_rows <= FDB_NOALLOC_MAX_ROWS, FDB_ERR_ARG,
                "fdb_linear_pass_noalloc handles at most %d rows per call (got %lld); larger batches need "
                "fdb_linear_pass_dev, which allocates its scratch through an fdb_ctx",
                FDB_NOALLOC_MAX_ROWS, (long long)n_rows);
    FDB_REQUIRE(n_rows == 0 || x != nullptr, FDB_ERR_ARG, "x is NULL");
    FDB_REQUIRE(mul != nullptr, FDB_ERR_ARG, "mul is NULL");
    FDB_REQUIRE(proba_cols == 1 || proba_cols == 2, FDB_ERR_ARG, "proba_cols must be 1 or 2");
    FDB_REQUIRE(fdb_aligned16(x) && fdb_aligned16(phi), FDB_ERR_ALIGN, "x and phi must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_rows == 0) {
        if (abs_sum) FDB_CUDA(cudaMemsetAsync(abs_sum, 0, 32 * sizeof(double), st));
        return FDB_OK;
    }
    LinearParams p;
    for (int k = 0; k < FDB_F; ++k) {
        p.sub[k] = sub ? sub[k] : 0.0;
        p.mul[k] = mul[k];
    }
    p.bias = bias;
    if (dtype == FDB_F32) return launch_noalloc<float>(st, n_rows, x, p, proba, proba_cols, logit, phi, cls, abs_sum);
    return launch_noalloc<double>(st, n_rows, x, p, proba, proba_cols, logit, phi, cls, abs_sum);
}
