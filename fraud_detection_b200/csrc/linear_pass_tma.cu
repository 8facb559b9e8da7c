// K1, streaming form - the fused linear pass for large batches, fed by the TMA engine (sm_100a).
//
// Same arithmetic and lane mapping as linear_pass.cu, but HBM traffic is moved by 1-D bulk
// copies (cp.async.bulk, SASS UBLKCP) instead of per-thread loads/stores:
//
//   warp 0 (one lane)  loader : 15,360-byte tiles (128 f32 rows / 64 f64 rows)  HBM -> smem ring
//   warps 2..9         compute: ONE THREAD PER ROW - 15 conflict-free LDS.64 (f32) / LDS.128 (f64) bring
//                               the row into registers, phi is written back in place, the dot product is a
//                               private register sum (no shuffles), parameters are uniform constant-bank
//                               operands, sigmoid runs on all 32 lanes and proba/logit/cls stores are coalesced;
//                               two groups of 4 warps alternate tiles
//   warp 1 (one lane)  storer : finished phi tiles  smem -> HBM (bulk async-group)
//
// An 8-stage ring keeps ~120 KB of loads in flight per SM (Little's law needs ~22 KB at
// 6.5 TB/s), so latency is hidden by the copy engine, not by occupancy, and the per-byte
// instruction count drops ~5x versus the load/store kernel.  One persistent CTA per SM; tiles
// are interleaved across CTAs (tile t -> CTA t mod grid) so the chip walks HBM sequentially.
// Rows beyond the last full tile (< 128) are handled by one warp with plain loads.

#include "linear_common.cuh"
#include "ptx_sm100.cuh"

using namespace k1;

namespace {

constexpr int kConsumerWarps = 8;
constexpr int kThreadsTma = (kConsumerWarps + 2) * 32;
constexpr int kTileBytes = 15360;                       // 32 groups of 480 B
constexpr int kStages = 8;

template <typename T> struct RowVec;
template <> struct RowVec<float> { using type = float2; };
template <> struct RowVec<double> { using type = double2; };

__device__ __forceinline__ float sigmoid_fast(float z) {
    return ptx::rcp_approx(1.0f + ptx::ex2_approx(-1.4426950408889634f * z));   // 2 MUFU, ~1e-7 abs error
}
__device__ __forceinline__ double sigmoid_fast(double z) { return 1.0 / (1.0 + exp(-z)); }

// all 32 lanes hold consecutive rows -> coalesced stores
template <typename T>
__device__ __forceinline__ void emit_row_fast(int64_t row, T z, T* __restrict__ proba, int proba_cols,
                                              T* __restrict__ logit, uint8_t* __restrict__ cls) {
    if (logit) logit[row] = z;
    if (proba) {
        const T p = sigmoid_fast(z);
        if (proba_cols == 2) {
            if constexpr (sizeof(T) == 4) reinterpret_cast<float2*>(proba)[row] = make_float2(1.0f - p, p);
            else reinterpret_cast<double2*>(proba)[row] = make_double2(1.0 - p, p);
        } else {
            proba[row] = p;
        }
    }
    if (cls) cls[row] = z > T(0) ? 1 : 0;
}

struct alignas(128) TmaSmem {
    unsigned char stage[kStages][kTileBytes];
    double acc[kConsumerWarps][32];
    double tail[32];
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t computed[kStages];
    int last;
};

template <typename T, bool DIVIDE, bool WRITE_PHI>
__global__ void __launch_bounds__(kThreadsTma, 1)
linear_pass_tma_kernel(const T* __restrict__ x, int64_t n_rows, const KParams<T> prm, T* __restrict__ proba,
                       int proba_cols, T* __restrict__ logit, T* __restrict__ phi, uint8_t* __restrict__ cls,
                       double* __restrict__ abs_sum, double* __restrict__ partials,
                       unsigned int* __restrict__ ticket) {
    constexpr int TR = kTileBytes / (FDB_F * (int)sizeof(T));   // rows per tile

    extern __shared__ __align__(128) unsigned char smem_raw[];
    TmaSmem& sm = *reinterpret_cast<TmaSmem*>(smem_raw);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int64_t n_tiles = n_rows / TR;
    const bool want_rows = (proba != nullptr) | (logit != nullptr) | (cls != nullptr);
    const bool want_abs = !DIVIDE && abs_sum != nullptr;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            ptx::mbar_init(&sm.full[s], 1);
            ptx::mbar_init(&sm.empty[s], WRITE_PHI ? 1 : TR / 32);
            ptx::mbar_init(&sm.computed[s], TR / 32);
        }
        ptx::mbar_fence_init();
    }
    __syncthreads();

    float col_acc32[FDB_F];      // per-thread |phi| column sums (compute warps only), flushed to f64
    double col_acc64[FDB_F];
#pragma unroll
    for (int q = 0; q < FDB_F; ++q) { col_acc32[q] = 0.f; col_acc64[q] = 0.0; }

    if (warp == 0) {
        // ------------------------------------------------------------- loader
        if (lane == 0) {
            int k = 0;
            for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
                const int s = k % kStages;
                const uint32_t ph = (k / kStages) & 1;
                ptx::mbar_wait(&sm.empty[s], ph ^ 1);
                ptx::mbar_arrive_expect_tx(&sm.full[s], kTileBytes);
                ptx::bulk_g2s(sm.stage[s], reinterpret_cast<const unsigned char*>(x) + t * kTileBytes, kTileBytes, &sm.full[s]);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------- storer
        if (WRITE_PHI && lane == 0) {
            int k = 0;
            for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
                const int s = k % kStages;
                const uint32_t ph = (k / kStages) & 1;
                ptx::mbar_wait(&sm.computed[s], ph);
                ptx::bulk_s2g(reinterpret_cast<unsigned char*>(phi) + t * kTileBytes, sm.stage[s], kTileBytes);
                ptx::bulk_commit();
                ptx::bulk_wait_read<1>();                 // the store committed one tile ago has left smem
                if (k >= 1) ptx::mbar_arrive(&sm.empty[(k - 1) % kStages]);
            }
            ptx::bulk_wait<0>();
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------- compute: one thread per row
        const int cw = warp - 2;
        const int group = cw >> 2;                 // warps 0-3 take even tiles of this CTA, 4-7 odd ones
        const int r = (cw & 3) * 32 + lane;        // row inside the tile
        const bool has_row = r < TR;
        constexpr int ROW_V = FDB_F / 2;           // vectors per row: 15 x float2 (f32) / 15 x double2 (f64)
        using RV = typename RowVec<T>::type;
        int k = 0;
        int since_flush = 0;
        for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
            if ((k & 1) != group) continue;
            const int s = k % kStages;
            const uint32_t ph = (k / kStages) & 1;
            ptx::mbar_wait(&sm.full[s], ph);
            if (has_row) {
                RV* row = reinterpret_cast<RV*>(sm.stage[s] + (size_t)r * FDB_F * sizeof(T));
                RV v[ROW_V];
#pragma unroll
                for (int i = 0; i < ROW_V; ++i) v[i] = row[i];
                T sum = T(0);
#pragma unroll
                for (int i = 0; i < ROW_V; ++i) {
                    T e0 = v[i].x - prm.sub[2 * i], e1 = v[i].y - prm.sub[2 * i + 1];
                    e0 = DIVIDE ? e0 / prm.mul[2 * i] : e0 * prm.mul[2 * i];
                    e1 = DIVIDE ? e1 / prm.mul[2 * i + 1] : e1 * prm.mul[2 * i + 1];
                    v[i].x = e0;
                    v[i].y = e1;
                    if constexpr (!DIVIDE) {
                        sum += e0 + e1;
                        if (want_abs) {
                            if constexpr (sizeof(T) == 4) {
                                col_acc32[2 * i] += fabsf((float)e0);
                                col_acc32[2 * i + 1] += fabsf((float)e1);
                            } else {
                                col_acc64[2 * i] += fabs((double)e0);
                                col_acc64[2 * i + 1] += fabs((double)e1);
                            }
                        }
                    }
                }
                if (WRITE_PHI) {
#pragma unroll
                    for (int i = 0; i < ROW_V; ++i) row[i] = v[i];
                }
                if constexpr (!DIVIDE) {
                    if (want_rows) emit_row_fast<T>(t * TR + r, sum + prm.bias, proba, proba_cols, logit, cls);
                }
            }
            if (WRITE_PHI) {
                ptx::fence_proxy_async_smem();     // phi written through the generic proxy, read by the bulk store
                __syncwarp();
                if (lane == 0 && (cw & 3) * 32 < TR) ptx::mbar_arrive(&sm.computed[s]);
            } else {
                __syncwarp();
                if (lane == 0 && (cw & 3) * 32 < TR) ptx::mbar_arrive(&sm.empty[s]);
            }
            if (sizeof(T) == 4 && want_abs && ++since_flush == 16) {
#pragma unroll
                for (int q = 0; q < FDB_F; ++q) { col_acc64[q] += (double)col_acc32[q]; col_acc32[q] = 0.f; }
                since_flush = 0;
            }
        }
        // rows past the last full tile: one warp, one feature per lane, plain loads
        double tail_abs = 0.0;
        if (blockIdx.x == gridDim.x - 1 && cw == 0) {
            for (int64_t rr = n_tiles * TR; rr < n_rows; ++rr) {
                T e = T(0);
                if (lane < FDB_F) {
                    const T d = x[rr * FDB_F + lane] - prm.sub[lane];
                    e = DIVIDE ? d / prm.mul[lane] : d * prm.mul[lane];
                    if (WRITE_PHI) phi[rr * FDB_F + lane] = e;
                    tail_abs += fabs((double)e);
                }
                if (!DIVIDE && want_rows) {
                    T sacc = e;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) sacc += __shfl_xor_sync(0xffffffffu, sacc, o);
                    if (lane == 0) emit_row<T>(rr, sacc + prm.bias, proba, proba_cols, logit, cls);
                }
            }
        }
        if (want_abs) {
            // fold the 32 lanes of this warp (fixed xor tree), lane f keeps column f
            double mine = 0.0;
#pragma unroll
            for (int q = 0; q < FDB_F; ++q) {
                double vq = col_acc64[q] + (double)col_acc32[q];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) vq += __shfl_xor_sync(0xffffffffu, vq, o);
                if (lane == q) mine = vq;
            }
            sm.acc[cw][lane] = mine;
            if (cw == 0) sm.tail[lane] = tail_abs;
        }
    }

    if constexpr (!DIVIDE) {
        if (!want_abs) return;
        // ---- |phi| column sums: fixed-order CTA reduction, then the last CTA folds all partials
        __syncthreads();
        if (threadIdx.x < 32) {
            double tot = 0.0;
            if (threadIdx.x < FDB_F) {
                const int f = threadIdx.x;
                for (int w = 0; w < kConsumerWarps; ++w) tot += sm.acc[w][f];
                tot += sm.tail[f];
            }
            partials[(size_t)blockIdx.x * 32 + threadIdx.x] = tot;
            __threadfence();
        }
        __syncthreads();
        if (threadIdx.x == 0) sm.last = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1 : 0;
        __syncthreads();
        if (sm.last && threadIdx.x < 32) {
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
    }
}

template <typename T, bool DIVIDE, bool WRITE_PHI>
int launch_inst(fdb_ctx* ctx, cudaStream_t st, int64_t n_rows, const void* x, const KParams<T>& kp, void* proba,
                int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum, double* partials,
                unsigned int* ticket) {
    static bool configured[FDB_MAX_DEVICES] = {};          // function attributes are per device
    auto kern = linear_pass_tma_kernel<T, DIVIDE, WRITE_PHI>;
    if (!configured[ctx->device]) {
        FDB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TmaSmem)));
        configured[ctx->device] = true;
    }
    constexpr int TR = kTileBytes / (FDB_F * (int)sizeof(T));
    const int64_t n_tiles = n_rows / TR;
    int64_t grid = ctx->sm_count < n_tiles ? ctx->sm_count : n_tiles;
    if (grid < 1) grid = 1;
    kern<<<(int)grid, kThreadsTma, sizeof(TmaSmem), st>>>((const T*)x, n_rows, kp, (T*)proba, proba_cols, (T*)logit,
                                                          (T*)phi, cls, abs_sum, partials, ticket);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}

template <typename T>
int launch_tma_typed(fdb_ctx* ctx, cudaStream_t st, int64_t n_rows, const void* x, const LinearParams& prm, bool divide,
                     void* proba, int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum,
                     double* partials, unsigned int* ticket) {
    const KParams<T> kp = make_kparams<T>(prm);
    if (divide)
        return launch_inst<T, true, true>(ctx, st, n_rows, x, kp, nullptr, 1, nullptr, phi, nullptr, nullptr, nullptr, nullptr);
    if (phi != nullptr)
        return launch_inst<T, false, true>(ctx, st, n_rows, x, kp, proba, proba_cols, logit, phi, cls, abs_sum, partials, ticket);
    return launch_inst<T, false, false>(ctx, st, n_rows, x, kp, proba, proba_cols, logit, nullptr, cls, abs_sum, partials, ticket);
}

}  // namespace

// Rows at which the streaming kernel takes over from the grid-stride kernel (two tiles per SM).
int64_t fdb_linear_pass_tma_min_rows(const fdb_ctx* ctx, int dtype) {
    const int tr = kTileBytes / (FDB_F * (dtype == FDB_F64 ? 8 : 4));
    return (int64_t)2 * ctx->sm_count * tr;
}

int fdb_launch_linear_pass_tma(fdb_ctx* ctx, cudaStream_t st, int dtype, int64_t n_rows, const void* x,
                               const LinearParams& prm, bool divide, void* proba, int proba_cols, void* logit,
                               void* phi, uint8_t* cls, double* abs_sum, double* partials, unsigned int* ticket) {
    if (dtype == FDB_F32)
        return launch_tma_typed<float>(ctx, st, n_rows, x, prm, divide, proba, proba_cols, logit, phi, cls, abs_sum, partials, ticket);
    return launch_tma_typed<double>(ctx, st, n_rows, x, prm, divide, proba, proba_cols, logit, phi, cls, abs_sum, partials, ticket);
}
