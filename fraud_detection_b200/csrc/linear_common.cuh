// Shared pieces of the K1 kernels (linear_pass.cu: grid-stride kernel for small batches and the
// no-alloc entry point; linear_pass_tma.cu: TMA-pipelined streaming kernel for large batches).
#pragma once

#include "fdb_internal.cuh"

namespace k1 {

template <typename T> struct Lane;
template <> struct Lane<float> {
    using V = float4;
    static constexpr int EPL = 4;   // elements per lane
    static constexpr int RPW = 4;   // rows per warp iteration
};
template <> struct Lane<double> {
    using V = double2;
    static constexpr int EPL = 2;
    static constexpr int RPW = 2;
};

template <typename T> struct KParams {
    T sub[FDB_F];
    T mul[FDB_F];
    T bias;
};

__device__ __forceinline__ void unpack(const float4& v, float (&e)[4]) { e[0] = v.x; e[1] = v.y; e[2] = v.z; e[3] = v.w; }
__device__ __forceinline__ void unpack(const double2& v, double (&e)[2]) { e[0] = v.x; e[1] = v.y; }
__device__ __forceinline__ float4 pack(const float (&e)[4]) { return make_float4(e[0], e[1], e[2], e[3]); }
__device__ __forceinline__ double2 pack(const double (&e)[2]) { return make_double2(e[0], e[1]); }

__device__ __forceinline__ float sigmoidT(float z) { return __frcp_rn(1.0f + __expf(-z)); }
__device__ __forceinline__ double sigmoidT(double z) { return 1.0 / (1.0 + exp(-z)); }

// proba_cols == 2 writes the sklearn layout [1 - p, p] (SK:linear_model/_base.py:438-440)
template <typename T>
__device__ __forceinline__ void emit_row(int64_t row, T z, T* __restrict__ proba, int proba_cols,
                                         T* __restrict__ logit, uint8_t* __restrict__ cls) {
    if (logit) logit[row] = z;
    if (proba) {
        const T p = sigmoidT(z);
        if (proba_cols == 2) {
            if constexpr (sizeof(T) == 4) reinterpret_cast<float2*>(proba)[row] = make_float2(1.0f - p, p);
            else reinterpret_cast<double2*>(proba)[row] = make_double2(1.0 - p, p);
        } else {
            proba[row] = p;
        }
    }
    if (cls) cls[row] = z > T(0) ? 1 : 0;
}


template <typename T>
inline KParams<T> make_kparams(const LinearParams& prm) {
    KParams<T> kp;
    for (int k = 0; k < FDB_F; ++k) {
        kp.sub[k] = (T)prm.sub[k];
        kp.mul[k] = (T)prm.mul[k];
    }
    kp.bias = (T)prm.bias;
    return kp;
}

}  // namespace k1
