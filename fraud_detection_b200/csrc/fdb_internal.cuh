// Internal declarations shared by the .cu files of libfdb200.so (not part of the C ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <stdio.h>

#include "../../include/fdb200.h"

#define FDB_F FDB_N_FEATURES
#define FDB_VERSION_NUM 100

// ---- error plumbing -------------------------------------------------------------------
void fdb_set_error(const char* fmt, ...);

#define FDB_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) {                                                         \
            fdb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),        \
                          __FILE__, __LINE__);                                           \
            return _e == cudaErrorMemoryAllocation ? FDB_ERR_NOMEM : FDB_ERR_CUDA;       \
        }                                                                                \
    } while (0)

#define FDB_REQUIRE(cond, code, ...)                                                     \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            fdb_set_error(__VA_ARGS__);                                                  \
            return (code);                                                               \
        }                                                                                \
    } while (0)

static inline bool fdb_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Makes `dev` the current device for the lifetime of the object and restores the caller's device afterwards (a
// process that drives several GPUs must not find its current device changed by a library call).
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; prev = -1; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// ---- context --------------------------------------------------------------------------
constexpr int FDB_MAX_DEVICES = 64;        // device indices a process may use
constexpr int FDB_PIPE_SLOTS = 3;          // host pipeline depth (H2D / kernel / D2H overlap)
constexpr int FDB_MAX_GRID = 148 * 16;     // upper bound on CTAs of any grid-stride kernel

struct fdb_pipe_slot {
    cudaStream_t stream = nullptr;
    void* d_in = nullptr;      // chunk of x
    void* d_out = nullptr;     // chunk of phi / transformed rows
    void* d_p = nullptr;       // chunk of proba
    void* d_z = nullptr;       // chunk of logits
    uint8_t* d_cls = nullptr;  // chunk of classes
};

struct fdb_ctx {
    std::mutex host_mu;                    // serialises the *_host entry points (they share streams and staging)
    int device = 0;
    int sm_count = 0;
    // scratch of the |phi| column-sum reduction: per-CTA partials + arrival ticket
    double* d_partials = nullptr;          // [FDB_MAX_GRID][32]
    unsigned int* d_ticket = nullptr;      // zero between launches
    double* d_abs_sum = nullptr;           // [32] result of the last host-pipeline call
    // host pipeline
    int64_t chunk_rows = 0;                // rows per pipeline chunk (0 = not allocated)
    size_t chunk_elem = 0;                 // element size the chunk buffers were sized for
    fdb_pipe_slot slot[FDB_PIPE_SLOTS];
    void* h_stage = nullptr;               // pinned staging for small (serving) batches
    size_t h_stage_bytes = 0;
    void* h_out[FDB_PIPE_SLOTS] = {};      // pinned landing buffers for phi chunks bound for PAGEABLE caller memory
    size_t h_out_bytes = 0;                // size of each
    void* h_in[FDB_PIPE_SLOTS] = {};       // pinned take-off buffers for row chunks coming from PAGEABLE caller memory
    size_t h_in_bytes = 0;
};

// ---- kernel launchers (device pointers, no synchronisation) ---------------------------
struct LinearParams {
    double sub[FDB_F];
    double mul[FDB_F];
    double bias;
};

int fdb_launch_linear_pass(fdb_ctx* ctx, cudaStream_t st, int dtype, int64_t n_rows, const void* x,
                           const LinearParams& prm, bool divide, void* proba, int proba_cols, void* logit,
                           void* phi, uint8_t* cls, double* abs_sum, double* partials, unsigned int* ticket);

// streaming (TMA bulk-copy) form of K1 for large batches - linear_pass_tma.cu
int64_t fdb_linear_pass_tma_min_rows(const fdb_ctx* ctx, int dtype);
int fdb_launch_linear_pass_tma(fdb_ctx* ctx, cudaStream_t st, int dtype, int64_t n_rows, const void* x,
                               const LinearParams& prm, bool divide, void* proba, int proba_cols, void* logit,
                               void* phi, uint8_t* cls, double* abs_sum, double* partials, unsigned int* ticket);
