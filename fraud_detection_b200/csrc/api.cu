// C ABI of libfdb200.so: context, error reporting, and the host-buffer pipelines around K1.
// Declarations and reference citations live in include/fdb200.h.

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <condition_variable>
#include <deque>
#include <memory>
#include <new>
#include <thread>
#include <vector>

#include "fdb_internal.cuh"

// ---------------------------------------------------------------- errors
static thread_local char g_err[512] = "";

void fdb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* fdb_last_error(void) { return g_err; }
extern "C" int fdb_version(void) { return FDB_VERSION_NUM; }

// ---------------------------------------------------------------- context
namespace {

constexpr int64_t kSmallRows = 16384;            // serving-size batches go through pinned staging
constexpr int64_t kDefaultChunkRows = 1 << 20;   // rows per pipeline chunk for large host batches
constexpr int kScratchSets = FDB_PIPE_SLOTS + 1; // set 0: _dev callers; sets 1..: the host entry points (one per pipeline slot)

size_t elem_size(int dtype) { return dtype == FDB_F64 ? 8 : 4; }

void free_pipeline(fdb_ctx* ctx) {
    for (auto& s : ctx->slot) {
        cudaFree(s.d_in); cudaFree(s.d_out); cudaFree(s.d_p); cudaFree(s.d_z); cudaFree(s.d_cls);
        s.d_in = s.d_out = s.d_p = s.d_z = nullptr;
        s.d_cls = nullptr;
    }
    ctx->chunk_rows = 0;
    ctx->chunk_elem = 0;
}

// batches up to this many rows are served without copy commands (kernel reads / writes mapped pinned memory)
int64_t zero_copy_rows() {
    static int64_t v = -1;
    if (v < 0) {
        const char* env = getenv("FDB_ZERO_COPY_ROWS");
        v = env ? atoll(env) : kSmallRows;      // measured: faster than copy commands up to the whole serving range
        if (v < 0) v = 0;
        if (v > kSmallRows) v = kSmallRows;
    }
    return v;
}

int64_t chunk_rows_cap() {
    const char* env = getenv("FDB_CHUNK_ROWS");
    if (env) {
        long long v = atoll(env);
        if (v >= 4) return (int64_t)(v & ~3LL);
    }
    return kDefaultChunkRows;
}

// Device chunk buffers for the host pipeline, grown on demand (never shrunk).
int ensure_pipeline(fdb_ctx* ctx, int64_t rows, size_t elem) {
    int64_t want = rows < chunk_rows_cap() ? rows : chunk_rows_cap();
    want = (want + 3) & ~int64_t(3);
    if (want < 4) want = 4;
    if (ctx->chunk_rows >= want && ctx->chunk_elem >= elem) return FDB_OK;
    if (ctx->chunk_rows > want) want = ctx->chunk_rows;
    if (ctx->chunk_elem > elem) elem = ctx->chunk_elem;
    if (ctx->chunk_rows > 0) {
        for (auto& s : ctx->slot) FDB_CUDA(cudaStreamSynchronize(s.stream));
        free_pipeline(ctx);
    }
    for (auto& s : ctx->slot) {
        FDB_CUDA(cudaMalloc(&s.d_in, (size_t)want * FDB_F * elem));
        FDB_CUDA(cudaMalloc(&s.d_out, (size_t)want * FDB_F * elem));
        FDB_CUDA(cudaMalloc(&s.d_p, (size_t)want * elem * 2));
        FDB_CUDA(cudaMalloc(&s.d_z, (size_t)want * elem));
        FDB_CUDA(cudaMalloc((void**)&s.d_cls, (size_t)want));
    }
    ctx->chunk_rows = want;
    ctx->chunk_elem = elem;
    return FDB_OK;
}

// Copy workers of one host call whose phi output is pageable memory.  A device-to-host copy into pageable memory is
// staged by the driver on the calling thread and takes the page faults of a freshly allocated result array one by one
// (measured: 68 MB in 15 ms, 4.5 GB/s, against 1.4 ms into page-locked memory).  Instead the chunks land in page-locked
// buffers of the context and these threads move them on - four slices of a chunk at a time, faults taken in parallel -
// while the GPU works on the next chunks.
class CopyWorkers {
public:
    explicit CopyWorkers(int n_threads) {
        for (int i = 0; i < n_threads; ++i) threads_.emplace_back([this] { run(); });
    }
    ~CopyWorkers() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    // split one chunk into a slice per thread; `slot` counts them so the landing buffer is not reused too early
    void post(int slot, char* dst, const char* src, size_t bytes) {
        const size_t n = threads_.size();
        const size_t per = ((bytes + n - 1) / n + 4095) & ~size_t(4095);
        std::lock_guard<std::mutex> lk(mu_);
        for (size_t off = 0; off < bytes; off += per) {
            q_.push_back({slot, dst + off, src + off, bytes - off < per ? bytes - off : per});
            ++pending_[slot];
        }
        cv_.notify_all();
    }
    void wait_slot(int slot) {
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [&] { return pending_[slot] == 0; });
    }
    void wait_all() {
        for (int s = 0; s < 2 * FDB_PIPE_SLOTS; ++s) wait_slot(s);
    }

private:
    struct Task { int slot; char* dst; const char* src; size_t n; };
    void run() {
        for (;;) {
            Task t;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
                if (q_.empty()) return;
                t = q_.front();
                q_.pop_front();
            }
            memcpy(t.dst, t.src, t.n);
            {
                std::lock_guard<std::mutex> lk(mu_);
                --pending_[t.slot];
            }
            done_.notify_all();
        }
    }
    std::vector<std::thread> threads_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    std::deque<Task> q_;
    int pending_[2 * FDB_PIPE_SLOTS] = {};      // [0, slots): chunks going out, [slots, 2 slots): chunks coming in
    bool stop_ = false;
};

bool is_device_accessible_host(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

int check_common(const fdb_ctx* ctx, int dtype, int64_t n_rows, const void* x) {
    FDB_REQUIRE(ctx != nullptr, FDB_ERR_ARG, "ctx is NULL");
    FDB_REQUIRE(dtype == FDB_F32 || dtype == FDB_F64, FDB_ERR_ARG, "unknown dtype %d", dtype);
    FDB_REQUIRE(n_rows >= 0, FDB_ERR_ARG, "n_rows is negative");
    FDB_REQUIRE(n_rows == 0 || x != nullptr, FDB_ERR_ARG, "x is NULL");
    return FDB_OK;
}

// Shared body of the *_host entry points: chunked 3-stream pipeline, or the pinned-staging
// fast path for serving-size batches.
int run_host(fdb_ctx* ctx, int dtype, int64_t n, const void* x, const LinearParams& prm, bool divide,
             void* proba, int proba_cols, void* logit, void* phi, uint8_t* cls, double* abs_sum) {
    std::lock_guard<std::mutex> lock(ctx->host_mu);
    DeviceGuard guard(ctx->device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", ctx->device);
    const size_t es = elem_size(dtype);
    if (abs_sum) memset(abs_sum, 0, 32 * sizeof(double));
    if (n == 0) return FDB_OK;
    int rc = ensure_pipeline(ctx, n, es);
    if (rc != FDB_OK) return rc;

    const char* xs = static_cast<const char*>(x);
    char* ph = static_cast<char*>(phi);
    char* pp = static_cast<char*>(proba);
    char* pz = static_cast<char*>(logit);
    const size_t row_b = FDB_F * es;

    if (n <= kSmallRows) {
        // serving path: one stream, everything through the context's pinned staging buffer
        fdb_pipe_slot& s = ctx->slot[0];
        // scratch set 1 (pipeline slot 0's): set 0 belongs to _dev callers, whose launches may overlap this one
        double* h_partials = ctx->d_partials + (size_t)FDB_MAX_GRID * 32;
        unsigned int* h_ticket = ctx->d_ticket + 1;
        char* st = static_cast<char*>(ctx->h_stage);
        const bool x_pinned = is_device_accessible_host(x);
        // output blocks inside the pinned staging buffer (offsets keep 16-byte alignment)
        char* o_phi = st + (size_t)kSmallRows * row_b;
        char* o_p = o_phi + (size_t)kSmallRows * row_b;
        char* o_z = o_p + (size_t)kSmallRows * es * 2;
        char* o_c = o_z + (size_t)kSmallRows * es;
        char* o_a = o_c + (size_t)kSmallRows;
        if (n <= zero_copy_rows()) {
            // request-size batches: no copy commands at all.  The staging buffer is page-locked and mapped, so the
            // kernel reads the rows and writes every output straight over PCIe; one launch, one stream wait.
            const void* src = x;
            if (!x_pinned || !fdb_aligned16(x)) { memcpy(st, x, (size_t)n * row_b); src = st; }
            rc = fdb_launch_linear_pass(ctx, s.stream, dtype, n, src, prm, divide, proba ? o_p : nullptr, proba_cols,
                                        logit ? o_z : nullptr, phi ? o_phi : nullptr, cls ? reinterpret_cast<uint8_t*>(o_c) : nullptr,
                                        abs_sum ? reinterpret_cast<double*>(o_a) : nullptr, h_partials, h_ticket);
            if (rc != FDB_OK) return rc;
            FDB_CUDA(cudaStreamSynchronize(s.stream));
            if (phi) memcpy(ph, o_phi, (size_t)n * row_b);
            if (proba) memcpy(pp, o_p, (size_t)n * es * proba_cols);
            if (logit) memcpy(pz, o_z, (size_t)n * es);
            if (cls) memcpy(cls, o_c, (size_t)n);
            if (abs_sum) memcpy(abs_sum, o_a, 32 * sizeof(double));
            return FDB_OK;
        }
        const void* src = x;
        if (!x_pinned) { memcpy(st, x, (size_t)n * row_b); src = st; }
        FDB_CUDA(cudaMemcpyAsync(s.d_in, src, (size_t)n * row_b, cudaMemcpyHostToDevice, s.stream));
        rc = fdb_launch_linear_pass(ctx, s.stream, dtype, n, s.d_in, prm, divide, proba ? s.d_p : nullptr, proba_cols,
                                    logit ? s.d_z : nullptr, phi ? s.d_out : nullptr, cls ? s.d_cls : nullptr,
                                    abs_sum ? ctx->d_abs_sum + 32 : nullptr, h_partials, h_ticket);
        if (rc != FDB_OK) return rc;
        if (phi) FDB_CUDA(cudaMemcpyAsync(o_phi, s.d_out, (size_t)n * row_b, cudaMemcpyDeviceToHost, s.stream));
        if (proba) FDB_CUDA(cudaMemcpyAsync(o_p, s.d_p, (size_t)n * es * proba_cols, cudaMemcpyDeviceToHost, s.stream));
        if (logit) FDB_CUDA(cudaMemcpyAsync(o_z, s.d_z, (size_t)n * es, cudaMemcpyDeviceToHost, s.stream));
        if (cls) FDB_CUDA(cudaMemcpyAsync(o_c, s.d_cls, (size_t)n, cudaMemcpyDeviceToHost, s.stream));
        if (abs_sum) FDB_CUDA(cudaMemcpyAsync(o_a, ctx->d_abs_sum + 32, 32 * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
        FDB_CUDA(cudaStreamSynchronize(s.stream));
        if (phi) memcpy(ph, o_phi, (size_t)n * row_b);
        if (proba) memcpy(pp, o_p, (size_t)n * es * proba_cols);
        if (logit) memcpy(pz, o_z, (size_t)n * es);
        if (cls) memcpy(cls, o_c, (size_t)n);
        if (abs_sum) memcpy(abs_sum, o_a, 32 * sizeof(double));
        return FDB_OK;
    }

    // bulk path: chunk c runs on slot c % 3; stream order protects each slot's buffers, and the
    // three streams overlap H2D(c+1), kernel(c) and D2H(c-1).  Pinned caller buffers DMA
    // directly; pageable INPUT is staged by the driver (14 GB/s measured, fine); a pageable phi OUTPUT goes
    // through the context's page-locked landing buffers and the copy workers above.
    // fill + drain cost one chunk's copy time, so aim for many chunks; the floor below keeps every copy large
    // enough (>= 7.8 MB) for the link to run at speed
    const bool use_workers = getenv("FDB_NO_COPY_WORKERS") == nullptr;
    const bool stage_phi = phi != nullptr && use_workers && !is_device_accessible_host(phi);
    // (input-only calls - predict, predict_proba - are better off with the driver's own pageable path: 4.7 against 7.0 ms
    //  for 68 MB measured; with phi staged the workers are there anyway and taking the rows too saves another 15-20 %)
    const bool stage_in = stage_phi && !is_device_accessible_host(x);
    int want_chunks = 32;
    if (const char* e = getenv("FDB_PIPE_CHUNKS")) { const int v = atoi(e); if (v >= 1 && v <= 4096) want_chunks = v; }
    int64_t chunk = (n + want_chunks - 1) / want_chunks;
    const int64_t floor_rows = (stage_phi || stage_in) ? 65536 : 262144;                // 30 MB copies (f32 rows); smaller hops when workers copy on
    if (chunk < floor_rows && !getenv("FDB_PIPE_CHUNKS")) chunk = floor_rows;
    if (chunk < 4 * kSmallRows) chunk = 4 * kSmallRows;
    chunk = (chunk + 3) & ~int64_t(3);
    if (chunk > ctx->chunk_rows) chunk = ctx->chunk_rows;
    const int64_t n_chunks = (n + chunk - 1) / chunk;
    double* h_abs = nullptr;     // per-chunk |phi| sums land in the pinned staging buffer
    if (abs_sum) {
        FDB_REQUIRE((size_t)n_chunks * 32 * sizeof(double) <= ctx->h_stage_bytes, FDB_ERR_ARG,
                    "too many chunks (%lld) for the |phi| summary; raise FDB_CHUNK_ROWS", (long long)n_chunks);
        h_abs = static_cast<double*>(ctx->h_stage);
    }
    if (stage_phi && ctx->h_out_bytes < (size_t)chunk * row_b) {
        for (auto& p : ctx->h_out) { if (p) cudaFreeHost(p); p = nullptr; }
        ctx->h_out_bytes = 0;
        for (auto& p : ctx->h_out) FDB_CUDA(cudaMallocHost(&p, (size_t)chunk * row_b));
        ctx->h_out_bytes = (size_t)chunk * row_b;
    }
    if (stage_in && ctx->h_in_bytes < (size_t)chunk * row_b) {
        for (auto& p : ctx->h_in) { if (p) cudaFreeHost(p); p = nullptr; }
        ctx->h_in_bytes = 0;
        for (auto& p : ctx->h_in) FDB_CUDA(cudaMallocHost(&p, (size_t)chunk * row_b));
        ctx->h_in_bytes = (size_t)chunk * row_b;
    }
    std::unique_ptr<CopyWorkers> workers;
    cudaEvent_t landed[FDB_PIPE_SLOTS] = {}, taken[FDB_PIPE_SLOTS] = {};
    if (stage_phi || stage_in) {
        // measured on the 16-vCPU B200 box, 68 MB in + 68 MB out: 2 threads 11.1 ms, 4 -> 7.2, 8 -> 6.0, 12 -> 6.1
        int n_workers = (int)(std::thread::hardware_concurrency() / 2);
        if (n_workers < 2) n_workers = 2;
        if (n_workers > 8) n_workers = 8;
        if (const char* e = getenv("FDB_COPY_THREADS")) { const int v = atoi(e); if (v >= 1 && v <= 64) n_workers = v; }
        workers.reset(new (std::nothrow) CopyWorkers(n_workers));
        FDB_REQUIRE(workers != nullptr, FDB_ERR_NOMEM, "out of host memory");
        for (auto& ev : landed) FDB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        for (auto& ev : taken) FDB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    }
    auto chunk_rows_of = [&](int64_t c) { return (n - c * chunk) < chunk ? (n - c * chunk) : chunk; };
    // hand chunk c (landed in its slot's page-locked buffer) to the copy workers
    auto forward = [&](int64_t c) -> cudaError_t {
        const int sl = (int)(c % FDB_PIPE_SLOTS);
        cudaError_t e2 = cudaEventSynchronize(landed[sl]);
        if (e2 == cudaSuccess)
            workers->post(sl, ph + (size_t)(c * chunk) * row_b, static_cast<const char*>(ctx->h_out[sl]), (size_t)chunk_rows_of(c) * row_b);
        return e2;
    };
    cudaError_t e = cudaSuccess;
    for (int64_t c = 0; c < n_chunks && e == cudaSuccess; ++c) {
        fdb_pipe_slot& s = ctx->slot[c % FDB_PIPE_SLOTS];
        const int set = (int)(c % FDB_PIPE_SLOTS);
        const int64_t r0 = c * chunk;
        const int64_t rows = chunk_rows_of(c);
        if (stage_phi) {
            if (c >= 2) e = forward(c - 2);              // chunk c - 2 has landed by now (or nearly): move it on
            workers->wait_slot(set);                     // chunk c - 3 left this slot's landing buffer (posted one trip ago)
            if (e != cudaSuccess) break;
        }
        const char* src = xs + (size_t)r0 * row_b;
        if (stage_in) {
            // pageable rows: the workers copy the chunk into this slot's page-locked buffer (once the DMA of chunk c - 3
            // has left it) while the GPU is busy with the chunks already queued; then a plain asynchronous DMA
            if (c >= FDB_PIPE_SLOTS) e = cudaEventSynchronize(taken[set]);
            if (e != cudaSuccess) break;
            workers->post(FDB_PIPE_SLOTS + set, static_cast<char*>(ctx->h_in[set]), src, (size_t)rows * row_b);
            workers->wait_slot(FDB_PIPE_SLOTS + set);
            src = static_cast<const char*>(ctx->h_in[set]);
        }
        e = cudaMemcpyAsync(s.d_in, src, (size_t)rows * row_b, cudaMemcpyHostToDevice, s.stream);
        if (e == cudaSuccess && stage_in) e = cudaEventRecord(taken[set], s.stream);
        if (e == cudaSuccess) {
            rc = fdb_launch_linear_pass(ctx, s.stream, dtype, rows, s.d_in, prm, divide, proba ? s.d_p : nullptr, proba_cols,
                                        logit ? s.d_z : nullptr, phi ? s.d_out : nullptr, cls ? s.d_cls : nullptr,
                                        abs_sum ? ctx->d_abs_sum + 32 * (set + 1) : nullptr,
                                        ctx->d_partials + (size_t)(set + 1) * FDB_MAX_GRID * 32, ctx->d_ticket + set + 1);
            if (rc != FDB_OK) break;
        }
        if (e == cudaSuccess && phi) {
            if (stage_phi) {
                e = cudaMemcpyAsync(ctx->h_out[set], s.d_out, (size_t)rows * row_b, cudaMemcpyDeviceToHost, s.stream);
                if (e == cudaSuccess) e = cudaEventRecord(landed[set], s.stream);
            } else {
                e = cudaMemcpyAsync(ph + (size_t)r0 * row_b, s.d_out, (size_t)rows * row_b, cudaMemcpyDeviceToHost, s.stream);
            }
        }
        if (e == cudaSuccess && proba) e = cudaMemcpyAsync(pp + (size_t)r0 * es * proba_cols, s.d_p, (size_t)rows * es * proba_cols, cudaMemcpyDeviceToHost, s.stream);
        if (e == cudaSuccess && logit) e = cudaMemcpyAsync(pz + (size_t)r0 * es, s.d_z, (size_t)rows * es, cudaMemcpyDeviceToHost, s.stream);
        if (e == cudaSuccess && cls) e = cudaMemcpyAsync(cls + r0, s.d_cls, (size_t)rows, cudaMemcpyDeviceToHost, s.stream);
        if (e == cudaSuccess && abs_sum) e = cudaMemcpyAsync(h_abs + c * 32, ctx->d_abs_sum + 32 * (set + 1), 32 * sizeof(double), cudaMemcpyDeviceToHost, s.stream);
    }
    if (stage_phi && rc == FDB_OK && e == cudaSuccess) {
        for (int64_t c = n_chunks >= 2 ? n_chunks - 2 : 0; c < n_chunks && e == cudaSuccess; ++c) e = forward(c);
        workers->wait_all();
    }
    for (auto& s : ctx->slot) {
        cudaError_t e2 = cudaStreamSynchronize(s.stream);
        if (e == cudaSuccess) e = e2;
    }
    workers.reset();
    for (auto& ev : landed)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : taken)
        if (ev) cudaEventDestroy(ev);
    if (rc != FDB_OK) return rc;
    if (e != cudaSuccess) {
        fdb_set_error("host pipeline failed: %s", cudaGetErrorString(e));
        return FDB_ERR_CUDA;
    }
    if (abs_sum) {
        for (int64_t c = 0; c < n_chunks; ++c)
            for (int k = 0; k < 31; ++k) abs_sum[k] += h_abs[c * 32 + k];
    }
    return FDB_OK;
}

void fill_params(LinearParams& p, const double* sub, const double* mul, double bias) {
    for (int k = 0; k < FDB_F; ++k) {
        p.sub[k] = sub ? sub[k] : 0.0;
        p.mul[k] = mul[k];
    }
    p.bias = bias;
}

}  // namespace

extern "C" int fdb_ctx_create(int device, fdb_ctx** out) {
    FDB_REQUIRE(out != nullptr, FDB_ERR_ARG, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    FDB_REQUIRE(e == cudaSuccess && count > 0, FDB_ERR_CUDA, "no CUDA device available (%s); libfdb200 has no CPU fallback",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    FDB_REQUIRE(device >= 0 && device < count && device < FDB_MAX_DEVICES, FDB_ERR_ARG, "device %d out of range (0..%d)", device, count - 1);
    DeviceGuard guard(device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", device);
    fdb_ctx* ctx = new (std::nothrow) fdb_ctx();
    FDB_REQUIRE(ctx != nullptr, FDB_ERR_NOMEM, "out of host memory");
    ctx->device = device;
    int rc = FDB_OK;
    do {
        cudaDeviceProp prop;
        if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) break;
        ctx->sm_count = prop.multiProcessorCount;
        if (prop.major < 10) {
            fdb_set_error("device %d is sm_%d%d; libfdb200 is built for sm_100a (B200) only", device, prop.major, prop.minor);
            rc = FDB_ERR_CUDA;
            break;
        }
        if ((e = cudaMalloc((void**)&ctx->d_partials, (size_t)kScratchSets * FDB_MAX_GRID * 32 * sizeof(double))) != cudaSuccess) break;
        if ((e = cudaMalloc((void**)&ctx->d_ticket, kScratchSets * sizeof(unsigned int))) != cudaSuccess) break;
        if ((e = cudaMemset(ctx->d_ticket, 0, kScratchSets * sizeof(unsigned int))) != cudaSuccess) break;
        if ((e = cudaMalloc((void**)&ctx->d_abs_sum, kScratchSets * 32 * sizeof(double))) != cudaSuccess) break;
        for (auto& s : ctx->slot)
            if ((e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking)) != cudaSuccess) break;
        if (e != cudaSuccess) break;
        ctx->h_stage_bytes = (size_t)kSmallRows * (2 * FDB_F * 8 + 3 * 8 + 1) + 32 * sizeof(double) + 256;
        if ((e = cudaMallocHost(&ctx->h_stage, ctx->h_stage_bytes)) != cudaSuccess) break;
    } while (0);
    if (e != cudaSuccess || rc != FDB_OK) {
        if (e != cudaSuccess) {
            fdb_set_error("context creation failed: %s", cudaGetErrorString(e));
            rc = e == cudaErrorMemoryAllocation ? FDB_ERR_NOMEM : FDB_ERR_CUDA;
        }
        fdb_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return FDB_OK;
}

extern "C" int fdb_ctx_destroy(fdb_ctx* ctx) {
    if (!ctx) return FDB_OK;
    DeviceGuard guard(ctx->device);
    cudaDeviceSynchronize();
    free_pipeline(ctx);
    for (auto& s : ctx->slot)
        if (s.stream) cudaStreamDestroy(s.stream);
    cudaFree(ctx->d_partials);
    cudaFree(ctx->d_ticket);
    cudaFree(ctx->d_abs_sum);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    for (auto& p : ctx->h_out)
        if (p) cudaFreeHost(p);
    for (auto& p : ctx->h_in)
        if (p) cudaFreeHost(p);
    delete ctx;
    return FDB_OK;
}

extern "C" int fdb_ctx_device(const fdb_ctx* ctx) { return ctx ? ctx->device : -1; }
extern "C" int fdb_ctx_sm_count(const fdb_ctx* ctx) { return ctx ? ctx->sm_count : 0; }

extern "C" int fdb_pinned_alloc(size_t bytes, void** out) {
    FDB_REQUIRE(out != nullptr, FDB_ERR_ARG, "out is NULL");
    FDB_CUDA(cudaMallocHost(out, bytes ? bytes : 1));
    return FDB_OK;
}

extern "C" int fdb_pinned_free(void* p) {
    if (p) FDB_CUDA(cudaFreeHost(p));
    return FDB_OK;
}

// ---------------------------------------------------------------- K1 entry points
extern "C" int fdb_transform_dev(fdb_ctx* ctx, void* stream, int dtype, int64_t n_rows, const void* x,
                                 const double* mean, const double* scale, void* out) {
    int rc = check_common(ctx, dtype, n_rows, x);
    if (rc != FDB_OK) return rc;
    FDB_REQUIRE(mean && scale, FDB_ERR_ARG, "mean/scale is NULL");
    FDB_REQUIRE(n_rows == 0 || out != nullptr, FDB_ERR_ARG, "out is NULL");
    FDB_REQUIRE(fdb_aligned16(x) && fdb_aligned16(out), FDB_ERR_ALIGN, "x and out must be 16-byte aligned");
    LinearParams p;
    fill_params(p, mean, scale, 0.0);
    return fdb_launch_linear_pass(ctx, (cudaStream_t)stream, dtype, n_rows, x, p, true, nullptr, 1, nullptr, out,
                                  nullptr, nullptr, nullptr, nullptr);
}

extern "C" int fdb_transform_host(fdb_ctx* ctx, int dtype, int64_t n_rows, const void* x, const double* mean,
                                  const double* scale, void* out) {
    int rc = check_common(ctx, dtype, n_rows, x);
    if (rc != FDB_OK) return rc;
    FDB_REQUIRE(mean && scale, FDB_ERR_ARG, "mean/scale is NULL");
    FDB_REQUIRE(n_rows == 0 || out != nullptr, FDB_ERR_ARG, "out is NULL");
    LinearParams p;
    fill_params(p, mean, scale, 0.0);
    return run_host(ctx, dtype, n_rows, x, p, true, nullptr, 1, nullptr, out, nullptr, nullptr);
}

extern "C" int fdb_linear_pass_dev(fdb_ctx* ctx, void* stream, int dtype, int64_t n_rows, const void* x,
                                   const double* sub, const double* mul, double bias, void* proba, int proba_cols,
                                   void* logit, void* phi, uint8_t* cls, double* abs_sum) {
    int rc = check_common(ctx, dtype, n_rows, x);
    if (rc != FDB_OK) return rc;
    FDB_REQUIRE(mul != nullptr, FDB_ERR_ARG, "mul is NULL");
    FDB_REQUIRE(proba_cols == 1 || proba_cols == 2, FDB_ERR_ARG, "proba_cols must be 1 or 2");
    FDB_REQUIRE(fdb_aligned16(x) && fdb_aligned16(phi), FDB_ERR_ALIGN, "x and phi must be 16-byte aligned");
    LinearParams p;
    fill_params(p, sub, mul, bias);
    return fdb_launch_linear_pass(ctx, (cudaStream_t)stream, dtype, n_rows, x, p, false, proba, proba_cols, logit, phi, cls,
                                  abs_sum, ctx->d_partials, ctx->d_ticket);
}

extern "C" int fdb_linear_pass_host(fdb_ctx* ctx, int dtype, int64_t n_rows, const void* x, const double* sub,
                                    const double* mul, double bias, void* proba, int proba_cols, void* logit, void* phi,
                                    uint8_t* cls, double* abs_sum) {
    int rc = check_common(ctx, dtype, n_rows, x);
    if (rc != FDB_OK) return rc;
    FDB_REQUIRE(mul != nullptr, FDB_ERR_ARG, "mul is NULL");
    FDB_REQUIRE(proba_cols == 1 || proba_cols == 2, FDB_ERR_ARG, "proba_cols must be 1 or 2");
    LinearParams p;
    fill_params(p, sub, mul, bias);
    return run_host(ctx, dtype, n_rows, x, p, false, proba, proba_cols, logit, phi, cls, abs_sum);
}
