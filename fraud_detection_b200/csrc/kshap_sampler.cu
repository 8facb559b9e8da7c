// On-device coalition sampler for Kernel SHAP.
//
// shap.KernelExplainer draws a fresh coalition sample for EVERY explained row [recalled, upstream
// _kernel.py::explain]: Shapley-kernel weights per coalition size, complete enumeration of the smallest /
// largest sizes while the budget allows, then random paired draws (a coalition and its complement) over the
// remaining sizes with duplicate merging, and renormalisation of the random part to the weight that is left.
// On the host that costs ~3.7 ms per row in Python; here one warp builds one row's set in a few microseconds,
// so the reference's per-row sampling becomes usable at 1e5+ rows.
//
// Equivalence is STATISTICAL, not bitwise: the schedule (enumerated block, size distribution, pairing,
// duplicate merging, 4 x budget cap on draws, weight renormalisation) is the reference's, the random stream is
// Philox4x32-10 (seed, subsequence = set * 32 + lane) instead of numpy's global Mersenne Twister.  The
// host sampler (fraud_detection_b200/sampler.py) remains the bit-compatible one.
//
// Per set (one warp, one CTA): every lane proposes one coalition per round (size from the remaining-size
// CDF, members by drawing the r-th still-unset feature with __fns), lane 0 then walks the 32 proposals in lane
// order through a shared-memory hash set, so a given (seed, set) always yields the same masks.

#include <curand_kernel.h>
#include <math.h>

#include <vector>

#include "kshap_internal.cuh"

namespace {

constexpr int kMaxSizes = 16;
// hash-set slots: >= 4 x the number of random primary draws (nsamples / 2), so probes stay short; the smaller
// table (24 KB) doubles the resident warps per SM for the common nsamples <= 2048

struct SamplerParams {
    uint32_t* bits;          // [S][m_pad]
    float* kw;               // [S][m_pad]
    int m_pad, nsamples, n_fixed;
    const uint32_t* fixed_bits;   // [n_fixed] enumerated block (same for every set)
    float fixed_w;                // weight of each enumerated coalition (all enumerated sizes share... see host)
    const float* fixed_kw;        // [n_fixed]
    int n_sizes;             // number of coalition sizes left for random draws
    int first_size;          // smallest of them
    int n_paired;            // sizes <= n_paired are drawn with their complement
    float cdf[kMaxSizes];    // cumulative distribution over the remaining sizes
    float weight_left;
    unsigned long long seed;
};

template <int kHashSlots, int kHashShift>
__global__ void __launch_bounds__(32) kshap_sample_sets(const SamplerParams p) {
    __shared__ uint32_t h_key[kHashSlots];
    __shared__ uint16_t h_idx[kHashSlots];
    const int s = blockIdx.x, lane = threadIdx.x;
    uint32_t* ob = p.bits + (size_t)s * p.m_pad;
    float* ow = p.kw + (size_t)s * p.m_pad;
    for (int i = lane; i < kHashSlots; i += 32) h_key[i] = 0u;        // 0 is never a drawn coalition (size >= 1)
    for (int i = lane; i < p.m_pad; i += 32) {
        ob[i] = i < p.n_fixed ? p.fixed_bits[i] : 0u;
        ow[i] = i < p.n_fixed ? p.fixed_kw[i] : 0.f;
    }
    __syncwarp();
    int samples_left = p.nsamples - p.n_fixed;
    if (samples_left <= 0 || p.n_sizes == 0) return;
    curandStatePhilox4_32_10_t rng;
    curand_init(p.seed, (unsigned long long)s * 32 + lane, 0, &rng);
    const uint32_t full = 0x3FFFFFFFu;
    int count = p.n_fixed;                       // next free slot (uniform across the warp)
    int draws_left = 4 * samples_left;           // the reference reads at most 4 x samples_left size draws
    while (samples_left > 0 && draws_left > 0) {
        // ---- every lane proposes one coalition
        const float u = curand_uniform(&rng);
        int si = 0;
        while (si < p.n_sizes - 1 && u > p.cdf[si]) ++si;
        const int size = p.first_size + si;
        uint32_t mask = 0u;
        for (int i = 0; i < size; ++i) {
            const uint32_t r = __umulhi(curand(&rng), (uint32_t)(FDB_F - i));       // uniform in [0, 30 - i)
            mask |= 1u << __fns(~mask & full, 0, (int)r + 1);
        }
        // ---- lane 0 commits the proposals in lane order (deterministic), the others idle
        const int take = draws_left < 32 ? draws_left : 32;
        for (int l = 0; l < take && samples_left > 0; ++l) {
            const uint32_t m = __shfl_sync(0xffffffffu, mask, l);
            const int msize = __popc(m);
            int slot = -1;
            bool is_new = false;
            if (lane == 0) {
                uint32_t h = (m * 2654435761u) >> kHashShift;
                while (true) {
                    const uint32_t k = h_key[h];
                    if (k == 0u) { h_key[h] = m; h_idx[h] = (uint16_t)count; slot = count; is_new = true; break; }
                    if (k == m) { slot = h_idx[h]; break; }
                    h = (h + 1) & (kHashSlots - 1);
                }
                if (is_new) {
                    ob[slot] = m;
                    ow[slot] = 1.0f;
                } else {
                    ow[slot] += 1.0f;
                }
            }
            is_new = __shfl_sync(0xffffffffu, (int)is_new, 0) != 0;
            slot = __shfl_sync(0xffffffffu, slot, 0);
            if (is_new) { ++count; --samples_left; }
            if (samples_left > 0 && msize <= p.n_paired) {                          // the complement rides along
                if (is_new) {
                    if (lane == 0) { ob[count] = full ^ m; ow[count] = 1.0f; }
                    ++count;
                    --samples_left;
                } else if (lane == 0) {
                    // the complement of a known coalition sits right after it - unless the budget ran out exactly
                    // when that coalition was added; a zero-weight slot there is still unused
                    if (ob[slot + 1] == (full ^ m)) ow[slot + 1] += 1.0f;
                }
            }
            __syncwarp();
        }
        draws_left -= take;
    }
    __syncwarp();
    // ---- random block weights sum to weight_left
    float tot = 0.f;
    for (int i = p.n_fixed + lane; i < count; i += 32) tot += ow[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    const float scale = tot > 0.f ? p.weight_left / tot : 0.f;
    for (int i = p.n_fixed + lane; i < count; i += 32) ow[i] *= scale;
}

double binom(int n, int k) {
    double r = 1.0;
    for (int i = 1; i <= k; ++i) r = r * (n - k + i) / i;
    return floor(r + 0.5);
}

void enumerate_size(int M, int size, std::vector<uint32_t>& out, bool paired) {
    std::vector<int> idx(size);
    for (int i = 0; i < size; ++i) idx[i] = i;
    const uint32_t full = (M >= 32) ? 0xFFFFFFFFu : ((1u << M) - 1u);
    while (true) {
        uint32_t b = 0;
        for (int i : idx) b |= 1u << i;
        out.push_back(b);
        if (paired) out.push_back(full ^ b);
        int i = size - 1;
        while (i >= 0 && idx[i] == M - size + i) --i;
        if (i < 0) break;
        ++idx[i];
        for (int j = i + 1; j < size; ++j) idx[j] = idx[j - 1] + 1;
    }
}

}  // namespace

int fdb_kshap_resolve_nsamples(int nsamples) {
    const long long cap = (1LL << FDB_F) - 2;
    if (nsamples <= 0) nsamples = 2 * FDB_F + 2048;      // "auto"
    return (long long)nsamples > cap ? (int)cap : nsamples;
}

// Fills plan->d_bits / d_kw for all sets.  nsamples must already be resolved; M is fixed at 30 features.
int fdb_kshap_sample_on_device(fdb_kshap_plan* pl, int nsamples, unsigned long long seed) {
    FDB_REQUIRE(nsamples >= 64 && nsamples <= 4096, FDB_ERR_ARG,
                "the on-device sampler supports 64 <= nsamples <= 4096 (got %d); use host-drawn masks otherwise", nsamples);
    const int M = FDB_F;
    const int n_sizes = (M - 1 + 1) / 2, n_paired = (M - 1) / 2;       // ceil((M-1)/2) = 15, floor = 14
    std::vector<double> wv(n_sizes);
    double sum = 0.0;
    for (int i = 1; i <= n_sizes; ++i) {
        wv[i - 1] = (M - 1.0) / (i * (double)(M - i));
        if (i <= n_paired) wv[i - 1] *= 2.0;
        sum += wv[i - 1];
    }
    for (double& w : wv) w /= sum;
    // complete enumeration of sizes while the budget allows (reference step 4)
    std::vector<uint32_t> fb;
    std::vector<float> fw;
    std::vector<double> rem = wv;
    double left = nsamples;
    int n_full = 0;
    for (int size = 1; size <= n_sizes; ++size) {
        double nsub = binom(M, size);
        const bool paired = size <= n_paired;
        if (paired) nsub *= 2.0;
        if (left * rem[size - 1] / nsub >= 1.0 - 1e-8) {
            ++n_full;
            left -= nsub;
            if (rem[size - 1] < 1.0) {
                const double f = 1.0 - rem[size - 1];
                for (double& r : rem) r /= f;
            }
            double w = wv[size - 1] / binom(M, size);
            if (paired) w /= 2.0;
            const size_t before = fb.size();
            enumerate_size(M, size, fb, paired);
            fw.insert(fw.end(), fb.size() - before, (float)w);
        } else {
            break;
        }
    }
    SamplerParams sp;
    sp.bits = pl->d_bits; sp.kw = pl->d_kw; sp.m_pad = pl->m_pad; sp.nsamples = nsamples;
    sp.n_fixed = (int)fb.size();
    sp.n_sizes = n_sizes - n_full; sp.first_size = n_full + 1; sp.n_paired = n_paired; sp.seed = seed;
    sp.fixed_w = 0.f;
    double weight_left = 0.0, csum = 0.0, tot = 0.0;
    std::vector<double> pr;
    for (int i = n_full; i < n_sizes; ++i) {
        weight_left += wv[i];
        double v = wv[i];
        if (i + 1 <= n_paired) v /= 2.0;                    // a paired draw yields two samples
        pr.push_back(v);
        tot += v;
    }
    for (size_t i = 0; i < pr.size() && i < (size_t)kMaxSizes; ++i) {
        csum += pr[i] / tot;
        sp.cdf[i] = (float)csum;
    }
    sp.weight_left = (float)weight_left;
    uint32_t* d_fb = nullptr;
    float* d_fw = nullptr;
    FDB_CUDA(cudaMalloc((void**)&d_fb, (fb.size() + 1) * 4));
    cudaError_t e = cudaMalloc((void**)&d_fw, (fw.size() + 1) * 4);
    if (e == cudaSuccess && !fb.empty()) e = cudaMemcpy(d_fb, fb.data(), fb.size() * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !fw.empty()) e = cudaMemcpy(d_fw, fw.data(), fw.size() * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        sp.fixed_bits = d_fb;
        sp.fixed_kw = d_fw;
        if (nsamples <= 2048) kshap_sample_sets<4096, 20><<<pl->n_sets, 32>>>(sp);
        else kshap_sample_sets<8192, 19><<<pl->n_sets, 32>>>(sp);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaFree(d_fb);
    cudaFree(d_fw);
    FDB_CUDA(e);
    return FDB_OK;
}
