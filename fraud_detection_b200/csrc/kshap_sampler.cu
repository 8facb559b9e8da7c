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
// Philox4x32-7 (key = seed, counter = (call, lane, set)) instead of numpy's global Mersenne Twister.  The
// host sampler (fraud_detection_b200/sampler.py) remains the bit-compatible one.
//
// Per set (one warp, one CTA): every lane proposes one coalition per round (size from the remaining-size
// CDF, members by Floyd's subset algorithm); the 32 proposals are then committed as if one after the other in
// lane order - parallel hash look-up, __match_any merging of equal proposals, slots from a prefix sum of the
// budget each proposal takes - so a given (seed, set) always yields the same masks.  The set is assembled in
// shared memory and written out once.

#include <math.h>

#include <new>
#include <vector>

#include "kshap_internal.cuh"

namespace {

constexpr int kMaxSizes = 16;

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

// Philox4x32 with 7 rounds (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3": 7 rounds pass BigCrush; the
// usual 10 are a safety margin), counter mode, stateless: two 32 x 32 -> 64 multiplies and three xors per round.
__device__ __forceinline__ uint4 philox4x32_7(uint32_t c0, uint32_t c1, uint32_t c2, unsigned long long key) {
    uint32_t c3 = 0u, k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
#pragma unroll
    for (int r = 0; r < 7; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0;
        c1 = lo1;
        c2 = hi0 ^ c3 ^ k1;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// Shared memory per set: the set itself (coalition words + draw counts) and an open-addressing hash table of slot
// indices (0xFFFF = empty; the key of an entry is s_bits[slot]; at most nsamples / 2 entries, load <= 1/2).
// 16 KB for nsamples <= 2048 (14 sets in flight per SM), 32 KB up to 4096.
template <int kHashSlots, int kHashShift, int kMaxSamples>
__global__ void __launch_bounds__(32) kshap_sample_sets(const SamplerParams p) {
    __shared__ __align__(16) uint32_t s_bits[kMaxSamples];
    __shared__ __align__(16) uint32_t s_cnt[kMaxSamples / 2];       // 16-bit draw counts, two per word (at most 4 x nsamples draws in all)
    __shared__ __align__(16) unsigned short h_slot[kHashSlots];
#ifdef FDB_SAMPLER_PAD
    __shared__ uint32_t s_pad[FDB_SAMPLER_PAD / 4];                  // occupancy experiment only (tools/variant_build.py)
    if (p.nsamples < 0) s_pad[threadIdx.x] = 1u;
    if (p.nsamples < -1) p.bits[0] = s_pad[(threadIdx.x + 1) & 31];
#endif
    const uint32_t allw = 0xffffffffu;
    const int s = blockIdx.x, lane = threadIdx.x;
    uint32_t* ob = p.bits + (size_t)s * p.m_pad;
    float* ow = p.kw + (size_t)s * p.m_pad;
    // 16-byte stores: the empty hash table (0xFFFF) and the zeroed draw counts; s_bits needs the enumerated block only
    // (entries beyond the final count are written out as zeros whatever shared memory holds)
    for (int i = lane; i < kHashSlots / 8; i += 32) reinterpret_cast<uint4*>(h_slot)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    for (int i = lane; i < kMaxSamples / 8; i += 32) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int i = lane; i < p.n_fixed; i += 32) s_bits[i] = p.fixed_bits[i];
    __syncwarp();
    int samples_left = p.nsamples - p.n_fixed;
    int count = p.n_fixed;                       // next free slot (uniform across the warp)
    const uint32_t full = 0x3FFFFFFFu;
    if (samples_left > 0 && p.n_sizes > 0) {
        uint32_t call = 0u;                      // Philox counter = (call, lane, set, 0), key = seed: one stream per (seed, set, lane)
        int draws_left = 4 * samples_left;       // the reference reads at most 4 x samples_left size draws
        while (samples_left > 0 && draws_left > 0) {
            // ---- every lane proposes one coalition: size from the remaining-size CDF, members by Floyd's
            //      algorithm (a uniform size-subset of the 30 features from `size` bounded draws)
            // Two Philox calls per proposal, always: word x of the first picks the size, the other words feed the members.
            // Three bounded draws per 32-bit word (t = hi(w n), w = lo(w n): the low word is as good as fresh for ranges
            // <= 30, the bias after three steps stays below 1e-4 relative), so five words cover the largest size (15).
            // Both the size look-up and Floyd's subset algorithm are fully unrolled and predicated - no loop whose trip
            // count is the largest size in the warp, no refill branch: ~120 instructions per proposal instead of ~600.
            const uint4 ra = philox4x32_7(call, (uint32_t)lane, (uint32_t)s, p.seed);
            const uint4 rb = philox4x32_7(call + 1u, (uint32_t)lane, (uint32_t)s, p.seed);
            call += 2u;
            const float u = (float)ra.x * 2.3283064365386963e-10f + 1.1641532182693481e-10f;   // (0, 1]
            int si = 0;
#pragma unroll
            for (int k = 0; k < kMaxSizes - 2; ++k) si += (u > p.cdf[k]) ? 1 : 0;
            si = si < p.n_sizes - 1 ? si : p.n_sizes - 1;
            const int size = p.first_size + si;
            const uint32_t words[5] = {ra.y, ra.z, ra.w, rb.x, rb.y};
            uint32_t mask = 0u;
#pragma unroll
            for (int w = 0; w < 5; ++w) {
                uint32_t cur = words[w];
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const int i = 3 * w + j;
                    const uint32_t n = (uint32_t)(FDB_F - size + i) + 1u;
                    const uint32_t t = __umulhi(cur, n);                                       // uniform in [0, n)
                    cur *= n;
                    const uint32_t pick = ((mask >> (t & 31u)) & 1u) ? (n - 1u) : t;
                    if (i < size) mask |= 1u << (pick & 31u);
                }
            }
            // ---- commit the 32 proposals AS IF one after the other in lane order (so (seed, set) is reproducible):
            //      look every proposal up, merge equal proposals of this round, hand out slots by a prefix sum
            const int take = draws_left < 32 ? draws_left : 32;
            const bool active = lane < take;
            const bool paired = __popc(mask) <= p.n_paired;
            int found = -1;
            if (active) {
                uint32_t h = (mask * 2654435761u) >> kHashShift;
                while (true) {
                    const uint32_t sl = h_slot[h];
                    if (sl == 0xFFFFu) break;
                    if (s_bits[sl] == mask) { found = (int)sl; break; }
                    h = (h + 1) & (kHashSlots - 1);
                }
            }
            const uint32_t peers = __match_any_sync(allw, active ? mask : (0x80000000u | (uint32_t)lane));
            const int leader = __ffs(peers) - 1;
            const bool is_new = active && found < 0 && leader == lane;
            const int inc = is_new ? (paired ? 2 : 1) : 0;
            int incl = inc;                                                               // inclusive prefix sum
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(allw, incl, o);
                if (lane >= o) incl += v;
            }
            const int before = incl - inc;
            const bool processed = active && samples_left - before > 0;                  // budget left when its turn comes
            int slot = is_new ? count + before : found;
            if (is_new && processed) {
                s_bits[slot] = mask;
                atomicAdd(&s_cnt[slot >> 1], 1u << ((slot & 1) * 16));
                if (paired && samples_left - before > 1) {                               // the complement rides along
                    s_bits[slot + 1] = full ^ mask;
                    atomicAdd(&s_cnt[(slot + 1) >> 1], 1u << (((slot + 1) & 1) * 16));
                }
            }
            {
                // Hash insert of this round's new coalitions without a 16-bit atomicCAS (an emulated loop around a 32-bit
                // CAS: it was a fifth of the kernel's instructions): every inserting lane walks to the first empty slot and
                // stores its index with a plain 16-bit store; where two lanes picked the same slot one store wins, the
                // other lane sees a foreign index after the warp barrier and walks on.  One trip in nearly every round.
                bool pending = is_new && processed;
                uint32_t h = (mask * 2654435761u) >> kHashShift;
                volatile unsigned short* hs = h_slot;
                while (__any_sync(allw, pending)) {
                    if (pending) {
                        while (hs[h] != 0xFFFFu) h = (h + 1) & (kHashSlots - 1);
                        hs[h] = (unsigned short)slot;
                    }
                    __syncwarp();
                    if (pending) {
                        if (hs[h] == (unsigned short)slot) pending = false;
                        else h = (h + 1) & (kHashSlots - 1);
                    }
                    __syncwarp();
                }
            }
            slot = __shfl_sync(allw, slot, leader);                                      // merged proposals follow their leader
            __syncwarp();
            if (processed && !is_new) {
                atomicAdd(&s_cnt[slot >> 1], 1u << ((slot & 1) * 16));
                // the complement of a known coalition sits right after it - unless the budget ran out exactly
                // when that coalition was added
                if (paired && slot + 1 < p.nsamples && s_bits[slot + 1] == (full ^ mask))
                    atomicAdd(&s_cnt[(slot + 1) >> 1], 1u << (((slot + 1) & 1) * 16));
            }
            int added = __shfl_sync(allw, incl, 31);
            if (added > samples_left) added = samples_left;
            count += added;
            samples_left -= added;
            draws_left -= take;
            __syncwarp();
        }
    }
    __syncwarp();
    // ---- write the set out; random block weights (draw counts) are scaled to sum to weight_left
    uint32_t tot = 0u;
    for (int i = p.n_fixed + lane; i < count; i += 32) tot += (s_cnt[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(allw, tot, o);
    const float scale = tot > 0u ? p.weight_left / (float)tot : 0.f;
    // four coalitions per lane and trip: LDS.128 + LDS.64 in, two STG.128 out (m_pad is a multiple of 128)
    for (int i4 = lane * 4; i4 < p.m_pad; i4 += 128) {
        uint4 b = make_uint4(0u, 0u, 0u, 0u);
        uint2 c = make_uint2(0u, 0u);
        if (i4 < kMaxSamples) {
            b = *reinterpret_cast<const uint4*>(s_bits + i4);
            c = *reinterpret_cast<const uint2*>(s_cnt + (i4 >> 1));
        }
        const uint32_t bb[4] = {b.x, b.y, b.z, b.w};
        const uint32_t cc[4] = {c.x & 0xFFFFu, c.x >> 16, c.y & 0xFFFFu, c.y >> 16};
        uint32_t ob4[4];
        float ow4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int i = i4 + j;
            const bool in_set = i < count;
            ob4[j] = in_set ? bb[j] : 0u;
            ow4[j] = !in_set ? 0.f : (i < p.n_fixed ? p.fixed_kw[i] : (float)cc[j] * scale);
        }
        *reinterpret_cast<uint4*>(ob + i4) = make_uint4(ob4[0], ob4[1], ob4[2], ob4[3]);
        *reinterpret_cast<float4*>(ow + i4) = make_float4(ow4[0], ow4[1], ow4[2], ow4[3]);
    }
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

// The sampling schedule (enumerated block, size distribution) depends only on nsamples: it is worked out and uploaded
// ONCE per plan and kept (plan->sampler_state), so that redrawing the sets is a single kernel launch with no
// allocation, copy or host synchronisation.
namespace {
struct SamplerState {
    SamplerParams sp;
    uint32_t* d_fb = nullptr;
    float* d_fw = nullptr;
};
}  // namespace

void fdb_kshap_sampler_free(fdb_kshap_plan* pl) {
    SamplerState* st = static_cast<SamplerState*>(pl->sampler_state);
    if (!st) return;
    cudaFree(st->d_fb);
    cudaFree(st->d_fw);
    delete st;
    pl->sampler_state = nullptr;
}

static int sampler_prepare(fdb_kshap_plan* pl, int nsamples) {
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
    SamplerState* st = new (std::nothrow) SamplerState();
    FDB_REQUIRE(st != nullptr, FDB_ERR_NOMEM, "out of host memory");
    SamplerParams& sp = st->sp;
    sp.bits = pl->d_bits; sp.kw = pl->d_kw; sp.m_pad = pl->m_pad; sp.nsamples = nsamples;
    sp.n_fixed = (int)fb.size();
    sp.n_sizes = n_sizes - n_full; sp.first_size = n_full + 1; sp.n_paired = n_paired; sp.seed = 0ULL;
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
    for (int i = 0; i < kMaxSizes; ++i) sp.cdf[i] = 1.0f;
    for (size_t i = 0; i < pr.size() && i < (size_t)kMaxSizes; ++i) {
        csum += pr[i] / tot;
        sp.cdf[i] = (float)csum;
    }
    sp.weight_left = (float)weight_left;
    cudaError_t e = cudaMalloc((void**)&st->d_fb, (fb.size() + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc((void**)&st->d_fw, (fw.size() + 1) * 4);
    if (e == cudaSuccess && !fb.empty()) e = cudaMemcpy(st->d_fb, fb.data(), fb.size() * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !fw.empty()) e = cudaMemcpy(st->d_fw, fw.data(), fw.size() * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(st->d_fb); cudaFree(st->d_fw);
        delete st;
        FDB_CUDA(e);
    }
    sp.fixed_bits = st->d_fb;
    sp.fixed_kw = st->d_fw;
    pl->sampler_state = st;
    return FDB_OK;
}

// Fills plan->d_bits / d_kw for all sets: one kernel on `stream`, no synchronisation.  nsamples must already be
// resolved (and stay the same for the life of the plan); M is fixed at 30 features.
int fdb_kshap_sample_on_device(fdb_kshap_plan* pl, cudaStream_t stream, int nsamples, unsigned long long seed) {
    if (pl->sampler_state == nullptr) {
        int rc = sampler_prepare(pl, nsamples);
        if (rc != FDB_OK) return rc;
    }
    SamplerState* st = static_cast<SamplerState*>(pl->sampler_state);
    FDB_REQUIRE(st->sp.nsamples == nsamples, FDB_ERR_STATE, "the plan was created for nsamples = %d", st->sp.nsamples);
    SamplerParams sp = st->sp;
    sp.seed = seed;
    if (nsamples <= 2048) kshap_sample_sets<2048, 21, 2048><<<pl->n_sets, 32, 0, stream>>>(sp);
    else kshap_sample_sets<4096, 20, 4096><<<pl->n_sets, 32, 0, stream>>>(sp);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
