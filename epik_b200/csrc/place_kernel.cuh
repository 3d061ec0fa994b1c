// The placement kernel: one warp per query sequence, persistent CTAs (one per SM).
//
// Replaces the body of the reference's OpenMP loop (epik/src/epik/place.cpp:230-268), i.e.
// place_seq (:320-440) + query_kmers (:278-316) + sum_scores (:164-184) +
// select_best_placements (:134-159) + the LWR loop (:241-264) + filter_by_ratio (:188-199).
//
// Per query a warp
//   1. stages the letter codes of a chunk of CHUNK windows in shared memory, builds the k-mer
//      codes (2 or 5 bits per letter) and probes the key table, lanes striding over windows;
//   2. compacts the hits in window order and gathers their posting lists with 128-bit loads
//      (two (branch, score) entries per lane), U lists in flight per lane, and adds them to
//      the per-branch accumulators {float S, uint32 C} held in shared memory. Lists are applied
//      in window order and a list never names a branch twice, so every S[b] sees exactly the
//      additions of place.cpp:358-367 in the same order: the sums are bit-identical to the
//      reference's and no atomics are needed;
//   3. (only for reads that have windows with one ambiguous letter) replays place.cpp:373-415;
//   4. sweeps the accumulators once: omega-threshold correction (place.cpp:417-422), FP64
//      10^S summed over all touched branches (sum_scores), a per-lane top-K kept in registers,
//      accumulators reset to zero on the way;
//   5. merges the per-lane lists with warp shuffles, computes the likelihood weight ratios in
//      FP64 and applies keep_factor.
// Float expressions that the reference evaluates without fused multiply-add are written with
// __fmul_rn/__fadd_rn/__fdiv_rn so that nvcc cannot contract them.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>
#include "device_db.cuh"

namespace epk
{
constexpr int CHUNK = 256;             // windows staged per pass
constexpr int ROUNDS = CHUNK / 32;     // windows per lane per pass
constexpr int CODE_BYTES = CHUNK + 32; // letter codes of one pass (k - 1 <= 31 extra letters)
constexpr int MAX_WARPS_PER_CTA = 16;
constexpr uint32_t NO_BRANCH = 0xFFFFFFFFu;

enum counter_slot { CTR_CURSOR = 0, CTR_WINDOWS = 1, CTR_HITS = 2, CTR_ENTRIES = 3, CTR_SHORT = 4, CTR_COUNT = 8 };

struct place_params
{
    device_db db;
    const char* seq;             // packed letters
    const uint64_t* seq_off;     // [n + 1]
    uint32_t n;
    uint32_t keep_at_most;
    double keep_factor;
    uint32_t* n_place;           // outputs, see epk_result
    uint32_t* branch;
    float* score;
    double* lwr;
    uint32_t* count;
    const uint32_t* letters;     // [256] letter table
    const uint32_t* fake_order;  // [keep_at_most] order of the zero-hit placements
    unsigned long long* counters; // [CTR_COUNT]
};

// bytes of shared memory one warp needs for a tree of n_nodes
__host__ __device__ inline size_t warp_smem_bytes(uint32_t n_nodes)
{
    size_t acc = (static_cast<size_t>(n_nodes) * 8 + 15) & ~static_cast<size_t>(15);
    size_t seen = ((static_cast<size_t>(n_nodes) + 31) / 32 * 4 + 15) & ~static_cast<size_t>(15);
    return acc + seen + CHUNK * 8 + CODE_BYTES;
}
constexpr size_t CTA_SMEM_FIXED = 1024; // letter table

struct window_class
{
    uint32_t key;      // code with 0 in the place of an ambiguous letter
    uint32_t gaps;     // letters that are gaps / unknown
    uint32_t ambs;     // ambiguous letters
    uint32_t amb_at;   // index of the (last) ambiguous letter within the window
};

// i2l::to_kmers<one_ambiguity_policy> for one window (call site place.cpp:294): leftmost letter
// in the most significant position; a window is an exact k-mer when gaps == 0 && ambs == 0, an
// ambiguous one when gaps == 0 && ambs == 1, skipped otherwise.
template<class CodePtr>
__device__ __forceinline__ window_class classify_window(CodePtr codes, uint32_t w, uint32_t k, uint32_t bits)
{
    window_class c{ 0, 0, 0, 0 };
    for (uint32_t j = 0; j < k; ++j)
    {
        const uint32_t l = codes[w + j];
        c.key = (c.key << bits) | (l & 31u);
        c.gaps += (l >> 6) & 1u;
        if (l & LETTER_AMBIGUOUS)
        {
            ++c.ambs;
            c.amb_at = j;
        }
    }
    return c;
}

__device__ __forceinline__ bool better(float s1, uint32_t b1, float s2, uint32_t b2)
{
    // higher score first (place.cpp:128-131); exact ties: lower branch id first (documented rule)
    return s1 > s2 || (s1 == s2 && b1 < b2);
}

// place.cpp:358-367 for one entry
__device__ __forceinline__ void add_entry(float2* acc, uint32_t branch, uint32_t score_bits)
{
    float2 a = acc[branch];
    a.x += __uint_as_float(score_bits);
    a.y = __uint_as_float(__float_as_uint(a.y) + 1u);
    acc[branch] = a;
}

// Gathers the posting lists of hits[0, nh) (packed offset|length, window order) and applies them
// in order. U segments of up to 64 entries are in flight per pass.
template<int U>
__device__ __forceinline__ uint64_t gather_hits(const device_db& db, float2* acc, const uint64_t* hits,
                                                uint32_t nh, uint32_t lane)
{
    uint64_t n_entries = 0;
    uint32_t h = 0, pos = 0;
    uint64_t cur = nh ? hits[0] : 0;
    while (h < nh)
    {
        uint4 v[U];
        uint32_t cnt[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            cnt[u] = 0;
            v[u] = make_uint4(0, 0, 0, 0);
            if (h < nh)
            {
                const uint64_t off = cur >> LEN_BITS;
                const uint32_t len = static_cast<uint32_t>(cur & LEN_MASK);
                const uint32_t c = min(len - pos, 64u);
                cnt[u] = c;
                if (2 * lane < c)
                {
                    v[u] = __ldg(reinterpret_cast<const uint4*>(db.postings + off + pos) + lane);
                }
                pos += 64;
                if (pos >= len)
                {
                    ++h;
                    pos = 0;
                    cur = h < nh ? hits[h] : 0;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            if (2 * lane < cnt[u])
            {
                add_entry(acc, v[u].x, v[u].y);
                if (2 * lane + 1 < cnt[u]) { add_entry(acc, v[u].z, v[u].w); }
            }
            n_entries += cnt[u];
            __syncwarp();
        }
    }
    return n_entries;
}

template<uint32_t TABLE, int KMAX, int U>
__global__ void __launch_bounds__(MAX_WARPS_PER_CTA * 32, 1) place_kernel(const place_params p)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const device_db& db = p.db;
    const uint32_t N = db.n_nodes;
    const uint32_t k = db.k;
    const uint32_t bits = db.bits;
    const uint32_t K = p.keep_at_most;

    uint32_t* letters = reinterpret_cast<uint32_t*>(smem);
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) { letters[i] = p.letters[i]; }

    unsigned char* base = smem + CTA_SMEM_FIXED + warp * warp_smem_bytes(N);
    const size_t acc_bytes = (static_cast<size_t>(N) * 8 + 15) & ~static_cast<size_t>(15);
    const size_t seen_words = (static_cast<size_t>(N) + 31) / 32;
    const size_t seen_bytes = (seen_words * 4 + 15) & ~static_cast<size_t>(15);
    float2* acc = reinterpret_cast<float2*>(base);
    uint32_t* seen = reinterpret_cast<uint32_t*>(base + acc_bytes);
    uint64_t* hits = reinterpret_cast<uint64_t*>(base + acc_bytes + seen_bytes);
    uint8_t* codes = reinterpret_cast<uint8_t*>(base + acc_bytes + seen_bytes + CHUNK * 8);

    for (uint32_t b = lane; b < N; b += 32) { acc[b] = make_float2(0.0f, 0.0f); }
    for (uint32_t i = lane; i < seen_words; i += 32) { seen[i] = 0; }
    __syncthreads();

    const float kf = static_cast<float>(k);
    const float logthr = db.log_threshold;
    uint64_t ctr_windows = 0, ctr_hits = 0, ctr_entries = 0;

    while (true)
    {
        uint32_t q = 0;
        if (lane == 0) { q = static_cast<uint32_t>(atomicAdd(p.counters + CTR_CURSOR, 1ull)); }
        q = __shfl_sync(0xFFFFFFFFu, q, 0);
        if (q >= p.n) { break; }

        const uint64_t s0 = p.seq_off[q];
        const uint64_t len = p.seq_off[q + 1] - s0;
        if (len + 1 < k)
        {
            // the reference dies here (size_t wrap at place.cpp:291); reported as EPK_E_SHORT_SEQ
            if (lane == 0)
            {
                atomicAdd(p.counters + CTR_SHORT, 1ull);
                p.n_place[q] = 0;
            }
            continue;
        }
        const uint64_t n_windows = len + 1 - k; // place.cpp:322, counts skipped windows too
        const char* seq = p.seq + s0;
        bool has_ambiguous = false;

        // ---- exact k-mers, window order (place.cpp:294-304, 349-371) ----
        for (uint64_t c0 = 0; c0 < n_windows; c0 += CHUNK)
        {
            const uint32_t nw = static_cast<uint32_t>(min(n_windows - c0, static_cast<uint64_t>(CHUNK)));
            const uint32_t ncode = nw + k - 1;
            for (uint32_t i = lane; i < ncode; i += 32)
            {
                codes[i] = static_cast<uint8_t>(letters[static_cast<uint8_t>(seq[c0 + i])]);
            }
            __syncwarp();

            uint64_t slot[ROUNDS];
#pragma unroll
            for (int r = 0; r < ROUNDS; ++r)
            {
                slot[r] = 0;
                const uint32_t w = r * 32 + lane;
                if (w < nw)
                {
                    const window_class c = classify_window(codes, w, k, bits);
                    if (c.gaps == 0)
                    {
                        if (c.ambs == 0) { slot[r] = db_search<TABLE>(db, c.key); }
                        else if (c.ambs == 1) { has_ambiguous = true; }
                    }
                }
            }
            uint32_t nh = 0;
#pragma unroll
            for (int r = 0; r < ROUNDS; ++r)
            {
                const bool hit = (slot[r] & LEN_MASK) != 0;
                const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, hit);
                if (hit) { hits[nh + __popc(ballot & ((1u << lane) - 1u))] = slot[r]; }
                nh += __popc(ballot);
            }
            __syncwarp();
            ctr_windows += nw;
            ctr_hits += nh;
            ctr_entries += gather_hits<U>(db, acc, hits, nh, lane);
        }

        // ---- windows with exactly one ambiguous letter (place.cpp:306-313, 373-415) ----
        has_ambiguous = __any_sync(0xFFFFFFFFu, has_ambiguous);
        if (has_ambiguous)
        {
            const float thr_rest = __fmul_rn(static_cast<float>(k - 1), db.threshold);
            for (uint64_t c0 = 0; c0 < n_windows; c0 += CHUNK)
            {
                const uint32_t nw = static_cast<uint32_t>(min(n_windows - c0, static_cast<uint64_t>(CHUNK)));
                const uint32_t ncode = nw + k - 1;
                __syncwarp();
                for (uint32_t i = lane; i < ncode; i += 32)
                {
                    codes[i] = static_cast<uint8_t>(letters[static_cast<uint8_t>(seq[c0 + i])]);
                }
                __syncwarp();
                for (uint32_t r = 0; r < ROUNDS && r * 32 < nw; ++r)
                {
                    const uint32_t w = r * 32 + lane;
                    window_class c{ 0, 1, 0, 0 };
                    if (w < nw) { c = classify_window(codes, w, k, bits); }
                    uint32_t todo = __ballot_sync(0xFFFFFFFFu, c.gaps == 0 && c.ambs == 1);
                    while (todo)
                    {
                        const int src = __ffs(todo) - 1;
                        todo &= todo - 1;
                        const uint32_t base_key = __shfl_sync(0xFFFFFFFFu, c.key, src);
                        const uint32_t amb_at = __shfl_sync(0xFFFFFFFFu, c.amb_at, src);
                        const uint32_t letter = letters[static_cast<uint8_t>(seq[c0 + r * 32 + src + amb_at])];
                        uint32_t resolutions = letter >> 8;
                        const uint32_t shift = bits * (k - 1 - amb_at);
                        // one key per resolution, each on its own (place.cpp:308-312)
                        while (resolutions)
                        {
                            const uint32_t code = __ffs(resolutions) - 1;
                            resolutions &= resolutions - 1;
                            const uint64_t found = db_search<TABLE>(db, base_key | (code << shift));
                            const uint32_t flen = static_cast<uint32_t>(found & LEN_MASK);
                            const uint2* list = db.postings + (found >> LEN_BITS);
                            ctr_hits += flen != 0;
                            ctr_entries += flen;
                            for (uint32_t e = lane; e < flen; e += 32)
                            {
                                const uint2 entry = __ldg(list + e);
                                const uint32_t bit = 1u << (entry.x & 31u);
                                const uint32_t old = atomicOr(seen + (entry.x >> 5), bit);
                                if (!(old & bit))
                                {
                                    // first ambiguous key that names this branch: (10^s + (k-1)·thr)/k is
                                    // added to the LOG score (place.cpp:391,400-410)
                                    const float prob = static_cast<float>(pow(10.0, static_cast<double>(__uint_as_float(entry.y))));
                                    const float mean = __fdiv_rn(__fadd_rn(prob, thr_rest), kf);
                                    float2 a = acc[entry.x];
                                    a.x = __fadd_rn(a.x, mean);
                                    a.y = __uint_as_float(__float_as_uint(a.y) + 1u);
                                    acc[entry.x] = a;
                                }
                            }
                            __syncwarp();
                        }
                    }
                }
            }
            for (uint32_t i = lane; i < seen_words; i += 32) { seen[i] = 0; }
        }

        // ---- correction, likelihood mass, per-lane top-K (place.cpp:417-422, 164-184, 134-159) ----
        float top_s[KMAX];
        uint32_t top_b[KMAX], top_c[KMAX];
#pragma unroll
        for (int i = 0; i < KMAX; ++i)
        {
            top_s[i] = -INFINITY;
            top_b[i] = NO_BRANCH;
            top_c[i] = 0;
        }
        double mass = 0.0;
        uint32_t touched = 0;
        for (uint32_t b = lane; b < N; b += 32)
        {
            const float2 a = acc[b];
            const uint32_t c = __float_as_uint(a.y);
            if (c != 0)
            {
                acc[b] = make_float2(0.0f, 0.0f);
                float s = __fadd_rn(a.x, __fmul_rn(static_cast<float>(n_windows - c), logthr));
                s = __fdiv_rn(s, kf);
                mass += exp10(static_cast<double>(s));
                ++touched;
                if (better(s, b, top_s[KMAX - 1], top_b[KMAX - 1]))
                {
                    top_s[KMAX - 1] = s;
                    top_b[KMAX - 1] = b;
                    top_c[KMAX - 1] = c;
#pragma unroll
                    for (int i = KMAX - 1; i > 0; --i)
                    {
                        if (better(top_s[i], top_b[i], top_s[i - 1], top_b[i - 1]))
                        {
                            const float ts = top_s[i]; top_s[i] = top_s[i - 1]; top_s[i - 1] = ts;
                            const uint32_t tb = top_b[i]; top_b[i] = top_b[i - 1]; top_b[i - 1] = tb;
                            const uint32_t tc = top_c[i]; top_c[i] = top_c[i - 1]; top_c[i - 1] = tc;
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
        {
            mass += __shfl_xor_sync(0xFFFFFFFFu, mass, d);
            touched += __shfl_xor_sync(0xFFFFFFFFu, touched, d);
        }

        // ---- merge: round i elects the best remaining head; lane i keeps it ----
        float my_s = 0.0f;
        uint32_t my_b = NO_BRANCH, my_c = 0;
        const uint32_t n_keep = min(K, touched);
        for (uint32_t i = 0; i < n_keep; ++i)
        {
            float bs = top_s[0];
            uint32_t bb = top_b[0];
            uint32_t bl = lane;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1)
            {
                const float os = __shfl_xor_sync(0xFFFFFFFFu, bs, d);
                const uint32_t ob = __shfl_xor_sync(0xFFFFFFFFu, bb, d);
                const uint32_t ol = __shfl_xor_sync(0xFFFFFFFFu, bl, d);
                if (better(os, ob, bs, bb)) { bs = os; bb = ob; bl = ol; }
            }
            const uint32_t bc = __shfl_sync(0xFFFFFFFFu, top_c[0], bl);
            if (lane == i) { my_s = bs; my_b = bb; my_c = bc; }
            if (lane == bl)
            {
#pragma unroll
                for (int j = 0; j < KMAX - 1; ++j)
                {
                    top_s[j] = top_s[j + 1];
                    top_b[j] = top_b[j + 1];
                    top_c[j] = top_c[j + 1];
                }
                top_s[KMAX - 1] = -INFINITY;
                top_b[KMAX - 1] = NO_BRANCH;
            }
        }

        // ---- likelihood weight ratios and keep_factor (place.cpp:164-184, 241-267) ----
        const float nwf = static_cast<float>(n_windows);
        const float absent_score = __fdiv_rn(__fmul_rn(nwf, logthr), kf);
        const double absent_mass = static_cast<double>(static_cast<float>(N) - static_cast<float>(touched))
                                   * exp10(static_cast<double>(absent_score));
        const double total = absent_mass + mass;
        uint32_t n_out = n_keep;
        if (touched == 0)
        {
            // no k-mer found: keep_at_most placements on the first branches (place.cpp:141-152), in the
            // order std::partial_sort leaves all-tied elements in (precomputed on the host)
            n_out = K;
            if (lane < K)
            {
                my_s = absent_score;
                my_b = p.fake_order[lane];
                my_c = 0;
            }
        }
        double ratio = 0.0;
        double keep_factor = p.keep_factor;
        if (total == 0.0) { keep_factor = 0.0; }
        else if (lane < n_out)
        {
            const double power = exp10(static_cast<double>(my_s));
            ratio = power == 0.0 ? 0.0 : power / total;
        }
        const double cut = __shfl_sync(0xFFFFFFFFu, ratio, 0) * keep_factor;
        const bool keep = lane < n_out && ratio >= cut;
        const uint32_t kept = __ballot_sync(0xFFFFFFFFu, keep);
        const uint32_t rank = __popc(kept & ((1u << lane) - 1u));
        const uint32_t n_kept = __popc(kept);
        const size_t row = static_cast<size_t>(q) * K;
        if (keep)
        {
            p.branch[row + rank] = my_b;
            p.score[row + rank] = my_s;
            p.lwr[row + rank] = ratio;
            p.count[row + rank] = my_c;
        }
        if (lane >= n_kept && lane < K)
        {
            p.branch[row + lane] = 0;
            p.score[row + lane] = 0.0f;
            p.lwr[row + lane] = 0.0;
            p.count[row + lane] = 0;
        }
        if (lane == 0) { p.n_place[q] = n_kept; }
        __syncwarp();
    }

    if (lane == 0)
    {
        atomicAdd(p.counters + CTR_WINDOWS, static_cast<unsigned long long>(ctr_windows));
        atomicAdd(p.counters + CTR_HITS, static_cast<unsigned long long>(ctr_hits));
        atomicAdd(p.counters + CTR_ENTRIES, static_cast<unsigned long long>(ctr_entries));
    }
}
} // namespace epk
