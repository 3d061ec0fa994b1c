// The placement kernel: one warp per query sequence, persistent CTAs (one per SM).
//
// Replaces the body of the reference's OpenMP loop (epik/src/epik/place.cpp:230-268), i.e.
// place_seq (:320-440) + query_kmers (:278-316) + sum_scores (:164-184) +
// select_best_placements (:134-159) + the LWR loop (:241-264) + filter_by_ratio (:188-199).
//
// Per query a warp
//   1. stages the letter codes of a chunk of CHUNK windows in shared memory; every lane rolls the
//      2-bit / 5-bit k-mer code over WPL consecutive windows and probes the key table (WPL
//      independent loads in flight per lane);
//   2. expands the hits, in window order, into segment descriptors (<= 64 postings each) and
//      gathers them with 128-bit loads (two (branch, score) postings per lane), U segments in
//      flight while the previous U are being applied to the per-branch accumulators
//      {float S, uint32 C} in shared memory. Segments are applied in window order and a list
//      never names a branch twice, so every S[b] sees exactly the additions of place.cpp:358-367
//      in the same order: the sums are bit-identical to the reference's, without atomics;
//   3. (only for reads that have windows with one ambiguous letter) replays place.cpp:373-415;
//   4. sweeps the accumulators twice. Pass A: omega-threshold correction (place.cpp:417-422),
//      FP64 10^S summed over all touched branches (sum_scores) and the per-lane maximum. The
//      keep_at_most-th largest lane maximum is a lower bound of the keep_at_most-th best score,
//      so pass B only has to collect the few branches at or above it (and clears the
//      accumulators on the way);
//   5. picks the best keep_at_most candidates with warp reductions, computes the likelihood
//      weight ratios in FP64 and applies keep_factor.
// Float expressions that the reference evaluates without fused multiply-add are written with
// __fmul_rn/__fadd_rn so that nvcc cannot contract them.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>
#include "device_db.cuh"

namespace epk
{
constexpr int WPL = 8;                 // consecutive windows per lane per chunk
constexpr int CHUNK = 32 * WPL;        // windows staged per pass
constexpr int CODE_BYTES = CHUNK + 32; // letter codes of one pass (k - 1 <= 31 extra letters)
constexpr int SEG_CAP = 256;           // segment descriptors buffered per warp
constexpr int SEG_LEN = 64;            // postings per segment: one 128-bit load per lane
constexpr int CAND_CAP = 128;          // top-k candidates buffered per warp (aliases the segments)
constexpr int MAX_WARPS_PER_CTA = 8;
constexpr uint32_t NO_BRANCH = 0xFFFFFFFFu;
constexpr uint32_t FULL = 0xFFFFFFFFu;

static_assert(CAND_CAP * 12 <= SEG_CAP * 8, "candidate arrays alias the segment buffer");

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
    return acc + seen + SEG_CAP * 8 + CODE_BYTES;
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

// ---- arithmetic helpers -------------------------------------------------------------------

// 10^s in FP64 for a float exponent (sum_scores and the LWR loop call std::pow(10.0, double(s)),
// place.cpp:44-47,175,254). 10^s = 2^n * 2^f with n = rint(s*log2(10)), |f| <= 1/2; 2^f by its
// Taylor polynomial of degree 12 (truncation 1.7e-16 relative), log2(10) carried in two doubles
// so that the reduction loses nothing. About 1.5 ulp; exponents near the FP64 range ends go to
// the library routine (gradual underflow is part of the reference's behaviour for long queries).
__device__ __forceinline__ double exp10_of_float(float s)
{
    if (!(fabsf(s) < 300.0f)) { return exp10(static_cast<double>(s)); }
    const double x = static_cast<double>(s);
    const double t = x * 3.321928094887362;
    const double t_lo = fma(x, 1.661617516973592e-16, fma(x, 3.321928094887362, -t));
    const double shifted = t + 6755399441055744.0; // 2^52 + 2^51: integer part lands in the low word
    const int n = __double2loint(shifted);
    const double f = (t - (shifted - 6755399441055744.0)) + t_lo;
    double p = 2.5678435993488206e-11;
    p = fma(p, f, 4.4455382718708116e-10);
    p = fma(p, f, 7.054911620801123e-09);
    p = fma(p, f, 1.01780860092397e-07);
    p = fma(p, f, 1.321548679014431e-06);
    p = fma(p, f, 1.5252733804059841e-05);
    p = fma(p, f, 0.0001540353039338161);
    p = fma(p, f, 0.0013333558146428443);
    p = fma(p, f, 0.009618129107628477);
    p = fma(p, f, 0.05550410866482158);
    p = fma(p, f, 0.24022650695910072);
    p = fma(p, f, 0.6931471805599453);
    p = fma(p, f, 1.0);
    return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}

// x / kf, correctly rounded (place.cpp:421 divides the corrected score by the k-mer size):
// q = x*r, one FMA residual step with r = RN(1/kf). Checked exhaustively against x / kf for
// every float with 1e-30 < |x| < 1e30 and kf = 1..32; the rest takes the IEEE routine.
__device__ __forceinline__ float div_by_k(float x, float kf, float rcp)
{
    const float ax = fabsf(x);
    if (!(ax > 1e-30f && ax < 1e30f)) { return __fdiv_rn(x, kf); }
    const float q = __fmul_rn(x, rcp);
    return __fmaf_rn(__fmaf_rn(-q, kf, x), rcp, q);
}

// order-preserving map float -> uint32 (and back); -0 and +0 are made one value first
__device__ __forceinline__ uint32_t score_key(float s)
{
    const uint32_t u = __float_as_uint(__fadd_rn(s, 0.0f));
    return (u >> 31) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_score(uint32_t key)
{
    return __uint_as_float((key >> 31) ? (key & 0x7FFFFFFFu) : ~key);
}

__device__ __forceinline__ uint32_t warp_inclusive_sum(uint32_t v, uint32_t lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1)
    {
        const uint32_t o = __shfl_up_sync(FULL, v, d);
        if (lane >= static_cast<uint32_t>(d)) { v += o; }
    }
    return v;
}

// ---- gathering posting lists ---------------------------------------------------------------

// segment descriptor: (index of the first posting) << 7 | number of postings (1..64)
template<int U>
__device__ __forceinline__ void load_segments(const uint2* __restrict__ postings, const uint64_t* segs, uint32_t first,
                                              uint32_t n, uint32_t lane, uint4 (&v)[U], uint32_t (&cnt)[U])
{
#pragma unroll
    for (int u = 0; u < U; ++u)
    {
        cnt[u] = 0;
        v[u] = make_uint4(0, 0, 0, 0);
        if (first + u < n)
        {
            const uint64_t d = segs[first + u];
            cnt[u] = static_cast<uint32_t>(d) & 127u;
            if (2 * lane < cnt[u])
            {
                v[u] = __ldg(reinterpret_cast<const uint4*>(postings + (d >> 7)) + lane);
            }
        }
    }
}

// place.cpp:358-367 for up to 64 postings of one list (two per lane; branches are distinct)
template<int U>
__device__ __forceinline__ void apply_segments(float2* acc, uint32_t lane, const uint4 (&v)[U], const uint32_t (&cnt)[U])
{
#pragma unroll
    for (int u = 0; u < U; ++u)
    {
        if (cnt[u] == 0) { break; }
        const bool p0 = 2 * lane < cnt[u];
        const bool p1 = 2 * lane + 1 < cnt[u];
        float2 a0 = make_float2(0.0f, 0.0f), a1 = make_float2(0.0f, 0.0f);
        if (p0) { a0 = acc[v[u].x]; }
        if (p1) { a1 = acc[v[u].z]; }
        a0.x = __fadd_rn(a0.x, __uint_as_float(v[u].y));
        a1.x = __fadd_rn(a1.x, __uint_as_float(v[u].w));
        a0.y = __uint_as_float(__float_as_uint(a0.y) + 1u);
        a1.y = __uint_as_float(__float_as_uint(a1.y) + 1u);
        if (p0) { acc[v[u].x] = a0; }
        if (p1) { acc[v[u].z] = a1; }
        __syncwarp();
    }
}

// Applies segs[0, n) in order; U segments are in flight while the previous U are applied.
template<int U>
__device__ __noinline__ void gather_segments(const uint2* __restrict__ postings, float2* acc, const uint64_t* segs,
                                             uint32_t n, uint32_t lane)
{
    uint4 va[U], vb[U];
    uint32_t ca[U], cb[U];
    load_segments<U>(postings, segs, 0, n, lane, va, ca);
    for (uint32_t first = 0; first < n; first += 2 * U)
    {
        if (first + U < n) { load_segments<U>(postings, segs, first + U, n, lane, vb, cb); }
        apply_segments<U>(acc, lane, va, ca);
        if (first + U >= n) { break; }
        if (first + 2 * U < n) { load_segments<U>(postings, segs, first + 2 * U, n, lane, va, ca); }
        apply_segments<U>(acc, lane, vb, cb);
    }
}

// One long list on its own (only when a single chunk's lists overflow the segment buffer).
__device__ __noinline__ void apply_list(const uint2* __restrict__ postings, float2* acc, uint64_t off, uint32_t len, uint32_t lane)
{
    for (uint32_t pos = 0; pos < len; pos += 4 * SEG_LEN)
    {
        uint4 v[4];
        uint32_t cnt[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            const uint32_t at = pos + u * SEG_LEN;
            cnt[u] = at < len ? min(len - at, static_cast<uint32_t>(SEG_LEN)) : 0u;
            v[u] = make_uint4(0, 0, 0, 0);
            if (2 * lane < cnt[u]) { v[u] = __ldg(reinterpret_cast<const uint4*>(postings + off + at) + lane); }
        }
        apply_segments<4>(acc, lane, v, cnt);
    }
}

// ---- top-k selection --------------------------------------------------------------------------

// Candidates (score key, branch, count) live in three arrays of CAND_CAP words. Selects the best
// `want` of the m buffered ones (higher score first, place.cpp:128-131; exact ties: lower branch
// id first — the documented tie rule), leaves them sorted in slots [0, n_sel) and in the registers
// of lanes [0, n_sel): lane i holds the i-th best. Returns n_sel = min(want, m).
__device__ __forceinline__ uint32_t select_best(uint32_t* cand_key, uint32_t* cand_branch, uint32_t* cand_count, uint32_t m,
                                                uint32_t want, uint32_t lane, uint32_t& my_key, uint32_t& my_branch,
                                                uint32_t& my_count)
{
    constexpr int PER_LANE = CAND_CAP / 32;
    uint32_t hi[PER_LANE], lo[PER_LANE], cn[PER_LANE];
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i)
    {
        const uint32_t j = lane + 32u * i;
        const bool have = j < m;
        hi[i] = have ? cand_key[j] : 0u;
        lo[i] = have ? ~cand_branch[j] : 0u;
        cn[i] = have ? cand_count[j] : 0u;
    }
    __syncwarp();
    my_key = 0;
    my_branch = NO_BRANCH;
    my_count = 0;
    uint32_t n_sel = 0;
    const uint32_t rounds = min(want, m);
    for (uint32_t round = 0; round < rounds; ++round)
    {
        // my best remaining candidate: largest (hi, lo)
        uint32_t bh = hi[0], bl = lo[0], bc = cn[0];
#pragma unroll
        for (int i = 1; i < PER_LANE; ++i)
        {
            const bool gt = hi[i] > bh || (hi[i] == bh && lo[i] > bl);
            bh = gt ? hi[i] : bh;
            bl = gt ? lo[i] : bl;
            bc = gt ? cn[i] : bc;
        }
        const uint32_t top_hi = __reduce_max_sync(FULL, bh);
        const uint32_t top_lo = __reduce_max_sync(FULL, bh == top_hi ? bl : 0u);
        const bool mine = bh == top_hi && bl == top_lo && bl != 0u; // branches are distinct: one lane
        const uint32_t src = __ffs(__ballot_sync(FULL, mine)) - 1u;
        const uint32_t top_count = __shfl_sync(FULL, bc, src);
        if (lane == round)
        {
            my_key = top_hi;
            my_branch = ~top_lo;
            my_count = top_count;
        }
        if (mine)
        {
#pragma unroll
            for (int i = 0; i < PER_LANE; ++i)
            {
                if (hi[i] == top_hi && lo[i] == top_lo) { hi[i] = 0u; lo[i] = 0u; }
            }
        }
        ++n_sel;
    }
    if (lane < n_sel)
    {
        cand_key[lane] = my_key;
        cand_branch[lane] = my_branch;
        cand_count[lane] = my_count;
    }
    __syncwarp();
    return n_sel;
}

// ---- the kernel ----------------------------------------------------------------------------------

template<uint32_t TABLE, int U>
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
    uint64_t* segs = reinterpret_cast<uint64_t*>(base + acc_bytes + seen_bytes);
    uint32_t* cand_key = reinterpret_cast<uint32_t*>(segs); // the segment buffer is idle during the sweeps
    uint32_t* cand_branch = cand_key + CAND_CAP;
    uint32_t* cand_count = cand_branch + CAND_CAP;
    uint8_t* codes = reinterpret_cast<uint8_t*>(base + acc_bytes + seen_bytes + SEG_CAP * 8);

    for (uint32_t b = lane; b < N; b += 32) { acc[b] = make_float2(0.0f, 0.0f); }
    for (uint32_t i = lane; i < seen_words; i += 32) { seen[i] = 0; }
    __syncthreads();

    const float kf = static_cast<float>(k);
    const float rcp_k = __frcp_rn(kf);
    const float logthr = db.log_threshold;
    const uint32_t key_mask = db.key_mask;
    uint64_t ctr_windows = 0, ctr_hits = 0, ctr_entries = 0; // per lane, folded once at the end

    while (true)
    {
        uint32_t q = 0;
        if (lane == 0) { q = static_cast<uint32_t>(atomicAdd(p.counters + CTR_CURSOR, 1ull)); }
        q = __shfl_sync(FULL, q, 0);
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
        uint32_t nseg = 0;
        if (lane == 0) { ctr_windows += n_windows; }

        // ---- exact k-mers, window order (place.cpp:294-304, 349-371) ----
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

            // lane L owns windows [L*WPL, L*WPL + WPL): one rolling code, WPL probes in flight
            uint64_t slot[WPL];
#pragma unroll
            for (int r = 0; r < WPL; ++r) { slot[r] = 0; }
            const uint32_t w0 = lane * WPL;
            if (w0 < nw)
            {
                uint32_t key = 0;
                int last_gap = -1, last_amb = -1, prev_amb = -1; // positions of the latest such letters
                for (uint32_t j = 0; j + 1 < k; ++j)
                {
                    const uint32_t l = codes[w0 + j];
                    const int at = static_cast<int>(w0 + j);
                    key = (key << bits) | (l & 31u);
                    last_gap = (l & LETTER_GAP) ? at : last_gap;
                    prev_amb = (l & LETTER_AMBIGUOUS) ? last_amb : prev_amb;
                    last_amb = (l & LETTER_AMBIGUOUS) ? at : last_amb;
                }
#pragma unroll
                for (int r = 0; r < WPL; ++r)
                {
                    const uint32_t w = w0 + r;
                    if (w < nw)
                    {
                        const uint32_t l = codes[w + k - 1];
                        const int at = static_cast<int>(w + k - 1);
                        key = ((key << bits) | (l & 31u)) & key_mask;
                        last_gap = (l & LETTER_GAP) ? at : last_gap;
                        prev_amb = (l & LETTER_AMBIGUOUS) ? last_amb : prev_amb;
                        last_amb = (l & LETTER_AMBIGUOUS) ? at : last_amb;
                        if (last_gap < static_cast<int>(w))
                        {
                            if (last_amb < static_cast<int>(w)) { slot[r] = db_search<TABLE>(db, key); }
                            else if (prev_amb < static_cast<int>(w)) { has_ambiguous = true; }
                        }
                    }
                }
            }

            // segments of this chunk's hits; window order = lane-major
            uint32_t my_segs = 0;
#pragma unroll
            for (int r = 0; r < WPL; ++r)
            {
                const uint32_t l = static_cast<uint32_t>(slot[r] & LEN_MASK);
                my_segs += (l + SEG_LEN - 1) / SEG_LEN;
                ctr_hits += l != 0;
                ctr_entries += l;
            }
            const uint32_t incl = warp_inclusive_sum(my_segs, lane);
            const uint32_t total = __shfl_sync(FULL, incl, 31);
            if (total == 0) { continue; }
            if (nseg + total > SEG_CAP)
            {
                __syncwarp();
                gather_segments<U>(db.postings, acc, segs, nseg, lane);
                nseg = 0;
                __syncwarp();
            }
            if (total <= SEG_CAP)
            {
                uint32_t at = nseg + incl - my_segs;
#pragma unroll
                for (int r = 0; r < WPL; ++r)
                {
                    const uint32_t l = static_cast<uint32_t>(slot[r] & LEN_MASK);
                    const uint64_t off = slot[r] >> LEN_BITS;
                    for (uint32_t pos = 0; pos < l; pos += SEG_LEN)
                    {
                        segs[at++] = ((off + pos) << 7) | min(l - pos, static_cast<uint32_t>(SEG_LEN));
                    }
                }
                nseg += total;
            }
            else
            {
                // more postings in one chunk than the buffer describes: list by list
                for (uint32_t src = 0; src < 32; ++src)
                {
#pragma unroll
                    for (int r = 0; r < WPL; ++r)
                    {
                        const uint64_t s = __shfl_sync(FULL, slot[r], src);
                        const uint32_t l = static_cast<uint32_t>(s & LEN_MASK);
                        if (l) { apply_list(db.postings, acc, s >> LEN_BITS, l, lane); }
                    }
                }
            }
        }
        __syncwarp();
        if (nseg) { gather_segments<U>(db.postings, acc, segs, nseg, lane); }
        __syncwarp();

        // ---- windows with exactly one ambiguous letter (place.cpp:306-313, 373-415) ----
        has_ambiguous = __any_sync(FULL, has_ambiguous);
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
                for (uint32_t r = 0; r * 32 < nw; ++r)
                {
                    const uint32_t w = r * 32 + lane;
                    window_class c{ 0, 1, 0, 0 };
                    if (w < nw) { c = classify_window(codes, w, k, bits); }
                    uint32_t todo = __ballot_sync(FULL, c.gaps == 0 && c.ambs == 1);
                    while (todo)
                    {
                        const int src = __ffs(todo) - 1;
                        todo &= todo - 1;
                        const uint32_t base_key = __shfl_sync(FULL, c.key, src);
                        const uint32_t amb_at = __shfl_sync(FULL, c.amb_at, src);
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
                            if (lane == 0)
                            {
                                ctr_hits += flen != 0;
                                ctr_entries += flen;
                            }
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
            __syncwarp();
        }

        // ---- pass A: correction (place.cpp:417-422), likelihood mass (:164-184), lane maxima ----
        const uint32_t nw32 = static_cast<uint32_t>(min(n_windows, static_cast<uint64_t>(0xFFFFFFFFu)));
        double mass = 0.0;
        uint32_t touched = 0;
        uint32_t lane_best = 0; // score key of my best branch, 0 = none
#pragma unroll 4
        for (uint32_t b = lane; b < N; b += 32)
        {
            const float2 a = acc[b];
            const uint32_t c = __float_as_uint(a.y);
            // the reference converts the size_t difference to float (place.cpp:420)
            const float absent = n_windows > 0xFFFFFFFFull ? static_cast<float>(n_windows - c) : static_cast<float>(nw32 - c);
            const float s = div_by_k(__fadd_rn(a.x, __fmul_rn(absent, logthr)), kf, rcp_k);
            const double e = exp10_of_float(s);
            if (c != 0)
            {
                acc[b].x = s;
                mass += e;
                ++touched;
                lane_best = max(lane_best, score_key(s));
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
        {
            mass += __shfl_xor_sync(FULL, mass, d);
            touched += __shfl_xor_sync(FULL, touched, d);
        }

        // the K-th largest lane maximum bounds the K-th best score from below
        uint32_t bar_key = 0, bar_branch = NO_BRANCH;
        {
            uint32_t mine = lane_best, kth = 0;
            for (uint32_t i = 0; i < K; ++i)
            {
                kth = __reduce_max_sync(FULL, mine);
                if (kth == 0) { break; }
                const uint32_t first = __ffs(__ballot_sync(FULL, mine == kth)) - 1u;
                if (lane == first) { mine = 0; }
            }
            bar_key = kth;
        }

        // ---- pass B: collect the branches at or above the bar, clear the accumulators ----
        __syncwarp();
        uint32_t n_cand = 0;
        uint32_t my_key = 0, my_b = NO_BRANCH, my_c = 0;
        for (uint32_t b0 = 0; b0 < N; b0 += 32)
        {
            const uint32_t b = b0 + lane;
            bool cand = false;
            uint32_t key = 0, c = 0;
            if (b < N)
            {
                const float2 a = acc[b];
                c = __float_as_uint(a.y);
                acc[b] = make_float2(0.0f, 0.0f);
                key = score_key(a.x);
                cand = c != 0 && (key > bar_key || (key == bar_key && b < bar_branch));
            }
            const uint32_t ballot = __ballot_sync(FULL, cand);
            if (ballot)
            {
                if (cand)
                {
                    const uint32_t at = n_cand + __popc(ballot & ((1u << lane) - 1u));
                    cand_key[at] = key;
                    cand_branch[at] = b;
                    cand_count[at] = c;
                }
                n_cand += __popc(ballot);
                __syncwarp();
                if (n_cand > CAND_CAP - 32)
                {
                    // keep the best K, raise the bar to the K-th of them
                    n_cand = select_best(cand_key, cand_branch, cand_count, n_cand, K, lane, my_key, my_b, my_c);
                    if (n_cand == K)
                    {
                        bar_key = __shfl_sync(FULL, my_key, K - 1);
                        bar_branch = __shfl_sync(FULL, my_b, K - 1);
                    }
                }
            }
        }
        __syncwarp();
        const uint32_t n_keep = select_best(cand_key, cand_branch, cand_count, n_cand, K, lane, my_key, my_b, my_c);
        float my_s = key_score(my_key);

        // ---- likelihood weight ratios and keep_factor (place.cpp:164-184, 241-267) ----
        const float nwf = static_cast<float>(n_windows);
        const float absent_score = __fdiv_rn(__fmul_rn(nwf, logthr), kf);
        const double absent_mass = static_cast<double>(static_cast<float>(N) - static_cast<float>(touched))
                                   * exp10_of_float(absent_score);
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
            const double power = exp10_of_float(my_s);
            ratio = power == 0.0 ? 0.0 : power / total;
        }
        const double cut = __shfl_sync(FULL, ratio, 0) * keep_factor;
        const bool keep = lane < n_out && ratio >= cut;
        const uint32_t kept = __ballot_sync(FULL, keep);
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

#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
    {
        ctr_windows += __shfl_xor_sync(FULL, ctr_windows, d);
        ctr_hits += __shfl_xor_sync(FULL, ctr_hits, d);
        ctr_entries += __shfl_xor_sync(FULL, ctr_entries, d);
    }
    if (lane == 0)
    {
        atomicAdd(p.counters + CTR_WINDOWS, static_cast<unsigned long long>(ctr_windows));
        atomicAdd(p.counters + CTR_HITS, static_cast<unsigned long long>(ctr_hits));
        atomicAdd(p.counters + CTR_ENTRIES, static_cast<unsigned long long>(ctr_entries));
    }
}
} // namespace epk
