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
constexpr int SEG_PAD = 8;             // zero descriptors after the last one (= segments in flight)
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
    return acc + seen + (SEG_CAP + SEG_PAD) * 8 + CODE_BYTES;
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

// Taylor coefficients ln(2)^i / i!, i = 12 .. 1, of 2^f (constant bank: DFMA takes them as operands)
__constant__ double EXP2_POLY[12] = {
    2.5678435993488206e-11, 4.4455382718708116e-10, 7.054911620801123e-09, 1.01780860092397e-07,
    1.321548679014431e-06,  1.5252733804059841e-05, 0.0001540353039338161, 0.0013333558146428443,
    0.009618129107628477,   0.05550410866482158,    0.24022650695910072,   0.6931471805599453 };

// 10^s in FP64 for a float exponent with |s| < 300 (sum_scores and the LWR loop call
// std::pow(10.0, double(s)), place.cpp:44-47,175,254). 10^s = 2^n * 2^f with
// n = rint(s*log2(10)), |f| <= 1/2, 2^f by its Taylor polynomial of degree 12. Relative error
// below 1e-13, dominated by the single-double log2(10) in the reduction; it is common to all
// terms of a ratio to first order. Straight-line code: callers interleave several of these.
__device__ __forceinline__ double exp10_small(float s)
{
    const double t = static_cast<double>(s) * 3.321928094887362;
    const double shifted = t + 6755399441055744.0; // 2^52 + 2^51: the integer part lands in the low word
    const int n = __double2loint(shifted);
    const double f = t - (shifted - 6755399441055744.0);
    double p = EXP2_POLY[0];
#pragma unroll
    for (int i = 1; i < 12; ++i) { p = fma(p, f, EXP2_POLY[i]); }
    p = fma(p, f, 1.0);
    return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}
__device__ __forceinline__ bool exp10_is_small(float s) { return fabsf(s) < 300.0f; }

// any exponent: near the ends of the FP64 range the library routine takes over (gradual underflow
// is part of the reference's behaviour for long queries)
__device__ __forceinline__ double exp10_of_float(float s)
{
    return exp10_is_small(s) ? exp10_small(s) : exp10(static_cast<double>(s));
}

// x / kf, correctly rounded (place.cpp:421 divides the corrected score by the k-mer size):
// q = x*r, one FMA residual step with r = RN(1/kf). Checked exhaustively against x / kf for
// every float with 1e-30 < |x| < 1e30 and kf = 1..32 (div_is_plain); the rest takes __fdiv_rn.
__device__ __forceinline__ float div_by_k_plain(float x, float kf, float rcp)
{
    const float q = __fmul_rn(x, rcp);
    return __fmaf_rn(__fmaf_rn(-q, kf, x), rcp, q);
}
__device__ __forceinline__ bool div_is_plain(float x)
{
    const float ax = fabsf(x);
    return ax > 1e-30f && ax < 1e30f;
}

// order-preserving map float -> uint32 (and back); -0 and +0 are made one value first
__device__ __forceinline__ uint32_t score_key(float s)
{
    const uint32_t u = __float_as_uint(__fadd_rn(s, 0.0f));
    return u ^ (static_cast<uint32_t>(static_cast<int32_t>(u) >> 31) | 0x80000000u);
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
//
// A segment descriptor is {index of the segment's first posting / 2, number of postings (1..64)};
// a zero count is padding. Lists start on 16-byte boundaries and segments on multiples of 64
// postings, so a lane fetches postings 2*lane and 2*lane + 1 of a segment with one 128-bit load.
// The hot loop is written in PTX: nvcc's version of it spent three quarters of its instructions
// on moves and branches.

struct seg_regs
{
    uint32_t b0, s0, b1, s1; // two postings: (branch, score bits) x 2
    uint32_t cnt;
};

__device__ __forceinline__ void load_segment(const unsigned char* lane_base, uint32_t segs_shared, uint32_t idx,
                                             uint32_t lane2, seg_regs& r)
{
#ifdef EPK_CXX_LOAD
    extern __shared__ __align__(16) unsigned char smem_base[];
    const uint32_t smem0 = static_cast<uint32_t>(__cvta_generic_to_shared(smem_base));
    const uint2 d = *reinterpret_cast<const uint2*>(smem_base + (segs_shared + idx * 8u - smem0));
    r.cnt = d.y;
    if (lane2 < d.y)
    {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(lane_base + 16ull * d.x));
        r.b0 = v.x; r.s0 = v.y; r.b1 = v.z; r.s1 = v.w;
    }
    return;
#endif
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b32 pair;\n\t"
        ".reg .b64 addr;\n\t"
        "ld.shared.v2.b32 {pair, %4}, [%5];\n\t"
        "setp.lt.u32 p, %7, %4;\n\t"
        "mad.wide.u32 addr, pair, 16, %6;\n\t"
        "@p ld.global.nc.v4.b32 {%0, %1, %2, %3}, [addr];\n\t"
        "}"
        : "+r"(r.b0), "+r"(r.s0), "+r"(r.b1), "+r"(r.s1), "=&r"(r.cnt)
        : "r"(segs_shared + idx * 8u), "l"(lane_base), "r"(lane2)
        : "memory");
}

// place.cpp:358-367 for the (up to) two postings a lane holds of one segment; branches of one
// list are distinct, so the two read-modify-writes of a lane and those of different lanes are
// independent.
__device__ __forceinline__ void apply_segment(uint32_t acc_shared, uint32_t lane2, const seg_regs& r)
{
#ifdef EPK_CXX_APPLY
    extern __shared__ __align__(16) unsigned char smem_base[];
    const uint32_t smem0 = static_cast<uint32_t>(__cvta_generic_to_shared(smem_base));
    float2* acc = reinterpret_cast<float2*>(smem_base + (acc_shared - smem0));
    const bool p0 = lane2 < r.cnt, p1 = lane2 + 1 < r.cnt;
    float2 a0 = acc[r.b0], a1 = acc[r.b1];
    a0.x = __fadd_rn(a0.x, __uint_as_float(r.s0));
    a1.x = __fadd_rn(a1.x, __uint_as_float(r.s1));
    a0.y = __uint_as_float(__float_as_uint(a0.y) + 1u);
    a1.y = __uint_as_float(__float_as_uint(a1.y) + 1u);
    if (p0) { acc[r.b0] = a0; }
    if (p1) { acc[r.b1] = a1; }
    __syncwarp();
    return;
#endif
    asm volatile(
        "{\n\t"
        ".reg .pred p0, p1;\n\t"
        ".reg .b32 a0, a1, c0, c1, x0, x1;\n\t"
        ".reg .f32 f0, f1;\n\t"
        "setp.lt.u32 p0, %5, %4;\n\t"
        "setp.lt.u32 p1, %6, %4;\n\t"
        "shl.b32 a0, %0, 3;\n\t"
        "shl.b32 a1, %2, 3;\n\t"
        "add.u32 a0, a0, %7;\n\t"
        "add.u32 a1, a1, %7;\n\t"
        "ld.shared.v2.b32 {x0, c0}, [a0];\n\t" // lanes past the end hold 0 or an older, valid branch
        "ld.shared.v2.b32 {x1, c1}, [a1];\n\t"
        "mov.b32 f0, x0;\n\t"
        "mov.b32 f1, x1;\n\t"
        "add.rn.f32 f0, f0, %1;\n\t"
        "add.rn.f32 f1, f1, %3;\n\t"
        "add.u32 c0, c0, 1;\n\t"
        "add.u32 c1, c1, 1;\n\t"
        "mov.b32 x0, f0;\n\t"
        "mov.b32 x1, f1;\n\t"
        "@p0 st.shared.v2.b32 [a0], {x0, c0};\n\t"
        "@p1 st.shared.v2.b32 [a1], {x1, c1};\n\t"
        "}"
        :
        : "r"(r.b0), "f"(__uint_as_float(r.s0)), "r"(r.b1), "f"(__uint_as_float(r.s1)), "r"(r.cnt), "r"(lane2),
          "r"(lane2 + 1u), "r"(acc_shared)
        : "memory");
    __syncwarp(); // the next segment may name the same branches from other lanes
}

// Applies segs[0, n) in order; U segments are in flight while the previous U are applied.
// segs[n, n + U) must hold zero descriptors.
#ifndef EPK_GATHER_INLINE
#define EPK_GATHER_INLINE __noinline__
#endif
template<int U>
__device__ EPK_GATHER_INLINE void gather_segments(const uint2* __restrict__ postings, uint32_t acc_shared, uint32_t segs_shared,
                                             uint32_t n, uint32_t lane)
{
    const unsigned char* lane_base = reinterpret_cast<const unsigned char*>(postings) + 16u * lane;
    const uint32_t lane2 = 2u * lane;
    seg_regs a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
    {
        a[u] = seg_regs{ 0, 0, 0, 0, 0 };
        b[u] = seg_regs{ 0, 0, 0, 0, 0 };
    }
#pragma unroll
    for (int u = 0; u < U; ++u) { load_segment(lane_base, segs_shared, u, lane2, a[u]); }
    for (uint32_t first = 0; first < n; first += 2 * U)
    {
        const bool more_b = first + U < n;
        if (more_b)
        {
#pragma unroll
            for (int u = 0; u < U; ++u) { load_segment(lane_base, segs_shared, first + U + u, lane2, b[u]); }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) { apply_segment(acc_shared, lane2, a[u]); }
        if (!more_b) { break; }
        if (first + 2 * U < n)
        {
#pragma unroll
            for (int u = 0; u < U; ++u) { load_segment(lane_base, segs_shared, first + 2 * U + u, lane2, a[u]); }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) { apply_segment(acc_shared, lane2, b[u]); }
    }
}

// One long list on its own (only when a single chunk's lists overflow the segment buffer).
__device__ __noinline__ void apply_list(const uint2* __restrict__ postings, uint32_t acc_shared, uint64_t off, uint32_t len,
                                        uint32_t lane)
{
    const uint32_t lane2 = 2u * lane;
    for (uint32_t pos = 0; pos < len; pos += SEG_LEN)
    {
        seg_regs r{ 0, 0, 0, 0, min(len - pos, static_cast<uint32_t>(SEG_LEN)) };
        if (lane2 < r.cnt)
        {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(postings + off + pos) + lane);
            r.b0 = v.x; r.s0 = v.y; r.b1 = v.z; r.s1 = v.w;
        }
        apply_segment(acc_shared, lane2, r);
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
    static_assert(U <= SEG_PAD, "zero padding of the segment buffer covers one batch of loads");

    uint32_t* letters = reinterpret_cast<uint32_t*>(smem);
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) { letters[i] = p.letters[i]; }

    unsigned char* base = smem + CTA_SMEM_FIXED + warp * warp_smem_bytes(N);
    const size_t acc_bytes = (static_cast<size_t>(N) * 8 + 15) & ~static_cast<size_t>(15);
    const size_t seen_words = (static_cast<size_t>(N) + 31) / 32;
    const size_t seen_bytes = (seen_words * 4 + 15) & ~static_cast<size_t>(15);
    float2* acc = reinterpret_cast<float2*>(base);
    uint32_t* seen = reinterpret_cast<uint32_t*>(base + acc_bytes);
    uint2* segs = reinterpret_cast<uint2*>(base + acc_bytes + seen_bytes);
    const uint32_t segs_shared = static_cast<uint32_t>(__cvta_generic_to_shared(segs));
    const uint32_t acc_shared = static_cast<uint32_t>(__cvta_generic_to_shared(acc));
    uint32_t* cand_key = reinterpret_cast<uint32_t*>(segs); // the segment buffer is idle during the sweeps
    uint32_t* cand_branch = cand_key + CAND_CAP;
    uint32_t* cand_count = cand_branch + CAND_CAP;
    uint8_t* codes = reinterpret_cast<uint8_t*>(base + acc_bytes + seen_bytes + (SEG_CAP + SEG_PAD) * 8);

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
                if (lane < SEG_PAD) { segs[nseg + lane] = make_uint2(0, 0); }
                __syncwarp();
                gather_segments<U>(db.postings, acc_shared, segs_shared, nseg, lane);
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
                    const uint32_t pair = static_cast<uint32_t>(slot[r] >> (LEN_BITS + 1));
                    for (uint32_t pos = 0; pos < l; pos += SEG_LEN)
                    {
                        segs[at++] = make_uint2(pair + pos / 2, min(l - pos, static_cast<uint32_t>(SEG_LEN)));
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
                        if (l) { apply_list(db.postings, acc_shared, s >> LEN_BITS, l, lane); }
                    }
                }
            }
        }
        if (nseg)
        {
            if (lane < SEG_PAD) { segs[nseg + lane] = make_uint2(0, 0); }
            __syncwarp();
            gather_segments<U>(db.postings, acc_shared, segs_shared, nseg, lane);
        }
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
        // Four branches per lane and iteration, straight-line, so that the four FP64 chains overlap.
        // The corrected score replaces S as its order-preserving key; C stays.
        const uint32_t nw32 = static_cast<uint32_t>(n_windows);
        double mass = 0.0;
        uint32_t touched = 0;
        uint32_t lane_best = 0; // score key of my best branch, 0 = none
        for (uint32_t base0 = 0; base0 < N; base0 += 128) // warp-uniform trip count
        {
            const uint32_t b0 = base0 + lane;
            float2 a[4];
            float x[4], s[4];
            double e[4];
            bool odd = false;
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                const uint32_t b = b0 + 32u * u;
                a[u] = b < N ? acc[b] : make_float2(0.0f, 0.0f);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                const uint32_t c = __float_as_uint(a[u].y);
                // the reference converts the size_t difference to float (place.cpp:420)
                x[u] = __fadd_rn(a[u].x, __fmul_rn(static_cast<float>(nw32 - c), logthr));
                s[u] = div_by_k_plain(x[u], kf, rcp_k);
                e[u] = exp10_small(s[u]);
                odd |= c != 0 && !(div_is_plain(x[u]) && exp10_is_small(s[u]));
            }
            if (odd)
            {
                // outside what the short forms were checked for (very long queries): the IEEE routines
#pragma unroll
                for (int u = 0; u < 4; ++u)
                {
                    s[u] = __fdiv_rn(x[u], kf);
                    e[u] = exp10(static_cast<double>(s[u]));
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                const uint32_t b = b0 + 32u * u;
                if (__float_as_uint(a[u].y) != 0)
                {
                    const uint32_t key = score_key(s[u]);
                    acc[b].x = __uint_as_float(key);
                    mass += e[u];
                    ++touched;
                    lane_best = max(lane_best, key);
                }
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
        for (uint32_t base0 = 0; base0 < N; base0 += 128) // warp-uniform trip count
        {
            const uint32_t b0 = base0 + lane;
            uint32_t key[4], cnt[4];
            bool cand[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                const uint32_t b = b0 + 32u * u;
                float2 a = make_float2(0.0f, 0.0f);
                if (b < N)
                {
                    a = acc[b];
                    acc[b] = make_float2(0.0f, 0.0f);
                }
                key[u] = __float_as_uint(a.x);
                cnt[u] = __float_as_uint(a.y);
                cand[u] = (cnt[u] != 0) & ((key[u] > bar_key) | ((key[u] == bar_key) & (b < bar_branch)));
            }
            if (__any_sync(FULL, cand[0] | cand[1] | cand[2] | cand[3]))
            {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                {
                    const uint32_t ballot = __ballot_sync(FULL, cand[u]);
                    if (cand[u])
                    {
                        const uint32_t at = n_cand + __popc(ballot & ((1u << lane) - 1u));
                        cand_key[at] = key[u];
                        cand_branch[at] = b0 + 32u * u;
                        cand_count[at] = cnt[u];
                    }
                    n_cand += __popc(ballot);
                    __syncwarp();
                    if (n_cand > CAND_CAP - 32)
                    {
                        // keep the best K, raise the bar to the K-th of them. Later branches of this
                        // group were tested against the old bar: a few extra candidates, never a miss.
                        n_cand = select_best(cand_key, cand_branch, cand_count, n_cand, K, lane, my_key, my_b, my_c);
                        if (n_cand == K)
                        {
                            bar_key = __shfl_sync(FULL, my_key, K - 1);
                            bar_branch = __shfl_sync(FULL, my_b, K - 1);
                        }
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
