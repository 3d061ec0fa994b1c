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
//      flight while the previous U are being applied to the per-branch accumulators in shared
//      memory: float S[N] and a count C[N] whose width follows the number of windows of the
//      query (uint8 up to 255 windows, uint16 up to 65535, uint32 beyond; one kernel
//      instantiation per width, so that 150-bp reads pay 5 bytes per branch and 8 warps fit one
//      SM at N = 5000). Segments are applied in window order and a list never names a branch
//      twice, so every S[b] sees exactly the additions of place.cpp:358-367 in the same order:
//      the sums are bit-identical to the reference's, without atomics;
//   3. (only for reads that have windows with one ambiguous letter) replays place.cpp:373-415;
//   4. sweeps the accumulators ONCE, four consecutive branches per lane and step: omega-threshold
//      correction (place.cpp:417-422), FP64 10^S summed over the touched branches (sum_scores),
//      collection of the top-k candidates against a bar that rises as candidates accumulate, and
//      clearing of the accumulators for the next query;
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
constexpr int CAND_CAP = 128;          // top-k candidates buffered per warp (aliases the segments)
constexpr int MAX_WARPS_PER_CTA = 8;
constexpr int MAX_OWNERS = 8;          // ranks of a database-sharded group (one NVSwitch domain)
constexpr uint32_t NO_BRANCH = 0xFFFFFFFFu;
constexpr uint32_t FULL = 0xFFFFFFFFu;

static_assert(CAND_CAP * 12 <= SEG_CAP * 8, "candidate arrays alias the segment buffer");

// One kernel instantiation per width of the hit counts: class c handles the queries with
// CLASS_MIN_WINDOWS[c] <= windows <= CLASS_MAX_WINDOWS[c].
constexpr int N_CLASSES = 3;
__host__ __device__ constexpr uint64_t class_max_windows(int c) { return c == 0 ? 255ull : c == 1 ? 65535ull : 0xFFFFFFFFull; }
__host__ __device__ constexpr uint32_t class_count_bytes(int c) { return c == 0 ? 1u : c == 1 ? 2u : 4u; }
__host__ __device__ inline int class_of_windows(uint64_t n_windows)
{
    return n_windows <= class_max_windows(0) ? 0 : n_windows <= class_max_windows(1) ? 1 : 2;
}

enum counter_slot { CTR_CURSOR = 0, CTR_WINDOWS = 1, CTR_HITS = 2, CTR_ENTRIES = 3, CTR_SHORT = 4, CTR_TOO_LONG = 5, CTR_UNSUPPORTED = 6, CTR_COUNT = 8 };

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
    unsigned long long* counters; // [CTR_COUNT] of this launch
    unsigned long long* totals;   // [CTR_COUNT] running totals of the placer, bumped with the counters (NULL: none)
    const uint32_t* class_reads; // [N_CLASSES] queries per class (classify_kernel); NULL = unknown
    uint32_t cls;                // class this launch handles
    // ---- database-sharded mode (partial rows out / partial rows in), see "partial rows" below ----
    unsigned char* rows_out[MAX_OWNERS]; // partial launch: where the rows of owner g's queries go
    uint32_t owner_first[MAX_OWNERS + 1]; // queries [owner_first[g], owner_first[g + 1]) are finalised by owner g
    uint32_t n_owners;
    const unsigned char* rows_in;        // finalize launch: n_shards slabs of n rows each
    uint64_t shard_stride;               // bytes from a row of shard g to the same row of shard g + 1
    uint32_t n_shards;
    uint32_t row_format;                 // partial rows: ROW_DENSE or ROW_COMPACT (both launches of a step agree on it)
    uint32_t team;                       // warps per accumulator set (TEAM launches)
    uint32_t prefetch;                   // L2 prefetch of a descriptor buffer's postings before it is gathered (PREFETCH_*)
    uint32_t* trace;                     // [4 n] per query: keys looked up, keys found, postings, checksum of the k-mer stream (NULL: none)
    // ---- accumulators in global memory (trees / reads too large for shared memory), see "accumulators in HBM" ----
    unsigned char* acc_scratch;          // per warp of the grid: float S[np] | uint32 C[np] | seen bitmap np / 8 bytes, all zero
    // ---- finalize launch: ambiguous windows of database-sharded mode, replayed by the owner ----
    const device_db* shard_views;        // [n_shards] the shards' tables and postings as seen from this device (NULL: none)
    uint32_t* seen_scratch;              // np / 32 words per warp of the finalize grid, zero between queries
    unsigned char* amb_scratch;          // float S[np] | uint16 C[np] per warp of the finalize grid, zero between queries
};


// Number of queries of each class (sequences shorter than k - 1 letters count for class 0, which
// reports them). Classes in unsupported_mask have no launch: their rows get n_place = 0 here.
__global__ void classify_kernel(const uint64_t* __restrict__ seq_off, uint32_t n, uint32_t k, uint32_t* class_reads,
                                uint32_t unsupported_mask, uint32_t* n_place, unsigned long long* totals)
{
    uint32_t mine[N_CLASSES] = { 0, 0, 0 };
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        const uint64_t len = seq_off[i + 1] - seq_off[i];
        const int c = len + 1 < k ? 0 : class_of_windows(len + 1 - k);
        if (n_place && ((unsupported_mask >> c) & 1u)) { n_place[i] = 0; } // no launch will write this row
        mine[0] += c == 0;
        mine[1] += c == 1;
        mine[2] += c == 2;
    }
#pragma unroll
    for (int c = 0; c < N_CLASSES; ++c)
    {
        const uint32_t total = __reduce_add_sync(FULL, mine[c]);
        if ((threadIdx.x & 31u) == 0 && total)
        {
            atomicAdd(class_reads + c, total);
            if (totals && ((unsupported_mask >> c) & 1u)) { atomicAdd(totals + CTR_UNSUPPORTED, static_cast<unsigned long long>(total)); }
        }
    }
}

// A work counter of the launch and, with it, the running total of the placer. (The totals used to be folded by a
// one-thread launch after the placement launch; with every SM's register file taken by the next placement launch
// of another stream, that little launch - and the copies out queued behind it - waited for a whole launch.)
__device__ __forceinline__ void bump(const place_params& p, int slot, unsigned long long v)
{
    atomicAdd(p.counters + slot, v);
    if (p.totals) { atomicAdd(p.totals + slot, v); }
}

// One term of the per-query checksum of the k-mer stream (epk_placer_set_trace): position of the window, code,
// number of keys of the window (1 = exact k-mer), length of the posting list found (0 = absent). The terms are
// summed modulo 2^32, so the order in which lanes meet the windows does not matter. Mirrored by tests/helpers.py.
__host__ __device__ inline uint32_t trace_term(uint32_t pos, uint32_t key, uint32_t group, uint32_t len)
{
    uint32_t h = pos * 0x9E3779B1u ^ key * 0x85EBCA6Bu ^ group * 0xC2B2AE35u ^ len * 0x27D4EB2Fu;
    h ^= h >> 15; h *= 0x2C1B3C6Du;
    h ^= h >> 12; h *= 0x297A2D39u;
    return h ^ (h >> 15);
}

// ---- phase clocks (kernel experiments only: -DEPK_PHASE_CLOCKS) ------------------------------------------
// Cycles the leader warps spent per phase of a query, summed over the launch: 0 claim + offsets, 1 letters + encode +
// probe (until the slots are in), 2 descriptors, 3 gather + accumulate, 4 ambiguous windows, 5 sweep, 6 weight ratios
// + rows out, 7 queries; inside the two-pass sweep: 5 first pass, 8 bar, 9 second pass, 10 final selection. Read and
// cleared by epk_debug_phase_clocks.
#ifndef EPK_ABLATE
#define EPK_ABLATE 0 // kernel experiments: 1 = no likelihood mass, 2 = no candidates, 4 = postings loaded but not accumulated
#endif
#ifdef EPK_PHASE_CLOCKS
__device__ unsigned long long g_phase_clocks[12];
#define EPK_PHASE_DECL uint32_t ph_acc[12] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 }; uint32_t ph_t = clock();
#define EPK_PHASE_PARAMS , uint32_t (&ph_acc)[12], uint32_t& ph_t
#define EPK_PHASE_ARGS , ph_acc, ph_t
#define EPK_PHASE_DUMMY uint32_t ph_acc[12] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 }; uint32_t ph_t = 0;
#define EPK_PHASE(i) { const uint32_t ph_now = clock(); ph_acc[i] += ph_now - ph_t; ph_t = ph_now; }
#define EPK_PHASE_COUNT(i) { ++ph_acc[i]; }
#define EPK_PHASE_FLUSH if (lane == 0) { for (int i = 0; i < 12; ++i) { atomicAdd(g_phase_clocks + i, static_cast<unsigned long long>(ph_acc[i])); } }
#else
#define EPK_PHASE_DECL
#define EPK_PHASE_PARAMS
#define EPK_PHASE_ARGS
#define EPK_PHASE_DUMMY
#define EPK_PHASE(i)
#define EPK_PHASE_COUNT(i)
#define EPK_PHASE_FLUSH
#endif

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

// Coefficients c8 .. c1 of a degree-8 near-minimax polynomial of 2^f on |f| <= 1/2 (c0 = 1);
// relative error 2.2e-12 (tools/exp2_poly.py). Constant bank: DFMA takes them as operands.
__constant__ double EXP2_POLY[8] = {
    1.325522879405974e-06, 1.5303702622244995e-05, 0.0001540345583542578, 0.001333347846654947,
    0.009618129159424775,  0.05550410906332551,    0.24022650695812914,   0.6931471805568318 };

// 10^s in FP64 for a float exponent with |s| < 300 (sum_scores and the LWR loop call
// std::pow(10.0, double(s)), place.cpp:44-47,175,254). 10^s = 2^n * 2^f with
// n = rint(s*log2(10)), |f| <= 1/2, 2^f by the polynomial above. Relative error below 3e-12:
// four orders of magnitude under what the float score itself resolves (an ulp of a score near
// -60 moves 10^s by 9e-6) and under the 1e-9 the parity tests ask of the weight ratios.
// Straight-line code: callers interleave several of these.
__device__ __forceinline__ double exp10_small(float s)
{
    const double t = static_cast<double>(s) * 3.321928094887362;
    const double shifted = t + 6755399441055744.0; // 2^52 + 2^51: the integer part lands in the low word
    const int n = __double2loint(shifted);
    const double f = t - (shifted - 6755399441055744.0);
    double p = EXP2_POLY[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) { p = fma(p, f, EXP2_POLY[i]); }
    p = fma(p, f, 1.0);
    return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}
constexpr float EXP10_SMALL_LIMIT = 290.0f;

// any exponent: near the ends of the FP64 range the library routine takes over (gradual underflow
// is part of the reference's behaviour for long queries)
__device__ __forceinline__ double exp10_of_float(float s)
{
    return fabsf(s) < EXP10_SMALL_LIMIT ? exp10_small(s) : exp10(static_cast<double>(s));
}

// x / kf, correctly rounded (place.cpp:421 divides the corrected score by the k-mer size):
// q = x*r, one FMA residual step with r = RN(1/kf). Checked exhaustively against x / kf for
// every float with 1e-30 < |x| < 1e30 and kf = 1..32 (tools/check_div.cu); +0 stays +0. Queries
// whose scores could leave that range take the sweep with __fdiv_rn (see fast_sweep_ok).
__device__ __forceinline__ float div_by_k_plain(float x, float kf, float rcp)
{
    const float q = __fmul_rn(x, rcp);
    return __fmaf_rn(__fmaf_rn(-q, kf, x), rcp, q);
}

// order-preserving map float -> uint32 (and back); -0 and +0 are made one value first
__device__ __forceinline__ uint32_t score_key(float s)
{
    const uint32_t u = __float_as_uint(__fadd_rn(s, 0.0f));
    return u ^ (static_cast<uint32_t>(static_cast<int32_t>(u) >> 31) | 0x80000000u);
}
// the same without folding -0 into +0 (for scores known not to be -0)
__device__ __forceinline__ uint32_t score_key_raw(float s)
{
    const uint32_t u = __float_as_uint(s);
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
    uint32_t cnt;            // postings of the segment, counted from the (even) posting the descriptor names
    uint32_t start;          // 1: the first of them belongs to the neighbouring branch range and is skipped
};

template<bool RANGE>
__device__ __forceinline__ void load_segment(const unsigned char* lane_base, uint32_t segs_shared, uint32_t idx,
                                             uint32_t lane2, seg_regs& r)
{
    if (RANGE)
    {
        // descriptors of a range team: bit 31 of the count word = skip the first posting
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            ".reg .b32 pair, y;\n\t"
            ".reg .b64 addr;\n\t"
            "ld.shared.v2.b32 {pair, y}, [%6];\n\t"
            "and.b32 %4, y, 0x7FFFFFFF;\n\t"
            "shr.u32 %5, y, 31;\n\t"
            "setp.lt.u32 p, %8, %4;\n\t"
            "mad.wide.u32 addr, pair, 16, %7;\n\t"
            "@p ld.global.nc.v4.b32 {%0, %1, %2, %3}, [addr];\n\t"
            "}"
            : "+r"(r.b0), "+r"(r.s0), "+r"(r.b1), "+r"(r.s1), "=&r"(r.cnt), "=&r"(r.start)
            : "r"(segs_shared + idx * 8u), "l"(lane_base), "r"(lane2)
            : "memory");
        return;
    }
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

// The accumulators of one warp: float S[np] and CT C[np] (np = padded_nodes(N)), as shared-window
// addresses for the PTX below.
struct acc_ref
{
    uint32_t s_shared; // S
    uint32_t c_shared; // C
};

// place.cpp:358-367 for the (up to) two postings a lane holds of one segment; branches of one
// list are distinct, so the two read-modify-writes of a lane and those of different lanes are
// independent. Lanes past the end of the segment hold 0 or an older, valid branch: they read it
// and store nothing.
#ifndef EPK_APPLY_SYNC
#define EPK_APPLY_SYNC __syncwarp();
#endif
#define EPK_APPLY_SEGMENT(NAME, CT, LD, ST, P0, EXTRA)                                                     \
    __device__ __forceinline__ void NAME(const acc_ref& acc, uint32_t lane2, const seg_regs& r, CT)          \
    {                                                                                                      \
        asm volatile(                                                                                      \
            "{\n\t"                                                                                        \
            ".reg .pred p0, p1;\n\t"                                                                       \
            ".reg .b32 a0, a1, d0, d1, c0, c1;\n\t"                                                        \
            ".reg .f32 f0, f1;\n\t"                                                                        \
            P0                                                                                             \
            "setp.lt.u32 p1, %6, %4;\n\t"                                                                  \
            "mad.lo.u32 a0, %0, 4, %7;\n\t"                                                                \
            "mad.lo.u32 a1, %2, 4, %7;\n\t"                                                                \
            "mad.lo.u32 d0, %0, %9, %8;\n\t"                                                               \
            "mad.lo.u32 d1, %2, %9, %8;\n\t"                                                               \
            "ld.shared.f32 f0, [a0];\n\t"                                                                  \
            "ld.shared.f32 f1, [a1];\n\t"                                                                  \
            "ld.shared." LD " c0, [d0];\n\t"                                                               \
            "ld.shared." LD " c1, [d1];\n\t"                                                               \
            "add.rn.f32 f0, f0, %1;\n\t"                                                                   \
            "add.rn.f32 f1, f1, %3;\n\t"                                                                   \
            "add.u32 c0, c0, 1;\n\t"                                                                       \
            "add.u32 c1, c1, 1;\n\t"                                                                       \
            "@p0 st.shared.f32 [a0], f0;\n\t"                                                              \
            "@p1 st.shared.f32 [a1], f1;\n\t"                                                              \
            "@p0 st.shared." ST " [d0], c0;\n\t"                                                           \
            "@p1 st.shared." ST " [d1], c1;\n\t"                                                           \
            "}"                                                                                            \
            :                                                                                              \
            : "r"(r.b0), "f"(__uint_as_float(r.s0)), "r"(r.b1), "f"(__uint_as_float(r.s1)), "r"(r.cnt),    \
              "r"(lane2), "r"(lane2 + 1u), "r"(acc.s_shared), "r"(acc.c_shared),                           \
              "n"(static_cast<int>(sizeof(CT))) EXTRA                                                      \
            : "memory");                                                                                   \
        EPK_APPLY_SYNC /* the next segment may name the same branches from other lanes */                   \
    }
#define EPK_P0_PLAIN "setp.lt.u32 p0, %5, %4;\n\t"
#define EPK_P0_RANGE "setp.ge.u32 p0, %5, %10;\n\tsetp.lt.and.u32 p0, %5, %4, p0;\n\t" /* lane2 in [start, cnt) */
#define EPK_NO_EXTRA
#define EPK_START_EXTRA , "r"(r.start)
EPK_APPLY_SEGMENT(apply_segment, uint8_t, "u8", "u8", EPK_P0_PLAIN, EPK_NO_EXTRA)
EPK_APPLY_SEGMENT(apply_segment, uint16_t, "u16", "u16", EPK_P0_PLAIN, EPK_NO_EXTRA)
EPK_APPLY_SEGMENT(apply_segment, uint32_t, "u32", "u32", EPK_P0_PLAIN, EPK_NO_EXTRA)
EPK_APPLY_SEGMENT(apply_segment_range, uint8_t, "u8", "u8", EPK_P0_RANGE, EPK_START_EXTRA)
EPK_APPLY_SEGMENT(apply_segment_range, uint16_t, "u16", "u16", EPK_P0_RANGE, EPK_START_EXTRA)
EPK_APPLY_SEGMENT(apply_segment_range, uint32_t, "u32", "u32", EPK_P0_RANGE, EPK_START_EXTRA)
#undef EPK_APPLY_SEGMENT

// ---- L2 prefetch of the postings a descriptor buffer names ------------------------------------------
//
// gather_segments keeps U segments (U * 512 bytes) of a warp in flight; with 8 warps per SM that is far too
// little to cover the latency of HBM. Before a buffer of descriptors is gathered, the warp therefore asks L2
// for every 128-byte line (PREFETCH_LINES) or 32-byte sector (PREFETCH_SECTORS) of all its segments at once -
// prefetch.global.L2 takes no register, no shared memory and no scoreboard, the two resources this kernel
// has none left of - and the in-order gather that follows finds its postings in L2.
enum prefetch_mode { PREFETCH_NONE = 0, PREFETCH_LINES = 1, PREFETCH_SECTORS = 2 };

template<uint32_t GRANULE, uint32_t LANES_PER_SEG>
__device__ __forceinline__ void prefetch_segments_as(const unsigned char* postings, uint32_t segs_shared, uint32_t n, uint32_t lane)
{
    constexpr uint32_t PER = 32u / LANES_PER_SEG; // segments per instruction
    static_assert(PER <= SEG_PAD, "descriptors read past the end are the zero padding");
    const uint32_t sub = lane % LANES_PER_SEG, which = lane / LANES_PER_SEG;
    for (uint32_t s0 = 0; s0 < n; s0 += PER)
    {
        uint32_t pair, cnt;
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(pair), "=r"(cnt) : "r"(segs_shared + (s0 + which) * 8u));
        cnt &= 0x7FFFFFFFu; // (bit 31: skip flag of a range team's descriptor)
        const uint64_t base = static_cast<uint64_t>(pair) * 16u;
        const uint32_t head = static_cast<uint32_t>(base) & (GRANULE - 1u); // the postings array itself is 256-byte aligned
        const uint32_t granules = (head + cnt * 8u + GRANULE - 1u) / GRANULE;
        if (sub < granules) { asm volatile("prefetch.global.L2 [%0];" ::"l"(postings + (base - head) + sub * GRANULE)); }
    }
}
__device__ __forceinline__ void prefetch_segments(uint32_t mode, const uint2* postings, uint32_t segs_shared, uint32_t n, uint32_t lane)
{
    const unsigned char* base = reinterpret_cast<const unsigned char*>(postings);
    if (mode == PREFETCH_LINES) { prefetch_segments_as<128, 8>(base, segs_shared, n, lane); }       // 512 unaligned bytes: up to 5 lines
    else if (mode == PREFETCH_SECTORS) { prefetch_segments_as<32, 32>(base, segs_shared, n, lane); } // up to 17 sectors
}

// Applies segs[0, n) in order; U segments are in flight while the previous U are applied.
// segs[n, n + U) must hold zero descriptors.
#ifndef EPK_GATHER_INLINE
#define EPK_GATHER_INLINE __noinline__
#endif
// load-only stand-in for apply_segment (MODE_GATHER_ONLY): the postings are consumed, nothing is stored
__device__ __forceinline__ void sink_segment(uint32_t& sink, uint32_t lane2, const seg_regs& r)
{
    if (lane2 < r.cnt) { sink ^= r.b0 ^ r.s0 ^ r.b1 ^ r.s1; }
    (void)r.start;
}

template<int U, class CT, bool SINK = false, bool RANGE = false>
__device__ EPK_GATHER_INLINE void gather_segments(const uint2* __restrict__ postings, acc_ref acc, uint32_t segs_shared,
                                             uint32_t n, uint32_t lane, uint32_t pf, uint32_t* sink_out = nullptr)
{
    uint32_t sink = 0;
    if (pf) { prefetch_segments(pf, postings, segs_shared, n, lane); }
    const unsigned char* lane_base = reinterpret_cast<const unsigned char*>(postings) + 16u * lane;
    const uint32_t lane2 = 2u * lane;
    seg_regs a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
    {
        a[u] = seg_regs{ 0, 0, 0, 0, 0, 0 };
        b[u] = seg_regs{ 0, 0, 0, 0, 0, 0 };
    }
#pragma unroll
    for (int u = 0; u < U; ++u) { load_segment<RANGE>(lane_base, segs_shared, u, lane2, a[u]); }
    for (uint32_t first = 0; first < n; first += 2 * U)
    {
        const bool more_b = first + U < n;
        if (more_b)
        {
#pragma unroll
            for (int u = 0; u < U; ++u) { load_segment<RANGE>(lane_base, segs_shared, first + U + u, lane2, b[u]); }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            if (SINK || (EPK_ABLATE & 4)) { sink_segment(sink, lane2, a[u]); }
            else if (RANGE) { apply_segment_range(acc, lane2, a[u], CT{}); }
            else { apply_segment(acc, lane2, a[u], CT{}); }
        }
        if (!more_b) { break; }
        if (first + 2 * U < n)
        {
#pragma unroll
            for (int u = 0; u < U; ++u) { load_segment<RANGE>(lane_base, segs_shared, first + 2 * U + u, lane2, a[u]); }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            if (SINK || (EPK_ABLATE & 4)) { sink_segment(sink, lane2, b[u]); }
            else if (RANGE) { apply_segment_range(acc, lane2, b[u], CT{}); }
            else { apply_segment(acc, lane2, b[u], CT{}); }
        }
    }
    if (SINK || (EPK_ABLATE & 4)) { if (sink_out) { *sink_out ^= sink; } }
}

// One long list on its own (only when a single chunk's lists overflow the segment buffer).
template<class CT>
__device__ __noinline__ void apply_list(const uint2* __restrict__ postings, acc_ref acc, uint64_t off, uint32_t len,
                                        uint32_t lane)
{
    const uint32_t lane2 = 2u * lane;
    for (uint32_t pos = 0; pos < len; pos += SEG_LEN)
    {
        seg_regs r{ 0, 0, 0, 0, min(len - pos, static_cast<uint32_t>(SEG_LEN)), 0 };
        if (lane2 < r.cnt)
        {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(postings + off + pos) + lane);
            r.b0 = v.x; r.s0 = v.y; r.b1 = v.z; r.s1 = v.w;
        }
        apply_segment(acc, lane2, r, CT{});
    }
}

// ---- accumulators in HBM ------------------------------------------------------------------------------
//
// The shared-memory accumulators bound the tree: 5.125 bytes per node for 150-bp reads, 8.125 for reads of more than
// 65535 windows - 44 k / 28 k nodes at most per SM. The reference has no such limit (its scratch is max_threads x
// node_count in host memory, place.cpp:92-96). Beyond it the same kernel runs with every warp's S / C / seen bitmap
// in a row of its own in global memory (L2-resident: a few hundred KB per warp): the same additions in the same
// order - results stay bit-identical to the reference's - through plain loads and stores, a __syncwarp() between
// segments as in shared memory. A fallback for shapes outside the design point, several times slower per posting.
__host__ __device__ inline size_t global_acc_bytes(uint32_t n_nodes)
{
    const size_t np = padded_nodes(n_nodes);
    return np * 8 + np / 8;
}
// what a warp of this form keeps in shared memory: segment descriptors and letter codes
constexpr size_t GLOBAL_WARP_SMEM = (SEG_CAP + SEG_PAD) * 8 + CODE_BYTES;

__device__ __forceinline__ void apply_segment_global(float* S, uint32_t* C, uint32_t lane2, const seg_regs& r)
{
    if (lane2 < r.cnt)
    {
        __stcg(S + r.b0, __fadd_rn(__ldcg(S + r.b0), __uint_as_float(r.s0)));
        __stcg(C + r.b0, __ldcg(C + r.b0) + 1u);
    }
    if (lane2 + 1u < r.cnt)
    {
        __stcg(S + r.b1, __fadd_rn(__ldcg(S + r.b1), __uint_as_float(r.s1)));
        __stcg(C + r.b1, __ldcg(C + r.b1) + 1u);
    }
    __syncwarp(); // the next segment may name the same branches from other lanes
}

template<int U>
__device__ __noinline__ void gather_segments_global(const uint2* __restrict__ postings, float* S, uint32_t* C, uint32_t segs_shared,
                                                    uint32_t n, uint32_t lane)
{
    const unsigned char* lane_base = reinterpret_cast<const unsigned char*>(postings) + 16u * lane;
    const uint32_t lane2 = 2u * lane;
    for (uint32_t first = 0; first < n; first += U)
    {
        seg_regs a[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            a[u] = seg_regs{ 0, 0, 0, 0, 0, 0 };
            load_segment<false>(lane_base, segs_shared, first + u, lane2, a[u]); // (descriptors past n are the zero padding)
        }
#pragma unroll
        for (int u = 0; u < U; ++u) { apply_segment_global(S, C, lane2, a[u]); }
    }
}

__device__ __noinline__ void apply_list_global(const uint2* __restrict__ postings, float* S, uint32_t* C, uint64_t off, uint32_t len,
                                               uint32_t lane)
{
    const uint32_t lane2 = 2u * lane;
    for (uint32_t pos = 0; pos < len; pos += SEG_LEN)
    {
        seg_regs r{ 0, 0, 0, 0, min(len - pos, static_cast<uint32_t>(SEG_LEN)), 0 };
        if (lane2 < r.cnt)
        {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(postings + off + pos) + lane);
            r.b0 = v.x; r.s0 = v.y; r.b1 = v.z; r.s1 = v.w;
        }
        apply_segment_global(S, C, lane2, r);
    }
}

// the sweep's view of such a row: four consecutive branches are taken (and cleared) at a time
struct global_acc
{
    float* S;
    uint32_t* C;
    __device__ __forceinline__ bool any(uint32_t j) const
    {
        const uint4 w = __ldcg(reinterpret_cast<const uint4*>(C + j));
        return (w.x | w.y | w.z | w.w) != 0u;
    }
    __device__ __forceinline__ float4 take(uint32_t j, uint32_t (&c)[4]) const
    {
        const float4 s4 = __ldcg(reinterpret_cast<const float4*>(S + j));
        const uint4 w = __ldcg(reinterpret_cast<const uint4*>(C + j));
        __stcg(reinterpret_cast<float4*>(S + j), make_float4(0.0f, 0.0f, 0.0f, 0.0f));
        __stcg(reinterpret_cast<uint4*>(C + j), make_uint4(0u, 0u, 0u, 0u));
        c[0] = w.x; c[1] = w.y; c[2] = w.z; c[3] = w.w;
        return s4;
    }
};

// ---- top-k selection --------------------------------------------------------------------------

// Candidates (score key, branch, count) live in three arrays of CAND_CAP words. Selects the best
// `want` of the m buffered ones (higher score first, place.cpp:128-131; exact ties: lower branch
// id first — the documented tie rule), leaves them sorted in slots [0, n_sel) and in the registers
// of lanes [0, n_sel): lane i holds the i-th best. Returns n_sel = min(want, m).
template<int PER_LANE>
__device__ __forceinline__ uint32_t select_best_n(uint32_t* cand_key, uint32_t* cand_branch, uint32_t* cand_count, uint32_t m,
                                                  uint32_t want, uint32_t lane, uint32_t& my_key, uint32_t& my_branch,
                                                  uint32_t& my_count)
{
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

// m <= 32 candidates, one per lane: every lane ranks its own against all the others - m independent pairs of
// shuffles instead of `want` dependent arg-max rounds (each a chain of two reductions, a vote and a shuffle).
__device__ __forceinline__ uint32_t select_best_ranked(uint32_t* cand_key, uint32_t* cand_branch, uint32_t* cand_count, uint32_t m,
                                                       uint32_t want, uint32_t lane, uint32_t& my_key, uint32_t& my_branch,
                                                       uint32_t& my_count)
{
    const bool have = lane < m;
    const uint32_t hi = have ? cand_key[lane] : 0u;
    const uint32_t lo = have ? ~cand_branch[lane] : 0u;
    const uint32_t cn = have ? cand_count[lane] : 0u;
    __syncwarp();
    uint32_t rank = 0; // candidates that come before mine: higher key, or the same key and a lower branch id
    for (uint32_t i = 0; i < m; ++i)
    {
        const uint32_t oh = __shfl_sync(FULL, hi, i);
        const uint32_t ol = __shfl_sync(FULL, lo, i);
        rank += (oh > hi) | ((oh == hi) & (ol > lo));
    }
    const uint32_t n_sel = min(want, m);
    if (have && rank < n_sel) // branches are distinct: so are the ranks
    {
        cand_key[rank] = hi;
        cand_branch[rank] = ~lo;
        cand_count[rank] = cn;
    }
    __syncwarp();
    my_key = lane < n_sel ? cand_key[lane] : 0u;
    my_branch = lane < n_sel ? cand_branch[lane] : NO_BRANCH;
    my_count = lane < n_sel ? cand_count[lane] : 0u;
    __syncwarp();
    return n_sel;
}

// m <= CAND_CAP candidates: the fewer there are, the fewer each lane has to look at per round
__device__ __noinline__ uint32_t select_best(uint32_t* cand_key, uint32_t* cand_branch, uint32_t* cand_count, uint32_t m,
                                             uint32_t want, uint32_t lane, uint32_t& my_key, uint32_t& my_branch,
                                             uint32_t& my_count)
{
    if (m <= 32) { return select_best_ranked(cand_key, cand_branch, cand_count, m, want, lane, my_key, my_branch, my_count); }
    if (m <= 64) { return select_best_n<2>(cand_key, cand_branch, cand_count, m, want, lane, my_key, my_branch, my_count); }
    return select_best_n<CAND_CAP / 32>(cand_key, cand_branch, cand_count, m, want, lane, my_key, my_branch, my_count);
}

// Where a warp keeps its candidates and the quad queue of the scan-and-queue sweep: CAP words each of
// key / branch / count, then 64 words of queue.
template<int CAP>
struct cand_buffers
{
    uint32_t* key;
    uint32_t* branch;
    uint32_t* count;
    uint32_t* pending;
    __device__ explicit cand_buffers(uint32_t* base) : key(base), branch(base + CAP), count(base + 2 * CAP), pending(base + 3 * CAP) {}
    static constexpr uint32_t WORDS = 3 * CAP + 64;
};

// ---- the sweep over the accumulators ---------------------------------------------------------------

// four consecutive hit counts, and their clearing, by width
__device__ __forceinline__ void load_counts4(const uint8_t* C, uint32_t j, uint32_t (&c)[4])
{
    const uint32_t w = *reinterpret_cast<const uint32_t*>(C + j);
    c[0] = w & 0xFFu; c[1] = (w >> 8) & 0xFFu; c[2] = (w >> 16) & 0xFFu; c[3] = w >> 24;
}
__device__ __forceinline__ void load_counts4(const uint16_t* C, uint32_t j, uint32_t (&c)[4])
{
    const uint2 w = *reinterpret_cast<const uint2*>(C + j);
    c[0] = w.x & 0xFFFFu; c[1] = w.x >> 16; c[2] = w.y & 0xFFFFu; c[3] = w.y >> 16;
}
__device__ __forceinline__ void load_counts4(const uint32_t* C, uint32_t j, uint32_t (&c)[4])
{
    const uint4 w = *reinterpret_cast<const uint4*>(C + j);
    c[0] = w.x; c[1] = w.y; c[2] = w.z; c[3] = w.w;
}
__device__ __forceinline__ bool any_count4(const uint8_t* C, uint32_t j) { return *reinterpret_cast<const uint32_t*>(C + j) != 0u; }
__device__ __forceinline__ bool any_count4(const uint16_t* C, uint32_t j)
{
    const uint2 w = *reinterpret_cast<const uint2*>(C + j);
    return (w.x | w.y) != 0u;
}
__device__ __forceinline__ bool any_count4(const uint32_t* C, uint32_t j)
{
    const uint4 w = *reinterpret_cast<const uint4*>(C + j);
    return (w.x | w.y | w.z | w.w) != 0u;
}
__device__ __forceinline__ void clear_counts4(uint8_t* C, uint32_t j) { *reinterpret_cast<uint32_t*>(C + j) = 0u; }
__device__ __forceinline__ void clear_counts4(uint16_t* C, uint32_t j) { *reinterpret_cast<uint2*>(C + j) = make_uint2(0u, 0u); }
__device__ __forceinline__ void clear_counts4(uint32_t* C, uint32_t j) { *reinterpret_cast<uint4*>(C + j) = make_uint4(0u, 0u, 0u, 0u); }

// Where the sweep finds the accumulators of its query. shared_acc: the warp's S / C in shared memory
// (taking a quad clears it for the next query). rows_acc (database-sharded mode, the owner's side): the
// partial rows of the shards in global memory, added in shard order as they are read - the sum never
// exists anywhere but in registers.
// take-and-clear of four consecutive counts through a shared-window address (the generic-pointer form made
// nvcc recompute the window base - S2UR SR_CgaCtaId + ULEA, ~35 cycles - in every step of the sweep)
__device__ __forceinline__ void take_counts4(uint32_t c_sh, uint32_t j, uint32_t (&c)[4], uint8_t)
{
    uint32_t w;
    asm volatile("ld.shared.b32 %0, [%1];\n\tst.shared.b32 [%1], %2;" : "=r"(w) : "r"(c_sh + j), "r"(0u));
    c[0] = w & 0xFFu; c[1] = (w >> 8) & 0xFFu; c[2] = (w >> 16) & 0xFFu; c[3] = w >> 24;
}
__device__ __forceinline__ void take_counts4(uint32_t c_sh, uint32_t j, uint32_t (&c)[4], uint16_t)
{
    uint32_t x, y;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];\n\tst.shared.v2.b32 [%2], {%3, %3};" : "=r"(x), "=r"(y) : "r"(c_sh + 2u * j), "r"(0u));
    c[0] = x & 0xFFFFu; c[1] = x >> 16; c[2] = y & 0xFFFFu; c[3] = y >> 16;
}
__device__ __forceinline__ void take_counts4(uint32_t c_sh, uint32_t j, uint32_t (&c)[4], uint32_t)
{
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];\n\tst.shared.v4.b32 [%4], {%5, %5, %5, %5};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]) : "r"(c_sh + 4u * j), "r"(0u));
}

template<class CT>
struct shared_acc
{
    float* S;
    CT* C;
    uint32_t s_sh, c_sh; // the same two arrays as shared-window addresses
    __device__ __forceinline__ shared_acc(float* s, CT* c)
        : S(s), C(c), s_sh(static_cast<uint32_t>(__cvta_generic_to_shared(s))), c_sh(static_cast<uint32_t>(__cvta_generic_to_shared(c))) {}
    __device__ __forceinline__ bool any(uint32_t j) const { return any_count4(C, j); }
    __device__ __forceinline__ float4 take(uint32_t j, uint32_t (&c)[4]) const
    {
        float4 s4;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];\n\tst.shared.v4.b32 [%4], {%5, %5, %5, %5};"
                     : "=f"(s4.x), "=f"(s4.y), "=f"(s4.z), "=f"(s4.w) : "r"(s_sh + 4u * j), "r"(0u));
        take_counts4(c_sh, j, c, CT{});
        return s4;
    }
};

// the layout of a compact partial row (what it is and why: "partial rows" below)
constexpr uint32_t ROW_HEADER_BYTES = 16;
__host__ __device__ inline uint64_t row_bitmap_offset() { return ROW_HEADER_BYTES; }
__host__ __device__ inline uint64_t row_steps_offset(uint32_t np) { return ROW_HEADER_BYTES + np / 8; }
__host__ __device__ inline uint64_t row_s_offset(uint32_t np) { return ROW_HEADER_BYTES + np / 8 + 4ull * (np / SWEEP_STEP); }
__host__ __device__ inline uint64_t row_c_offset(uint32_t np) { return row_s_offset(np) + 4ull * np; }

enum row_format_kind { ROW_DENSE = 0, ROW_COMPACT = 1 };

struct rows_acc
{
    const unsigned char* row0;   // the query's row of shard 0; shard g is g * stride further
    uint64_t stride;
    uint32_t n_shards;
    uint32_t np;
    uint32_t format;             // ROW_DENSE: float S[np] | uint16 C[np] behind the header; ROW_COMPACT: bitmap, step counts, values
    float* amb_s;                // dense contributions of the ambiguous windows the owner replayed (NULL: none);
    uint16_t* amb_c;             // taken - and cleared - with the rows

    __device__ __forceinline__ bool any(uint32_t j) const
    {
        uint32_t seen = 0;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            if (format == ROW_DENSE)
            {
                const uint2 u = __ldg(reinterpret_cast<const uint2*>(row0 + g * stride + ROW_HEADER_BYTES + 4ull * np + 2ull * j));
                seen |= u.x | u.y;
                continue;
            }
            const uint32_t word = __ldg(reinterpret_cast<const uint32_t*>(row0 + g * stride + ROW_HEADER_BYTES) + j / 32u);
            seen |= (word >> (j % 32u)) & 0xFu;
        }
        if (amb_c)
        {
            const uint2 u = __ldcg(reinterpret_cast<const uint2*>(amb_c + j));
            seen |= u.x | u.y;
        }
        return seen != 0u;
    }
    __device__ __forceinline__ float4 take(uint32_t j, uint32_t (&c)[4]) const
    {
        float sv[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
        c[0] = c[1] = c[2] = c[3] = 0u;
        const uint32_t step = j / SWEEP_STEP, in_step = j % SWEEP_STEP;
        const uint32_t which = in_step / 32u, shift = in_step % 32u;
        constexpr uint32_t GROUP = 4; // shards whose loads are in flight together (a row's bitmap, its step count, its values)
        if (format == ROW_DENSE)
        {
            const unsigned char* s_rows = row0 + ROW_HEADER_BYTES + 4ull * j;
            const unsigned char* c_rows = row0 + ROW_HEADER_BYTES + 4ull * np + 2ull * j;
            float4 s4 = __ldg(reinterpret_cast<const float4*>(s_rows));
            uint2 c2 = __ldg(reinterpret_cast<const uint2*>(c_rows));
            for (uint32_t g = 1; g < n_shards; ++g)
            {
                const float4 t = __ldg(reinterpret_cast<const float4*>(s_rows + g * stride));
                const uint2 u = __ldg(reinterpret_cast<const uint2*>(c_rows + g * stride));
                s4.x = __fadd_rn(s4.x, t.x); s4.y = __fadd_rn(s4.y, t.y);
                s4.z = __fadd_rn(s4.z, t.z); s4.w = __fadd_rn(s4.w, t.w);
                c2.x += u.x; c2.y += u.y; // 16-bit halves cannot carry: a branch is hit at most n_windows <= 65535 times
            }
            sv[0] = s4.x; sv[1] = s4.y; sv[2] = s4.z; sv[3] = s4.w;
            c[0] = c2.x & 0xFFFFu; c[1] = c2.x >> 16; c[2] = c2.y & 0xFFFFu; c[3] = c2.y >> 16;
        }
        for (uint32_t g0 = 0; format == ROW_COMPACT && g0 < n_shards; g0 += GROUP)
        {
            uint4 w4[GROUP];
            uint32_t base[GROUP];
#pragma unroll
            for (uint32_t i = 0; i < GROUP; ++i)
            {
                w4[i] = make_uint4(0u, 0u, 0u, 0u);
                base[i] = 0u;
                if (g0 + i < n_shards)
                {
                    const unsigned char* row = row0 + (g0 + i) * stride;
                    w4[i] = __ldg(reinterpret_cast<const uint4*>(row + ROW_HEADER_BYTES) + step);
                    base[i] = __ldg(reinterpret_cast<const uint32_t*>(row + row_steps_offset(np)) + step);
                }
            }
            uint32_t nib[GROUP];
            float vs[GROUP][4];
            uint32_t vc[GROUP][4];
#pragma unroll
            for (uint32_t i = 0; i < GROUP; ++i)
            {
                const uint32_t w[4] = { w4[i].x, w4[i].y, w4[i].z, w4[i].w };
                uint32_t before = 0, word = 0;
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k)
                {
                    before += k < which ? __popc(w[k]) : 0u;
                    word = k == which ? w[k] : word;
                }
                nib[i] = (word >> shift) & 0xFu;
                uint32_t at = base[i] + before + __popc(word & ((1u << shift) - 1u));
                const unsigned char* row = row0 + (g0 + i) * stride;
                const float* s_in = reinterpret_cast<const float*>(row + row_s_offset(np));
                const uint16_t* c_in = reinterpret_cast<const uint16_t*>(row + row_c_offset(np));
#pragma unroll
                for (int u = 0; u < 4; ++u)
                {
                    vs[i][u] = 0.0f;
                    vc[i][u] = 0u;
                    if ((nib[i] >> u) & 1u)
                    {
                        vs[i][u] = __ldg(s_in + at);
                        vc[i][u] = __ldg(c_in + at);
                        ++at;
                    }
                }
            }
            // shard order: a fixed association, the same on every run
#pragma unroll
            for (uint32_t i = 0; i < GROUP; ++i)
            {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                {
                    if ((nib[i] >> u) & 1u)
                    {
                        sv[u] = __fadd_rn(sv[u], vs[i][u]);
                        c[u] += vc[i][u]; // a branch is hit at most n_windows <= 65535 times
                    }
                }
            }
        }
        if (amb_c)
        {
            const float4 t = __ldcg(reinterpret_cast<const float4*>(amb_s + j));
            const uint2 u = __ldcg(reinterpret_cast<const uint2*>(amb_c + j));
            if ((u.x | u.y) != 0u)
            {
                const uint32_t ac[4] = { u.x & 0xFFFFu, u.x >> 16, u.y & 0xFFFFu, u.y >> 16 };
                const float as[4] = { t.x, t.y, t.z, t.w };
#pragma unroll
                for (int v = 0; v < 4; ++v)
                {
                    if (ac[v]) { sv[v] = __fadd_rn(sv[v], as[v]); c[v] += ac[v]; }
                }
                __stcg(reinterpret_cast<float4*>(amb_s + j), make_float4(0.0f, 0.0f, 0.0f, 0.0f));
                __stcg(reinterpret_cast<uint2*>(amb_c + j), make_uint2(0u, 0u));
            }
        }
        return make_float4(sv[0], sv[1], sv[2], sv[3]);
    }
};

// Terms of the likelihood mass more than this many decades under the best score the warp has seen
// are left out: each is below 1e-20 of the total, all of them together below N * 1e-20.
constexpr float MASS_CUTOFF_DECADES = 20.0f;

// a scan step with at least this many touched quads (of 32) is processed in place rather than queued
constexpr uint32_t DENSE_QUADS = 24;
static_assert(3 * CAND_CAP * 4 + 64 * 4 <= (SEG_CAP + SEG_PAD) * 8, "candidates and the quad queue alias the segment buffer");

struct sweep_result
{
    double mass;       // sum over the touched branches of 10^score            (place.cpp:174-175)
    uint32_t touched;  // |L|: branches with at least one hit                   (place.cpp:166)
    uint32_t n_keep;   // min(keep_at_most, touched); lane i < n_keep holds the i-th best in key/branch/count
    uint32_t key, branch, count;
};

// One pass over S[0, np), C[0, np) of a query with nw windows; leaves both arrays zeroed.
// Per branch b with C[b] != 0 (place.cpp:417-422):  S[b] += float(nw - C[b]) * log10thr;  S[b] /= k.
// FAST: the short forms of the division and of 10^s (valid for the score range established by
// fast_sweep_ok) and the mass cut-off; otherwise the IEEE division and the library exp10, every term.
// The warp looks at the steps scan0, scan0 + stride, ... (the warps of a team take interleaved steps).
template<class ACC, bool FAST, bool DENSE, int CAP>
__device__ __forceinline__ sweep_result sweep(const ACC acc, uint32_t np, uint32_t nw32, float logthr, float kf,
                                              float rcp_k, uint32_t K, uint32_t lane, const cand_buffers<CAP>& cb,
                                              uint32_t scan0, uint32_t stride)
{
    uint32_t* const cand_key = cb.key;
    uint32_t* const cand_branch = cb.branch;
    uint32_t* const cand_count = cb.count;
    sweep_result out{ 0.0, 0u, 0u, 0u, NO_BRANCH, 0u };
    double mass = 0.0;
    uint32_t touched = 0, n_cand = 0;
    // A candidate beats (bar_key, bar_branch): higher key, or the same key and a lower branch id. The
    // bar is a lower bound of the K-th best score: the K-th largest of the lanes' best keys so far
    // (refreshed after steps 1, 2, 4, 8 ...), or the K-th best candidate when the buffer was squeezed.
    uint32_t bar_key = 0, bar_branch = 0; // nothing beats key 0 / branch 0: untouched branches have key 0
    uint32_t cut_key = 1;                 // mass terms below this key are negligible; 1 admits every touched branch
    uint32_t lane_max = 0;                // best score key this lane has seen
    uint32_t step = 0;
    // The accumulators are looked at in runs of 4 consecutive branches ("quads"), 32 quads per scan step.
    // `dense` (the query gathered at least a sixth as many postings as there are branches): every step is
    // processed as it stands, lane l takes quad l. Otherwise the count words are scanned first: a step
    // with many touched quads is still processed in place, the touched quads of sparse steps are queued
    // and processed 32 at a time, so that a query that touched few branches pays for those only (a
    // 100-residue protein read touches a few dozen of the 2000 branches of its tree).
    uint32_t* pending = cb.pending; // [64] first branch of queued quads
    uint32_t n_pending = 0, scan = scan0;
    while (true)
    {
        // ---- next 32 quads to process: j = first branch of this lane's quad ----
        uint32_t j = 0;
        bool active = DENSE, have = false;
        if (DENSE)
        {
            // a query with many postings touched most quads: every step as it stands, no scan
            if (scan >= np) { break; }
            j = scan + 4u * lane;
            scan += stride;
            have = true;
        }
        while (!DENSE && !have && scan < np)
        {
            const uint32_t mine = scan + 4u * lane;
            const bool hit = acc.any(mine);
            const uint32_t ballot = __ballot_sync(FULL, hit);
            const uint32_t n_hit = __popc(ballot);
            scan += stride;
            if (n_hit >= DENSE_QUADS)
            {
                j = mine;
                active = true;
                have = true;
            }
            else if (n_hit)
            {
                if (hit) { pending[n_pending + __popc(ballot & ((1u << lane) - 1u))] = mine; }
                n_pending += n_hit;
                __syncwarp();
                if (n_pending >= 32)
                {
                    j = pending[lane];
                    const uint32_t rest = n_pending - 32;
                    const uint32_t moved = lane < rest ? pending[32 + lane] : 0u;
                    __syncwarp();
                    if (lane < rest) { pending[lane] = moved; }
                    __syncwarp();
                    n_pending = rest;
                    active = true;
                    have = true;
                }
            }
        }
        if (!DENSE && !have)
        {
            if (n_pending == 0) { break; }
            active = lane < n_pending;
            j = active ? pending[lane] : 0u;
            n_pending = 0;
        }

        // ---- the quad: correction, keys ----
        float4 s4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        uint32_t c[4] = { 0u, 0u, 0u, 0u };
        if (DENSE || active) { s4 = acc.take(j, c); }
        const float sum[4] = { s4.x, s4.y, s4.z, s4.w };
        float s[4];
        uint32_t key[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            // the reference converts the size_t difference to float (place.cpp:420)
            const float x = __fadd_rn(sum[u], __fmul_rn(static_cast<float>(nw32 - c[u]), logthr));
            s[u] = FAST ? div_by_k_plain(x, kf, rcp_k) : __fdiv_rn(x, kf);
            // FAST: no -0 can come out of the division (x is 0 or |x| > 1e-30), the key needs no folding
            key[u] = c[u] != 0 ? (FAST ? score_key_raw(s[u]) : score_key(s[u])) : 0u;
            touched += c[u] != 0;
            lane_max = max(lane_max, key[u]);
        }
        if ((step & (step + 1)) == 0)
        {
            // K-th largest of the lanes' maxima (0 while fewer than K lanes have a touched branch). A round takes
            // the maximum and strikes every lane that holds it: lanes that tie are struck together, which can only
            // lower the result - still a lower bound of the K-th best score. (Striking one lane per round through
            // ballot + ffs cost 130 cycles a round, 7 % of a 150-bp query.)
            uint32_t mine = lane_max, kth = 0, top = 0;
            for (uint32_t i = 0; i < K; ++i)
            {
                kth = __reduce_max_sync(FULL, mine);
                top = i == 0 ? kth : top;
                mine = mine == kth ? 0u : mine;
            }
            if (kth > bar_key)
            {
                bar_key = kth;
                bar_branch = NO_BRANCH;
            }
            if (FAST && top != 0) { cut_key = max(cut_key, score_key_raw(key_score(top) - MASS_CUTOFF_DECADES)); }
        }
        ++step;
        // ---- likelihood mass (place.cpp:164-184) ----
        const uint32_t best4 = max(max(key[0], key[1]), max(key[2], key[3])); // best key of the quad
        if (!(EPK_ABLATE & 1) && __any_sync(FULL, best4 >= cut_key))
        {
            double e[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { e[u] = FAST ? exp10_small(s[u]) : exp10(static_cast<double>(s[u])); }
#pragma unroll
            for (int u = 0; u < 4; ++u) { e[u] = key[u] >= cut_key ? e[u] : 0.0; }
            mass += (e[0] + e[1]) + (e[2] + e[3]); // a tree: four chained additions waited 8 cycles each
        }
        // ---- top-k candidates (place.cpp:134-159) ----
        // (a candidate has key >= bar_key: one compare of the quad's best key keeps the four exact tests, ties
        // and all, out of the common path)
        if (!(EPK_ABLATE & 2) && __any_sync(FULL, best4 >= max(bar_key, 1u)))
        {
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                const bool cand = (key[u] > bar_key) | ((key[u] == bar_key) & (j + u < bar_branch));
                const uint32_t ballot = __ballot_sync(FULL, cand);
                if (ballot == 0) { continue; }
                if (cand)
                {
                    const uint32_t at = n_cand + __popc(ballot & ((1u << lane) - 1u));
                    cand_key[at] = key[u];
                    cand_branch[at] = j + u;
                    cand_count[at] = c[u];
                }
                n_cand += __popc(ballot);
                __syncwarp();
                if (n_cand > CAP - 32)
                {
                    // keep the best K, raise the bar to the K-th of them
                    n_cand = select_best(cand_key, cand_branch, cand_count, n_cand, K, lane, out.key, out.branch, out.count);
                    if (n_cand == K)
                    {
                        bar_key = __shfl_sync(FULL, out.key, K - 1);
                        bar_branch = __shfl_sync(FULL, out.branch, K - 1);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
    {
        mass += __shfl_xor_sync(FULL, mass, d);
        touched += __shfl_xor_sync(FULL, touched, d);
    }
    __syncwarp();
    out.n_keep = select_best(cand_key, cand_branch, cand_count, n_cand, K, lane, out.key, out.branch, out.count);
    out.mass = mass;
    out.touched = touched;
    return out;
}

// ---- the dense sweep over shared-memory accumulators: two passes ---------------------------------------
//
// Phase clocks of the one-pass sweep on C2 (150-bp reads, 4999 branches; gpurun r3b): 36 k cycles per query, of
// which 12 k the likelihood mass - four dependent DFMA chains per lane against ~20 cycles of FP64 latency - and
// 13 k the candidate bookkeeping against a bar that is refreshed six times per query: votes, ballots and branches
// in 15 of the 40 steps. Here
//   pass 1 corrects NQ quads (4 * NQ branches) per lane and iteration - 4 * NQ independent 10^s chains -, adds the
//          mass, tracks the lane's best key and leaves the KEYS in place of the sums (counts stay);
//   then   the bar is the K-th largest of the lanes' best keys over the WHOLE query - at most K - 1 lanes, hence
//          typically a dozen branches, beat it;
//   pass 2 re-reads the keys (LDS.128 per quad), clears sums and counts, and hands the few quads that reach the
//          bar to the same tie-aware candidate code as the one-pass sweep.
// Same results as sweep<..., FAST = true, DENSE = true>: scores, counts, ties; the mass differs in the order of
// its FP64 additions only.
#ifndef EPK_SWEEP_NQ
#define EPK_SWEEP_NQ 2
#endif
#ifndef EPK_SWEEP_NQ_WIDE
#define EPK_SWEEP_NQ_WIDE 1 // the 16-warp launch (128 registers per thread): C4 75.2 -> 77.4 M reads/s, C1p 8.19 -> 8.30 M (profiles/r4w)
#endif
#ifndef EPK_WIDE_AHEAD
#define EPK_WIDE_AHEAD 1 // kernel experiments: 0 = the 16-warp launch claims its queries one at a time
#endif
#ifndef EPK_SWEEP_P2
#define EPK_SWEEP_P2 4
#endif
__device__ __forceinline__ void load_counts4_sh(uint32_t c_sh, uint32_t j, uint32_t (&c)[4], uint8_t)
{
    uint32_t w;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(c_sh + j));
    c[0] = w & 0xFFu; c[1] = (w >> 8) & 0xFFu; c[2] = (w >> 16) & 0xFFu; c[3] = w >> 24;
}
__device__ __forceinline__ void load_counts4_sh(uint32_t c_sh, uint32_t j, uint32_t (&c)[4], uint16_t)
{
    uint32_t x, y;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(c_sh + 2u * j));
    c[0] = x & 0xFFFFu; c[1] = x >> 16; c[2] = y & 0xFFFFu; c[3] = y >> 16;
}
__device__ __forceinline__ void load_counts4_sh(uint32_t c_sh, uint32_t j, uint32_t (&c)[4], uint32_t)
{
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]) : "r"(c_sh + 4u * j));
}

// single words through the shared window, as volatile asm like every other access of the two-pass sweep (the
// compiler keeps volatile asm statements in program order; it knows nothing about the memory they touch)
__device__ __forceinline__ uint32_t lds_b32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t load_count_sh(uint32_t c_sh, uint32_t b, uint8_t)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(c_sh + b));
    return v;
}
__device__ __forceinline__ uint32_t load_count_sh(uint32_t c_sh, uint32_t b, uint16_t)
{
    uint32_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(c_sh + 2u * b));
    return v;
}
__device__ __forceinline__ uint32_t load_count_sh(uint32_t c_sh, uint32_t b, uint32_t) { return lds_b32(c_sh + 4u * b); }
__device__ __forceinline__ void clear_counts4_sh(uint32_t c_sh, uint32_t j, uint8_t) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(c_sh + j), "r"(0u)); }
__device__ __forceinline__ void clear_counts4_sh(uint32_t c_sh, uint32_t j, uint16_t) { asm volatile("st.shared.v2.b32 [%0], {%1, %1};" ::"r"(c_sh + 2u * j), "r"(0u)); }
__device__ __forceinline__ void clear_counts4_sh(uint32_t c_sh, uint32_t j, uint32_t) { asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(c_sh + 4u * j), "r"(0u)); }

template<class CT, int CAP, int NQ>
__device__ __forceinline__ sweep_result sweep_dense_two_pass(const shared_acc<CT> acc, uint32_t np, uint32_t nw32, float logthr,
                                                             float kf, float rcp_k, uint32_t K, uint32_t lane,
                                                             const cand_buffers<CAP>& cb, uint32_t scan0, uint32_t stride
                                                             EPK_PHASE_PARAMS)
{
    uint32_t* const cand_key = cb.key;
    uint32_t* const cand_branch = cb.branch;
    uint32_t* const cand_count = cb.count;
    sweep_result out{ 0.0, 0u, 0u, 0u, NO_BRANCH, 0u };
    double mass = 0.0;
    uint32_t touched = 0, lane_max = 0, cut_key = 1, step = 0;
    const uint32_t lane4 = 4u * lane;

    // ---- pass 1: correction (place.cpp:417-422), likelihood mass (place.cpp:164-184), keys in place of the sums ----
    for (uint32_t scan = scan0; scan < np; scan += NQ * stride)
    {
        float sc[NQ][4];
        uint32_t key[NQ][4];
        uint32_t best = 0;
#pragma unroll
        for (int q = 0; q < NQ; ++q)
        {
            const uint32_t j = scan + q * stride + lane4;
            float4 s4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            uint32_t c[4] = { 0u, 0u, 0u, 0u };
            if (q == 0 || scan + q * stride < np) // (warp-uniform)
            {
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(s4.x), "=f"(s4.y), "=f"(s4.z), "=f"(s4.w) : "r"(acc.s_sh + 4u * j));
                load_counts4_sh(acc.c_sh, j, c, CT{});
            }
            const float sum[4] = { s4.x, s4.y, s4.z, s4.w };
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                // the reference converts the size_t difference to float (place.cpp:420)
                const float x = __fadd_rn(sum[u], __fmul_rn(static_cast<float>(nw32 - c[u]), logthr));
                sc[q][u] = div_by_k_plain(x, kf, rcp_k);
                key[q][u] = c[u] != 0 ? score_key_raw(sc[q][u]) : 0u; // (no -0 can come out of the division)
                touched += c[u] != 0;
            }
            best = max(best, max(max(key[q][0], key[q][1]), max(key[q][2], key[q][3])));
        }
        lane_max = max(lane_max, best);
        if ((step & (step + 1)) == 0)
        {
            // mass terms 20 decades under the best score seen so far are negligible
            const uint32_t top = __reduce_max_sync(FULL, lane_max);
            if (top != 0) { cut_key = max(cut_key, score_key_raw(key_score(top) - MASS_CUTOFF_DECADES)); }
        }
        ++step;
        if (!(EPK_ABLATE & 1) && __any_sync(FULL, best >= cut_key))
        {
            double e[NQ][4];
#pragma unroll
            for (int q = 0; q < NQ; ++q)
            {
#pragma unroll
                for (int u = 0; u < 4; ++u) { e[q][u] = exp10_small(sc[q][u]); }
            }
#pragma unroll
            for (int q = 0; q < NQ; ++q)
            {
#pragma unroll
                for (int u = 0; u < 4; ++u) { e[q][u] = key[q][u] >= cut_key ? e[q][u] : 0.0; }
            }
            double part = 0.0;
#pragma unroll
            for (int q = 0; q < NQ; ++q) { part += (e[q][0] + e[q][1]) + (e[q][2] + e[q][3]); }
            mass += part;
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q)
        {
            if (q == 0 || scan + q * stride < np)
            {
                const uint32_t j = scan + q * stride + lane4;
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(acc.s_sh + 4u * j), "r"(key[q][0]), "r"(key[q][1]),
                             "r"(key[q][2]), "r"(key[q][3]));
            }
        }
    }

    EPK_PHASE(5)
    // ---- the bar: K-th largest of the lanes' best keys (0 while fewer than K lanes have a touched branch). A round
    // takes the maximum and strikes every lane that holds it; lanes that tie are struck together, which can only
    // lower the result - still a lower bound of the K-th best key of the query. ----
    uint32_t bar_key = 0, bar_branch = NO_BRANCH;
    {
        uint32_t mine = lane_max, kth = 0;
        for (uint32_t i = 0; i < K; ++i)
        {
            kth = __reduce_max_sync(FULL, mine);
            mine = mine == kth ? 0u : mine;
        }
        bar_key = kth;
    }

    EPK_PHASE(8)
    // ---- pass 2: the candidates (place.cpp:134-159) ----
    // Every key that reaches the bar is a candidate; there are few (K - 1 lanes at most have a best key above the
    // bar). The lanes append theirs on their own - a counter in shared memory hands out the slots, no vote, no
    // prefix sum: the collective version of this loop cost 6 k cycles per C2 query (gpurun r3e), ten times its
    // loads. Only when more candidates turn up than the buffer holds (masses of exact ties) is the scan repeated
    // with the collective append-and-squeeze of the one-pass sweep. The accumulators are cleared afterwards.
    constexpr int P2 = EPK_SWEEP_P2;
    uint32_t* const counter = cb.pending; // (the quad queue of the scan-and-queue sweep: idle here)
    const uint32_t floor_key = max(bar_key, 1u);
    if (lane == 0) { *counter = 0u; }
    __syncwarp();
    if (!(EPK_ABLATE & 2))
    {
        for (uint32_t scan = scan0; scan < np; scan += P2 * stride)
        {
            uint32_t key[P2][4];
            uint32_t best = 0;
#pragma unroll
            for (int i = 0; i < P2; ++i)
            {
                key[i][0] = key[i][1] = key[i][2] = key[i][3] = 0u;
                if (i == 0 || scan + i * stride < np) // (warp-uniform)
                {
                    const uint32_t j = scan + i * stride + lane4;
                    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(key[i][0]), "=r"(key[i][1]), "=r"(key[i][2]), "=r"(key[i][3]) : "r"(acc.s_sh + 4u * j));
                }
                best = max(best, max(max(key[i][0], key[i][1]), max(key[i][2], key[i][3])));
            }
            if (best >= floor_key)
            {
                uint32_t mask = 0; // bit 4 i + u: branch u of the lane's quad in step i
#pragma unroll
                for (int i = 0; i < P2; ++i)
                {
#pragma unroll
                    for (int u = 0; u < 4; ++u) { mask |= key[i][u] >= floor_key ? 1u << (4 * i + u) : 0u; }
                }
                uint32_t at = atomicAdd(counter, static_cast<uint32_t>(__popc(mask)));
                while (mask && at < static_cast<uint32_t>(CAP))
                {
                    const uint32_t b = __ffs(mask) - 1u;
                    mask &= mask - 1u;
                    const uint32_t branch = scan + (b >> 2) * stride + lane4 + (b & 3u);
                    cand_key[at] = lds_b32(acc.s_sh + 4u * branch); // (the key pass 1 left there)
                    cand_branch[at] = branch;
                    cand_count[at] = load_count_sh(acc.c_sh, branch, CT{});
                    ++at;
                }
            }
        }
    }
    __syncwarp();
    uint32_t n_cand = *counter;
    __syncwarp();
    if (n_cand > static_cast<uint32_t>(CAP))
    {
        // too many for the buffer: position by position with votes, squeezing to the best K (and raising the bar to
        // the K-th of them, ties by branch id) whenever the buffer fills up
        n_cand = 0;
        for (uint32_t scan = scan0; scan < np; scan += stride)
        {
            for (uint32_t u = 0; u < 4; ++u)
            {
                const uint32_t branch = scan + lane4 + u;
                const uint32_t key = lds_b32(acc.s_sh + 4u * branch);
                const bool cand = (key > bar_key) | ((key == bar_key) & (branch < bar_branch) & (key != 0u));
                const uint32_t ballot = __ballot_sync(FULL, cand);
                if (ballot == 0) { continue; }
                if (cand)
                {
                    const uint32_t at = n_cand + __popc(ballot & ((1u << lane) - 1u));
                    cand_key[at] = key;
                    cand_branch[at] = branch;
                    cand_count[at] = load_count_sh(acc.c_sh, branch, CT{});
                }
                n_cand += __popc(ballot);
                __syncwarp();
                if (n_cand > CAP - 32)
                {
                    n_cand = select_best(cand_key, cand_branch, cand_count, n_cand, K, lane, out.key, out.branch, out.count);
                    if (n_cand == K)
                    {
                        bar_key = __shfl_sync(FULL, out.key, K - 1);
                        bar_branch = __shfl_sync(FULL, out.branch, K - 1);
                    }
                }
            }
        }
    }
    // ---- sums and counts cleared for the next query ----
    for (uint32_t scan = scan0; scan < np; scan += stride)
    {
        const uint32_t j = scan + lane4;
        asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(acc.s_sh + 4u * j), "r"(0u));
        clear_counts4_sh(acc.c_sh, j, CT{});
    }
    EPK_PHASE(9)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
    {
        mass += __shfl_xor_sync(FULL, mass, d);
        touched += __shfl_xor_sync(FULL, touched, d);
    }
    __syncwarp();
    out.n_keep = select_best(cand_key, cand_branch, cand_count, n_cand, K, lane, out.key, out.branch, out.count);
    EPK_PHASE(10)
    out.mass = mass;
    out.touched = touched;
    return out;
}

// The forms off the beaten track, out of line (they are long, and cold for 150-bp reads on a 5000-branch
// tree): IEEE arithmetic; scan-and-queue for queries that touched few branches.
template<class ACC, int CAP>
__device__ __noinline__ sweep_result sweep_ieee(const ACC acc, uint32_t np, uint32_t nw32, float logthr, float kf,
                                                float rcp_k, uint32_t K, uint32_t lane, const cand_buffers<CAP>& cb,
                                                uint32_t scan0, uint32_t stride)
{
    return sweep<ACC, false, false, CAP>(acc, np, nw32, logthr, kf, rcp_k, K, lane, cb, scan0, stride);
}
template<class ACC, int CAP>
__device__ __noinline__ sweep_result sweep_sparse(const ACC acc, uint32_t np, uint32_t nw32, float logthr, float kf,
                                                  float rcp_k, uint32_t K, uint32_t lane, const cand_buffers<CAP>& cb,
                                                  uint32_t scan0, uint32_t stride)
{
    return sweep<ACC, true, false, CAP>(acc, np, nw32, logthr, kf, rcp_k, K, lane, cb, scan0, stride);
}

// Whether every score a query with nwf windows can end up with stays inside what the short forms
// of the sweep were checked for: |score| < EXP10_SMALL_LIMIT, and a corrected sum that is 0 or has
// 1e-30 < |x| < 1e30 (the database guarantees the latter through fast_ok: finite scores <= 0 with
// magnitude 0 or >= 1e-20).
__device__ __forceinline__ bool fast_sweep_ok(const device_db& db, float nwf, float kf)
{
    const float lo = nwf * fminf(db.score_lo, db.log_threshold) / kf;
    const float hi = (nwf * fmaxf(db.log_threshold, 0.0f) + 1.0f) / kf; // ambiguous windows add at most 1 (a probability)
    return db.fast_ok != 0 && lo > -EXP10_SMALL_LIMIT && hi < EXP10_SMALL_LIMIT;
}

// ---- the end of a query ---------------------------------------------------------------------------
//
// The accumulators S, C of one query with n_windows windows are complete (place.cpp:349-415 done):
// correction + likelihood mass + top-k (the sweep), weight ratios, keep_factor, the rows of query q.
// `gathered` = postings the query gathered (warp-uniform; only steers dense / scan-and-queue).

// the dense form by kind of accumulators: two passes over shared memory, one pass over the partial rows in HBM
template<int CAP, int NQ, class CT>
__device__ __forceinline__ sweep_result sweep_dense_form(const shared_acc<CT> acc, uint32_t np, uint32_t nw32, float logthr, float kf,
                                                         float rcp_k, uint32_t K, uint32_t lane, const cand_buffers<CAP>& cb,
                                                         uint32_t scan0, uint32_t stride EPK_PHASE_PARAMS)
{
    return sweep_dense_two_pass<CT, CAP, NQ>(acc, np, nw32, logthr, kf, rcp_k, K, lane, cb, scan0, stride EPK_PHASE_ARGS);
}
template<int CAP, int NQ, class ACC>
__device__ __forceinline__ sweep_result sweep_dense_form(const ACC acc, uint32_t np, uint32_t nw32, float logthr, float kf,
                                                         float rcp_k, uint32_t K, uint32_t lane, const cand_buffers<CAP>& cb,
                                                         uint32_t scan0, uint32_t stride EPK_PHASE_PARAMS)
{
    return sweep<ACC, true, true, CAP>(acc, np, nw32, logthr, kf, rcp_k, K, lane, cb, scan0, stride);
}

// the sweep form that fits the query
template<class ACC, int CAP, int NQ = EPK_SWEEP_NQ>
__device__ __forceinline__ sweep_result run_sweep(const device_db& db, const ACC acc, uint32_t np, uint32_t nw32,
                                                  uint32_t gathered, uint32_t K, uint32_t lane, const cand_buffers<CAP>& cb,
                                                  uint32_t scan0, uint32_t stride EPK_PHASE_PARAMS)
{
    const float kf = static_cast<float>(db.k);
    const float rcp_k = __frcp_rn(kf);
    const float logthr = db.log_threshold;
    const float nwf = static_cast<float>(nw32);
    const bool dense = 6ull * gathered >= np;
    return !fast_sweep_ok(db, nwf, kf) ? sweep_ieee<ACC, CAP>(acc, np, nw32, logthr, kf, rcp_k, K, lane, cb, scan0, stride)
           : dense                     ? sweep_dense_form<CAP, NQ>(acc, np, nw32, logthr, kf, rcp_k, K, lane, cb, scan0, stride EPK_PHASE_ARGS)
                                       : sweep_sparse<ACC, CAP>(acc, np, nw32, logthr, kf, rcp_k, K, lane, cb, scan0, stride);
}

// likelihood weight ratios and keep_factor (place.cpp:164-184, 241-267), rows of query q
__device__ __forceinline__ void emit_query(const place_params& p, const sweep_result& sw, uint32_t nw32, uint32_t q, uint32_t lane)
{
    const device_db& db = p.db;
    const uint32_t N = db.n_nodes;
    const uint32_t K = p.keep_at_most;
    const float kf = static_cast<float>(db.k);
    const float logthr = db.log_threshold;
    const float nwf = static_cast<float>(nw32);
    const uint32_t touched = sw.touched;
    uint32_t my_b = sw.branch, my_c = sw.count;
    float my_s = key_score(sw.key);

    const float absent_score = __fdiv_rn(__fmul_rn(nwf, logthr), kf);
    const double absent_mass = static_cast<double>(static_cast<float>(N) - static_cast<float>(touched))
                               * exp10_of_float(absent_score);
    const double total = absent_mass + sw.mass;
    uint32_t n_out = sw.n_keep;
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

// one warp does it all; cand_base: CAND_CAP * 3 + 64 words (the idle segment buffer)
template<class ACC, int NQ = EPK_SWEEP_NQ>
__device__ __forceinline__ void finish_query(const place_params& p, const ACC acc, uint32_t np, uint64_t n_windows,
                                             uint32_t gathered, uint32_t q, uint32_t lane, uint32_t* cand_base)
{
    const cand_buffers<CAND_CAP> cb(cand_base);
    const uint32_t nw32 = static_cast<uint32_t>(n_windows);
    EPK_PHASE_DUMMY
    const sweep_result sw = run_sweep<ACC, CAND_CAP, NQ>(p.db, acc, np, nw32, gathered, p.keep_at_most, lane, cb, 0u, SWEEP_STEP EPK_PHASE_ARGS);
    emit_query(p, sw, nw32, q, lane);
}

// ---- teams: several warps sweep one query ---------------------------------------------------------
//
// A large tree leaves room for few accumulator sets per SM (N = 20 000: two), and a lone warp per set
// then spends two thirds of its time in the sweep with the SM idle around it (ncu, profiles/r1w). In
// team mode every set has T warps (T * sets = 8): the leader (member 0) encodes, probes and gathers as
// before - the order of the additions per branch is untouched - then all T members sweep interleaved
// 128-branch steps through ONE out-of-line copy of the sweep (the two-warps experiment of DESIGN.md 3.1
// died of instruction-cache misses with two inlined copies), each with its own candidate buffer; the
// leader adds the members' masses in member order and merges their best keep_at_most candidates. The
// team meets on a named barrier (id 1 + set): A = accumulators complete, B = members done.
constexpr int TEAM_CAND_CAP = 64;
constexpr uint32_t TEAM_MEMBER_WORDS = cand_buffers<TEAM_CAND_CAP>::WORDS;     // candidate buffer of one member
struct team_mailbox { uint32_t q, nw32, gathered, pad; };                      // leader -> members (q == NO_BRANCH: leave)
struct team_result { uint32_t mass_lo, mass_hi, touched, n_keep; };            // member -> leader
__host__ __device__ inline size_t team_smem_bytes(uint32_t team)
{
    return team <= 1 ? 0 : team * (TEAM_MEMBER_WORDS * 4 + sizeof(team_result)) + sizeof(team_mailbox);
}

__device__ __forceinline__ void team_sync(uint32_t set, uint32_t team)
{
    asm volatile("bar.sync %0, %1;" ::"r"(set + 1u), "r"(team * 32u) : "memory");
}

// one member's share of the sweep; out of line so that all members of all teams run the same code
template<class CT>
__device__ __noinline__ sweep_result team_sweep(const device_db& db, float* S, CT* C, uint32_t np, uint32_t nw32,
                                                uint32_t gathered, uint32_t K, uint32_t lane, uint32_t* member_base,
                                                uint32_t member, uint32_t team)
{
    const cand_buffers<TEAM_CAND_CAP> cb(member_base);
    EPK_PHASE_DUMMY
    return run_sweep<shared_acc<CT>, TEAM_CAND_CAP>(db, shared_acc<CT>(S, C), np, nw32, gathered, K, lane, cb,
                                                    member * SWEEP_STEP, team * SWEEP_STEP EPK_PHASE_ARGS);
}

// mailbox.pad: 0 = sweep; GATHER | n_hits = gather the published hits (range teams); DUMP = write this member's part of
// the query's partial row (database-sharded mode; nw32 carries the row flags)
constexpr uint32_t TEAM_CMD_GATHER = 0x80000000u;
constexpr uint32_t TEAM_CMD_DUMP = 0x40000000u;

template<class CT>
__device__ __forceinline__ void dump_partial_row(const place_params& p, float* S, CT* C, uint32_t np, uint32_t q,
                                                 uint32_t flags, uint32_t gathered, uint32_t lane, uint32_t* stage, uint32_t part = 0,
                                                 uint32_t parts = 1, uint32_t* step_counts = nullptr, uint32_t set = 0);

template<int U, class CT, bool SINK>
__device__ __noinline__ void range_gather(const device_db& db, acc_ref acc, const uint64_t* hits, uint32_t n_hits, uint2* my_segs,
                                          uint32_t member, uint32_t team, uint32_t lane, uint32_t pf, uint32_t* sink_out);

template<int U, class CT, bool SINK>
__device__ __forceinline__ void team_helper_loop(const place_params& p, float* S, CT* C, uint32_t np, uint32_t lane,
                                                 unsigned char* team_area, const uint64_t* hits, uint32_t set,
                                                 uint32_t member, uint32_t team)
{
    const acc_ref acc{ static_cast<uint32_t>(__cvta_generic_to_shared(S)), static_cast<uint32_t>(__cvta_generic_to_shared(C)) };
    const team_mailbox* box = reinterpret_cast<const team_mailbox*>(team_area);
    team_result* results = reinterpret_cast<team_result*>(team_area + sizeof(team_mailbox));
    uint32_t* member_base = reinterpret_cast<uint32_t*>(team_area + sizeof(team_mailbox) + team * sizeof(team_result))
                            + member * TEAM_MEMBER_WORDS;
    while (true)
    {
        team_sync(set, team); // A
        const uint32_t q = box->q, nw32 = box->nw32, gathered = box->gathered, cmd = box->pad;
        if (q == NO_BRANCH) { break; }
        if (cmd & TEAM_CMD_GATHER)
        {
            uint32_t sink = 0;
            range_gather<U, CT, SINK>(p.db, acc, hits, cmd & ~TEAM_CMD_GATHER, reinterpret_cast<uint2*>(member_base), member, team,
                                      lane, p.prefetch, &sink);
            if (SINK)
            {
                sink = __reduce_xor_sync(FULL, sink);
                if (lane == 0) { atomicXor(p.n_place + q, sink); } // the ceiling launch: one checksum word per query
            }
            team_sync(set, team); // G: this chunk's postings are in the accumulators
            continue;
        }
        if (cmd & TEAM_CMD_DUMP)
        {
            dump_partial_row<CT>(p, S, C, np, q, nw32, gathered, lane, member_base, member, team,
                                 reinterpret_cast<uint32_t*>(const_cast<uint64_t*>(hits)), set);
            team_sync(set, team); // the row is out, the accumulators are clear
            continue;
        }
        const sweep_result sw = team_sweep<CT>(p.db, S, C, np, nw32, gathered, p.keep_at_most, lane, member_base, member, team);
        if (lane == 0)
        {
            results[member] = team_result{ static_cast<uint32_t>(__double2loint(sw.mass)),
                                           static_cast<uint32_t>(__double2hiint(sw.mass)), sw.touched, sw.n_keep };
        }
        __syncwarp();
        team_sync(set, team); // B
    }
}

template<class CT>
__device__ __forceinline__ void finish_query_leader(const place_params& p, float* S, CT* C, uint32_t np, uint32_t nw32,
                                                    uint32_t gathered, uint32_t q, uint32_t lane, unsigned char* team_area,
                                                    uint32_t set, uint32_t team)
{
    const uint32_t K = p.keep_at_most;
    team_mailbox* box = reinterpret_cast<team_mailbox*>(team_area);
    const team_result* results = reinterpret_cast<const team_result*>(team_area + sizeof(team_mailbox));
    uint32_t* bases = reinterpret_cast<uint32_t*>(team_area + sizeof(team_mailbox) + team * sizeof(team_result));
    if (lane == 0) { *box = team_mailbox{ q, nw32, gathered, 0u }; }
    __syncwarp();
    team_sync(set, team); // A: the accumulators are complete
    sweep_result sw = team_sweep<CT>(p.db, S, C, np, nw32, gathered, K, lane, bases, 0u, team);
    __syncwarp();
    team_sync(set, team); // B: every member's share is in its buffer and its result slot
    const cand_buffers<TEAM_CAND_CAP> mine(bases);
    for (uint32_t m = 1; m < team; ++m)
    {
        const team_result r = results[m];
        sw.mass += __hiloint2double(static_cast<int>(r.mass_hi), static_cast<int>(r.mass_lo));
        sw.touched += r.touched;
        if (r.n_keep == 0) { continue; }
        const cand_buffers<TEAM_CAND_CAP> theirs(bases + m * TEAM_MEMBER_WORDS);
        if (lane < r.n_keep)
        {
            mine.key[sw.n_keep + lane] = theirs.key[lane];
            mine.branch[sw.n_keep + lane] = theirs.branch[lane];
            mine.count[sw.n_keep + lane] = theirs.count[lane];
        }
        __syncwarp();
        sw.n_keep = select_best(mine.key, mine.branch, mine.count, sw.n_keep + r.n_keep, K, lane, sw.key, sw.branch, sw.count);
    }
    emit_query(p, sw, nw32, q, lane);
}

// ---- range teams: every member gathers its own branch range ------------------------------------------
//
// With branch-sorted lists behind range headers (db_build.cuh) the gather splits too: the leader
// publishes the packed slots of a chunk's hit windows, in window order, and member m of T turns every hit
// into descriptors of ITS sub-list - the postings with branch in [m * np / T, (m + 1) * np / T), found
// through the list's header - and applies them to the shared accumulators. The ranges are disjoint, so
// members never touch the same S[b], and each member walks the hits in window order: every S[b] still
// receives the additions of place.cpp:358-367 in the reference's order. A sub-list may start on an odd
// posting: its first descriptor then carries the skip flag (bit 31 of the count word).
constexpr uint32_t RANGE_DESC_CAP = TEAM_MEMBER_WORDS / 2 - SEG_PAD; // descriptors a member buffers (aliases its candidates)
constexpr uint32_t RANGE_HEADER_POSTINGS = 4;                        // = RANGE_HEADER_ENTRIES of db_build.cuh

template<int U, class CT, bool SINK>
__device__ __noinline__ void range_gather(const device_db& db, acc_ref acc, const uint64_t* hits, uint32_t n_hits,
                                          uint2* my_segs, uint32_t member, uint32_t team, uint32_t lane, uint32_t pf, uint32_t* sink_out)
{
    const uint32_t segs_shared = static_cast<uint32_t>(__cvta_generic_to_shared(my_segs));
    const uint32_t part_lo = member * (8u / team); // header entries are eighths of the padded branch range
    const uint32_t part_hi = (member + 1u) * (8u / team);
    uint32_t n_desc = 0;
    for (uint32_t h0 = 0; h0 < n_hits; h0 += 32)
    {
        uint32_t pair0 = 0, count = 0, first_skip = 0;
        if (h0 + lane < n_hits)
        {
            const uint64_t slot = hits[h0 + lane];
            const uint64_t off = slot >> LEN_BITS;
            const uint32_t* header = reinterpret_cast<const uint32_t*>(db.postings + (off - RANGE_HEADER_POSTINGS));
            const uint32_t lo = part_lo == 0 ? 0u : __ldg(header + part_lo - 1);
            const uint32_t hi = __ldg(header + part_hi - 1);
            if (hi > lo)
            {
                first_skip = lo & 1u;
                pair0 = static_cast<uint32_t>((off + (lo - first_skip)) >> 1);
                count = hi - (lo - first_skip);
            }
        }
        const uint32_t my_total = (count + SEG_LEN - 1) / SEG_LEN;
        uint32_t emitted = 0; // of my_total
        while (true)
        {
            const uint32_t remaining = my_total - emitted;
            const uint32_t incl = warp_inclusive_sum(remaining, lane);
            const uint32_t total = __shfl_sync(FULL, incl, 31);
            if (total == 0) { break; }
            const uint32_t space = RANGE_DESC_CAP - n_desc;
            const uint32_t excl = incl - remaining;
            const uint32_t can = excl >= space ? 0u : min(remaining, space - excl);
            for (uint32_t i = 0; i < can; ++i)
            {
                const uint32_t sidx = emitted + i;
                const uint32_t pos = sidx * SEG_LEN;
                my_segs[n_desc + excl + i] = make_uint2(pair0 + pos / 2, min(count - pos, static_cast<uint32_t>(SEG_LEN))
                                                                             | ((sidx == 0 && first_skip) ? 0x80000000u : 0u));
            }
            emitted += can;
            n_desc += min(total, space);
            if (total >= space)
            {
                if (lane < SEG_PAD) { my_segs[n_desc + lane] = make_uint2(0, 0); }
                __syncwarp();
                gather_segments<U, CT, SINK, true>(db.postings, acc, segs_shared, n_desc, lane, pf, sink_out);
                n_desc = 0;
                __syncwarp();
            }
        }
    }
    if (n_desc)
    {
        if (lane < SEG_PAD) { my_segs[n_desc + lane] = make_uint2(0, 0); }
        __syncwarp();
        gather_segments<U, CT, SINK, true>(db.postings, acc, segs_shared, n_desc, lane, pf, sink_out);
    }
    __syncwarp();
}

// ---- partial rows (database-sharded mode) ---------------------------------------------------------
//
// When the database is split by key over the GPUs of a group (second row), every GPU
// runs the encode / probe / gather phases of EVERY query against its shard and leaves, per query, a COMPACT row
//   { uint32 flags, postings gathered, touched branches, 0
//   | bitmap of the touched branches, np / 8 bytes
//   | touched branches before each 128-branch step, np / 128 uint32
//   | float S of the touched branches, in branch order | uint16 C of the touched branches, in branch order }
// in the memory of the GPU that finalises that query (its own, a send buffer, or a peer's slab mapped over NVLink) - in
// a slot of fixed size (partial_row_bytes: room for every branch), of which only the used prefix of each part is
// written and read: a shard of G touches about 1 / G of what the query touches, so the bytes that cross NVLink and
// that the owner reads scale with the touched branches, not with G x N (round 1 wrote dense 6-byte-per-branch rows:
// 105 KB per query over NVLink at G = 8 on a 20 k-branch tree, 1.46 M reads/s; gpurun r3o). The owner walks the
// branches in order, finds a branch's values through the bitmap and the step counts, adds the shards in shard order
// and runs finish_query on the sums. The additions of one branch are no longer in window order: scores agree with the
// reference to float rounding of a differently associated sum (asserted at 1e-5 relative), not bit for bit.
constexpr uint32_t ROW_FLAG_AMBIGUOUS = 1u; // the query has windows with one ambiguous letter: the first-key-wins rule of
                                            // place.cpp:391 needs the global key order, which a shard does not have ->
                                            // replayed by the owner (replay_ambiguous), or reported unsupported
__host__ __device__ inline uint64_t partial_row_bytes(uint32_t n_nodes)
{
    const uint32_t np = padded_nodes(n_nodes);
    return (row_c_offset(np) + 2ull * np + 15ull) & ~15ull;
}

__device__ __forceinline__ void store_counts4_u16(unsigned char* at, const uint32_t (&c)[4])
{
    *reinterpret_cast<uint2*>(at) = make_uint2(c[0] | (c[1] << 16), c[2] | (c[3] << 16));
}

// The dense form of a row { header | float S[np] | uint16 C[np] }: what a shard writes when it touches a large part of the
// tree anyway (two shards: half of what the query touches) - a straight copy, no counting, no compaction.
// S, C of query q -> its row at the owner; leaves both arrays zeroed
// (a team dumps a row together: member `part` of `parts` takes every parts-th 128-branch step)
template<class CT>
__device__ __forceinline__ void dump_dense_row(const place_params& p, float* S, CT* C, uint32_t np, uint32_t q,
                                                 uint32_t flags, uint32_t gathered, uint32_t lane, uint32_t part,
                                                 uint32_t parts)
{
    uint32_t g = 0;
    while (g + 1 < p.n_owners && q >= p.owner_first[g + 1]) { ++g; }
    unsigned char* row = p.rows_out[g] + static_cast<uint64_t>(q - p.owner_first[g]) * partial_row_bytes(p.db.n_nodes);
    if (lane == 0 && part == 0) { *reinterpret_cast<uint4*>(row) = make_uint4(flags, gathered, 0u, 0u); }
    unsigned char* s_out = row + ROW_HEADER_BYTES;
    unsigned char* c_out = s_out + 4ull * np;
    for (uint32_t j = 4u * lane + part * SWEEP_STEP; j < np; j += parts * SWEEP_STEP)
    {
        const float4 s4 = *reinterpret_cast<const float4*>(S + j);
        uint32_t c[4];
        load_counts4(C, j, c);
        *reinterpret_cast<float4*>(S + j) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        clear_counts4(C, j);
        *reinterpret_cast<float4*>(s_out + 4ull * j) = s4;
        store_counts4_u16(c_out + 2ull * j, c);
    }
    __syncwarp();
}

// S, C of query q -> its row at the owner; leaves both arrays zeroed.
// `stage`: 192 words of shared memory of this warp (128 floats + 128 uint16).
// A team dumps a row together: member `part` of `parts` takes every parts-th 128-branch step; the position of a step's
// values in the row depends on all the steps before it, so the members first count (step_counts: shared memory of
// the set, np / 128 words), the leader turns the counts into positions, then everybody writes.
template<class CT>
__device__ __forceinline__ void dump_partial_row(const place_params& p, float* S, CT* C, uint32_t np, uint32_t q,
                                                 uint32_t flags, uint32_t gathered, uint32_t lane, uint32_t* stage, uint32_t part,
                                                 uint32_t parts, uint32_t* step_counts, uint32_t set)
{
    if (p.row_format == ROW_DENSE)
    {
        dump_dense_row<CT>(p, S, C, np, q, flags, gathered, lane, part, parts);
        return;
    }
    uint32_t g = 0;
    while (g + 1 < p.n_owners && q >= p.owner_first[g + 1]) { ++g; }
    unsigned char* row = p.rows_out[g] + static_cast<uint64_t>(q - p.owner_first[g]) * partial_row_bytes(p.db.n_nodes);
    uint32_t* bitmap = reinterpret_cast<uint32_t*>(row + row_bitmap_offset());
    uint32_t* steps = reinterpret_cast<uint32_t*>(row + row_steps_offset(np));
    float* s_out = reinterpret_cast<float*>(row + row_s_offset(np));
    uint16_t* c_out = reinterpret_cast<uint16_t*>(row + row_c_offset(np));
    const uint32_t n_steps = np / SWEEP_STEP;
    uint32_t n_touched = 0;
    if (parts > 1)
    {
        for (uint32_t t = part; t < n_steps; t += parts)
        {
            uint32_t c[4];
            load_counts4(C, t * SWEEP_STEP + 4u * lane, c);
            const uint32_t mine = (c[0] != 0) + (c[1] != 0) + (c[2] != 0) + (c[3] != 0);
            const uint32_t total = __reduce_add_sync(FULL, mine);
            if (lane == 0) { step_counts[t] = total; }
        }
        __syncwarp();
        team_sync(set, parts); // every step is counted
        if (part == 0)
        {
            uint32_t running = 0;
            for (uint32_t t0 = 0; t0 < n_steps; t0 += 32)
            {
                const uint32_t v = t0 + lane < n_steps ? step_counts[t0 + lane] : 0u;
                const uint32_t incl = warp_inclusive_sum(v, lane);
                if (t0 + lane < n_steps) { step_counts[t0 + lane] = running + incl - v; }
                running += __shfl_sync(FULL, incl, 31);
            }
            n_touched = running;
            __syncwarp();
        }
        team_sync(set, parts); // the positions are there
    }
    uint32_t base = 0;
    for (uint32_t t = part; t < n_steps; t += parts)
    {
        const uint32_t j = t * SWEEP_STEP + 4u * lane;
        const float4 s4 = *reinterpret_cast<const float4*>(S + j);
        uint32_t c[4];
        load_counts4(C, j, c);
        *reinterpret_cast<float4*>(S + j) = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        clear_counts4(C, j);
        const uint32_t nib = (c[0] != 0 ? 1u : 0u) | (c[1] != 0 ? 2u : 0u) | (c[2] != 0 ? 4u : 0u) | (c[3] != 0 ? 8u : 0u);
        const uint32_t mine = __popc(nib);
        const uint32_t incl = warp_inclusive_sum(mine, lane);
        if (parts > 1) { base = step_counts[t]; }
        // the step's values are compacted in shared memory first and leave as runs of consecutive words: a lane writing
        // its own one to four values straight to the row made 8 partly filled store instructions per step - over NVLink
        // that was slower than the dense row (gpurun r3r)
        const uint32_t step_total = __shfl_sync(FULL, incl, 31);
        uint32_t at = incl - mine;
        const float sv[4] = { s4.x, s4.y, s4.z, s4.w };
        float* stage_s = reinterpret_cast<float*>(stage);
        uint16_t* stage_c = reinterpret_cast<uint16_t*>(stage + SWEEP_STEP);
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            if ((nib >> u) & 1u)
            {
                stage_s[at] = sv[u];
                stage_c[at] = static_cast<uint16_t>(c[u]);
                ++at;
            }
        }
        __syncwarp();
        for (uint32_t i = lane; i < step_total; i += 32)
        {
            s_out[base + i] = stage_s[i];
            c_out[base + i] = stage_c[i];
        }
        __syncwarp();
        // the bitmap word of 8 lanes: their nibbles side by side
        uint32_t word = nib << (4u * (lane & 7u));
        word |= __shfl_xor_sync(FULL, word, 1);
        word |= __shfl_xor_sync(FULL, word, 2);
        word |= __shfl_xor_sync(FULL, word, 4);
        if ((lane & 7u) == 0) { bitmap[t * 4u + (lane >> 3)] = word; }
        if (lane == 0) { steps[t] = base; }
        if (parts == 1) { base += step_total; }
    }
    if (parts == 1) { n_touched = base; }
    if (lane == 0 && part == 0) { *reinterpret_cast<uint4*>(row) = make_uint4(flags, gathered, n_touched, 0u); }
    __syncwarp();
}

// the segment descriptors of one lane's hit windows, in window order, from `at` on
__device__ __forceinline__ void emit_lane_descriptors(uint2* at, const uint64_t (&slot)[WPL])
{
#pragma unroll
    for (int r = 0; r < WPL; ++r)
    {
        const uint32_t l = static_cast<uint32_t>(slot[r] & LEN_MASK);
        const uint32_t pair = static_cast<uint32_t>(slot[r] >> (LEN_BITS + 1));
        for (uint32_t pos = 0; pos < l; pos += SEG_LEN) { *at++ = make_uint2(pair + pos / 2, min(l - pos, static_cast<uint32_t>(SEG_LEN))); }
    }
}

// ---- the kernel ----------------------------------------------------------------------------------

// MODE_PARTIAL: database-sharded mode, the accumulators of a query leave as a partial row instead of
// being swept (queries with ambiguous windows are flagged, their ambiguous windows left out).
// MODE_GATHER_ONLY: the measuring stick of the roofline - the same encode / probe / segment / gather
// instruction stream with the read-modify-write of the accumulators replaced by an XOR into registers
// and no sweep: what the memory system delivers for this access pattern (one word per query comes out).
enum kernel_mode { MODE_FULL = 0, MODE_PARTIAL = 1, MODE_GATHER_ONLY = 2 };

// TEAM: p.team warps per accumulator set (see "teams"); blockDim = 32 * p.team * sets.
// MAXW: warps per CTA the launch may use (small trees: 16 accumulator sets fit an SM, and 16 warps at 128 registers
// hide the latency of the key-table probes better than 8 at 255).
// GLOBAL: the accumulators live in p.acc_scratch ("accumulators in HBM"; CT = uint32_t, one warp per query).
template<uint32_t TABLE, int U, class CT, int MODE = MODE_FULL, bool TEAM = false, int MAXW = MAX_WARPS_PER_CTA, bool GLOBAL = false>
__global__ void __launch_bounds__(MAXW * 32, 1) place_kernel(const place_params p)
{
    static_assert(!GLOBAL || (!TEAM && MODE == MODE_FULL && sizeof(CT) == 4), "accumulators in HBM: the plain fused launch, 32-bit counts");
    constexpr bool PARTIAL = MODE == MODE_PARTIAL;
    constexpr bool GATHER_ONLY = MODE == MODE_GATHER_ONLY;
    extern __shared__ __align__(16) unsigned char smem[];
    if (p.class_reads && p.class_reads[p.cls] == 0) { return; } // no query of this class in the batch
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t team = TEAM ? p.team : 1u;
    const uint32_t warp = (threadIdx.x >> 5) / team;   // accumulator set
    const uint32_t member = (threadIdx.x >> 5) % team; // 0: the leader
    const device_db& db = p.db;
    const uint32_t N = db.n_nodes;
    const uint32_t np = padded_nodes(N);
    const uint32_t k = db.k;
    const uint32_t bits = db.bits;
    static_assert(U <= SEG_PAD, "zero padding of the segment buffer covers one batch of loads");

    uint32_t* letters = reinterpret_cast<uint32_t*>(smem);
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) { letters[i] = p.letters[i]; }

    // per set: float S[np] | CT C[np] | seen bitmap np/8 bytes | segment descriptors | letter codes | team area
    // (GLOBAL: the first three in this warp's row of p.acc_scratch, zero from the start and after every query)
    unsigned char* base = GLOBAL ? p.acc_scratch + (static_cast<size_t>(blockIdx.x) * (blockDim.x >> 5) + warp) * global_acc_bytes(N)
                                 : smem + CTA_SMEM_FIXED + warp * (warp_smem_bytes(N, sizeof(CT)) + team_smem_bytes(team));
    unsigned char* team_area = base + warp_smem_bytes(N, sizeof(CT));
    float* S = reinterpret_cast<float*>(base);
    CT* C = reinterpret_cast<CT*>(base + static_cast<size_t>(np) * 4);
    uint32_t* seen = reinterpret_cast<uint32_t*>(base + static_cast<size_t>(np) * (4 + sizeof(CT)));
    const uint32_t seen_words = np / 32;
    uint2* segs = GLOBAL ? reinterpret_cast<uint2*>(smem + CTA_SMEM_FIXED + warp * GLOBAL_WARP_SMEM)
                         : reinterpret_cast<uint2*>(reinterpret_cast<unsigned char*>(seen) + np / 8);
    const uint32_t segs_shared = static_cast<uint32_t>(__cvta_generic_to_shared(segs));
    const acc_ref acc{ static_cast<uint32_t>(__cvta_generic_to_shared(S)), static_cast<uint32_t>(__cvta_generic_to_shared(C)) };
    uint32_t* cand_base = reinterpret_cast<uint32_t*>(segs); // the segment buffer is idle during the sweep
    uint8_t* codes = reinterpret_cast<uint8_t*>(segs + SEG_CAP + SEG_PAD);

    if (member == 0 && !GLOBAL)
    {
        for (uint32_t b = lane; b < np; b += 32)
        {
            S[b] = 0.0f;
            C[b] = 0;
        }
        for (uint32_t i = lane; i < seen_words; i += 32) { seen[i] = 0; }
    }
    __syncthreads();
    if (TEAM && member != 0)
    {
        team_helper_loop<U, CT, GATHER_ONLY>(p, S, C, np, lane, team_area, reinterpret_cast<const uint64_t*>(segs), warp, member, team);
        return;
    }

    const float kf = static_cast<float>(k);
    const uint32_t key_mask = db.key_mask;
    uint64_t ctr_windows = 0, ctr_hits = 0, ctr_entries = 0; // per lane, folded once at the end
    EPK_PHASE_DECL

    // Looking one query ahead: a query used to start with two dependent round trips to memory - the cursor, then its
    // two offsets - before a single letter could be asked for (1.7 k of the 65 k cycles of a C2 query; phase clocks,
    // gpurun r3e). Now the cursor is bumped for the NEXT query at the top of this one (A), and the next query's offsets
    // are asked for before this one's last gather (B): both are there when the sweep is over.
    // (Carrying more than that across the sweep - the next query's letters, its key-table slots for an L2 prefetch
    // of its lists - cost more in spilled registers than the latencies it hid: 28 instead of 38 M reads/s, gpurun r3g.)
    // (In the 16-warp, 128-register launch of small trees too: 8.12 against 8.08 M reads/s on C1p, 75.0 against 74.4 M on C4; gpurun r4b.)
    constexpr bool AHEAD = MAXW <= MAX_WARPS_PER_CTA || EPK_WIDE_AHEAD;
    uint32_t q_next = 0;
    if (AHEAD && lane == 0) { q_next = static_cast<uint32_t>(atomicAdd(p.counters + CTR_CURSOR, 1ull)); }
    q_next = __shfl_sync(FULL, q_next, 0);
    uint64_t next_s0 = 0, next_len = 0; // offsets of query q_next ...
    bool next_known = false;            // ... when they have been asked for

    while (true)
    {
        if (!AHEAD)
        {
            if (lane == 0) { q_next = static_cast<uint32_t>(atomicAdd(p.counters + CTR_CURSOR, 1ull)); }
            q_next = __shfl_sync(FULL, q_next, 0);
        }
        const uint32_t q = q_next;
        if (q >= p.n)
        {
            if (TEAM)
            {
                if (lane == 0) { reinterpret_cast<team_mailbox*>(team_area)->q = NO_BRANCH; }
                __syncwarp();
                team_sync(warp, team); // A: tells the members to leave
            }
            break;
        }
        uint32_t claim = 0; // A (read at B, or where this query is left early)
        if (AHEAD && lane == 0) { claim = static_cast<uint32_t>(atomicAdd(p.counters + CTR_CURSOR, 1ull)); }
        const uint64_t s0 = AHEAD && next_known ? next_s0 : p.seq_off[q];
        const uint64_t len = AHEAD && next_known ? next_len : p.seq_off[q + 1] - s0;
        next_known = false;
        if (len + 1 < k)
        {
            // the reference dies here (size_t wrap at place.cpp:291); reported as EPK_E_SHORT_SEQ
            if (lane == 0 && p.cls == 0 && !PARTIAL) // (sharded mode: the finalising launch reports it)
            {
                bump(p, CTR_SHORT, 1ull);
                p.n_place[q] = 0;
            }
            q_next = __shfl_sync(FULL, claim, 0);
            continue;
        }
        const uint64_t entries_before = ctr_entries;
        const uint64_t n_windows = len + 1 - k; // place.cpp:322, counts skipped windows too
        if (class_of_windows(n_windows) != static_cast<int>(p.cls)) // another launch's query
        {
            q_next = __shfl_sync(FULL, claim, 0);
            continue;
        }
        if (n_windows > class_max_windows(N_CLASSES - 1))
        {
            if (lane == 0 && !PARTIAL)
            {
                bump(p, CTR_TOO_LONG, 1ull);
                p.n_place[q] = 0;
            }
            q_next = __shfl_sync(FULL, claim, 0);
            continue;
        }
        const char* seq = p.seq + s0;
        EPK_PHASE(0)
        EPK_PHASE_COUNT(7)
        bool has_ambiguous = false;
        uint32_t nseg = 0, sink = 0;
        uint32_t tr_keys = 0, tr_hits = 0, tr_entries = 0, tr_sum = 0; // this lane's share of the query's trace record
        if (GATHER_ONLY && TEAM && lane == 0) { p.n_place[q] = 0; } // every member XORs its part in (after a team barrier)
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

            uint64_t slot[WPL];
#pragma unroll
            for (int r = 0; r < WPL; ++r) { slot[r] = 0; }
            // lane L owns windows [L*wpl, L*wpl + wpl): short queries spread over all the lanes
            const uint32_t wpl = min(static_cast<uint32_t>(WPL), (nw + 31u) / 32u);
            const uint32_t w0 = lane * wpl;
            uint32_t keys[WPL];
            uint32_t look = 0; // bit r: window w0 + r is an exact k-mer to look up
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
                    keys[r] = 0;
                    if (static_cast<uint32_t>(r) < wpl && w < nw)
                    {
                        const uint32_t l = codes[w + k - 1];
                        const int at = static_cast<int>(w + k - 1);
                        key = ((key << bits) | (l & 31u)) & key_mask;
                        keys[r] = key;
                        last_gap = (l & LETTER_GAP) ? at : last_gap;
                        prev_amb = (l & LETTER_AMBIGUOUS) ? last_amb : prev_amb;
                        last_amb = (l & LETTER_AMBIGUOUS) ? at : last_amb;
                        if (last_gap < static_cast<int>(w))
                        {
                            if (last_amb < static_cast<int>(w)) { look |= 1u << r; }
                            else if (prev_amb < static_cast<int>(w)) { has_ambiguous = true; }
                        }
                    }
                }
            }
            // all the look-ups of the lane go out before any of them is waited for
            if (TABLE == TABLE_DIRECT)
            {
#pragma unroll
                for (int r = 0; r < WPL; ++r)
                {
                    if ((look >> r) & 1u) { slot[r] = __ldg(db.direct + keys[r]); }
                }
            }
            else
            {
                uint4 first[WPL];
#pragma unroll
                for (int r = 0; r < WPL; ++r)
                {
                    first[r] = make_uint4(0u, HASH_EMPTY, 0u, 0u);
                    if ((look >> r) & 1u) { first[r] = __ldg(db.hashed + hash_slot(keys[r], db.hash_shift)); }
                }
#pragma unroll
                for (int r = 0; r < WPL; ++r)
                {
                    if ((look >> r) & 1u) { slot[r] = db_search_from(db, keys[r], first[r]); }
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
            EPK_PHASE(1)
            if (p.trace)
            {
#pragma unroll
                for (int r = 0; r < WPL; ++r)
                {
                    if (!((look >> r) & 1u)) { continue; }
                    const uint32_t l = static_cast<uint32_t>(slot[r] & LEN_MASK);
                    ++tr_keys;
                    tr_hits += l != 0;
                    tr_entries += l;
                    tr_sum += trace_term(static_cast<uint32_t>(c0) + w0 + r, keys[r], 1u, l);
                }
            }
            if (TEAM && db.range_headers)
            {
                // range teams: publish the hits in window order, every member gathers its branch range of them
                uint32_t my_hits = 0;
#pragma unroll
                for (int r = 0; r < WPL; ++r) { my_hits += (slot[r] & LEN_MASK) != 0; }
                const uint32_t hits_incl = warp_inclusive_sum(my_hits, lane);
                const uint32_t n_hits = __shfl_sync(FULL, hits_incl, 31);
                if (n_hits == 0) { continue; }
                uint64_t* hits = reinterpret_cast<uint64_t*>(segs);
                uint32_t at = hits_incl - my_hits;
#pragma unroll
                for (int r = 0; r < WPL; ++r)
                {
                    if (slot[r] & LEN_MASK) { hits[at++] = slot[r]; }
                }
                team_mailbox* box = reinterpret_cast<team_mailbox*>(team_area);
                if (lane == 0) { *box = team_mailbox{ q, 0u, 0u, TEAM_CMD_GATHER | n_hits }; }
                __syncwarp();
                team_sync(warp, team); // command: gather
                uint2* my_desc = reinterpret_cast<uint2*>(team_area + sizeof(team_mailbox) + team * sizeof(team_result));
                range_gather<U, CT, GATHER_ONLY>(db, acc, hits, n_hits, my_desc, 0u, team, lane, p.prefetch, &sink);
                team_sync(warp, team); // G
                continue;
            }
            const uint32_t incl = warp_inclusive_sum(my_segs, lane);
            const uint32_t total = __shfl_sync(FULL, incl, 31);
            if (total == 0) { continue; }
            // Descriptors are emitted lane by lane (= in window order), as many lanes at a time as the
            // buffer has room for; a full buffer is gathered and the emission goes on. Only a lane whose own
            // windows need more than SEG_CAP descriptors (lists of tens of thousands of postings) is
            // walked list by list.
            const uint32_t excl = incl - my_segs;
            uint32_t done = 0; // lanes [0, done) have been emitted
            while (done < 32)
            {
                const uint32_t start = __shfl_sync(FULL, excl, done);
                const uint32_t space = SEG_CAP - nseg;
                const bool fit = lane >= done && incl - start <= space; // monotone in the lane: a run from `done`
                const uint32_t n_fit = __popc(__ballot_sync(FULL, fit));
                if (n_fit == 0)
                {
                    if (nseg)
                    {
                        if (lane < SEG_PAD) { segs[nseg + lane] = make_uint2(0, 0); }
                        __syncwarp();
                        EPK_PHASE(2)
                        if (GLOBAL) { gather_segments_global<U>(db.postings, S, reinterpret_cast<uint32_t*>(C), segs_shared, nseg, lane); }
                        else { gather_segments<U, CT, GATHER_ONLY>(db.postings, acc, segs_shared, nseg, lane, p.prefetch, &sink); }
                        EPK_PHASE(3)
                        nseg = 0;
                        __syncwarp();
                        continue;
                    }
#pragma unroll
                    for (int r = 0; r < WPL; ++r)
                    {
                        const uint64_t s = __shfl_sync(FULL, slot[r], done);
                        const uint32_t l = static_cast<uint32_t>(s & LEN_MASK);
                        if (l && GLOBAL) { apply_list_global(db.postings, S, reinterpret_cast<uint32_t*>(C), s >> LEN_BITS, l, lane); }
                        else if (l) { apply_list<CT>(db.postings, acc, s >> LEN_BITS, l, lane); }
                    }
                    ++done;
                    continue;
                }
                if (fit) { emit_lane_descriptors(segs + (nseg + excl - start), slot); }
                nseg += __shfl_sync(FULL, incl, done + n_fit - 1) - start;
                done += n_fit;
            }
        }
        // B: the next query's index is there; its offsets are asked for
        if (AHEAD) { q_next = __shfl_sync(FULL, claim, 0); }
        uint64_t off0 = 0, off1 = 0;
        if (AHEAD && q_next < p.n)
        {
            off0 = __ldg(p.seq_off + q_next);
            off1 = __ldg(p.seq_off + q_next + 1);
        }
        if (nseg)
        {
            if (lane < SEG_PAD) { segs[nseg + lane] = make_uint2(0, 0); }
            __syncwarp();
            EPK_PHASE(2)
            if (GLOBAL) { gather_segments_global<U>(db.postings, S, reinterpret_cast<uint32_t*>(C), segs_shared, nseg, lane); }
            else { gather_segments<U, CT, GATHER_ONLY>(db.postings, acc, segs_shared, nseg, lane, p.prefetch, &sink); }
            EPK_PHASE(3)
        }
        __syncwarp();
        if (AHEAD && q_next < p.n)
        {
            next_s0 = off0;
            next_len = off1 - off0;
            next_known = true;
        }
        if (GATHER_ONLY)
        {
            sink = __reduce_xor_sync(FULL, sink);
            if (lane == 0)
            {
                if (TEAM) { atomicXor(p.n_place + q, sink); } // the members have added theirs
                else { p.n_place[q] = sink; }
            }
            continue;
        }

        // ---- windows with exactly one ambiguous letter (place.cpp:306-313, 373-415) ----
        has_ambiguous = __any_sync(FULL, has_ambiguous);
        if (PARTIAL)
        {
            const uint32_t gathered_here = __reduce_add_sync(FULL, static_cast<uint32_t>(min(static_cast<unsigned long long>(ctr_entries - entries_before), 0xFFFFFFull)));
            const uint32_t row_flags = has_ambiguous ? ROW_FLAG_AMBIGUOUS : 0u;
            if (TEAM)
            {
                if (lane == 0) { *reinterpret_cast<team_mailbox*>(team_area) = team_mailbox{ q, row_flags, gathered_here, TEAM_CMD_DUMP }; }
                __syncwarp();
                team_sync(warp, team); // command: dump
                dump_partial_row<CT>(p, S, C, np, q, row_flags, gathered_here, lane,
                                     reinterpret_cast<uint32_t*>(team_area + sizeof(team_mailbox) + team * sizeof(team_result)), 0u, team,
                                     reinterpret_cast<uint32_t*>(segs), warp);
                team_sync(warp, team);
            }
            else { dump_partial_row<CT>(p, S, C, np, q, row_flags, gathered_here, lane, reinterpret_cast<uint32_t*>(segs)); }
            continue;
        }
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
                                if (p.trace)
                                {
                                    ++tr_keys;
                                    tr_hits += flen != 0;
                                    tr_entries += flen;
                                    tr_sum += trace_term(static_cast<uint32_t>(c0) + r * 32u + src, base_key | (code << shift),
                                                         __popc(letter >> 8), flen);
                                }
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
                                    if (GLOBAL) // (through L2 like every other access of a row in HBM: no stale L1 line)
                                    {
                                        __stcg(S + entry.x, __fadd_rn(__ldcg(S + entry.x), mean));
                                        __stcg(C + entry.x, static_cast<CT>(__ldcg(C + entry.x) + 1u));
                                    }
                                    else
                                    {
                                        S[entry.x] = __fadd_rn(S[entry.x], mean);
                                        C[entry.x] = static_cast<CT>(C[entry.x] + 1u);
                                    }
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

        EPK_PHASE(4)
        if (p.trace)
        {
            tr_keys = __reduce_add_sync(FULL, tr_keys);
            tr_hits = __reduce_add_sync(FULL, tr_hits);
            tr_entries = __reduce_add_sync(FULL, tr_entries);
            tr_sum = __reduce_add_sync(FULL, tr_sum);
            if (lane == 0) { *reinterpret_cast<uint4*>(p.trace + 4ull * q) = make_uint4(tr_keys, tr_hits, tr_entries, tr_sum); }
        }
        // ---- correction, likelihood mass, top-k, weight ratios (place.cpp:417-422, 164-184, 134-159, 241-267) ----
        const uint32_t gathered = __reduce_add_sync(FULL, static_cast<uint32_t>(min(static_cast<unsigned long long>(ctr_entries - entries_before), 0xFFFFFFull)));
        if (TEAM) { finish_query_leader<CT>(p, S, C, np, static_cast<uint32_t>(n_windows), gathered, q, lane, team_area, warp, team); }
        else if (GLOBAL) { finish_query(p, global_acc{ S, reinterpret_cast<uint32_t*>(C) }, np, n_windows, gathered, q, lane, cand_base); }
        else
        {
#ifdef EPK_PHASE_CLOCKS
            const cand_buffers<CAND_CAP> cb(cand_base);
            const sweep_result sw = run_sweep<shared_acc<CT>, CAND_CAP>(p.db, shared_acc<CT>(S, C), np, static_cast<uint32_t>(n_windows), gathered, p.keep_at_most, lane, cb, 0u, SWEEP_STEP EPK_PHASE_ARGS);
            EPK_PHASE(10)
            emit_query(p, sw, static_cast<uint32_t>(n_windows), q, lane);
            EPK_PHASE(6)
#else
            finish_query<shared_acc<CT>, (MAXW > MAX_WARPS_PER_CTA ? EPK_SWEEP_NQ_WIDE : EPK_SWEEP_NQ)>(p, shared_acc<CT>(S, C), np, n_windows, gathered, q, lane, cand_base);
#endif
        }
    }
    EPK_PHASE_FLUSH

#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
    {
        ctr_windows += __shfl_xor_sync(FULL, ctr_windows, d);
        ctr_hits += __shfl_xor_sync(FULL, ctr_hits, d);
        ctr_entries += __shfl_xor_sync(FULL, ctr_entries, d);
    }
    if (lane == 0)
    {
        bump(p, CTR_WINDOWS, static_cast<unsigned long long>(ctr_windows));
        bump(p, CTR_HITS, static_cast<unsigned long long>(ctr_hits));
        bump(p, CTR_ENTRIES, static_cast<unsigned long long>(ctr_entries));
    }
}
// Database-sharded mode, the owner's side: the sweep of finish_query runs straight over the partial rows
// of the n_shards shards (rows_acc adds them in shard order as it reads: a fixed association, the same
// on every run). One warp per query; shared memory only holds the candidate buffers, so the launch is
// not limited by the size of the tree.
constexpr size_t FINALIZE_WARP_SMEM = cand_buffers<CAND_CAP>::WORDS * 4 + CODE_BYTES; // candidates, letter codes of a chunk
constexpr size_t FINALIZE_CTA_SMEM = 1024;                                               // letter table
constexpr int FINALIZE_WARPS = 8;

// phylo_kmer_db::search in a shard seen through its view (the table kind is the view's: a run-time branch)
__device__ __forceinline__ uint64_t view_search(const device_db& view, uint32_t key)
{
    return view.table_kind == TABLE_DIRECT ? db_search<TABLE_DIRECT>(view, key) : db_search<TABLE_HASH>(view, key);
}

// place.cpp:373-415 for one query of database-sharded mode, by the warp that finalises it: every key an ambiguous window
// resolves to lives in exactly one shard, so walking the windows in order and looking each key up in its shard keeps
// the reference's global "first ambiguous key to name a branch" order. The contributions go to a dense scratch row of
// the warp (S float, C uint16, zero between queries), which the sweep adds to the shards' sums and clears.
__device__ __noinline__ void replay_ambiguous(const place_params& p, const char* seq, uint64_t n_windows, float* S, uint16_t* C,
                                              uint32_t np, uint32_t* seen, uint8_t* codes, const uint32_t* letters,
                                              uint32_t lane, uint64_t& hits, uint64_t& entries)
{
    const device_db& db = p.db;
    const uint32_t k = db.k, bits = db.bits;
    const float kf = static_cast<float>(k);
    const float thr_rest = __fmul_rn(static_cast<float>(k - 1), db.threshold);
    for (uint64_t c0 = 0; c0 < n_windows; c0 += CHUNK)
    {
        const uint32_t nw = static_cast<uint32_t>(min(n_windows - c0, static_cast<uint64_t>(CHUNK)));
        const uint32_t ncode = nw + k - 1;
        __syncwarp();
        for (uint32_t i = lane; i < ncode; i += 32) { codes[i] = static_cast<uint8_t>(letters[static_cast<uint8_t>(seq[c0 + i])]); }
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
                while (resolutions) // one key per resolution, each on its own (place.cpp:308-312)
                {
                    const uint32_t code = __ffs(resolutions) - 1;
                    resolutions &= resolutions - 1;
                    const uint32_t key = base_key | (code << shift);
                    const device_db& view = p.shard_views[shard_of_key(key, p.n_shards)];
                    const uint64_t found = view_search(view, key);
                    const uint32_t flen = static_cast<uint32_t>(found & LEN_MASK);
                    const uint2* list = view.postings + (found >> LEN_BITS);
                    hits += flen != 0;
                    entries += flen;
                    for (uint32_t e = lane; e < flen; e += 32)
                    {
                        const uint2 entry = __ldg(list + e);
                        const uint32_t bit = 1u << (entry.x & 31u);
                        const uint32_t old = atomicOr(seen + (entry.x >> 5), bit);
                        if (!(old & bit))
                        {
                            // first ambiguous key that names this branch: (10^s + (k-1)·thr)/k is added to the LOG
                            // score (place.cpp:391,400-410)
                            const float prob = static_cast<float>(pow(10.0, static_cast<double>(__uint_as_float(entry.y))));
                            const float mean = __fdiv_rn(__fadd_rn(prob, thr_rest), kf);
                            // (a branch gets ONE such contribution per query - the seen bit - so this is a plain store
                            // into the warp's zeroed scratch, which the sweep adds to the shards' sums and clears)
                            __stcg(S + entry.x, mean);
                            __stcg(C + entry.x, static_cast<uint16_t>(1u));
                        }
                    }
                    __syncwarp();
                }
            }
        }
    }
    for (uint32_t i = lane; i < np / 32; i += 32) { seen[i] = 0; }
    __threadfence_block();
    __syncwarp();
}

__global__ void __launch_bounds__(FINALIZE_WARPS * 32, 2) finalize_kernel(const place_params p)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const device_db& db = p.db;
    const uint32_t N = db.n_nodes;
    const uint32_t np = padded_nodes(N);
    const uint32_t k = db.k;
    uint32_t* letters = reinterpret_cast<uint32_t*>(smem);
    const bool can_replay = p.shard_views != nullptr && p.seq != nullptr;
    if (can_replay)
    {
        for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) { letters[i] = p.letters[i]; }
        __syncthreads();
    }
    uint32_t* cand_base = reinterpret_cast<uint32_t*>(smem + FINALIZE_CTA_SMEM + warp * FINALIZE_WARP_SMEM);
    uint8_t* codes = reinterpret_cast<uint8_t*>(cand_base + cand_buffers<CAND_CAP>::WORDS);
    uint32_t* seen = p.seen_scratch ? p.seen_scratch + (static_cast<size_t>(blockIdx.x) * FINALIZE_WARPS + warp) * (np / 32) : nullptr;
    const uint64_t row_bytes = partial_row_bytes(N);

    while (true)
    {
        uint32_t q = 0;
        if (lane == 0) { q = static_cast<uint32_t>(atomicAdd(p.counters + CTR_CURSOR, 1ull)); }
        q = __shfl_sync(FULL, q, 0);
        if (q >= p.n) { break; }
        const uint64_t s0 = p.seq_off[q];
        const uint64_t len = p.seq_off[q + 1] - s0;
        const unsigned char* row0 = p.rows_in + static_cast<uint64_t>(q) * row_bytes;
        uint32_t flags = 0, gathered = 0;
        if (lane < p.n_shards)
        {
            const uint4 h = *reinterpret_cast<const uint4*>(row0 + lane * p.shard_stride);
            flags = h.x;
            gathered = h.y;
        }
        flags = __reduce_or_sync(FULL, flags);
        gathered = __reduce_add_sync(FULL, gathered);
        const bool too_short = len + 1 < k;
        const uint64_t n_windows = too_short ? 0 : len + 1 - k;
        const bool ambiguous = (flags & ROW_FLAG_AMBIGUOUS) != 0;
        if (too_short || n_windows > class_max_windows(1) || (ambiguous && !can_replay))
        {
            if (lane == 0)
            {
                atomicAdd(p.counters + (too_short ? CTR_SHORT : CTR_UNSUPPORTED), 1ull);
                p.n_place[q] = 0;
            }
            continue;
        }
        uint64_t amb_hits = 0, amb_entries = 0;
        float* amb_s = nullptr;
        uint16_t* amb_c = nullptr;
        if (ambiguous)
        {
            unsigned char* mine = p.amb_scratch + (static_cast<size_t>(blockIdx.x) * FINALIZE_WARPS + warp) * (6ull * np);
            amb_s = reinterpret_cast<float*>(mine);
            amb_c = reinterpret_cast<uint16_t*>(mine + 4ull * np);
            replay_ambiguous(p, p.seq + s0, n_windows, amb_s, amb_c, np, seen, codes, letters, lane, amb_hits, amb_entries);
            gathered = static_cast<uint32_t>(min(static_cast<unsigned long long>(gathered) + amb_entries, 0xFFFFFFull));
        }
        if (lane == 0)
        {
            atomicAdd(p.counters + CTR_WINDOWS, static_cast<unsigned long long>(n_windows));
            atomicAdd(p.counters + CTR_ENTRIES, static_cast<unsigned long long>(gathered));
            if (amb_hits) { atomicAdd(p.counters + CTR_HITS, static_cast<unsigned long long>(amb_hits)); }
        }
        const rows_acc acc{ row0, p.shard_stride, p.n_shards, np, p.row_format, amb_s, amb_c };
        finish_query(p, acc, np, n_windows, gathered, q, lane, cand_base);
    }
}
} // namespace epk
