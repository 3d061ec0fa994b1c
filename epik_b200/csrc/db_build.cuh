// Load-time kernels: the device side of i2l::load(db_file, mu, omega, max_entries)
// (call site epik/src/epik/main.cpp:277) after the file has been parsed into CSR arrays.
//
//   1. informative-k-mer filter: k-mers ordered by decreasing filter value (stable radix sort;
//      with no filter values the given order is the file order), whole k-mers kept while the
//      entries taken so far are below min(mu * total, max_entries); of a kept k-mer only the
//      entries scoring >= log10(score_threshold(user omega, k)) survive.
//      [UNVERIFIED vs upstream i2l — the submodule is absent; SURVEY.md Appendix B]
//   2. layout: surviving lists are packed at 16-byte boundaries (even entry counts) by an
//      exclusive scan + a warp-per-list compaction;
//   3. key table: direct-address slots or open addressing (see device_db.cuh).
//   4. validation: branch ids < n_nodes and distinct within a list (the placement kernel
//      relies on it to add a list without atomics).
#pragma once

#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>
#include "device_db.cuh"

namespace epk
{
// len[i] = offsets[i + 1] - offsets[i]
__global__ void list_lengths_kernel(const uint64_t* offsets, uint64_t n_keys, uint64_t* len)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i < n_keys) { len[i] = offsets[i + 1] - offsets[i]; }
}

__global__ void iota_kernel(uint32_t* idx, uint64_t n)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i < n) { idx[i] = static_cast<uint32_t>(i); }
}

// sorted_len[j] = len[order[j]]
__global__ void gather_lengths_kernel(const uint64_t* len, const uint32_t* order, uint64_t n_keys, uint64_t* sorted_len)
{
    const uint64_t j = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (j < n_keys) { sorted_len[j] = len[order ? order[j] : j]; }
}

// keep[order[j]] = entries taken before k-mer j < budget
__global__ void prefix_cut_kernel(const uint64_t* before, const uint32_t* order, uint64_t n_keys, uint64_t budget,
                                  uint8_t* keep)
{
    const uint64_t j = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (j < n_keys) { keep[order ? order[j] : j] = before[j] < budget ? 1 : 0; }
}


// keep[i] &= key i belongs to shard `index` (keep_in == nullptr: every key was kept so far)
__global__ void shard_keep_kernel(const uint32_t* keys, uint64_t n_keys, uint32_t index, uint32_t n_shards,
                                  const uint8_t* keep_in, uint8_t* keep)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i < n_keys) { keep[i] = ((keep_in ? keep_in[i] : 1) && shard_of_key(keys[i], n_shards) == index) ? 1 : 0; }
}

// One warp per list: surviving entry count (and its even-padded size).
__global__ void surviving_lengths_kernel(const uint64_t* offsets, const uint2* entries, const uint8_t* keep,
                                         uint64_t n_keys, float min_score, bool use_min_score,
                                         uint64_t* new_len, uint64_t* padded_len, uint32_t header_entries)
{
    const uint64_t i = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (i >= n_keys) { return; }
    uint64_t n = 0;
    if (!keep || keep[i])
    {
        const uint64_t b = offsets[i], e = offsets[i + 1];
        if (!use_min_score) { n = e - b; }
        else
        {
            uint32_t mine = 0;
            for (uint64_t j = b + lane; j < e; j += 32)
            {
                mine += __uint_as_float(entries[j].y) >= min_score ? 1u : 0u;
            }
            for (int d = 16; d > 0; d >>= 1) { mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d); }
            n = mine;
        }
    }
    if (lane == 0)
    {
        new_len[i] = n;
        padded_len[i] = n ? ((n + 1) & ~1ull) + header_entries : 0;
    }
}

// One warp per list: copies the surviving entries, in order, to their packed position.
__global__ void pack_lists_kernel(const uint64_t* offsets, const uint2* entries, const uint64_t* new_len,
                                  const uint64_t* new_off, uint64_t n_keys, float min_score, bool use_min_score,
                                  uint2* postings)
{
    const uint64_t i = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (i >= n_keys) { return; }
    const uint64_t n = new_len[i];
    if (n == 0) { return; }
    const uint64_t b = offsets[i], e = offsets[i + 1];
    uint2* dst = postings + new_off[i];
    uint64_t written = 0;
    for (uint64_t j0 = b; j0 < e; j0 += 32)
    {
        const uint64_t j = j0 + lane;
        uint2 v = make_uint2(0, 0);
        bool ok = false;
        if (j < e)
        {
            v = entries[j];
            ok = !use_min_score || __uint_as_float(v.y) >= min_score;
        }
        const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, ok);
        if (ok) { dst[written + __popc(ballot & ((1u << lane) - 1u))] = v; }
        written += __popc(ballot);
    }
    if (lane == 0 && (n & 1)) { dst[n] = make_uint2(0, 0); } // pad entry, never applied
}

// status bits
constexpr uint32_t BUILD_DUPLICATE_KEY = 1, BUILD_KEY_RANGE = 2, BUILD_BAD_BRANCH = 4, BUILD_DUPLICATE_BRANCH = 8,
                   BUILD_LIST_TOO_LONG = 16;

__global__ void fill_direct_kernel(const uint32_t* keys, const uint64_t* new_len, const uint64_t* new_off,
                                   uint64_t n_keys, uint64_t n_slots, unsigned long long* slots, uint32_t* status)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n_keys) { return; }
    const uint64_t n = new_len[i];
    if (keys[i] >= n_slots) { atomicOr(status, BUILD_KEY_RANGE); return; }
    if (n > LEN_MASK) { atomicOr(status, BUILD_LIST_TOO_LONG); return; }
    if (n == 0) { return; }
    const unsigned long long packed = (new_off[i] << LEN_BITS) | n;
    if (atomicExch(slots + keys[i], packed) != 0ull) { atomicOr(status, BUILD_DUPLICATE_KEY); }
}

__global__ void clear_hash_kernel(uint4* slots, uint64_t n_slots)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i < n_slots) { slots[i] = make_uint4(0, HASH_EMPTY, 0, 0); }
}

// Linear probing; a slot is claimed by a compare-and-swap on its length word. Keys are unique, so
// builders never need to compare keys (a half-written slot is simply "taken").
__global__ void fill_hash_kernel(const uint32_t* keys, const uint64_t* new_len, const uint64_t* new_off,
                                 uint64_t n_keys, uint4* slots, uint64_t mask, uint32_t shift, uint32_t* status)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n_keys) { return; }
    const uint64_t n = new_len[i];
    if (n == 0) { return; }
    if (n >= HASH_EMPTY || n > LEN_MASK) { atomicOr(status, BUILD_LIST_TOO_LONG); return; }
    const uint32_t key = keys[i];
    uint64_t h = hash_slot(key, shift);
    while (true)
    {
        uint32_t* words = reinterpret_cast<uint32_t*>(slots + h);
        if (atomicCAS(words + 1, HASH_EMPTY, static_cast<uint32_t>(n)) == HASH_EMPTY)
        {
            words[0] = key;
            words[2] = static_cast<uint32_t>(new_off[i]);
            words[3] = static_cast<uint32_t>(new_off[i] >> 32);
            return;
        }
        h = (h + 1) & mask;
    }
}

// After the hash table is complete: every key must be found exactly where a probe ends
// (detects duplicate keys: the second copy stops at the first).
__global__ void check_hash_kernel(const uint32_t* keys, const uint64_t* new_len, const uint64_t* new_off,
                                  uint64_t n_keys, const uint4* slots, uint64_t mask, uint32_t shift, uint32_t* status)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n_keys || new_len[i] == 0) { return; }
    const uint32_t key = keys[i];
    uint64_t h = hash_slot(key, shift);
    while (true)
    {
        const uint4 s = slots[h];
        if (s.y == HASH_EMPTY) { atomicOr(status, BUILD_DUPLICATE_KEY); return; }
        if (s.x == key)
        {
            const uint64_t off = (static_cast<uint64_t>(s.w) << 32) | s.z;
            if (off != new_off[i]) { atomicOr(status, BUILD_DUPLICATE_KEY); }
            return;
        }
        h = (h + 1) & mask;
    }
}

// One warp per list, a bitmap of n_nodes bits per warp in shared memory.
__global__ void validate_lists_kernel(const uint2* postings, const uint64_t* new_len, const uint64_t* new_off,
                                      uint64_t n_keys, uint32_t n_nodes, uint32_t* status)
{
    extern __shared__ uint32_t bitmap_all[];
    const uint32_t words = (n_nodes + 31) / 32;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps_per_block = blockDim.x >> 5;
    uint32_t* bitmap = bitmap_all + static_cast<size_t>(warp) * words;
    for (uint32_t w = lane; w < words; w += 32) { bitmap[w] = 0; }
    __syncwarp();
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * warps_per_block + warp; i < n_keys;
         i += static_cast<uint64_t>(gridDim.x) * warps_per_block)
    {
        const uint64_t n = new_len[i];
        const uint2* list = postings + new_off[i];
        for (uint64_t j = lane; j < n; j += 32)
        {
            const uint32_t b = list[j].x;
            if (b >= n_nodes) { atomicOr(status, BUILD_BAD_BRANCH); continue; }
            const uint32_t bit = 1u << (b & 31u);
            if (atomicOr(bitmap + (b >> 5), bit) & bit) { atomicOr(status, BUILD_DUPLICATE_BRANCH); }
        }
        __syncwarp();
        for (uint64_t j = lane; j < n; j += 32)
        {
            const uint32_t b = list[j].x;
            if (b < n_nodes) { bitmap[b >> 5] = 0; }
        }
        __syncwarp();
    }
}

// ---- branch-sorted lists with range headers (large trees) -----------------------------------------
//
// For trees so large that few accumulator sets fit an SM, the placement kernel splits a query over the
// warps of a team by BRANCH RANGE (place_kernel.cuh, "range teams"): member m of T applies, of every
// posting list the query hits, only the postings whose branch lies in [m * Np / T, (m + 1) * Np / T).
// For that a list is stored sorted by branch (a list names a branch once, so the order inside a list
// does not change any sum) behind a 32-byte header h[0..7]: h[r] = number of postings with branch
// < (r + 1) * Np / 8 (h[7] = the length), i.e. the split positions for T = 2, 4 or 8.
constexpr uint32_t RANGE_HEADER_ENTRIES = 4; // 32 bytes, in units of postings; the header ends where the list begins
constexpr uint32_t RANGE_PARTS = 8;

__global__ void add_header_room_kernel(uint64_t* new_off, const uint64_t* new_len, uint64_t n_keys)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i < n_keys && new_len[i] != 0) { new_off[i] += RANGE_HEADER_ENTRIES; }
}

// One warp per list: counting sort by branch through a bitmap of np bits (shared memory, per warp) and the
// header. `in` and `out` have the same layout; np = n_nodes rounded up to 128. Lists with an invalid or
// repeated branch are copied as they are (validate_lists_kernel reports them).
__global__ void sort_lists_kernel(const uint2* in, uint2* out, const uint64_t* new_len, const uint64_t* new_off,
                                  uint64_t n_keys, uint32_t n_nodes, uint32_t np)
{
    extern __shared__ uint32_t sort_smem[];
    const uint32_t words = np / 32;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps_per_block = blockDim.x >> 5;
    uint32_t* bitmap = sort_smem + static_cast<size_t>(warp) * 2 * words; // bits | exclusive prefix of the word popcounts
    uint32_t* before = bitmap + words;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * warps_per_block + warp; i < n_keys;
         i += static_cast<uint64_t>(gridDim.x) * warps_per_block)
    {
        const uint64_t n = new_len[i];
        if (n == 0) { continue; }
        const uint2* list = in + new_off[i];
        uint2* dst = out + new_off[i];
        for (uint32_t w = lane; w < words; w += 32) { bitmap[w] = 0; }
        __syncwarp();
        bool bad = false;
        for (uint64_t j = lane; j < n; j += 32)
        {
            const uint32_t b = list[j].x;
            if (b >= n_nodes) { bad = true; continue; }
            const uint32_t bit = 1u << (b & 31u);
            if (atomicOr(bitmap + (b >> 5), bit) & bit) { bad = true; }
        }
        bad = __any_sync(0xFFFFFFFFu, bad);
        __syncwarp();
        uint32_t running = 0;
        for (uint32_t w0 = 0; w0 < words; w0 += 32)
        {
            const uint32_t w = w0 + lane;
            const uint32_t c = w < words ? __popc(bitmap[w]) : 0u;
            uint32_t incl = c;
            for (int d = 1; d < 32; d <<= 1)
            {
                const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                if (lane >= static_cast<uint32_t>(d)) { incl += o; }
            }
            if (w < words) { before[w] = running + incl - c; }
            running += __shfl_sync(0xFFFFFFFFu, incl, 31);
        }
        __syncwarp();
        for (uint64_t j = lane; j < ((n + 1) & ~1ull); j += 32)
        {
            const uint2 v = list[j]; // the pad entry of an odd list travels with it
            uint64_t at = j;
            if (!bad && j < n) { at = before[v.x >> 5] + __popc(bitmap[v.x >> 5] & ((1u << (v.x & 31u)) - 1u)); }
            dst[at] = v;
        }
        if (lane < RANGE_PARTS)
        {
            uint32_t h = static_cast<uint32_t>(n);
            if (lane + 1 < RANGE_PARTS && !bad)
            {
                const uint32_t bound = (lane + 1) * (np / RANGE_PARTS); // a multiple of 16
                h = before[bound >> 5] + __popc(bitmap[bound >> 5] & ((1u << (bound & 31u)) - 1u));
            }
            reinterpret_cast<uint32_t*>(dst - RANGE_HEADER_ENTRIES)[lane] = h;
        }
        __syncwarp();
    }
}

// Range of the loaded scores, as order-preserving keys: stats[0] = min, stats[1] = max,
// stats[2] = smallest non-zero magnitude (raw bits), stats[3] = 1 when a score is not finite.
// stats must start as {0xFFFFFFFF, 0, 0x7F800000, 0}.
__device__ __forceinline__ uint32_t ordered_key(float s)
{
    const uint32_t u = __float_as_uint(s);
    return u ^ (static_cast<uint32_t>(static_cast<int32_t>(u) >> 31) | 0x80000000u);
}
__host__ __device__ inline float ordered_key_value(uint32_t key)
{
    const uint32_t u = (key >> 31) ? (key & 0x7FFFFFFFu) : ~key;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}
__global__ void score_stats_kernel(const uint2* postings, const uint64_t* new_len, const uint64_t* new_off,
                                   uint64_t n_keys, uint32_t* stats)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t warp = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) >> 5;
    const uint64_t n_warps = (static_cast<uint64_t>(gridDim.x) * blockDim.x) >> 5;
    uint32_t lo = 0xFFFFFFFFu, hi = 0u, tiny = 0x7F800000u, bad = 0u;
    for (uint64_t i = warp; i < n_keys; i += n_warps)
    {
        const uint64_t n = new_len[i];
        const uint2* list = postings + new_off[i];
        for (uint64_t j = lane; j < n; j += 32)
        {
            const float s = __uint_as_float(list[j].y);
            const uint32_t mag = list[j].y & 0x7FFFFFFFu;
            if (mag >= 0x7F800000u) { bad = 1u; continue; }
            const uint32_t key = ordered_key(s);
            lo = min(lo, key);
            hi = max(hi, key);
            if (mag != 0u) { tiny = min(tiny, mag); }
        }
    }
    lo = __reduce_min_sync(0xFFFFFFFFu, lo);
    hi = __reduce_max_sync(0xFFFFFFFFu, hi);
    tiny = __reduce_min_sync(0xFFFFFFFFu, tiny);
    bad = __reduce_max_sync(0xFFFFFFFFu, bad);
    if (lane == 0)
    {
        atomicMin(stats + 0, lo);
        atomicMax(stats + 1, hi);
        atomicMin(stats + 2, tiny);
        atomicMax(stats + 3, bad);
    }
}

// totals[0] = sum of new_len, totals[1] = number of non-empty lists
__global__ void count_loaded_kernel(const uint64_t* new_len, uint64_t n_keys, unsigned long long* totals)
{
    uint64_t entries = 0, lists = 0;
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < n_keys;
         i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        entries += new_len[i];
        lists += new_len[i] != 0;
    }
    for (int d = 16; d > 0; d >>= 1)
    {
        entries += __shfl_xor_sync(0xFFFFFFFFu, entries, d);
        lists += __shfl_xor_sync(0xFFFFFFFFu, lists, d);
    }
    if ((threadIdx.x & 31u) == 0)
    {
        atomicAdd(totals, static_cast<unsigned long long>(entries));
        atomicAdd(totals + 1, static_cast<unsigned long long>(lists));
    }
}
} // namespace epk
