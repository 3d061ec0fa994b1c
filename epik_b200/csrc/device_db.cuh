// HBM-resident form of the phylo-k-mer database and the letter tables, shared by every kernel.
//
// What the reference keeps as i2l::phylo_kmer_db — a hash map key -> std::vector<pkdb_value>
// searched at epik/src/epik/place.cpp:300,311 — lives here as
//   * a key table: direct-address (one 8-byte slot per possible code) when the code space is
//     small (alphabet_bits * k <= 26), otherwise open addressing with 16-byte slots, load <= 0.5;
//   * one array of (branch, score) postings, each list starting on a 16-byte boundary (lists are
//     padded to an even number of entries) so a lane can fetch two entries with one 128-bit load
//     and a whole list can be moved by one bulk copy.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace epk
{
constexpr uint32_t TABLE_DIRECT = 0;
constexpr uint32_t TABLE_HASH = 1;

// packed direct slot: offset (in entries) << LEN_BITS | length; length 0 = key absent
constexpr int LEN_BITS = 24;
constexpr uint64_t LEN_MASK = (1ull << LEN_BITS) - 1;
constexpr uint32_t HASH_EMPTY = 0xFFFFFFFFu; // .y (length) of a free hash slot

// letter table entry: bits 0-4 code, bit 5 ambiguous, bit 6 gap/unknown, bits 8-27 resolution set
constexpr uint32_t LETTER_AMBIGUOUS = 1u << 5;
constexpr uint32_t LETTER_GAP = 1u << 6;

struct device_db
{
    const uint64_t* direct;  // [1 << (bits*k)] packed slots            (TABLE_DIRECT)
    const uint4* hashed;     // [hash_mask + 1] {key, len, off_lo, off_hi} (TABLE_HASH)
    const uint2* postings;   // (branch, score bits), lists 16-byte aligned
    uint64_t hash_mask;
    uint32_t hash_shift;     // 32 - log2(slots)
    uint32_t table_kind;
    uint32_t k;              // letters per k-mer
    uint32_t bits;           // bits per letter: 2 (DNA) or 5 (amino acids)
    uint32_t key_mask;
    uint32_t n_nodes;
    uint32_t alphabet;
    float threshold;         // i2l::score_threshold(omega, k), place.cpp:87
    float log_threshold;     // std::log10 of it, place.cpp:88
    float score_lo;          // smallest / largest score of the loaded postings (0 when there are none)
    float score_hi;
    uint32_t range_headers;  // 1: lists are sorted by branch and preceded by a 32-byte header of split positions
                             // (db_build.cuh "branch-sorted lists with range headers"); large trees only
    uint32_t fast_ok;        // 1 when every score is finite, <= 0, and 0 or >= 1e-20 in magnitude:
                             // the short arithmetic forms of the sweep then hold (place_kernel.cuh)
};

__host__ __device__ inline uint32_t hash_slot(uint32_t key, uint32_t shift)
{
    return (key * 0x9E3779B1u) >> shift;
}

// Database-sharded mode: the shard a key belongs to is a range of a 16-bit multiplicative hash of the
// key (shard g of G owns hashes [g * 65536 / G, (g + 1) * 65536 / G)).
__host__ __device__ inline uint32_t shard_of_key(uint32_t key, uint32_t n_shards)
{
    return (((key * 0x9E3779B1u) >> 16) * n_shards) >> 16;
}

// phylo_kmer_db::search (place.cpp:300,311): returns the packed (offset << LEN_BITS | length)
// of the key's posting list, 0 when the key is absent (or its list is empty).
template<uint32_t TABLE>
__device__ __forceinline__ uint64_t db_search(const device_db& db, uint32_t key)
{
    if (TABLE == TABLE_DIRECT)
    {
        return __ldg(db.direct + key);
    }
    else
    {
        uint64_t h = hash_slot(key, db.hash_shift);
        while (true)
        {
            const uint4 s = __ldg(db.hashed + h);
            if (s.y == HASH_EMPTY) { return 0; }
            if (s.x == key)
            {
                const uint64_t off = (static_cast<uint64_t>(s.w) << 32) | s.z;
                return (off << LEN_BITS) | s.y;
            }
            h = (h + 1) & db.hash_mask;
        }
    }
}

// The same for the hashed table when the first slot of the probe sequence has already been fetched
// (the placement kernel issues the first probes of all the windows of a lane before resolving any).
__device__ __forceinline__ uint64_t db_search_from(const device_db& db, uint32_t key, uint4 s)
{
    uint64_t h = hash_slot(key, db.hash_shift);
    while (true)
    {
        if (s.y == HASH_EMPTY) { return 0; }
        if (s.x == key)
        {
            const uint64_t off = (static_cast<uint64_t>(s.w) << 32) | s.z;
            return (off << LEN_BITS) | s.y;
        }
        h = (h + 1) & db.hash_mask;
        s = __ldg(db.hashed + h);
    }
}

// ---- shapes of the placement kernel that the database build also needs ---------------------------------
constexpr int WPL = 8;                 // consecutive windows per lane per chunk
constexpr int CHUNK = 32 * WPL;        // windows staged per pass
constexpr int CODE_BYTES = CHUNK + 32; // letter codes of one pass (k - 1 <= 31 extra letters)
constexpr int SEG_CAP = 256;           // segment descriptors buffered per warp
constexpr int SEG_LEN = 64;            // postings per segment: one 128-bit load per lane
constexpr int SEG_PAD = 16;            // zero descriptors after the last one (>= segments in flight)
constexpr int SWEEP_STEP = 128;        // branches per warp and run of the sweep (4 consecutive ones per lane)

__host__ __device__ inline uint32_t padded_nodes(uint32_t n_nodes)
{
    return (n_nodes + SWEEP_STEP - 1) / SWEEP_STEP * SWEEP_STEP;
}

// bytes of shared memory one warp needs for a tree of n_nodes and counts of count_bytes
__host__ __device__ inline size_t warp_smem_bytes(uint32_t n_nodes, uint32_t count_bytes)
{
    const size_t np = padded_nodes(n_nodes);
    return np * 4 + np * count_bytes + np / 8 + (SEG_CAP + SEG_PAD) * 8 + CODE_BYTES;
}
constexpr size_t CTA_SMEM_FIXED = 1024; // letter table

// Letter tables: the 2-bit / 5-bit codes of i2l (absent submodule; SURVEY.md Appendix B, the same
// tables the oracle shim states with switch statements). Filled by make_letter_table on the host.
void make_letter_table(uint32_t alphabet, uint32_t* table /*[256]*/);
} // namespace epk
