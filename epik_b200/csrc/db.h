// Internal definition of the opaque epk_db handle (include/epik_b200.h).
#pragma once

#include <memory>
#include <string>
#include <vector>
#include "common.cuh"
#include "device_db.cuh"

struct epk_db
{
    int device = 0;
    epk::device_db d{};
    uint32_t table_kind = 0;
    uint64_t table_slots = 0;
    epk::device_buffer<uint64_t> direct;
    epk::device_buffer<uint4> hashed;
    epk::device_buffer<uint2> postings;
    epk::device_buffer<uint32_t> letters;
    std::vector<double> distal, pendant; // place.cpp:98-125, by post-order id
    float omega = 0.0f;
    uint64_t n_entries_total = 0, n_entries_loaded = 0, n_keys_loaded = 0;
    uint32_t shard_index = 0, shard_count = 1; // database-sharded mode: which part of the keys this is
    std::shared_ptr<void> pool;                // pooled database (pool.cu): the common address range d.postings points into
    std::shared_ptr<void> exported;            // epk_db_pool_share: this shard's postings in exportable memory
    size_t pool_local_bytes = 0;               // ... and how much of it sits in this GPU's HBM
};

namespace epk
{
const std::string& last_error();
float score_threshold(uint32_t alphabet, float omega, uint32_t k);
}
