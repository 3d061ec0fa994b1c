// epk_db_*: builds the HBM-resident database from the CSR description (include/epik_b200.h).
// Stands in for i2l::load + what placer::placer precomputes from the tree
// (epik/src/epik/main.cpp:277, epik/src/epik/place.cpp:83-126).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <limits>
#include <vector>
#include <cub/cub.cuh>

#include "common.cuh"
#include "db.h"
#include "db_build.cuh"
#include "synth_gen.cuh"

namespace epk
{
namespace
{
    thread_local std::string g_last_error;

    // DNA: A C G T -> 0..3 (U as T), IUPAC sets; amino acids: the RAPPAS order
    // R H K D E S T N Q C G P A I L M F W Y V -> 0..19 (U as C, O as L), B Z J X sets.
    // Everything else (gaps, '*', digits ...) is LETTER_GAP. [i2l absent: SURVEY.md Appendix B]
    void set_letter(uint32_t* table, char c, std::initializer_list<int> codes)
    {
        uint32_t v = 0;
        if (codes.size() == 1) { v = static_cast<uint32_t>(*codes.begin()); }
        else
        {
            v = LETTER_AMBIGUOUS;
            for (int code : codes) { v |= 1u << (8 + code); }
        }
        table[static_cast<unsigned char>(c)] = v;
        table[static_cast<unsigned char>(c) | 0x20u] = v; // lower case
    }

    // Host -> device through two pinned staging buffers, copies overlapped with the host memcpy.
    void upload(void* dst, const void* src, size_t bytes, cudaStream_t stream)
    {
        if (bytes == 0) { return; }
        cudaPointerAttributes attr{};
        if (cudaPointerGetAttributes(&attr, src) == cudaSuccess && attr.type == cudaMemoryTypeHost)
        {
            EPK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream));
            EPK_CUDA(cudaStreamSynchronize(stream));
            return;
        }
        cudaGetLastError();
        const size_t piece = std::min<size_t>(bytes, 32u << 20);
        pinned_buffer<unsigned char> stage[2];
        cudaEvent_t done[2];
        for (int i = 0; i < 2; ++i)
        {
            stage[i].resize(piece);
            EPK_CUDA(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
        }
        try
        {
            int slot = 0;
            for (size_t at = 0; at < bytes; at += piece, slot ^= 1)
            {
                const size_t n = std::min(piece, bytes - at);
                EPK_CUDA(cudaEventSynchronize(done[slot]));
                std::memcpy(stage[slot].get(), static_cast<const unsigned char*>(src) + at, n);
                EPK_CUDA(cudaMemcpyAsync(static_cast<unsigned char*>(dst) + at, stage[slot].get(), n,
                                         cudaMemcpyHostToDevice, stream));
                EPK_CUDA(cudaEventRecord(done[slot], stream));
            }
            EPK_CUDA(cudaStreamSynchronize(stream));
        }
        catch (...)
        {
            cudaStreamSynchronize(stream);
            for (auto& e : done) { cudaEventDestroy(e); }
            throw;
        }
        for (auto& e : done) { cudaEventDestroy(e); }
    }

    uint64_t exclusive_scan(const uint64_t* in, uint64_t* out, uint64_t n, cudaStream_t stream)
    {
        // returns the total; out[i] = sum of in[0, i)
        if (n == 0) { return 0; }
        size_t tmp_bytes = 0;
        EPK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, out, n, stream));
        device_buffer<unsigned char> tmp(tmp_bytes);
        uint64_t last_in = 0, last_out = 0;
        EPK_CUDA(cudaMemcpyAsync(&last_in, in + n - 1, 8, cudaMemcpyDeviceToHost, stream));
        EPK_CUDA(cub::DeviceScan::ExclusiveSum(tmp.get(), tmp_bytes, in, out, n, stream));
        EPK_CUDA(cudaMemcpyAsync(&last_out, out + n - 1, 8, cudaMemcpyDeviceToHost, stream));
        EPK_CUDA(cudaStreamSynchronize(stream));
        return last_in + last_out;
    }
} // namespace

void set_last_error(const std::string& message) { g_last_error = message; }
const std::string& last_error() { return g_last_error; }

void make_letter_table(uint32_t alphabet, uint32_t* table)
{
    for (int i = 0; i < 256; ++i) { table[i] = LETTER_GAP; }
    if (alphabet == EPK_DNA)
    {
        set_letter(table, 'A', { 0 }); set_letter(table, 'C', { 1 }); set_letter(table, 'G', { 2 });
        set_letter(table, 'T', { 3 }); set_letter(table, 'U', { 3 });
        set_letter(table, 'R', { 0, 2 }); set_letter(table, 'Y', { 1, 3 }); set_letter(table, 'S', { 1, 2 });
        set_letter(table, 'W', { 0, 3 }); set_letter(table, 'K', { 2, 3 }); set_letter(table, 'M', { 0, 1 });
        set_letter(table, 'B', { 1, 2, 3 }); set_letter(table, 'D', { 0, 2, 3 }); set_letter(table, 'H', { 0, 1, 3 });
        set_letter(table, 'V', { 0, 1, 2 }); set_letter(table, 'N', { 0, 1, 2, 3 });
    }
    else
    {
        const char* order = "RHKDESTNQCGPAILMFWYV";
        for (int i = 0; i < 20; ++i) { set_letter(table, order[i], { i }); }
        set_letter(table, 'U', { 9 }); set_letter(table, 'O', { 14 });
        set_letter(table, 'B', { 3, 7 }); set_letter(table, 'Z', { 4, 8 }); set_letter(table, 'J', { 13, 14 });
        set_letter(table, 'X', { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19 });
    }
}

float score_threshold(uint32_t alphabet, float omega, uint32_t k)
{
    // i2l::score_threshold(omega, k) (call site place.cpp:87): (omega / |alphabet|)^k in float
    return std::pow(omega / static_cast<float>(alphabet == EPK_AA ? 20 : 4), static_cast<float>(k));
}
} // namespace epk

using namespace epk;

// Where the raw CSR of a build comes from: host arrays (epk_db_create) or the on-device generator
// (epk_db_create_synthetic). Keys and offsets are small (12 bytes per key) and live on the device for the whole build;
// the entries - the bulk: 8 bytes each, 32 GB for BASELINE config 5 - are only ever on the device one key range at a
// time: the build streams them through a chunk buffer into their packed (and, for large trees, branch-sorted) place,
// so that the raw and the packed database never coexist in HBM.
struct raw_source
{
    uint64_t n_keys = 0, n_total = 0;
    const uint64_t* h_offsets = nullptr;    // [n_keys + 1], host
    std::vector<uint64_t> h_offsets_owned;  // (when the source made them itself)
    device_buffer<uint32_t> keys;           // [n_keys], device
    device_buffer<uint64_t> offsets;        // [n_keys + 1], device
    const float* h_filter_value = nullptr;  // [n_keys] host, or NULL
    // entries of keys [k0, k1), i.e. raw entries [h_offsets[k0], h_offsets[k1]), to dst on the stream; complete on return
    std::function<void(uint64_t k0, uint64_t k1, uint2* dst, cudaStream_t stream)> fill;
};

struct db_shape
{
    uint32_t alphabet, kmer_size;
    float omega;
    uint32_t n_nodes;
    const double* branch_length;
    const uint64_t* subtree_num_nodes;
    const double* subtree_total_length;
};

constexpr uint64_t BUILD_CHUNK_ENTRIES = 64ull << 20; // raw entries streamed per step (512 MB)

static void check_shape(const db_shape& d, uint32_t shard_index, uint32_t shard_count)
{
    if (shard_count == 0 || shard_count > 65536 || shard_index >= shard_count) { throw error(EPK_E_INVALID, "bad shard index / count"); }
    if (d.alphabet != EPK_DNA && d.alphabet != EPK_AA) { throw error(EPK_E_INVALID, "unknown alphabet"); }
    const uint32_t bits = d.alphabet == EPK_AA ? 5 : 2;
    if (d.kmer_size == 0 || d.kmer_size * bits > 32) { throw error(EPK_E_INVALID, "k-mer size does not fit a 32-bit key"); }
    if (d.n_nodes == 0) { throw error(EPK_E_INVALID, "empty tree"); }
    if (!d.branch_length || !d.subtree_num_nodes || !d.subtree_total_length) { throw error(EPK_E_INVALID, "null array in the database description"); }
}

static void select_device(int device, epk_db& db)
{
    int n_devices = 0;
    if (cudaGetDeviceCount(&n_devices) != cudaSuccess || n_devices == 0)
    {
        cudaGetLastError();
        throw error(EPK_E_CUDA, "no CUDA device: epik_b200 has no CPU path");
    }
    if (device < 0 || device >= n_devices) { throw error(EPK_E_INVALID, "no such CUDA device"); }
    EPK_CUDA(cudaSetDevice(device));
    db.device = device;
}

static void build_db(const db_shape& d, raw_source& src, const epk_load_filter* filter, uint32_t shard_index,
                     uint32_t shard_count, epk_db& db, cudaStream_t stream)
{
    const uint32_t bits = d.alphabet == EPK_AA ? 5 : 2;
    const uint64_t n_keys = src.n_keys;
    const uint64_t n_total = src.n_total;
    const uint64_t* h_off = src.h_offsets;

    float omega = d.omega;
    bool use_min_score = false;
    float min_score = 0.0f;
    uint64_t budget = std::numeric_limits<uint64_t>::max();
    if (filter)
    {
        if (!(filter->mu >= 0.0 && filter->mu <= 1.0)) { throw error(EPK_E_INVALID, "mu must be in [0, 1]"); } // main.cpp:196-202
        omega = filter->omega;
        use_min_score = true;
        min_score = std::log10(score_threshold(d.alphabet, filter->omega, d.kmer_size));
        if (filter->mu < 1.0) { budget = static_cast<uint64_t>(filter->mu * static_cast<double>(n_total)); }
        budget = std::min(budget, filter->max_entries);
    }

    // ---- the key ranges the entries are streamed in: whole lists, about BUILD_CHUNK_ENTRIES raw entries each ----
    std::vector<std::pair<uint64_t, uint64_t>> chunks;
    uint64_t max_chunk_raw = 0;
    uint64_t chunk_entries = BUILD_CHUNK_ENTRIES;
    if (const char* text = std::getenv("EPK_BUILD_CHUNK_ENTRIES")) // (tests: many small ranges)
    {
        const unsigned long long v = std::strtoull(text, nullptr, 10);
        if (v) { chunk_entries = v; }
    }
    for (uint64_t k0 = 0; k0 < n_keys;)
    {
        const uint64_t* stop = std::upper_bound(h_off + k0 + 1, h_off + n_keys + 1, h_off[k0] + chunk_entries);
        const uint64_t k1 = std::max<uint64_t>(k0 + 1, static_cast<uint64_t>(stop - h_off) - 1);
        chunks.emplace_back(k0, k1);
        max_chunk_raw = std::max(max_chunk_raw, h_off[k1] - h_off[k0]);
        k0 = k1;
    }
    device_buffer<uint2> raw(max_chunk_raw);
    static_assert(sizeof(epk_entry) == sizeof(uint2), "posting layout");

    const unsigned T = 256;
    // large tree (at most four accumulator sets of the placement kernel fit an SM): lists sorted by branch
    // behind range headers, for the range teams of place_kernel.cuh
    const bool with_ranges = 5 * epk::warp_smem_bytes(d.n_nodes, 1) + epk::CTA_SMEM_FIXED > 232448;
    device_buffer<uint64_t> new_len(n_keys), padded_len(n_keys), new_off(n_keys), region_off;
    uint64_t n_padded = 0;
    const uint32_t* keys = src.keys.get();
    const uint64_t* offsets = src.offsets.get();
    if (n_keys)
    {
        // ---- informative-k-mer prefix (mu / max_entries) ----
        device_buffer<uint8_t> keep;
        if (budget < n_total)
        {
            device_buffer<uint64_t> len(n_keys), sorted_len(n_keys), before(n_keys);
            list_lengths_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(offsets, n_keys, len.get());
            device_buffer<uint32_t> order;
            if (src.h_filter_value)
            {
                device_buffer<float> fv(n_keys), fv_sorted(n_keys);
                device_buffer<uint32_t> idx(n_keys);
                order.resize(n_keys);
                upload(fv.get(), src.h_filter_value, n_keys * 4, stream);
                iota_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(idx.get(), n_keys);
                size_t tmp_bytes = 0;
                EPK_CUDA(cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp_bytes, fv.get(), fv_sorted.get(),
                                                                   idx.get(), order.get(), n_keys, 0, 32, stream));
                device_buffer<unsigned char> tmp(tmp_bytes);
                EPK_CUDA(cub::DeviceRadixSort::SortPairsDescending(tmp.get(), tmp_bytes, fv.get(), fv_sorted.get(),
                                                                   idx.get(), order.get(), n_keys, 0, 32, stream));
                EPK_CUDA(cudaStreamSynchronize(stream));
            }
            gather_lengths_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(len.get(), order.get(), n_keys, sorted_len.get());
            exclusive_scan(sorted_len.get(), before.get(), n_keys, stream);
            keep.resize(n_keys);
            prefix_cut_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(before.get(), order.get(), n_keys, budget, keep.get());
            EPK_CUDA(cudaStreamSynchronize(stream));
        }
        // ---- database-sharded mode: after the (global) informative prefix, only this shard's keys stay ----
        if (shard_count > 1)
        {
            device_buffer<uint8_t> mine(n_keys);
            shard_keep_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(keys, n_keys, shard_index, shard_count, keep.get(), mine.get());
            EPK_CUDA(cudaStreamSynchronize(stream));
            keep = std::move(mine);
        }
        // ---- per-entry omega threshold, packed layout ----
        const uint32_t header = with_ranges ? RANGE_HEADER_ENTRIES : 0u;
        if (!use_min_score)
        {
            // the surviving length of a kept list is its length: no entry has to be looked at
            surviving_lengths_kernel<<<blocks_for(n_keys * 32, T), T, 0, stream>>>(offsets, nullptr, keep.get(), n_keys, min_score,
                                                                                   false, new_len.get(), padded_len.get(), header);
        }
        else
        {
            for (const auto& [k0, k1] : chunks)
            {
                src.fill(k0, k1, raw.get(), stream);
                surviving_lengths_kernel<<<blocks_for((k1 - k0) * 32, T), T, 0, stream>>>(
                    offsets + k0, raw.get() - h_off[k0], keep.get() ? keep.get() + k0 : nullptr, k1 - k0, min_score, true,
                    new_len.get() + k0, padded_len.get() + k0, header);
                EPK_CUDA(cudaStreamSynchronize(stream)); // the chunk buffer is refilled next
            }
        }
        n_padded = exclusive_scan(padded_len.get(), new_off.get(), n_keys, stream);
        if (with_ranges)
        {
            region_off.resize(n_keys);
            EPK_CUDA(cudaMemcpyAsync(region_off.get(), new_off.get(), n_keys * 8, cudaMemcpyDeviceToDevice, stream));
            add_header_room_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(new_off.get(), new_len.get(), n_keys);
        }
    }
    db.postings.resize(std::max<uint64_t>(n_padded, 2));
    EPK_CUDA(cudaMemsetAsync(db.postings.get(), 0, 16, stream));
    device_buffer<unsigned long long> totals(2);
    device_buffer<uint32_t> status(1);
    EPK_CUDA(cudaMemsetAsync(totals.get(), 0, 16, stream));
    EPK_CUDA(cudaMemsetAsync(status.get(), 0, 4, stream));
    if (n_keys)
    {
        // ---- the entries, one key range at a time: raw chunk -> packed (-> sorted by branch behind range headers) ----
        const uint64_t* region = with_ranges ? region_off.get() : new_off.get();
        std::vector<uint64_t> bound(chunks.size() + 1, n_padded); // packed position where each chunk's region starts
        for (size_t c = 0; c < chunks.size(); ++c)
        {
            EPK_CUDA(cudaMemcpyAsync(&bound[c], region + chunks[c].first, 8, cudaMemcpyDeviceToHost, stream));
        }
        EPK_CUDA(cudaStreamSynchronize(stream));
        uint64_t max_region = 0;
        for (size_t c = 0; c < chunks.size(); ++c) { max_region = std::max(max_region, bound[c + 1] - bound[c]); }
        device_buffer<uint2> packed_chunk;
        unsigned sort_warps = 8;
        size_t sort_smem = 0;
        if (with_ranges)
        {
            packed_chunk.resize(std::max<uint64_t>(max_region, 2));
            const uint32_t np = epk::padded_nodes(d.n_nodes);
            while (sort_warps > 1 && static_cast<size_t>(sort_warps) * 2 * (np / 32) * 4 > 96 * 1024) { sort_warps >>= 1; }
            sort_smem = static_cast<size_t>(sort_warps) * 2 * (np / 32) * 4;
            if (sort_smem > 200 * 1024) { throw error(EPK_E_UNSUPPORTED, "tree too large"); }
            EPK_CUDA(cudaFuncSetAttribute(sort_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sort_smem)));
        }
        for (size_t c = 0; c < chunks.size(); ++c)
        {
            const auto [k0, k1] = chunks[c];
            const uint64_t nk = k1 - k0;
            src.fill(k0, k1, raw.get(), stream);
            // (absolute offsets, biased base pointers: the kernels index as if the whole arrays were there)
            uint2* pack_to = with_ranges ? packed_chunk.get() - bound[c] : db.postings.get();
            if (with_ranges) { EPK_CUDA(cudaMemsetAsync(packed_chunk.get(), 0, (bound[c + 1] - bound[c]) * 8, stream)); }
            pack_lists_kernel<<<blocks_for(nk * 32, T), T, 0, stream>>>(offsets + k0, raw.get() - h_off[k0], new_len.get() + k0,
                                                                       new_off.get() + k0, nk, min_score, use_min_score, pack_to);
            if (with_ranges)
            {
                const uint32_t np = epk::padded_nodes(d.n_nodes);
                sort_lists_kernel<<<1184, sort_warps * 32, sort_smem, stream>>>(pack_to, db.postings.get(), new_len.get() + k0,
                                                                                new_off.get() + k0, nk, d.n_nodes, np);
            }
            EPK_CUDA(cudaStreamSynchronize(stream)); // the chunk buffers are reused by the next range
            EPK_CUDA(cudaGetLastError());
        }
        count_loaded_kernel<<<std::min(blocks_for(n_keys, T), 1024u), T, 0, stream>>>(new_len.get(), n_keys, totals.get());
    }
    raw.release();
    unsigned long long h_totals[2] = { 0, 0 };
    EPK_CUDA(cudaMemcpyAsync(h_totals, totals.get(), 16, cudaMemcpyDeviceToHost, stream));
    EPK_CUDA(cudaStreamSynchronize(stream));
    EPK_CUDA(cudaGetLastError());

    // ---- key table ----
    device_db& dd = db.d;
    std::memset(&dd, 0, sizeof(dd));
    const uint32_t key_bits = bits * d.kmer_size;
    if (key_bits <= 26)
    {
        db.table_kind = TABLE_DIRECT;
        db.table_slots = 1ull << key_bits;
        db.direct.resize(db.table_slots);
        EPK_CUDA(cudaMemsetAsync(db.direct.get(), 0, db.direct.bytes(), stream));
        if (n_keys)
        {
            fill_direct_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(
                keys, new_len.get(), new_off.get(), n_keys, db.table_slots,
                reinterpret_cast<unsigned long long*>(db.direct.get()), status.get());
        }
    }
    else
    {
        db.table_kind = TABLE_HASH;
        uint64_t slots = 1024;
        while (slots < 2 * h_totals[1]) { slots <<= 1; }
        if (slots > (1ull << 32)) { throw error(EPK_E_UNSUPPORTED, "hash table too large"); }
        uint32_t log2 = 0;
        while ((1ull << log2) < slots) { ++log2; }
        db.table_slots = slots;
        db.hashed.resize(slots);
        dd.hash_mask = slots - 1;
        dd.hash_shift = 32 - log2;
        clear_hash_kernel<<<blocks_for(slots, T), T, 0, stream>>>(db.hashed.get(), slots);
        if (n_keys)
        {
            fill_hash_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(
                keys, new_len.get(), new_off.get(), n_keys, db.hashed.get(), dd.hash_mask, dd.hash_shift, status.get());
            check_hash_kernel<<<blocks_for(n_keys, T), T, 0, stream>>>(
                keys, new_len.get(), new_off.get(), n_keys, db.hashed.get(), dd.hash_mask, dd.hash_shift, status.get());
        }
    }
    // ---- validation of the lists ----
    if (n_keys)
    {
        const uint32_t words = (d.n_nodes + 31) / 32;
        unsigned warps = 8;
        while (warps > 1 && static_cast<size_t>(warps) * words * 4 > 96 * 1024) { warps >>= 1; }
        const size_t smem = static_cast<size_t>(warps) * words * 4;
        if (smem > 200 * 1024) { throw error(EPK_E_UNSUPPORTED, "tree too large"); }
        EPK_CUDA(cudaFuncSetAttribute(validate_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        validate_lists_kernel<<<1184, warps * 32, smem, stream>>>(db.postings.get(), new_len.get(), new_off.get(), n_keys,
                                                                  d.n_nodes, status.get());
    }
    // ---- range of the loaded scores (which arithmetic the sweep of the placement kernel may use) ----
    device_buffer<uint32_t> stats(4);
    uint32_t h_stats[4] = { 0xFFFFFFFFu, 0u, 0x7F800000u, 0u };
    EPK_CUDA(cudaMemcpyAsync(stats.get(), h_stats, sizeof(h_stats), cudaMemcpyHostToDevice, stream));
    if (n_keys)
    {
        score_stats_kernel<<<1184, 256, 0, stream>>>(db.postings.get(), new_len.get(), new_off.get(), n_keys, stats.get());
    }
    EPK_CUDA(cudaMemcpyAsync(h_stats, stats.get(), sizeof(h_stats), cudaMemcpyDeviceToHost, stream));
    uint32_t h_status = 0;
    EPK_CUDA(cudaMemcpyAsync(&h_status, status.get(), 4, cudaMemcpyDeviceToHost, stream));
    EPK_CUDA(cudaStreamSynchronize(stream));
    EPK_CUDA(cudaGetLastError());
    if (h_status & BUILD_DUPLICATE_KEY) { throw error(EPK_E_INVALID, "duplicate k-mer key in the database"); }
    if (h_status & BUILD_KEY_RANGE) { throw error(EPK_E_INVALID, "k-mer key out of range for this k"); }
    if (h_status & BUILD_BAD_BRANCH) { throw error(EPK_E_INVALID, "branch id >= node count in a posting list"); }
    if (h_status & BUILD_DUPLICATE_BRANCH) { throw error(EPK_E_INVALID, "a posting list names the same branch twice"); }
    if (h_status & BUILD_LIST_TOO_LONG) { throw error(EPK_E_UNSUPPORTED, "posting list longer than 2^24 - 1 entries"); }

    // ---- letters, thresholds, tree tables ----
    uint32_t table[256];
    make_letter_table(d.alphabet, table);
    db.letters.resize(256);
    EPK_CUDA(cudaMemcpy(db.letters.get(), table, sizeof(table), cudaMemcpyHostToDevice));

    dd.direct = db.direct.get();
    dd.hashed = db.hashed.get();
    dd.postings = db.postings.get();
    dd.table_kind = db.table_kind;
    dd.k = d.kmer_size;
    dd.bits = bits;
    dd.key_mask = key_bits >= 32 ? 0xFFFFFFFFu : ((1u << key_bits) - 1u);
    dd.n_nodes = d.n_nodes;
    dd.range_headers = with_ranges ? 1u : 0u;
    dd.alphabet = d.alphabet;
    dd.threshold = score_threshold(d.alphabet, omega, d.kmer_size);
    dd.log_threshold = std::log10(dd.threshold); // place.cpp:88, float log10
    {
        const bool any = h_stats[0] <= h_stats[1];
        dd.score_lo = any ? ordered_key_value(h_stats[0]) : 0.0f;
        dd.score_hi = any ? ordered_key_value(h_stats[1]) : 0.0f;
        float tiny;
        std::memcpy(&tiny, &h_stats[2], 4);
        dd.fast_ok = (h_stats[3] == 0 && dd.score_hi <= 0.0f && tiny >= 1e-20f) ? 1u : 0u;
    }

    // place.cpp:98-125: distal = branch_length / 2, pendant = mean subtree branch length + distal
    db.distal.resize(d.n_nodes);
    db.pendant.resize(d.n_nodes);
    for (uint32_t i = 0; i < d.n_nodes; ++i)
    {
        const double distal = d.branch_length[i] / 2;
        double mean = 0.0;
        if (d.subtree_num_nodes[i] > 1) { mean = d.subtree_total_length[i] / static_cast<double>(d.subtree_num_nodes[i]); }
        db.distal[i] = distal;
        db.pendant[i] = mean + distal;
    }

    db.omega = omega;
    db.shard_index = shard_index;
    db.shard_count = shard_count;
    db.n_entries_total = n_total;
    db.n_entries_loaded = h_totals[0];
    db.n_keys_loaded = h_totals[1];
}

// epk_db_create / epk_db_create_shard: the raw CSR is the caller's host arrays
static void create_from_host(const epk_db_desc& d, const epk_load_filter* filter, int device, uint32_t shard_index,
                             uint32_t shard_count, epk_db& db)
{
    const db_shape shape{ d.alphabet, d.kmer_size, d.omega, d.n_nodes, d.branch_length, d.subtree_num_nodes, d.subtree_total_length };
    check_shape(shape, shard_index, shard_count);
    if (d.n_keys > 0xFFFFFFFFull) { throw error(EPK_E_INVALID, "too many keys"); }
    if (d.n_keys && (!d.keys || !d.offsets)) { throw error(EPK_E_INVALID, "null array in the database description"); }
    if (d.n_keys && d.offsets[0] != 0) { throw error(EPK_E_INVALID, "offsets[0] must be 0"); }
    // every kernel of the build trusts e - b of neighbouring offsets: a decreasing pair would wrap into a list of 2^64 postings
    for (uint64_t i = 0; i < d.n_keys; ++i)
    {
        if (d.offsets[i + 1] < d.offsets[i]) { throw error(EPK_E_INVALID, "offsets must be non-decreasing"); }
    }
    select_device(device, db);
    cudaStream_t stream;
    EPK_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    struct stream_guard { cudaStream_t s; ~stream_guard() { cudaStreamDestroy(s); } } guard{ stream };

    raw_source src;
    static const uint64_t no_offsets[1] = { 0 };
    src.n_keys = d.n_keys;
    src.n_total = d.n_keys ? d.offsets[d.n_keys] : 0;
    src.h_offsets = d.n_keys ? d.offsets : no_offsets;
    src.h_filter_value = d.filter_value;
    if (src.n_total && !d.entries) { throw error(EPK_E_INVALID, "null entries"); }
    src.keys.resize(d.n_keys);
    src.offsets.resize(d.n_keys + 1);
    if (d.n_keys)
    {
        upload(src.keys.get(), d.keys, d.n_keys * 4, stream);
        upload(src.offsets.get(), d.offsets, (d.n_keys + 1) * 8, stream);
    }
    else { EPK_CUDA(cudaMemsetAsync(src.offsets.get(), 0, 8, stream)); }
    const epk_entry* entries = d.entries;
    const uint64_t* h_off = src.h_offsets;
    src.fill = [entries, h_off](uint64_t k0, uint64_t k1, uint2* dst, cudaStream_t s) {
        upload(dst, entries + h_off[k0], (h_off[k1] - h_off[k0]) * 8, s);
    };
    build_db(shape, src, filter, shard_index, shard_count, db, stream);
}

extern "C" epk_status epk_db_create(const epk_db_desc* desc, const epk_load_filter* filter, int device, epk_db** out)
{
    return guarded([&] {
        if (!desc || !out) { throw error(EPK_E_INVALID, "null argument"); }
        *out = nullptr;
        auto db = std::make_unique<epk_db>();
        create_from_host(*desc, filter, device, 0, 1, *db);
        *out = db.release();
    });
}

extern "C" epk_status epk_db_create_shard(const epk_db_desc* desc, const epk_load_filter* filter, int device,
                                          uint32_t shard_index, uint32_t shard_count, epk_db** out)
{
    return guarded([&] {
        if (!desc || !out) { throw error(EPK_E_INVALID, "null argument"); }
        *out = nullptr;
        auto db = std::make_unique<epk_db>();
        create_from_host(*desc, filter, device, shard_index, shard_count, *db);
        *out = db.release();
    });
}

// epk_db_create_synthetic: the raw CSR is generated on the device (synth_gen.cuh), keys and offsets up front, the
// entries one key range at a time into the build's chunk buffer
extern "C" epk_status epk_db_create_synthetic(const epk_synth_desc* sd, const epk_load_filter* filter, int device,
                                              uint32_t shard_index, uint32_t shard_count, epk_db** out)
{
    return guarded([&] {
        if (!sd || !out) { throw error(EPK_E_INVALID, "null argument"); }
        *out = nullptr;
        const db_shape shape{ sd->alphabet, sd->kmer_size, sd->omega, sd->n_nodes, sd->branch_length, sd->subtree_num_nodes,
                              sd->subtree_total_length };
        check_shape(shape, shard_index, shard_count);
        if (!sd->length_table || sd->length_table_bits == 0 || sd->length_table_bits > 20) { throw error(EPK_E_INVALID, "bad length table"); }
        const uint32_t bits = sd->alphabet == EPK_AA ? 5 : 2;
        uint64_t n_candidates = 1;
        for (uint32_t i = 0; i < sd->kmer_size; ++i) { n_candidates *= sd->alphabet == EPK_AA ? 20u : 4u; }
        if (sd->key_space_limit && sd->key_space_limit < n_candidates) { n_candidates = sd->key_space_limit; }
        (void)bits;
        for (uint32_t i = 0; i < (1u << sd->length_table_bits); ++i)
        {
            // distinct branches need start + 7 * length < start + n_nodes
            if (sd->length_table[i] == 0 || 8ull * sd->length_table[i] > sd->n_nodes) { throw error(EPK_E_INVALID, "list length outside [1, n_nodes / 8]"); }
        }
        auto db = std::make_unique<epk_db>();
        select_device(device, *db);
        cudaStream_t stream;
        EPK_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        struct stream_guard { cudaStream_t s; ~stream_guard() { cudaStreamDestroy(s); } } guard{ stream };

        device_buffer<uint32_t> table(1ull << sd->length_table_bits);
        EPK_CUDA(cudaMemcpyAsync(table.get(), sd->length_table, table.bytes(), cudaMemcpyHostToDevice, stream));
        synth_params sp{};
        sp.base_present = mix64((sd->seed << 8) ^ 3u);
        sp.base_length = mix64((sd->seed << 8) ^ 4u);
        sp.base_stride = mix64((sd->seed << 8) ^ 6u);
        sp.base_start = mix64((sd->seed << 8) ^ 7u);
        sp.base_score = mix64((sd->seed << 8) ^ 8u);
        sp.presence_threshold = sd->presence_threshold;
        sp.length_table = table.get();
        sp.length_shift = 64 - sd->length_table_bits;
        sp.n_nodes = sd->n_nodes;
        sp.alphabet = sd->alphabet;
        sp.k = sd->kmer_size;
        sp.score_lo = sd->score_lo;
        sp.score_span = -0.3 - sd->score_lo;

        raw_source src;
        const unsigned T = 256;
        {
            device_buffer<uint64_t> present(n_candidates), length(n_candidates), at(n_candidates), before(n_candidates);
            synth_lengths_kernel<<<blocks_for(n_candidates, T), T, 0, stream>>>(sp, n_candidates, present.get(), length.get());
            src.n_keys = exclusive_scan(present.get(), at.get(), n_candidates, stream);
            src.n_total = exclusive_scan(length.get(), before.get(), n_candidates, stream);
            src.keys.resize(src.n_keys);
            src.offsets.resize(src.n_keys + 1);
            synth_scatter_kernel<<<blocks_for(n_candidates, T), T, 0, stream>>>(sp, n_candidates, present.get(), at.get(), before.get(),
                                                                               src.keys.get(), src.offsets.get());
            EPK_CUDA(cudaMemcpyAsync(src.offsets.get() + src.n_keys, &src.n_total, 8, cudaMemcpyHostToDevice, stream));
            EPK_CUDA(cudaStreamSynchronize(stream));
            EPK_CUDA(cudaGetLastError());
        }
        src.h_offsets_owned.resize(src.n_keys + 1);
        EPK_CUDA(cudaMemcpy(src.h_offsets_owned.data(), src.offsets.get(), (src.n_keys + 1) * 8, cudaMemcpyDeviceToHost));
        src.h_offsets = src.h_offsets_owned.data();
        const uint32_t* d_keys = src.keys.get();
        const uint64_t* d_offsets = src.offsets.get();
        const uint64_t* h_off = src.h_offsets;
        src.fill = [sp, d_keys, d_offsets, h_off](uint64_t k0, uint64_t k1, uint2* dst, cudaStream_t s) {
            synth_entries_kernel<<<blocks_for((k1 - k0) * 32, 256), 256, 0, s>>>(sp, d_keys + k0, d_offsets + k0, k1 - k0, dst - h_off[k0]);
            EPK_CUDA(cudaGetLastError());
        };
        build_db(shape, src, filter, shard_index, shard_count, *db, stream);
        *out = db.release();
    });
}

extern "C" uint32_t epk_shard_of_key(uint32_t key, uint32_t shard_count)
{
    return shard_count ? epk::shard_of_key(key, shard_count) : 0;
}

extern "C" epk_status epk_db_get_info(const epk_db* db, epk_db_info* info)
{
    return guarded([&] {
        if (!db || !info) { throw error(EPK_E_INVALID, "null argument"); }
        info->alphabet = db->d.alphabet;
        info->kmer_size = db->d.k;
        info->n_nodes = db->d.n_nodes;
        info->table_kind = db->table_kind;
        info->omega = db->omega;
        info->threshold = db->d.threshold;
        info->log10_threshold = db->d.log_threshold;
        info->n_keys_loaded = db->n_keys_loaded;
        info->n_entries_loaded = db->n_entries_loaded;
        info->n_entries_total = db->n_entries_total;
        info->table_slots = db->table_slots;
        info->device_bytes = db->direct.bytes() + db->hashed.bytes() + db->postings.bytes() + db->letters.bytes() + db->pool_local_bytes;
    });
}

extern "C" void epk_db_destroy(epk_db* db)
{
    if (db)
    {
        cudaSetDevice(db->device);
        delete db;
    }
}

extern "C" const char* epk_last_error(void) { return epk::last_error().c_str(); }
extern "C" uint32_t epk_abi_version(void) { return EPK_ABI_VERSION; }

extern "C" int epk_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" void* epk_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess)
    {
        cudaGetLastError();
        epk::set_last_error("cudaMallocHost failed");
        return nullptr;
    }
    return p;
}

extern "C" void epk_host_free(void* p)
{
    if (p) { cudaFreeHost(p); }
}
