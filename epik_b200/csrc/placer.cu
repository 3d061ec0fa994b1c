// epk_placer_* / epk_place*: the host side of epik::placer (reference
// epik/include/epik/place.h:81-140, epik/src/epik/place.cpp:83-126,201-275) in front of the
// sm_100a placement kernel (place_kernel.cuh).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <future>
#include <memory>
#include <mutex>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "db.h"
#include "place_kernel.cuh"

namespace epk
{
namespace
{
    constexpr int N_SLOTS = 3; // chunks in flight in epk_place: copy in / compute / copy out

    // letter codes straight from the sequence bytes (the placement kernel stages them in shared memory)
    struct letter_codes
    {
        const uint32_t* letters;
        const char* seq;
        __device__ uint32_t operator[](uint64_t i) const { return letters[static_cast<uint8_t>(seq[i])] & 0xFFu; }
    };

    // One thread, one sequence: the (window, key) stream in the order place.cpp:294-313 consumes it,
    // through the same classify_window / db_search the placement kernel runs.
    template<uint32_t TABLE>
    __global__ void kmers_kernel(device_db db, const uint32_t* letters, const char* seq, uint64_t len, uint64_t cap,
                                 uint32_t* pos, uint32_t* key, uint32_t* group, long long* hit_len,
                                 unsigned long long* n_out)
    {
        if (threadIdx.x != 0) { return; }
        uint64_t n = 0;
        for (uint64_t w = 0; w + db.k <= len; ++w)
        {
            const window_class c = classify_window(letter_codes{ letters, seq + w }, 0u, db.k, db.bits);
            if (c.gaps != 0 || c.ambs > 1) { continue; }
            uint32_t resolutions = 1, shift = 0, n_keys = 1;
            if (c.ambs == 1)
            {
                resolutions = letters[static_cast<uint8_t>(seq[w + c.amb_at])] >> 8;
                shift = db.bits * (db.k - 1 - c.amb_at);
                n_keys = __popc(resolutions);
            }
            while (resolutions)
            {
                const uint32_t code = c.ambs ? __ffs(resolutions) - 1 : 0;
                resolutions &= resolutions - 1;
                const uint32_t kk = c.key | (code << shift);
                if (n < cap)
                {
                    const uint64_t found = db_search<TABLE>(db, kk);
                    pos[n] = static_cast<uint32_t>(w);
                    key[n] = kk;
                    group[n] = n_keys;
                    hit_len[n] = (found & LEN_MASK) ? static_cast<long long>(found & LEN_MASK) : -1;
                }
                ++n;
            }
        }
        *n_out = n;
    }

    struct out_arrays
    {
        device_buffer<uint32_t> n_place, branch, count;
        device_buffer<float> score;
        device_buffer<double> lwr;
    };

    struct chunk_slot
    {
        cudaStream_t stream = nullptr;
        device_buffer<char> seq;
        device_buffer<uint64_t> off;
        out_arrays out;
        device_buffer<unsigned long long> counters; // [N_CLASSES][CTR_COUNT]
        device_buffer<uint32_t> class_reads;         // [N_CLASSES]
        pinned_buffer<char> h_seq;        // staging when the caller's memory is pageable
        pinned_buffer<uint64_t> h_off;
        pinned_buffer<unsigned char> h_out;
        size_t reads_cap = 0, bytes_cap = 0;
        // pending copy-out of the previous chunk that used this slot (pageable destination)
        bool pending = false;
        uint32_t pending_first = 0, pending_n = 0;
        // rows of the caller's arrays this slot's work in flight will have filled once its stream is idle
        bool rows_in_flight = false;
        uint32_t rows_first = 0, rows_n = 0;
    };
} // namespace
} // namespace epk

using namespace epk;

struct epk_placer
{
    const epk_db* db = nullptr;
    int device = 0; // copy: the destructor must not touch the database (it may already be gone)
    uint32_t keep_at_most = 0;
    double keep_factor = 0.0;
    uint32_t variant = 0;
    uint32_t prefetch = 0;   // prefetch_mode of the gathers
    bool prefetch_default = true; // not set through the variant word: lines, except in the 16-warp launch of small trees
    uint32_t* trace = nullptr; // epk_placer_set_trace: DEVICE array of the caller, [4 n] of the next epk_place_device calls
    uint32_t chunk_reads = 0;
    bool auto_chunk = true;  // chunk_reads was not given: small passes are cut finer so that their copies overlap too
    int n_sms = 0;
    uint32_t warps_per_cta[N_CLASSES] = { 0, 0, 0 }; // 0: the accumulators of this class do not fit shared memory
    size_t smem_bytes[N_CLASSES] = { 0, 0, 0 };
    uint32_t unsupported_mask = 0;
    uint32_t team[N_CLASSES] = { 1, 1, 1 };           // warps per accumulator set of the fused launch
    bool wide[N_CLASSES] = { false, false, false };   // 16 one-warp sets per CTA (small trees)
    uint32_t sets_fused[N_CLASSES] = { 0, 0, 0 };     // accumulator sets per CTA of the fused launch
    size_t smem_fused[N_CLASSES] = { 0, 0, 0 };
    device_buffer<uint32_t> fake_order;
    device_buffer<unsigned long long> totals;          // running work counters (bumped by the launches)
    device_buffer<unsigned long long> device_counters; // per-launch counters of epk_place_device
    device_buffer<uint32_t> device_class_reads;
    uint32_t global_mask = 0;                        // classes placed with their accumulators in HBM (place_kernel<..., GLOBAL>)
    device_buffer<unsigned char> acc_scratch;         // N_SLOTS + 1 regions of n_sms * 8 rows (one region per stream that may launch)
    size_t acc_region_bytes = 0;
    device_buffer<device_db> shard_views;   // epk_placer_set_shard_views: the shards of a database-sharded group, from this device
    uint32_t n_shard_views = 0;
    device_buffer<uint32_t> seen_scratch;   // bitmaps of the finalising launch's ambiguous-window replay (one per warp)
    device_buffer<unsigned char> amb_scratch; // and the dense rows its contributions go to (float S | uint16 C per warp)
    chunk_slot slots[N_SLOTS];
    uint64_t n_launches = 0, n_reads = 0;
    cudaStream_t last_stream = nullptr;
    bool totals_dirty = false;
};

#ifndef EPK_U_MAIN
#define EPK_U_MAIN 8 // segments in flight per buffer of the one-warp-per-query launch (kernel experiments: 12, 16)
#endif

namespace
{
    using kernel_fn = void (*)(const place_params);

    // the launch of database-sharded mode that leaves partial rows (classes 0 and 1 only)
    kernel_fn pick_partial_kernel(const epk_placer& pl, int cls)
    {
        if (pl.team[cls] > 1)
        {
            if (pl.db->table_kind == TABLE_DIRECT)
            {
                return cls == 0 ? place_kernel<TABLE_DIRECT, 8, uint8_t, MODE_PARTIAL, true> : place_kernel<TABLE_DIRECT, 8, uint16_t, MODE_PARTIAL, true>;
            }
            return cls == 0 ? place_kernel<TABLE_HASH, 8, uint8_t, MODE_PARTIAL, true> : place_kernel<TABLE_HASH, 8, uint16_t, MODE_PARTIAL, true>;
        }
        if (pl.db->table_kind == TABLE_DIRECT)
        {
            return cls == 0 ? place_kernel<TABLE_DIRECT, 8, uint8_t, MODE_PARTIAL> : place_kernel<TABLE_DIRECT, 8, uint16_t, MODE_PARTIAL>;
        }
        return cls == 0 ? place_kernel<TABLE_HASH, 8, uint8_t, MODE_PARTIAL> : place_kernel<TABLE_HASH, 8, uint16_t, MODE_PARTIAL>;
    }

    // the load-only form of the class-0 launch (roofline ceiling, epk_gather_ceiling_device)
    kernel_fn pick_gather_only_kernel(const epk_placer& pl)
    {
        if (pl.team[0] > 1)
        {
            return pl.db->table_kind == TABLE_DIRECT ? place_kernel<TABLE_DIRECT, 8, uint8_t, MODE_GATHER_ONLY, true>
                                                     : place_kernel<TABLE_HASH, 8, uint8_t, MODE_GATHER_ONLY, true>;
        }
        return pl.db->table_kind == TABLE_DIRECT ? place_kernel<TABLE_DIRECT, 8, uint8_t, MODE_GATHER_ONLY>
                                                 : place_kernel<TABLE_HASH, 8, uint8_t, MODE_GATHER_ONLY>;
    }

    // Partial rows of database-sharded mode: dense when a shard touches a large part of what the query touches (two
    // shards: half), compact from three shards on - measured on BASELINE config 5: 2 GPUs 5.6 M reads/s dense / 4.1 M
    // compact, 8 GPUs 1.5 M dense / 3.1 M compact (gpurun r3n, r3s, r3o, r3t). EPK_ROW_FORMAT=dense|compact overrides.
    uint32_t row_format_of(const epk_placer& pl)
    {
        if (const char* text = std::getenv("EPK_ROW_FORMAT"))
        {
            if (std::strcmp(text, "dense") == 0) { return ROW_DENSE; }
            if (std::strcmp(text, "compact") == 0) { return ROW_COMPACT; }
        }
        return pl.db->shard_count >= 3 ? ROW_COMPACT : ROW_DENSE;
    }

    // the launch for classes whose accumulators do not fit shared memory: rows in HBM, 32-bit counts, 8 warps per CTA
    kernel_fn pick_global_kernel(const epk_placer& pl)
    {
        return pl.db->table_kind == TABLE_DIRECT ? place_kernel<TABLE_DIRECT, 4, uint32_t, MODE_FULL, false, MAX_WARPS_PER_CTA, true>
                                                 : place_kernel<TABLE_HASH, 4, uint32_t, MODE_FULL, false, MAX_WARPS_PER_CTA, true>;
    }

    kernel_fn pick_kernel(const epk_placer& pl, int cls)
    {
        if (pl.wide[cls])
        {
            // small tree: 16 accumulator sets and 16 warps per SM
            if (pl.db->table_kind == TABLE_DIRECT)
            {
                return cls == 0 ? place_kernel<TABLE_DIRECT, 4, uint8_t, MODE_FULL, false, 16> : place_kernel<TABLE_DIRECT, 4, uint16_t, MODE_FULL, false, 16>;
            }
            return cls == 0 ? place_kernel<TABLE_HASH, 4, uint8_t, MODE_FULL, false, 16> : place_kernel<TABLE_HASH, 4, uint16_t, MODE_FULL, false, 16>;
        }
        if (pl.team[cls] > 1)
        {
            // large tree, few accumulator sets per SM: several warps per set (place_kernel.cuh, "teams")
            if (pl.db->table_kind == TABLE_DIRECT)
            {
                return cls == 0 ? place_kernel<TABLE_DIRECT, 8, uint8_t, MODE_FULL, true> : place_kernel<TABLE_DIRECT, 8, uint16_t, MODE_FULL, true>;
            }
            return cls == 0 ? place_kernel<TABLE_HASH, 8, uint8_t, MODE_FULL, true> : place_kernel<TABLE_HASH, 8, uint16_t, MODE_FULL, true>;
        }
        if (pl.db->table_kind == TABLE_DIRECT)
        {
            return cls == 0 ? place_kernel<TABLE_DIRECT, EPK_U_MAIN, uint8_t> : cls == 1 ? place_kernel<TABLE_DIRECT, EPK_U_MAIN, uint16_t>
                                                                                          : place_kernel<TABLE_DIRECT, EPK_U_MAIN, uint32_t>;
        }
        return cls == 0 ? place_kernel<TABLE_HASH, EPK_U_MAIN, uint8_t> : cls == 1 ? place_kernel<TABLE_HASH, EPK_U_MAIN, uint16_t>
                                                                                    : place_kernel<TABLE_HASH, EPK_U_MAIN, uint32_t>;
    }

    // One batch of n sequences already on the device: a count of the queries per class, then one
    // launch of the placement kernel per class (uint8 / uint16 / uint32 hit counts; a launch whose
    // class is empty returns at once). The launches add their work counters to pl.totals themselves:
    // nothing small follows a placement launch on its stream. `class_mask` names
    // the classes the caller knows to be present (all of them when it does not know).
    void launch(epk_placer& pl, const char* d_seq, const uint64_t* d_off, uint32_t n, const epk_result& d_out,
                unsigned long long* counters, uint32_t* class_reads, uint32_t class_mask, cudaStream_t stream,
                uint32_t* trace = nullptr, uint32_t acc_region = N_SLOTS)
    {
        place_params p{};
        p.db = pl.db->d;
        p.seq = d_seq;
        p.seq_off = d_off;
        p.n = n;
        p.keep_at_most = pl.keep_at_most;
        p.prefetch = pl.prefetch;
        p.keep_factor = pl.keep_factor;
        p.n_place = d_out.n_place;
        p.branch = d_out.branch;
        p.score = d_out.score;
        p.lwr = d_out.lwr;
        p.count = d_out.count;
        p.letters = pl.db->letters.get();
        p.fake_order = pl.fake_order.get();
        p.class_reads = class_reads;
        p.totals = pl.totals.get();
        p.trace = trace;
        if (trace && n) { EPK_CUDA(cudaMemsetAsync(trace, 0, 16ull * n, stream)); } // queries that are not placed keep zeros
        EPK_CUDA(cudaMemsetAsync(counters, 0, N_CLASSES * CTR_COUNT * sizeof(unsigned long long), stream));
        EPK_CUDA(cudaMemsetAsync(class_reads, 0, N_CLASSES * sizeof(uint32_t), stream));
        {
            const cudaError_t stale = cudaGetLastError();
            if (stale != cudaSuccess) { throw error(EPK_E_CUDA, std::string("stale CUDA error before launch: ") + cudaGetErrorString(stale)); }
        }
        if (n)
        {
            const unsigned cgrid = static_cast<unsigned>(std::min<uint64_t>(4ull * pl.n_sms, (static_cast<uint64_t>(n) + 255) / 256));
            classify_kernel<<<cgrid, 256, 0, stream>>>(d_off, n, pl.db->d.k, class_reads, pl.unsupported_mask, d_out.n_place, pl.totals.get());
            EPK_CUDA(cudaGetLastError());
            ++pl.n_launches;
            for (int c = 0; c < N_CLASSES; ++c)
            {
                if (!((class_mask >> c) & 1u)) { continue; }
                p.cls = static_cast<uint32_t>(c);
                p.counters = counters + c * CTR_COUNT;
                if ((pl.global_mask >> c) & 1u)
                {
                    // accumulators in HBM: this stream's region of rows, one per warp of the grid
                    p.acc_scratch = pl.acc_scratch.get() + static_cast<size_t>(acc_region) * pl.acc_region_bytes;
                    p.team = 1;
                    const unsigned grid = static_cast<unsigned>(
                        std::min<uint64_t>(pl.n_sms, (static_cast<uint64_t>(n) + MAX_WARPS_PER_CTA - 1) / MAX_WARPS_PER_CTA));
                    pick_global_kernel(pl)<<<grid, MAX_WARPS_PER_CTA * 32, CTA_SMEM_FIXED + MAX_WARPS_PER_CTA * GLOBAL_WARP_SMEM, stream>>>(p);
                    EPK_CUDA(cudaGetLastError());
                    ++pl.n_launches;
                    continue;
                }
                if (pl.warps_per_cta[c] == 0) { continue; }
                const unsigned grid = static_cast<unsigned>(
                    std::min<uint64_t>(pl.n_sms, (static_cast<uint64_t>(n) + pl.sets_fused[c] - 1) / pl.sets_fused[c]));
                p.team = pl.team[c];
                // (prefetching a descriptor buffer's lines gains 2 % at 8 warps per SM and loses 1 - 4 % at 16: gpurun r4a)
                p.prefetch = pl.prefetch_default && pl.wide[c] ? static_cast<uint32_t>(PREFETCH_NONE) : pl.prefetch;
                pick_kernel(pl, c)<<<grid, pl.sets_fused[c] * pl.team[c] * 32, pl.smem_fused[c], stream>>>(p);
                EPK_CUDA(cudaGetLastError());
                ++pl.n_launches;
            }
        }
        pl.n_reads += n;
        pl.totals_dirty = true;
        pl.last_stream = stream;
    }

    // classes present among sequences [first, last) (host offsets)
    uint32_t classes_of(const uint64_t* seq_off, uint32_t first, uint32_t last, uint32_t k)
    {
        uint32_t mask = 0;
        for (uint32_t i = first; i < last && mask != (1u << N_CLASSES) - 1u; ++i)
        {
            const uint64_t len = seq_off[i + 1] - seq_off[i];
            mask |= 1u << (len + 1 < k ? 0 : class_of_windows(len + 1 - k));
        }
        return mask;
    }

    void reserve_slot(epk_placer& pl, chunk_slot& s, size_t reads, size_t bytes, bool stage_in, bool stage_out)
    {
        const size_t K = pl.keep_at_most;
        if (reads > s.reads_cap)
        {
            EPK_CUDA(cudaStreamSynchronize(s.stream));
            const size_t cap = std::max(reads, static_cast<size_t>(pl.chunk_reads));
            s.off.resize(cap + 1);
            s.out.n_place.resize(cap);
            s.out.branch.resize(cap * K);
            s.out.count.resize(cap * K);
            s.out.score.resize(cap * K);
            s.out.lwr.resize(cap * K);
            s.h_off.release();
            s.h_out.release();
            s.reads_cap = cap;
        }
        if (bytes > s.bytes_cap)
        {
            EPK_CUDA(cudaStreamSynchronize(s.stream));
            const size_t cap = bytes + bytes / 4 + 4096;
            s.seq.resize(cap);
            s.h_seq.release();
            s.bytes_cap = cap;
        }
        if (stage_in)
        {
            if (s.h_seq.size() < s.bytes_cap) { s.h_seq.resize(s.bytes_cap); }
            if (s.h_off.size() < s.reads_cap + 1) { s.h_off.resize(s.reads_cap + 1); }
        }
        if (stage_out && s.h_out.size() < s.reads_cap * (4 + K * 20))
        {
            s.h_out.resize(s.reads_cap * (4 + K * 20));
        }
    }

    bool is_pinned(const void* p)
    {
        cudaPointerAttributes attr{};
        if (cudaPointerGetAttributes(&attr, p) != cudaSuccess)
        {
            cudaGetLastError();
            return false;
        }
        return attr.type == cudaMemoryTypeHost;
    }

    // layout of the pinned result staging of a slot: n_place | branch | count | score | lwr
    struct staged_out
    {
        uint32_t *n_place, *branch, *count;
        float* score;
        double* lwr;
        staged_out(unsigned char* base, size_t cap, size_t K)
        {
            lwr = reinterpret_cast<double*>(base);
            n_place = reinterpret_cast<uint32_t*>(base + cap * K * 8);
            branch = n_place + cap;
            count = branch + cap * K;
            score = reinterpret_cast<float*>(count + cap * K);
        }
    };

    void finish_slot(epk_placer& pl, chunk_slot& s, const epk_result& out)
    {
        if (!s.pending) { return; }
        EPK_CUDA(cudaStreamSynchronize(s.stream));
        const size_t K = pl.keep_at_most;
        const staged_out st(s.h_out.get(), s.reads_cap, K);
        const size_t first = s.pending_first, n = s.pending_n;
        std::memcpy(out.n_place + first, st.n_place, n * 4);
        std::memcpy(out.branch + first * K, st.branch, n * K * 4);
        std::memcpy(out.count + first * K, st.count, n * K * 4);
        std::memcpy(out.score + first * K, st.score, n * K * 4);
        std::memcpy(out.lwr + first * K, st.lwr, n * K * 8);
        s.pending = false;
    }
} // namespace

extern "C" epk_status epk_placer_create(const epk_db* db, uint64_t keep_at_most, double keep_factor,
                                        const epk_placer_opts* opts, epk_placer** out)
{
    return guarded([&] {
        if (!db || !out) { throw error(EPK_E_INVALID, "null argument"); }
        *out = nullptr;
        if (keep_at_most == 0 || keep_at_most > 32) { throw error(EPK_E_UNSUPPORTED, "keep_at_most must be in [1, 32]"); }
        EPK_CUDA(cudaSetDevice(db->device));
        auto pl = std::make_unique<epk_placer>();
        pl->db = db;
        pl->device = db->device;
        pl->keep_at_most = static_cast<uint32_t>(keep_at_most);
        pl->keep_factor = keep_factor;
        pl->variant = opts ? opts->variant & 0xFu : 0;
        // bits 4-7 of the variant word: 0 = this build's default, otherwise prefetch_mode + 1 (kernel experiments)
        pl->prefetch = opts && (opts->variant >> 4) ? ((opts->variant >> 4) & 0xFu) - 1u : PREFETCH_LINES;
        pl->prefetch_default = !(opts && (opts->variant >> 4));
        // A pooled database (pool.cu) is read over NVLink for the most part, and prefetch.global.L2 of a PEER address is
        // pathological: 87 k instead of 7.4 M reads/s on two GPUs (gpurun r4d / r4e; peer lines are not kept in the
        // local L2 - the instruction seems to serialise on the link). Never there, whatever the variant word says.
        if (db->pool) { pl->prefetch = PREFETCH_NONE; }
        pl->chunk_reads = opts && opts->chunk_reads ? opts->chunk_reads : (1u << 18);
        pl->auto_chunk = !(opts && opts->chunk_reads);

        int max_smem = 0;
        EPK_CUDA(cudaDeviceGetAttribute(&pl->n_sms, cudaDevAttrMultiProcessorCount, db->device));
        EPK_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, db->device));
        // warps per CTA by class of hit-count width; a class whose accumulators do not fit one SM's
        // shared memory is unsupported (its queries are reported, not placed)
        for (int c = 0; c < N_CLASSES; ++c)
        {
            const size_t per_warp = warp_smem_bytes(db->d.n_nodes, class_count_bytes(c));
            if (CTA_SMEM_FIXED + per_warp > static_cast<size_t>(max_smem))
            {
                pl->global_mask |= 1u << c; // too large for shared memory: this class keeps its accumulators in HBM
                continue;
            }
            pl->warps_per_cta[c] = static_cast<uint32_t>(
                std::min<size_t>(MAX_WARPS_PER_CTA, (static_cast<size_t>(max_smem) - CTA_SMEM_FIXED) / per_warp));
            pl->smem_bytes[c] = CTA_SMEM_FIXED + pl->warps_per_cta[c] * per_warp;
            // teams: when at most four sets fit, every set gets 8 / sets warps (a power of two) for its sweep
            pl->sets_fused[c] = pl->warps_per_cta[c];
            pl->smem_fused[c] = pl->smem_bytes[c];
            if (pl->warps_per_cta[c] <= 4 && c < 2 && pl->variant != 1)
            {
                uint32_t team = pl->warps_per_cta[c] == 1 ? 8 : pl->warps_per_cta[c] == 2 ? 4 : 2;
                const size_t per_set = per_warp + team_smem_bytes(team);
                const uint32_t sets = static_cast<uint32_t>(
                    std::min<size_t>(pl->warps_per_cta[c], (static_cast<size_t>(max_smem) - CTA_SMEM_FIXED) / per_set));
                if (sets >= 1)
                {
                    pl->team[c] = team;
                    pl->sets_fused[c] = sets;
                    pl->smem_fused[c] = CTA_SMEM_FIXED + sets * per_set;
                }
            }
            if (c < 2 && pl->variant != 1 && CTA_SMEM_FIXED + 16 * per_warp <= static_cast<size_t>(max_smem))
            {
                pl->wide[c] = true;
                pl->sets_fused[c] = 16;
                pl->smem_fused[c] = CTA_SMEM_FIXED + 16 * per_warp;
            }
            // the attribute belongs to the function, not to this placer: always the device's limit, so that
            // placers for trees of different sizes can live side by side
            EPK_CUDA(cudaFuncSetAttribute(pick_kernel(*pl, c), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
            if (c == 0)
            {
                EPK_CUDA(cudaFuncSetAttribute(pick_gather_only_kernel(*pl), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
            }
            if (c < 2)
            {
                EPK_CUDA(cudaFuncSetAttribute(pick_partial_kernel(*pl, c), cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
            }
        }
        if (pl->global_mask)
        {
            pl->acc_region_bytes = static_cast<size_t>(pl->n_sms) * MAX_WARPS_PER_CTA * global_acc_bytes(db->d.n_nodes);
            pl->acc_scratch.resize((N_SLOTS + 1) * pl->acc_region_bytes);
            EPK_CUDA(cudaMemset(pl->acc_scratch.get(), 0, pl->acc_scratch.bytes()));
            EPK_CUDA(cudaFuncSetAttribute(pick_global_kernel(*pl), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          static_cast<int>(CTA_SMEM_FIXED + MAX_WARPS_PER_CTA * GLOBAL_WARP_SMEM)));
        }

        // zero-hit queries: keep_at_most placements on branches 0..K-1, all tied, pushed through
        // std::partial_sort (place.cpp:141-156). The order it leaves them in is the library's; ask it.
        struct tied { uint32_t branch; float score; };
        std::vector<tied> fake(pl->keep_at_most);
        for (uint32_t i = 0; i < pl->keep_at_most; ++i) { fake[i] = { i, 0.0f }; }
        std::partial_sort(fake.begin(), fake.end(), fake.end(), [](const tied& a, const tied& b) { return a.score > b.score; });
        std::vector<uint32_t> order(pl->keep_at_most);
        for (uint32_t i = 0; i < pl->keep_at_most; ++i) { order[i] = fake[i].branch; }
        pl->fake_order.resize(order.size());
        EPK_CUDA(cudaMemcpy(pl->fake_order.get(), order.data(), order.size() * 4, cudaMemcpyHostToDevice));

        pl->totals.resize(CTR_COUNT);
        pl->device_counters.resize(N_CLASSES * CTR_COUNT);
        pl->device_class_reads.resize(N_CLASSES);
        EPK_CUDA(cudaMemset(pl->totals.get(), 0, pl->totals.bytes()));
        for (auto& s : pl->slots)
        {
            EPK_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
            s.counters.resize(N_CLASSES * CTR_COUNT);
            s.class_reads.resize(N_CLASSES);
        }
        *out = pl.release();
    });
}

extern "C" void epk_placer_destroy(epk_placer* pl)
{
    if (!pl) { return; }
    cudaSetDevice(pl->device);
    for (auto& s : pl->slots)
    {
        if (s.stream)
        {
            cudaStreamSynchronize(s.stream);
            cudaStreamDestroy(s.stream);
        }
    }
    delete pl;
}

extern "C" epk_status epk_place_device(epk_placer* pl, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                       const epk_result* d_out, void* stream)
{
    return guarded([&] {
        if (!pl || !d_out || (n && (!d_seq_off || !d_out->n_place || !d_out->branch || !d_out->score || !d_out->lwr || !d_out->count)))
        {
            throw error(EPK_E_INVALID, "null argument");
        }
        if (d_out->distal || d_out->pendant) { throw error(EPK_E_INVALID, "distal/pendant must be NULL for device results"); }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        const cudaStream_t s = static_cast<cudaStream_t>(stream);
        EPK_CUDA(cudaMemsetAsync(pl->totals.get(), 0, pl->totals.bytes(), s));
        pl->n_launches = 0;
        pl->n_reads = 0;
        launch(*pl, d_seq, d_seq_off, n, *d_out, pl->device_counters.get(), pl->device_class_reads.get(),
               (1u << N_CLASSES) - 1u, s, pl->trace);
    });
}

// ---- placement without any allocation (no counterpart in the reference) ----
//
// The accumulators-in-HBM form of the placement kernel (place_kernel<..., GLOBAL>) with its scratch - the launch
// counters and one accumulator row per warp - in a file-scope array of fixed size instead of cudaMalloc'ed memory.
// The array's name is a host-side stand-in in host code: its device address is asked for once per device, with that
// device current (cudaGetSymbolAddress), and that pointer is what the kernel gets.
constexpr size_t NOALLOC_SCRATCH_BYTES = static_cast<size_t>(EPK_NOALLOC_SCRATCH_BYTES);
constexpr size_t NOALLOC_COUNTER_BYTES = 256; // [N_CLASSES][CTR_COUNT] launch counters in front of the rows
static_assert(N_CLASSES * CTR_COUNT * sizeof(unsigned long long) <= NOALLOC_COUNTER_BYTES, "counters fit their corner of the scratch");
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

namespace
{
    constexpr int NOALLOC_MAX_DEVICES = 64;
    std::mutex g_noalloc_lock;
    unsigned char* g_noalloc_address[NOALLOC_MAX_DEVICES] = {}; // device address of g_noalloc_scratch, per device

    unsigned char* noalloc_scratch_of(int device)
    {
        if (device < 0 || device >= NOALLOC_MAX_DEVICES) { throw error(EPK_E_UNSUPPORTED, "device ordinal beyond 63"); }
        const std::lock_guard<std::mutex> hold(g_noalloc_lock);
        if (!g_noalloc_address[device])
        {
            void* address = nullptr; // (the device is current: the caller has set it)
            EPK_CUDA(cudaGetSymbolAddress(&address, g_noalloc_scratch));
            if (!address) { throw error(EPK_E_CUDA, "cudaGetSymbolAddress gave no address for the static scratch"); }
            g_noalloc_address[device] = static_cast<unsigned char*>(address);
        }
        return g_noalloc_address[device];
    }
}

extern "C" epk_status epk_place_device_noalloc(epk_placer* pl, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                               const epk_result* d_out, void* stream)
{
    return guarded([&] {
        if (!pl || !d_out || (n && (!d_seq_off || !d_out->n_place || !d_out->branch || !d_out->score || !d_out->lwr || !d_out->count)))
        {
            throw error(EPK_E_INVALID, "null argument");
        }
        if (d_out->distal || d_out->pendant) { throw error(EPK_E_INVALID, "distal/pendant must be NULL for device results"); }
        const uint32_t n_nodes = pl->db->d.n_nodes;
        const size_t row_bytes = global_acc_bytes(n_nodes);
        if (n_nodes > EPK_NOALLOC_MAX_NODES || NOALLOC_COUNTER_BYTES + MAX_WARPS_PER_CTA * row_bytes > NOALLOC_SCRATCH_BYTES)
        {
            throw error(EPK_E_UNSUPPORTED, "epk_place_device_noalloc: a tree of " + std::to_string(n_nodes) + " nodes is beyond the "
                                               + std::to_string(EPK_NOALLOC_MAX_NODES) + " nodes its static scratch ("
                                               + std::to_string(NOALLOC_SCRATCH_BYTES >> 20) + " MiB) is sized for");
        }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        const cudaStream_t s = static_cast<cudaStream_t>(stream);
        unsigned char* scratch = noalloc_scratch_of(pl->db->device);
        if (n == 0) { return; }

        // as many CTAs of 8 warps as there are rows for, one per SM at most
        const size_t rows_that_fit = (NOALLOC_SCRATCH_BYTES - NOALLOC_COUNTER_BYTES) / row_bytes;
        const unsigned grid = static_cast<unsigned>(std::min<uint64_t>(
            std::min<uint64_t>(pl->n_sms, rows_that_fit / MAX_WARPS_PER_CTA), (static_cast<uint64_t>(n) + MAX_WARPS_PER_CTA - 1) / MAX_WARPS_PER_CTA));
        // the array keeps what the last call left: the counters and the rows of this launch start at zero
        EPK_CUDA(cudaMemsetAsync(scratch, 0, NOALLOC_COUNTER_BYTES + static_cast<size_t>(grid) * MAX_WARPS_PER_CTA * row_bytes, s));

        place_params p{};
        p.db = pl->db->d;
        p.seq = d_seq;
        p.seq_off = d_seq_off;
        p.n = n;
        p.keep_at_most = pl->keep_at_most;
        p.prefetch = pl->prefetch;
        p.keep_factor = pl->keep_factor;
        p.n_place = d_out->n_place;
        p.branch = d_out->branch;
        p.score = d_out->score;
        p.lwr = d_out->lwr;
        p.count = d_out->count;
        p.letters = pl->db->letters.get();
        p.fake_order = pl->fake_order.get();
        p.class_reads = nullptr; // (not counted: every launch walks the batch and takes the queries of its class)
        p.totals = nullptr;      // (the placer's running counters are not touched: epk_placer_counters does not see this call)
        p.trace = nullptr;
        p.team = 1;
        p.acc_scratch = scratch + NOALLOC_COUNTER_BYTES;
        {
            const cudaError_t stale = cudaGetLastError();
            if (stale != cudaSuccess) { throw error(EPK_E_CUDA, std::string("stale CUDA error before launch: ") + cudaGetErrorString(stale)); }
        }
        // one launch per width class of the hit counts, back to back on the stream: the rows hold 32-bit counts, so every
        // class goes through the same instantiation; each launch leaves its rows zeroed for the next
        for (int c = 0; c < N_CLASSES; ++c)
        {
            p.cls = static_cast<uint32_t>(c);
            p.counters = reinterpret_cast<unsigned long long*>(scratch) + c * CTR_COUNT;
            pick_global_kernel(*pl)<<<grid, MAX_WARPS_PER_CTA * 32, CTA_SMEM_FIXED + MAX_WARPS_PER_CTA * GLOBAL_WARP_SMEM, s>>>(p);
            EPK_CUDA(cudaGetLastError());
        }
    });
}

extern "C" uint32_t epk_trace_term(uint32_t pos, uint32_t key, uint32_t group, uint32_t len)
{
    return trace_term(pos, key, group, len);
}

extern "C" epk_status epk_placer_set_trace(epk_placer* pl, uint32_t* d_trace)
{
    return guarded([&] {
        if (!pl) { throw error(EPK_E_INVALID, "null argument"); }
        pl->trace = d_trace;
    });
}

extern "C" epk_status epk_place(epk_placer* pl, const char* seq, const uint64_t* seq_off, uint32_t n, const epk_result* out)
{
    return guarded([&] {
        if (!pl || !out || (n && (!seq_off || !out->n_place || !out->branch || !out->score || !out->lwr || !out->count)))
        {
            throw error(EPK_E_INVALID, "null argument");
        }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        const size_t K = pl->keep_at_most;
        const uint32_t k = pl->db->d.k;
        // a call that failed half way (an allocation in reserve_slot, say) may have left results staged for ITS caller's
        // arrays: nothing of them is ever copied into this caller's
        struct pending_guard
        {
            epk_placer& pl;
            void clear()
            {
                for (auto& s : pl.slots)
                {
                    if (s.pending || s.rows_in_flight) { cudaStreamSynchronize(s.stream); }
                    s.pending = false;
                    s.rows_in_flight = false;
                }
            }
            ~pending_guard() { clear(); }
        } guard{ *pl };
        guard.clear();
        EPK_CUDA(cudaMemset(pl->totals.get(), 0, pl->totals.bytes()));
        pl->n_launches = 0;
        pl->n_reads = 0;
        if (n == 0) { return; }

        const bool in_pinned = is_pinned(seq) && is_pinned(seq_off);
        const bool out_pinned = is_pinned(out->n_place) && is_pinned(out->branch) && is_pinned(out->score)
                                && is_pinned(out->lwr) && is_pinned(out->count);
        const size_t max_chunk_bytes = 256u << 20;

        // distal / pendant are look-ups by branch (place.cpp:434-437; zero for the zero-hit placements): 2 x 8 bytes per
        // placement row - 1.1 GB for 10 M reads, 0.2 s on one thread. The rows of a chunk are filled on a thread of their
        // own as soon as the chunk has landed in the caller's arrays, while the GPU works on the chunks behind it.
        const bool want_lengths = out->distal || out->pendant;
        std::vector<std::future<void>> filling;
        auto fill_lengths = [pl, out, K](size_t row0, size_t rows) {
            const auto& db = *pl->db;
            for (size_t i = row0; i < row0 + rows; ++i)
            {
                for (size_t j = 0; j < K; ++j)
                {
                    const size_t at = i * K + j;
                    const bool real = j < out->n_place[i] && out->count[at] != 0;
                    if (out->distal) { out->distal[at] = real ? db.distal[out->branch[at]] : 0.0; }
                    if (out->pendant) { out->pendant[at] = real ? db.pendant[out->branch[at]] : 0.0; }
                }
            }
        };
        auto landed = [&](chunk_slot& s) { // the slot's stream is idle and its rows are in the caller's arrays
            if (s.rows_in_flight && want_lengths)
            {
                filling.push_back(std::async(std::launch::async, fill_lengths, static_cast<size_t>(s.rows_first), static_cast<size_t>(s.rows_n)));
            }
            s.rows_in_flight = false;
        };
        struct fill_guard // (an exception on the way out must not leave threads writing into the caller's arrays)
        {
            std::vector<std::future<void>>& f;
            ~fill_guard() { for (auto& x : f) { if (x.valid()) { x.wait(); } } }
        } wait_for_fills{ filling };

        uint32_t first = 0;
        int turn = 0;
        while (first < n)
        {
            // chunk = up to chunk_reads sequences and about max_chunk_bytes letters; a pass of fewer than eight such chunks
            // is cut into about eight (of 16 k reads or 32 MB at least) so that its copies overlap its launches too: one
            // chunk of 100 k 1.5-kb reads spent as long copying as computing (C1p: 4.4 M reads/s end to end, 7.8 M resident)
            uint32_t per_chunk = pl->chunk_reads;
            size_t chunk_bytes = max_chunk_bytes;
            if (pl->auto_chunk)
            {
                per_chunk = std::min<uint32_t>(per_chunk, std::max<uint32_t>(16384u, n / 8u));
                chunk_bytes = std::min<size_t>(chunk_bytes, std::max<size_t>(32u << 20, (seq_off[n] - seq_off[0]) / 8));
            }
            uint32_t last = static_cast<uint32_t>(std::min<uint64_t>(n, static_cast<uint64_t>(first) + per_chunk));
            while (last > first + 1 && seq_off[last] - seq_off[first] > chunk_bytes)
            {
                last = first + (last - first) / 2;
            }
            const uint32_t m = last - first;
            const size_t bytes = seq_off[last] - seq_off[first];
            chunk_slot& s = pl->slots[turn % N_SLOTS];
            ++turn;
            if (s.rows_in_flight)
            {
                EPK_CUDA(cudaStreamSynchronize(s.stream)); // (two younger chunks are in flight behind it)
                finish_slot(*pl, s, *out);
                landed(s);
            }
            finish_slot(*pl, s, *out);
            reserve_slot(*pl, s, m, bytes, !in_pinned, !out_pinned);

            const char* h_seq = seq + seq_off[first];
            const uint64_t* h_off = seq_off + first;
            if (!in_pinned)
            {
                EPK_CUDA(cudaStreamSynchronize(s.stream)); // staging buffers free again
                std::memcpy(s.h_seq.get(), h_seq, bytes);
                std::memcpy(s.h_off.get(), h_off, (static_cast<size_t>(m) + 1) * 8);
                h_seq = s.h_seq.get();
                h_off = s.h_off.get();
            }
            if (bytes) { EPK_CUDA(cudaMemcpyAsync(s.seq.get(), h_seq, bytes, cudaMemcpyHostToDevice, s.stream)); }
            EPK_CUDA(cudaMemcpyAsync(s.off.get(), h_off, (static_cast<size_t>(m) + 1) * 8, cudaMemcpyHostToDevice, s.stream));

            epk_result d{};
            d.n_place = s.out.n_place.get();
            d.branch = s.out.branch.get();
            d.score = s.out.score.get();
            d.lwr = s.out.lwr.get();
            d.count = s.out.count.get();
            // offsets stay absolute: bias the base pointer instead of rewriting them
            launch(*pl, s.seq.get() - seq_off[first], s.off.get(), m, d, s.counters.get(), s.class_reads.get(),
                   classes_of(seq_off, first, last, k), s.stream, nullptr, static_cast<uint32_t>((turn - 1) % N_SLOTS));

            if (out_pinned)
            {
                EPK_CUDA(cudaMemcpyAsync(out->n_place + first, d.n_place, m * 4ull, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(out->branch + first * K, d.branch, m * K * 4, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(out->count + first * K, d.count, m * K * 4, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(out->score + first * K, d.score, m * K * 4, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(out->lwr + first * K, d.lwr, m * K * 8, cudaMemcpyDeviceToHost, s.stream));
            }
            else
            {
                const staged_out st(s.h_out.get(), s.reads_cap, K);
                EPK_CUDA(cudaMemcpyAsync(st.n_place, d.n_place, m * 4ull, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(st.branch, d.branch, m * K * 4, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(st.count, d.count, m * K * 4, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(st.score, d.score, m * K * 4, cudaMemcpyDeviceToHost, s.stream));
                EPK_CUDA(cudaMemcpyAsync(st.lwr, d.lwr, m * K * 8, cudaMemcpyDeviceToHost, s.stream));
                s.pending = true;
                s.pending_first = first;
                s.pending_n = m;
            }
            s.rows_in_flight = true;
            s.rows_first = first;
            s.rows_n = m;
            first = last;
        }
        for (auto& s : pl->slots)
        {
            finish_slot(*pl, s, *out);
            EPK_CUDA(cudaStreamSynchronize(s.stream));
            landed(s);
        }
        for (auto& f : filling) { f.get(); }
        unsigned long long totals[CTR_COUNT];
        EPK_CUDA(cudaMemcpy(totals, pl->totals.get(), sizeof(totals), cudaMemcpyDeviceToHost));
        // every other row of the pass is complete (lengths included) before the rows left with n_place = 0 are reported
        if (totals[CTR_SHORT] != 0)
        {
            throw error(EPK_E_SHORT_SEQ, std::to_string(totals[CTR_SHORT]) + " sequence(s) shorter than k - 1 = "
                                             + std::to_string(k - 1) + " letters");
        }
        if (totals[CTR_UNSUPPORTED] + totals[CTR_TOO_LONG] != 0)
        {
            throw error(EPK_E_UNSUPPORTED, std::to_string(totals[CTR_UNSUPPORTED] + totals[CTR_TOO_LONG])
                                               + " sequence(s) too long for the shared-memory accumulators of a tree of "
                                               + std::to_string(pl->db->d.n_nodes) + " nodes");
        }
    });
}

// ---- one host process, several GPUs ---------------------------------------------------------------

static epk_status check_many_args(epk_placer* const* placers, uint32_t n_placers, const uint64_t* seq_off, uint32_t n,
                                  const epk_result* out)
{
    return guarded([&] {
        if (!placers || n_placers == 0 || !out) { throw error(EPK_E_INVALID, "null argument"); }
        for (uint32_t g = 0; g < n_placers; ++g)
        {
            if (!placers[g]) { throw error(EPK_E_INVALID, "null placer"); }
            if (placers[g]->keep_at_most != placers[0]->keep_at_most) { throw error(EPK_E_INVALID, "the placers must share keep_at_most"); }
            for (uint32_t h = 0; h < g; ++h)
            {
                if (placers[h] == placers[g]) { throw error(EPK_E_INVALID, "the same placer twice"); }
            }
        }
        if (n_placers == 1) { return; }
        if (n && (!seq_off || !out->n_place || !out->branch || !out->score || !out->lwr || !out->count)) { throw error(EPK_E_INVALID, "null argument"); }
    });
}

namespace
{
    // rows [first, ...) of the caller's result arrays
    epk_result result_slice(const epk_result& out, size_t first, size_t K)
    {
        epk_result r{};
        r.n_place = out.n_place + first;
        r.branch = out.branch + first * K;
        r.score = out.score + first * K;
        r.lwr = out.lwr + first * K;
        r.count = out.count + first * K;
        r.distal = out.distal ? out.distal + first * K : nullptr;
        r.pendant = out.pendant ? out.pendant + first * K : nullptr;
        return r;
    }
}

// The records of a pass are independent (the reference's only parallelism is an OpenMP loop over them,
// place.cpp:218-230): contiguous slices of the pass go to the placers - one per GPU, each over its own replica of the
// database - on one host thread each, and land in the caller's arrays in input order.
extern "C" epk_status epk_place_many(epk_placer* const* placers, uint32_t n_placers, const char* seq, const uint64_t* seq_off,
                                     uint32_t n, const epk_result* out)
{
    const epk_status args = check_many_args(placers, n_placers, seq_off, n, out);
    if (args != EPK_OK) { return args; }
    if (n_placers == 1) { return epk_place(placers[0], seq, seq_off, n, out); }
    const size_t K = placers[0]->keep_at_most;
    std::vector<epk_status> status(n_placers, EPK_OK);
    std::vector<std::string> message(n_placers);
    std::vector<std::thread> workers;
    for (uint32_t g = 0; g < n_placers; ++g)
    {
        workers.emplace_back([&, g] {
            const uint32_t first = static_cast<uint32_t>(static_cast<uint64_t>(n) * g / n_placers);
            const uint32_t last = static_cast<uint32_t>(static_cast<uint64_t>(n) * (g + 1) / n_placers);
            const epk_result part = result_slice(*out, first, K);
            // (offsets stay absolute: sequence i of the slice is seq[seq_off[first + i] ..))
            status[g] = epk_place(placers[g], seq, seq_off + first, last - first, &part);
            if (status[g] != EPK_OK) { message[g] = epk_last_error(); }
        });
    }
    for (auto& w : workers) { w.join(); }
    // the most serious condition wins; EPK_E_SHORT_SEQ / EPK_E_UNSUPPORTED leave every other row complete
    epk_status worst = EPK_OK;
    for (uint32_t g = 0; g < n_placers; ++g)
    {
        const bool soft = status[g] == EPK_E_SHORT_SEQ || status[g] == EPK_E_UNSUPPORTED;
        const bool worst_soft = worst == EPK_E_SHORT_SEQ || worst == EPK_E_UNSUPPORTED;
        if (status[g] != EPK_OK && (worst == EPK_OK || (worst_soft && !soft)))
        {
            worst = status[g];
            set_last_error("device " + std::to_string(placers[g]->device) + ": " + message[g]);
        }
    }
    return worst;
}

// ---- the random-gather ceiling of the roofline -----------------------------------------------------

extern "C" epk_status epk_gather_ceiling_device(epk_placer* pl, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                                uint32_t* d_checksum, void* stream)
{
    return guarded([&] {
        if (!pl || (n && (!d_seq_off || !d_checksum))) { throw error(EPK_E_INVALID, "null argument"); }
        if (pl->warps_per_cta[0] == 0) { throw error(EPK_E_UNSUPPORTED, "the ceiling launch exists for trees whose accumulators fit shared memory"); }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        const cudaStream_t s = static_cast<cudaStream_t>(stream);
        place_params p{};
        p.db = pl->db->d;
        p.seq = d_seq;
        p.seq_off = d_seq_off;
        p.n = n;
        p.keep_at_most = pl->keep_at_most;
        p.prefetch = pl->prefetch;
        p.n_place = d_checksum;
        p.letters = pl->db->letters.get();
        p.cls = 0;
        p.counters = pl->device_counters.get();
        p.totals = pl->totals.get();
        EPK_CUDA(cudaMemsetAsync(pl->totals.get(), 0, pl->totals.bytes(), s));
        EPK_CUDA(cudaMemsetAsync(p.counters, 0, N_CLASSES * CTR_COUNT * sizeof(unsigned long long), s));
        EPK_CUDA(cudaMemsetAsync(pl->device_class_reads.get(), 0, N_CLASSES * sizeof(uint32_t), s));
        pl->n_launches = 0;
        pl->n_reads = n;
        if (n)
        {
            const bool teams = pl->team[0] > 1; // (the 16-set form of small trees exists for the fused launch only)
            const uint32_t sets = teams ? pl->sets_fused[0] : pl->warps_per_cta[0];
            const unsigned grid = static_cast<unsigned>(std::min<uint64_t>(pl->n_sms, (static_cast<uint64_t>(n) + sets - 1) / sets));
            p.team = pl->team[0];
            pick_gather_only_kernel(*pl)<<<grid, sets * pl->team[0] * 32, teams ? pl->smem_fused[0] : pl->smem_bytes[0], s>>>(p);
            EPK_CUDA(cudaGetLastError());
            pl->n_launches = 1;
        }
        pl->totals_dirty = true;
        pl->last_stream = s;
    });
}

// ---- database-sharded mode -----------------------------------------------------------------------

extern "C" uint64_t epk_partial_row_bytes(const epk_placer* pl)
{
    return pl ? partial_row_bytes(pl->db->d.n_nodes) : 0;
}

extern "C" epk_status epk_place_partial_device(epk_placer* pl, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                               uint32_t n_owners, const uint32_t* owner_first, void* const* d_rows,
                                               void* stream)
{
    return guarded([&] {
        if (!pl || (n && (!d_seq_off || !owner_first || !d_rows))) { throw error(EPK_E_INVALID, "null argument"); }
        if (n_owners == 0 || n_owners > static_cast<uint32_t>(MAX_OWNERS)) { throw error(EPK_E_INVALID, "n_owners must be in [1, 8]"); }
        if (pl->warps_per_cta[1] == 0)
        {
            throw error(EPK_E_UNSUPPORTED, "tree too large for the shared-memory accumulators of database-sharded mode");
        }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        const cudaStream_t s = static_cast<cudaStream_t>(stream);
        place_params p{};
        p.db = pl->db->d;
        p.seq = d_seq;
        p.seq_off = d_seq_off;
        p.n = n;
        p.keep_at_most = pl->keep_at_most;
        p.prefetch = pl->prefetch;
        p.keep_factor = pl->keep_factor;
        p.letters = pl->db->letters.get();
        p.fake_order = pl->fake_order.get();
        p.n_owners = n_owners;
        p.row_format = row_format_of(*pl);
        for (uint32_t g = 0; g < n_owners; ++g)
        {
            if (owner_first[g] > owner_first[g + 1]) { throw error(EPK_E_INVALID, "owner_first must be non-decreasing"); }
            if (owner_first[g] < owner_first[g + 1] && !d_rows[g]) { throw error(EPK_E_INVALID, "null row buffer"); }
            p.owner_first[g] = owner_first[g];
            p.rows_out[g] = static_cast<unsigned char*>(d_rows[g]);
        }
        p.owner_first[n_owners] = owner_first[n_owners];
        if (owner_first[0] != 0 || owner_first[n_owners] != n) { throw error(EPK_E_INVALID, "owner_first must span [0, n]"); }

        unsigned long long* counters = pl->device_counters.get();
        uint32_t* class_reads = pl->device_class_reads.get();
        EPK_CUDA(cudaMemsetAsync(pl->totals.get(), 0, pl->totals.bytes(), s));
        EPK_CUDA(cudaMemsetAsync(counters, 0, N_CLASSES * CTR_COUNT * sizeof(unsigned long long), s));
        EPK_CUDA(cudaMemsetAsync(class_reads, 0, N_CLASSES * sizeof(uint32_t), s));
        pl->n_launches = 0;
        pl->n_reads = n;
        if (n)
        {
            const unsigned cgrid = static_cast<unsigned>(std::min<uint64_t>(4ull * pl->n_sms, (static_cast<uint64_t>(n) + 255) / 256));
            classify_kernel<<<cgrid, 256, 0, s>>>(d_seq_off, n, pl->db->d.k, class_reads, 0u, nullptr, nullptr);
            EPK_CUDA(cudaGetLastError());
            ++pl->n_launches;
            p.class_reads = class_reads;
            p.totals = pl->totals.get();
            for (int c = 0; c < 2; ++c)
            {
                p.cls = static_cast<uint32_t>(c);
                p.counters = counters + c * CTR_COUNT;
                const bool teams = pl->team[c] > 1;
                const uint32_t sets = teams ? pl->sets_fused[c] : pl->warps_per_cta[c];
                const unsigned grid = static_cast<unsigned>(std::min<uint64_t>(pl->n_sms, (static_cast<uint64_t>(n) + sets - 1) / sets));
                p.team = pl->team[c];
                pick_partial_kernel(*pl, c)<<<grid, sets * pl->team[c] * 32, teams ? pl->smem_fused[c] : pl->smem_bytes[c], s>>>(p);
                EPK_CUDA(cudaGetLastError());
                ++pl->n_launches;
            }
        }
        pl->totals_dirty = true;
        pl->last_stream = s;
    });
}

extern "C" epk_status epk_db_get_view(const epk_db* db, epk_db_view* out)
{
    return guarded([&] {
        if (!db || !out) { throw error(EPK_E_INVALID, "null argument"); }
        out->table = reinterpret_cast<uint64_t>(db->table_kind == TABLE_DIRECT ? static_cast<const void*>(db->direct.get())
                                                                               : static_cast<const void*>(db->hashed.get()));
        out->postings = reinterpret_cast<uint64_t>(db->pool ? static_cast<const void*>(db->d.postings) : static_cast<const void*>(db->postings.get()));
        out->hash_mask = db->d.hash_mask;
        out->hash_shift = db->d.hash_shift;
        out->table_kind = db->table_kind;
    });
}

extern "C" epk_status epk_placer_set_shard_views(epk_placer* pl, uint32_t n_shards, const epk_db_view* views)
{
    return guarded([&] {
        if (!pl || (n_shards && !views)) { throw error(EPK_E_INVALID, "null argument"); }
        if (n_shards > 32) { throw error(EPK_E_INVALID, "n_shards must be in [0, 32]"); }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        EPK_CUDA(cudaDeviceSynchronize()); // no launch may still be reading the old views
        pl->n_shard_views = 0;
        pl->shard_views.release();
        if (n_shards == 0) { return; }
        std::vector<device_db> host(n_shards, pl->db->d); // k, alphabet, thresholds are the group's; the tables are each shard's
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            if (!views[g].table || !views[g].postings) { throw error(EPK_E_INVALID, "null address in a shard view"); }
            if (views[g].table_kind != pl->db->table_kind) { throw error(EPK_E_INVALID, "the shards of a group share one table kind"); }
            host[g].direct = views[g].table_kind == TABLE_DIRECT ? reinterpret_cast<const uint64_t*>(views[g].table) : nullptr;
            host[g].hashed = views[g].table_kind == TABLE_HASH ? reinterpret_cast<const uint4*>(views[g].table) : nullptr;
            host[g].postings = reinterpret_cast<const uint2*>(views[g].postings);
            host[g].hash_mask = views[g].hash_mask;
            host[g].hash_shift = views[g].hash_shift;
            host[g].table_kind = views[g].table_kind;
        }
        pl->shard_views.resize(n_shards);
        EPK_CUDA(cudaMemcpy(pl->shard_views.get(), host.data(), n_shards * sizeof(device_db), cudaMemcpyHostToDevice));
        const size_t warps = static_cast<size_t>(2) * pl->n_sms * FINALIZE_WARPS, np = padded_nodes(pl->db->d.n_nodes);
        if (pl->seen_scratch.size() < warps * (np / 32))
        {
            pl->seen_scratch.resize(warps * (np / 32));
            EPK_CUDA(cudaMemset(pl->seen_scratch.get(), 0, pl->seen_scratch.bytes()));
            pl->amb_scratch.resize(warps * 6 * np);
            EPK_CUDA(cudaMemset(pl->amb_scratch.get(), 0, pl->amb_scratch.bytes()));
        }
        pl->n_shard_views = n_shards;
    });
}

extern "C" epk_status epk_place_finalize_seq_device(epk_placer* pl, void* d_rows, uint32_t n_shards, uint64_t shard_stride_bytes,
                                                    const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                                    const epk_result* d_out, void* stream)
{
    return guarded([&] {
        if (!pl || !d_out || (n && (!d_rows || !d_seq_off || !d_out->n_place || !d_out->branch || !d_out->score || !d_out->lwr || !d_out->count)))
        {
            throw error(EPK_E_INVALID, "null argument");
        }
        if (n_shards == 0 || n_shards > 32) { throw error(EPK_E_INVALID, "n_shards must be in [1, 32]"); }
        if (d_out->distal || d_out->pendant) { throw error(EPK_E_INVALID, "distal/pendant must be NULL for device results"); }
        if (shard_stride_bytes % 16 != 0 || (reinterpret_cast<uintptr_t>(d_rows) & 15u)) { throw error(EPK_E_INVALID, "rows must be 16-byte aligned"); }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        const cudaStream_t s = static_cast<cudaStream_t>(stream);
        place_params p{};
        p.db = pl->db->d;
        p.seq_off = d_seq_off;
        p.n = n;
        p.keep_at_most = pl->keep_at_most;
        p.prefetch = pl->prefetch;
        p.keep_factor = pl->keep_factor;
        p.n_place = d_out->n_place;
        p.branch = d_out->branch;
        p.score = d_out->score;
        p.lwr = d_out->lwr;
        p.count = d_out->count;
        p.fake_order = pl->fake_order.get();
        p.rows_in = static_cast<const unsigned char*>(d_rows);
        p.n_shards = n_shards;
        p.row_format = row_format_of(*pl);
        p.shard_stride = shard_stride_bytes;
        // ambiguous windows are replayed here when the owner can see every shard and has the letters
        if (d_seq && pl->n_shard_views == n_shards)
        {
            p.seq = d_seq;
            p.letters = pl->db->letters.get();
            p.shard_views = pl->shard_views.get();
            p.seen_scratch = pl->seen_scratch.get();
            p.amb_scratch = pl->amb_scratch.get();
        }
        // the counters of this call replace those of the partial launch: they now describe whole queries
        EPK_CUDA(cudaMemsetAsync(pl->totals.get(), 0, pl->totals.bytes(), s));
        p.counters = pl->totals.get();
        pl->n_launches = 0;
        pl->n_reads = n;
        if (n)
        {
            const unsigned grid = static_cast<unsigned>(
                std::min<uint64_t>(2ull * pl->n_sms, (static_cast<uint64_t>(n) + FINALIZE_WARPS - 1) / FINALIZE_WARPS));
            finalize_kernel<<<grid, FINALIZE_WARPS * 32, FINALIZE_CTA_SMEM + FINALIZE_WARPS * FINALIZE_WARP_SMEM, s>>>(p);
            EPK_CUDA(cudaGetLastError());
            ++pl->n_launches;
        }
        pl->totals_dirty = true;
        pl->last_stream = s;
    });
}

extern "C" epk_status epk_place_finalize_device(epk_placer* pl, const void* d_rows, uint32_t n_shards,
                                                uint64_t shard_stride_bytes, const uint64_t* d_seq_off, uint32_t n,
                                                const epk_result* d_out, void* stream)
{
    return epk_place_finalize_seq_device(pl, const_cast<void*>(d_rows), n_shards, shard_stride_bytes, nullptr, d_seq_off, n, d_out, stream);
}

// ---- peer memory for the fused exchange of database-sharded mode -------------------------------------

extern "C" epk_status epk_device_alloc(int device, uint64_t bytes, void** out)
{
    return guarded([&] {
        if (!out) { throw error(EPK_E_INVALID, "null argument"); }
        *out = nullptr;
        EPK_CUDA(cudaSetDevice(device));
        EPK_CUDA(cudaMalloc(out, bytes ? bytes : 1));
    });
}

extern "C" void epk_device_free(int device, void* p)
{
    if (p)
    {
        cudaSetDevice(device);
        cudaFree(p);
    }
}

extern "C" epk_status epk_ipc_export(const void* d_ptr, unsigned char* handle_out)
{
    return guarded([&] {
        if (!d_ptr || !handle_out) { throw error(EPK_E_INVALID, "null argument"); }
        static_assert(sizeof(cudaIpcMemHandle_t) == EPK_IPC_HANDLE_BYTES, "handle size");
        cudaIpcMemHandle_t h;
        EPK_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
        std::memcpy(handle_out, &h, sizeof(h));
    });
}

extern "C" epk_status epk_ipc_import(int device, const unsigned char* handle, void** out)
{
    return guarded([&] {
        if (!handle || !out) { throw error(EPK_E_INVALID, "null argument"); }
        *out = nullptr;
        EPK_CUDA(cudaSetDevice(device));
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, sizeof(h));
        EPK_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    });
}

extern "C" void epk_ipc_close(int device, void* p)
{
    if (p)
    {
        cudaSetDevice(device);
        cudaIpcCloseMemHandle(p);
    }
}

extern "C" epk_status epk_placer_counters(epk_placer* pl, epk_counters* out)
{
    return guarded([&] {
        if (!pl || !out) { throw error(EPK_E_INVALID, "null argument"); }
        EPK_CUDA(cudaSetDevice(pl->db->device));
        if (pl->last_stream || pl->totals_dirty) { EPK_CUDA(cudaStreamSynchronize(pl->last_stream)); }
        unsigned long long totals[CTR_COUNT];
        EPK_CUDA(cudaMemcpy(totals, pl->totals.get(), sizeof(totals), cudaMemcpyDeviceToHost));
        out->n_reads = pl->n_reads;
        out->n_windows = totals[CTR_WINDOWS];
        out->n_hits = totals[CTR_HITS];
        out->n_entries = totals[CTR_ENTRIES];
        out->n_short = totals[CTR_SHORT];
        out->n_unsupported = totals[CTR_UNSUPPORTED] + totals[CTR_TOO_LONG];
        out->n_launches = pl->n_launches;
    });
}

namespace
{
    // place.cpp:73-81 for records [0, n) of a batch: group_rep[i] = index of the first record with the same letters.
    // An open-addressing table of (hash, first index) reused from batch to batch: the node-per-key std::unordered_map
    // this replaces grouped 6 M reads/s on one thread, less than one GPU places (gpurun r3o: end to end on 8 GPUs).
    struct batch_grouper
    {
        std::vector<uint64_t> slot_hash;
        std::vector<uint32_t> slot_first; // 0xFFFFFFFF = free
        static uint64_t hash_of(const char* p, size_t len)
        {
            uint64_t h = 0x9E3779B97F4A7C15ull ^ (len * 0xD6E8FEB86659FD93ull);
            size_t i = 0;
            for (; i + 8 <= len; i += 8)
            {
                uint64_t w;
                std::memcpy(&w, p + i, 8);
                h = (h ^ w) * 0xBF58476D1CE4E5B9ull;
                h ^= h >> 29;
            }
            uint64_t w = 0;
            std::memcpy(&w, p + i, len - i);
            h = (h ^ w) * 0x94D049BB133111EBull;
            return h ^ (h >> 32);
        }
        void group(const char* seq, const uint64_t* seq_off, uint32_t n, uint32_t* group_rep)
        {
            size_t slots = 64;
            while (slots < 2ull * n) { slots <<= 1; }
            if (slot_first.size() != slots)
            {
                slot_hash.assign(slots, 0);
                slot_first.assign(slots, 0xFFFFFFFFu);
            }
            else { std::fill(slot_first.begin(), slot_first.end(), 0xFFFFFFFFu); }
            const size_t mask = slots - 1;
            for (uint32_t i = 0; i < n; ++i)
            {
                const char* p = seq + seq_off[i];
                const size_t len = seq_off[i + 1] - seq_off[i];
                const uint64_t h = hash_of(p, len);
                size_t at = h & mask;
                while (true)
                {
                    const uint32_t first = slot_first[at];
                    if (first == 0xFFFFFFFFu)
                    {
                        slot_first[at] = i;
                        slot_hash[at] = h;
                        group_rep[i] = i;
                        break;
                    }
                    if (slot_hash[at] == h && seq_off[first + 1] - seq_off[first] == len && std::memcmp(seq + seq_off[first], p, len) == 0)
                    {
                        group_rep[i] = first;
                        break;
                    }
                    at = (at + 1) & mask;
                }
            }
        }
    };
}

extern "C" epk_status epk_group_batch(const char* seq, const uint64_t* seq_off, uint32_t n, uint32_t* group_rep)
{
    return guarded([&] {
        if (n && (!seq_off || !group_rep)) { throw error(EPK_E_INVALID, "null argument"); }
        batch_grouper g;
        g.group(seq, seq_off, n, group_rep);
    });
}

// The grouping of a whole pass, --batch-size records at a time (main.cpp:332-344: identical sequences merge only
// within one batch), the batches spread over host threads.
extern "C" epk_status epk_group_batches(const char* seq, const uint64_t* seq_off, uint32_t n, uint32_t batch_size,
                                        uint32_t n_threads, uint32_t* group_rep)
{
    return guarded([&] {
        if (n && (!seq_off || !group_rep)) { throw error(EPK_E_INVALID, "null argument"); }
        if (batch_size == 0) { batch_size = std::max<uint32_t>(n, 1); }
        const uint32_t n_batches = (n + batch_size - 1) / batch_size;
        const uint32_t threads = std::max<uint32_t>(1, std::min<uint32_t>(std::min<uint32_t>(n_threads, n_batches), 256));
        auto work = [&](uint32_t t) {
            batch_grouper g;
            for (uint32_t b = t; b < n_batches; b += threads)
            {
                const uint32_t first = b * batch_size, last = std::min<uint64_t>(n, static_cast<uint64_t>(first) + batch_size);
                g.group(seq, seq_off + first, last - first, group_rep + first); // (indices within the batch)
            }
        };
        if (threads == 1) { work(0); return; }
        std::vector<std::thread> pool;
        for (uint32_t t = 0; t < threads; ++t) { pool.emplace_back(work, t); }
        for (auto& th : pool) { th.join(); }
    });
}

extern "C" epk_status epk_kmers(const epk_db* db, const char* seq, uint64_t len, uint64_t cap, uint32_t* pos,
                                uint32_t* key, uint32_t* group, int64_t* hit_len, uint64_t* n_out)
{
    return guarded([&] {
        if (!db || !n_out || (len && !seq) || (cap && (!pos || !key || !group || !hit_len))) { throw error(EPK_E_INVALID, "null argument"); }
        EPK_CUDA(cudaSetDevice(db->device));
        device_buffer<char> d_seq(len + 1);
        device_buffer<uint32_t> d_pos(cap + 1), d_key(cap + 1), d_group(cap + 1);
        device_buffer<long long> d_hit(cap + 1);
        device_buffer<unsigned long long> d_n(1);
        if (len) { EPK_CUDA(cudaMemcpy(d_seq.get(), seq, len, cudaMemcpyHostToDevice)); }
        if (db->table_kind == TABLE_DIRECT)
        {
            kmers_kernel<TABLE_DIRECT><<<1, 32>>>(db->d, db->letters.get(), d_seq.get(), len, cap, d_pos.get(), d_key.get(),
                                                  d_group.get(), d_hit.get(), d_n.get());
        }
        else
        {
            kmers_kernel<TABLE_HASH><<<1, 32>>>(db->d, db->letters.get(), d_seq.get(), len, cap, d_pos.get(), d_key.get(),
                                                d_group.get(), d_hit.get(), d_n.get());
        }
        EPK_CUDA(cudaGetLastError());
        unsigned long long n = 0;
        EPK_CUDA(cudaMemcpy(&n, d_n.get(), 8, cudaMemcpyDeviceToHost));
        const size_t m = static_cast<size_t>(std::min<uint64_t>(n, cap));
        if (m)
        {
            EPK_CUDA(cudaMemcpy(pos, d_pos.get(), m * 4, cudaMemcpyDeviceToHost));
            EPK_CUDA(cudaMemcpy(key, d_key.get(), m * 4, cudaMemcpyDeviceToHost));
            EPK_CUDA(cudaMemcpy(group, d_group.get(), m * 4, cudaMemcpyDeviceToHost));
            EPK_CUDA(cudaMemcpy(hit_len, d_hit.get(), m * 8, cudaMemcpyDeviceToHost));
        }
        *n_out = n;
    });
}

#ifdef EPK_PHASE_CLOCKS
// kernel experiments only (not part of the ABI): cycles per phase summed over the launches since the last call
extern "C" int epk_debug_phase_clocks(unsigned long long* out12)
{
    if (cudaDeviceSynchronize() != cudaSuccess) { return 1; }
    if (cudaMemcpyFromSymbol(out12, epk::g_phase_clocks, sizeof(unsigned long long) * 12) != cudaSuccess) { return 1; }
    unsigned long long zero[12] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 };
    return cudaMemcpyToSymbol(epk::g_phase_clocks, zero, sizeof(zero)) != cudaSuccess;
}
#endif
