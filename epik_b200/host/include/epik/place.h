// epik::placer on B200: the same public interface as the reference's epik/include/epik/place.h:81-140
// (constructor arguments, place(seq_records, num_threads), result types of place.h:45-75), implemented on
// the C ABI of libepik_b200.so (include/epik_b200.h). Driver code written against the reference's header
// compiles against this one unchanged; there is no CPU path behind it.
#ifndef EPIK_B200_HOST_PLACE_H
#define EPIK_B200_HOST_PLACE_H

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <exception>
#include <limits>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>

#include <epik_b200.h>
#include <epik/flat_db.h>

namespace epik::impl
{
    /// "sequence content -> headers of the records that carry it" (place.h:42)
    using sequence_map_t = std::unordered_map<std::string_view, std::vector<std::string_view>>;

    /// One placement of one sequence; field for field place.h:45-56 with i2l's types spelled out
    /// (branch_type = uint32_t, score_type = float, branch_length_type = double).
    struct placement
    {
        using weight_ratio_type = double;

        uint32_t branch_id;
        float score;
        weight_ratio_type weight_ratio;
        size_t count;
        double distal_length;
        double pendant_length;
    };

    /// Move-only like place.h:59-68.
    struct placed_sequence
    {
        std::string_view sequence;
        std::vector<placement> placements;

        placed_sequence() = default;
        placed_sequence(const placed_sequence&) = delete;
        placed_sequence(placed_sequence&&) noexcept = default;
        placed_sequence& operator=(const placed_sequence&) = delete;
        placed_sequence& operator=(placed_sequence&&) noexcept = default;
    };

    /// place.h:72-75. The string_views point into the caller's batch, which must outlive the result.
    struct placed_collection
    {
        sequence_map_t sequence_map;
        std::vector<placed_sequence> placed_seqs;
    };

    /// fn(i) for i in [0, n) on up to `threads` host threads (the -j / --jobs of the driver, main.cpp:212): the
    /// host side of a batch - grouping, result objects, jplace text - is independent per reference-sized batch
    template<class F>
    void parallel_for(size_t n, size_t threads, F&& fn)
    {
        threads = std::max<size_t>(1, std::min<size_t>(std::min(threads, n), 256));
        if (threads == 1)
        {
            for (size_t i = 0; i < n; ++i) { fn(i); }
            return;
        }
        std::vector<std::thread> pool;
        std::exception_ptr error;
        std::mutex error_lock;
        for (size_t t = 0; t < threads; ++t)
        {
            pool.emplace_back([&, t] {
                try
                {
                    for (size_t i = t; i < n; i += threads) { fn(i); }
                }
                catch (...)
                {
                    const std::lock_guard<std::mutex> hold(error_lock);
                    if (!error) { error = std::current_exception(); }
                }
            });
        }
        for (auto& th : pool) { th.join(); }
        if (error) { std::rethrow_exception(error); }
    }

    /// Pinned host memory (epk_host_alloc) that only ever grows.
    template<class T>
    class pinned_array
    {
    public:
        pinned_array() = default;
        pinned_array(const pinned_array&) = delete;
        pinned_array& operator=(const pinned_array&) = delete;
        ~pinned_array() { epk_host_free(_data); }
        T* reserve(size_t count)
        {
            if (count > _capacity)
            {
                epk_host_free(_data);
                _capacity = count + count / 4 + 16;
                _data = static_cast<T*>(epk_host_alloc(_capacity * sizeof(T)));
                if (!_data)
                {
                    _capacity = 0;
                    throw std::runtime_error(std::string("Could not allocate pinned memory: ") + epk_last_error());
                }
            }
            return _data;
        }
        T* data() const noexcept { return _data; }
    private:
        T* _data = nullptr;
        size_t _capacity = 0;
    };

    inline void check(epk_status status)
    {
        if (status != EPK_OK) { throw std::runtime_error(epk_last_error()); }
    }
}

namespace epik
{
    /// The --mu / --omega / --max-ram arguments of i2l::load (main.cpp:244-277).
    struct load_filter
    {
        double mu = 1.0;
        float omega = 1.5f;
        size_t max_entries = std::numeric_limits<size_t>::max();
    };

    /// HBM-resident database (epk_db): what `const auto db = i2l::load(...)` is to main.cpp:277.
    class device_database
    {
    public:
        device_database(const flat_db& db, const load_filter* filter = nullptr, int device = 0)
        {
            const epk_db_desc desc = db.describe();
            epk_load_filter f{};
            if (filter)
            {
                f.mu = filter->mu;
                f.omega = filter->omega;
                f.max_entries = filter->max_entries;
            }
            impl::check(epk_db_create(&desc, filter ? &f : nullptr, device, &_handle));
            impl::check(epk_db_get_info(_handle, &_info));
        }
        /// Shard `shard_index` of `shard_count` of the keys (database-sharded forms; epk_db_create_shard).
        device_database(const flat_db& db, const load_filter* filter, int device, uint32_t shard_index, uint32_t shard_count)
        {
            const epk_db_desc desc = db.describe();
            epk_load_filter f{};
            if (filter)
            {
                f.mu = filter->mu;
                f.omega = filter->omega;
                f.max_entries = filter->max_entries;
            }
            impl::check(epk_db_create_shard(&desc, filter ? &f : nullptr, device, shard_index, shard_count, &_handle));
            impl::check(epk_db_get_info(_handle, &_info));
        }

        /// The database POOLED over the memory of `devices` (epk_db_pool): shard g is built on devices[g], the shards are
        /// mapped into one address range every GPU reads, and each returned database - a whole database to a placer -
        /// sits on one of the devices. For databases that exceed one GPU; results are the unsharded database's.
        static std::vector<std::unique_ptr<device_database>> pooled(const flat_db& db, const load_filter* filter,
                                                                    const std::vector<int>& devices)
        {
            const uint32_t G = static_cast<uint32_t>(devices.size());
            std::vector<std::unique_ptr<device_database>> shards(G);
            impl::parallel_for(G, G, [&](size_t g) {
                shards[g].reset(new device_database(db, filter, devices[g], static_cast<uint32_t>(g), G));
            });
            std::vector<epk_db*> in(G), out(G, nullptr);
            for (uint32_t g = 0; g < G; ++g) { in[g] = shards[g]->_handle; }
            impl::check(epk_db_pool(in.data(), G, out.data()));
            std::vector<std::unique_ptr<device_database>> made(G);
            for (uint32_t g = 0; g < G; ++g)
            {
                made[g].reset(new device_database(out[g]));
            }
            return made; // (the shards go out of scope here: the pool holds copies)
        }
        device_database(const device_database&) = delete;
        device_database& operator=(const device_database&) = delete;
        ~device_database() noexcept { epk_db_destroy(_handle); }

        const epk_db* handle() const noexcept { return _handle; }
        /// the getters main.cpp:285-292 prints
        size_t kmer_size() const noexcept { return _info.kmer_size; }
        float omega() const noexcept { return _info.omega; }
        size_t get_num_entries_loaded() const noexcept { return _info.n_entries_loaded; }
        size_t get_num_entries_total() const noexcept { return _info.n_entries_total; }
        std::string_view sequence_type() const noexcept { return _info.alphabet == EPK_AA ? "Proteins" : "DNA"; }
        const epk_db_info& info() const noexcept { return _info; }

    private:
        explicit device_database(epk_db* handle) : _handle(handle) { impl::check(epk_db_get_info(_handle, &_info)); }
        epk_db* _handle = nullptr;
        epk_db_info _info{};
    };

    /// \brief Places a collection of fasta sequences (place.h:81-140).
    class placer
    {
        using placed_collection = impl::placed_collection;
        using placed_sequence = impl::placed_sequence;

    public:
        /// The reference's constructor (place.h:94-95) for i2l-shaped arguments: flattens the database
        /// and uploads it. `max_threads` sized per-thread scratch in the reference; it is accepted and unused.
        template<class Db, class Tree>
        placer(const Db& db, const Tree& original_tree, size_t keep_at_most, double keep_factor, size_t max_threads,
               uint32_t alphabet = EPK_DNA)
            : _owned_flat(new flat_db(flatten(db, original_tree, alphabet))),
              _owned_db(new device_database(*_owned_flat)), _db(_owned_db), _host_threads(std::max<size_t>(1, max_threads))
        {
            _owned_flat->entries.clear(); // the device holds them now
            _owned_flat->entries.shrink_to_fit();
            create(keep_at_most, keep_factor);
        }

        /// Over a database that is already on the device (it must outlive the placer, as in place.h:91-93).
        /// `max_threads` (the reference sized per-thread scratch with it) sizes the host side here: the records of a
        /// pass are packed on that many threads.
        placer(const device_database& db, size_t keep_at_most, double keep_factor, size_t max_threads = 1)
            : _db(&db), _host_threads(std::max<size_t>(1, max_threads))
        {
            create(keep_at_most, keep_factor);
        }

        /// Over replicas of one database on several GPUs: a pass is cut into one contiguous slice per replica
        /// (epk_place_many), the results come back in input order. One process, one FASTA stream, one jplace, as in
        /// main.cpp:332-366 - with every GPU of the node behind it.
        placer(const std::vector<const device_database*>& replicas, size_t keep_at_most, double keep_factor, size_t max_threads = 1)
            : _db(replicas.empty() ? nullptr : replicas.front()), _host_threads(std::max<size_t>(1, max_threads))
        {
            if (replicas.empty()) { throw std::runtime_error("No database replica given"); }
            create(keep_at_most, keep_factor);
            for (size_t g = 1; g < replicas.size(); ++g)
            {
                epk_placer* more = nullptr;
                impl::check(epk_placer_create(replicas[g]->handle(), keep_at_most, keep_factor, nullptr, &more));
                _replicas.push_back(more);
            }
        }

        placer(const placer&) = delete;
        placer(placer&&) = delete;
        placer& operator=(const placer&) = delete;
        placer& operator=(placer&&) = delete;
        ~placer() noexcept
        {
            for (epk_placer* more : _replicas) { epk_placer_destroy(more); }
            epk_placer_destroy(_handle);
            delete _owned_db;
            delete _owned_flat;
        }

        /// \brief Places a collection of fasta sequences (place.h:103, place.cpp:201-275).
        /// Record needs sequence() and header() convertible to std::string_view, as i2l::seq_record
        /// (place.cpp:73-79). Identical sequences of the batch share one placed_sequence (place.cpp:207);
        /// placed_seqs come in order of first occurrence (the reference: unordered_map order).
        template<class Record>
        placed_collection place(const std::vector<Record>& seq_records, size_t num_threads)
        {
            (void)num_threads;
            const epk_result out = run(seq_records);
            return collect(seq_records, out, 0, seq_records.size());
        }

        /// Many reference-sized batches in one device pass: the records are placed together (one
        /// epk_place call keeps the GPU busy), then grouped batch by batch exactly as consecutive
        /// place() calls on slices of `batch_size` records would (main.cpp:332-344: identical sequences
        /// merge only within one --batch-size batch).
        template<class Record>
        std::vector<placed_collection> place_batched(const std::vector<Record>& seq_records, size_t batch_size,
                                                     size_t num_threads = 1)
        {
            const epk_result out = run(seq_records);
            if (batch_size == 0) { batch_size = std::max<size_t>(1, seq_records.size()); }
            const size_t n_batches = (seq_records.size() + batch_size - 1) / batch_size;
            std::vector<placed_collection> batches(n_batches);
            impl::parallel_for(n_batches, num_threads, [&](size_t b) {
                batches[b] = collect(seq_records, out, b * batch_size, std::min(seq_records.size(), (b + 1) * batch_size));
            });
            return batches;
        }

        /// What place_batched computes, left in arrays instead of objects: rows [n][keep_at_most] of the C ABI's epk_result
        /// and, per record, the index within its --batch-size batch of the first record with the same sequence
        /// (place.cpp:73-81). Valid until the placement after the next one (two result sets take turns). A pass of a
        /// million reads costs more as 10^6 small objects than on the device; the driver formats its jplace straight
        /// from these arrays (jplace_writer::format_arrays).
        struct placed_arrays
        {
            epk_result rows;
            const uint32_t* group_rep;
            size_t keep_at_most, batch_size, n;
        };
        template<class Record>
        placed_arrays place_arrays(const std::vector<Record>& seq_records, size_t batch_size)
        {
            const epk_result out = run(seq_records);
            result_set& set = _sets[_turn % 2];
            const size_t n = seq_records.size();
            if (batch_size == 0) { batch_size = std::max<size_t>(1, n); }
            set.group_rep.resize(n + 1);
            impl::check(epk_group_batches(_letters.data(), _offsets.data(), static_cast<uint32_t>(n), static_cast<uint32_t>(batch_size),
                                          static_cast<uint32_t>(_host_threads), set.group_rep.data()));
            return placed_arrays{ out, set.group_rep.data(), _keep_at_most, batch_size, n };
        }

        /// Work counters of the last place() (windows looked up, hits, postings gathered, launches).
        epk_counters counters()
        {
            epk_counters c{};
            impl::check(epk_placer_counters(_handle, &c));
            return c;
        }

    private:
        /// packs the records into pinned memory and runs epk_place; the result arrays are the placer's own
        template<class Record>
        epk_result run(const std::vector<Record>& seq_records)
        {
            const size_t n = seq_records.size();
            if (n > std::numeric_limits<uint32_t>::max()) { throw std::runtime_error("Batch too large"); }
            const size_t K = _keep_at_most;
            uint64_t* offsets = _offsets.reserve(n + 1);
            size_t total = 0;
            for (size_t i = 0; i < n; ++i)
            {
                offsets[i] = total;
                total += std::string_view(seq_records[i].sequence()).size();
            }
            offsets[n] = total;
            char* letters = _letters.reserve(total + 1);
            // (a pass is hundreds of MB of letters: one thread copies them at a few GB/s, the GPUs eat 5 GB/s each)
            const size_t pieces = std::min<size_t>(_host_threads, std::max<size_t>(1, n / 4096));
            impl::parallel_for(pieces, pieces, [&](size_t t) {
                for (size_t i = n * t / pieces; i < n * (t + 1) / pieces; ++i)
                {
                    const std::string_view s(seq_records[i].sequence());
                    std::memcpy(letters + offsets[i], s.data(), s.size());
                }
            });

            result_set& set = _sets[++_turn % 2];
            epk_result out{};
            out.n_place = set.n_place.reserve(n + 1);
            out.branch = set.branch.reserve(n * K + 1);
            out.score = set.score.reserve(n * K + 1);
            out.lwr = set.lwr.reserve(n * K + 1);
            out.count = set.count.reserve(n * K + 1);
            out.distal = set.distal.reserve(n * K + 1);
            out.pendant = set.pendant.reserve(n * K + 1);
            if (_replicas.empty()) { impl::check(epk_place(_handle, letters, offsets, static_cast<uint32_t>(n), &out)); }
            else
            {
                std::vector<epk_placer*> all{ _handle };
                all.insert(all.end(), _replicas.begin(), _replicas.end());
                impl::check(epk_place_many(all.data(), static_cast<uint32_t>(all.size()), letters, offsets, static_cast<uint32_t>(n), &out));
            }
            return out;
        }

        /// group_by_sequence_content + the result types of place.h for records [first, last)
        template<class Record>
        placed_collection collect(const std::vector<Record>& seq_records, const epk_result& out, size_t first, size_t last)
        {
            const size_t K = _keep_at_most, n = last - first;
            std::vector<uint32_t> rep(n + 1); // (called concurrently for different batches)
            impl::check(epk_group_batch(_letters.data(), _offsets.data() + first, static_cast<uint32_t>(n), rep.data()));
            placed_collection result;
            result.sequence_map.reserve(n);
            for (size_t i = first; i < last; ++i)
            {
                const std::string_view seq(seq_records[i].sequence());
                result.sequence_map[seq].push_back(std::string_view(seq_records[i].header()));
                if (rep[i - first] != i - first) { continue; }
                placed_sequence placed;
                placed.sequence = seq;
                placed.placements.reserve(out.n_place[i]);
                for (size_t j = 0; j < out.n_place[i]; ++j)
                {
                    const size_t at = i * K + j;
                    placed.placements.push_back({ out.branch[at], out.score[at], out.lwr[at], out.count[at],
                                                  out.distal[at], out.pendant[at] });
                }
                result.placed_seqs.push_back(std::move(placed));
            }
            return result;
        }

        void create(size_t keep_at_most, double keep_factor)
        {
            _keep_at_most = keep_at_most;
            impl::check(epk_placer_create(_db->handle(), keep_at_most, keep_factor, nullptr, &_handle));
        }

        flat_db* _owned_flat = nullptr;
        device_database* _owned_db = nullptr;
        const device_database* _db = nullptr;
        epk_placer* _handle = nullptr;
        std::vector<epk_placer*> _replicas; // placers on the other GPUs (database replicated), epk_place_many
        size_t _keep_at_most = 0;
        size_t _host_threads = 1;

        impl::pinned_array<char> _letters;
        impl::pinned_array<uint64_t> _offsets;
        /// two sets of result arrays take turns: the results of a pass stay valid while the NEXT pass is placed (the
        /// driver formats the text of one pass while the next one is on the device)
        struct result_set
        {
            impl::pinned_array<uint32_t> n_place, branch, count;
            impl::pinned_array<float> score;
            impl::pinned_array<double> lwr, distal, pendant;
            std::vector<uint32_t> group_rep;
        };
        result_set _sets[2];
        size_t _turn = 0;
    };
}

#endif
