// TEST INFRASTRUCTURE ONLY. C wrapper around the reference's own placer.
//
// This translation unit is linked with /root/reference/epik/src/epik/place.cpp, compiled UNMODIFIED
// where it lies (see oracle/Makefile), against the i2l shim under oracle/shim/. The resulting
// libepik_ref_{dna,aa}.so goes to oracle/_ref/ (git-ignored). Nothing of the reference is copied here:
// this file only drives epik::placer the way epik/src/epik/main.cpp:332-365 does.
#include <chrono>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <i2l/phylo_kmer_db.h>
#include <i2l/phylo_tree.h>
#include <i2l/kmer_iterator.h>
#include <i2l/seq_record.h>
#include <epik/place.h>

#include "oracle_capi.h"

namespace
{
    thread_local std::string g_error;

    struct ref_handle
    {
        std::unique_ptr<i2l::phylo_kmer_db> db;
        std::unique_ptr<i2l::phylo_tree> tree;
        std::unique_ptr<epik::placer> placer;
        uint64_t keep_at_most;
        uint64_t max_threads;
    };
}

extern "C" int orc_alphabet(void)
{
#ifdef SEQ_TYPE_AA
    return 1;
#else
    return 0;
#endif
}

extern "C" const char* orc_kind(void) { return "reference"; }
extern "C" const char* orc_last_error(void) { return g_error.c_str(); }

extern "C" void* orc_create(const orc_db_desc* desc, uint64_t keep_at_most, double keep_factor, uint64_t max_threads)
{
    try
    {
        if (static_cast<int>(desc->alphabet) != orc_alphabet())
        {
            throw std::runtime_error("this reference build handles the other alphabet");
        }
        auto h = std::make_unique<ref_handle>();
        h->db = std::make_unique<i2l::phylo_kmer_db>(desc->kmer_size, desc->omega, "");
        for (uint64_t i = 0; i < desc->n_keys; ++i)
        {
            const auto begin = desc->offsets[i], end = desc->offsets[i + 1];
            static_assert(sizeof(orc_entry) == sizeof(i2l::pkdb_value));
            h->db->insert_list(desc->keys[i], reinterpret_cast<const i2l::pkdb_value*>(desc->entries + begin), end - begin);
        }
        h->db->set_num_entries_total(desc->offsets[desc->n_keys]);
        auto& index = h->db->tree_index();
        std::vector<double> lengths(desc->n_nodes);
        for (uint32_t i = 0; i < desc->n_nodes; ++i)
        {
            index.push_back({ static_cast<size_t>(desc->subtree_num_nodes[i]), desc->subtree_total_length[i] });
            lengths[i] = desc->branch_length[i];
        }
        h->tree = std::make_unique<i2l::phylo_tree>(lengths);
        h->keep_at_most = keep_at_most;
        h->max_threads = max_threads;
        /// main.cpp:295
        h->placer = std::make_unique<epik::placer>(*h->db, *h->tree, keep_at_most, keep_factor, max_threads);
        return h.release();
    }
    catch (const std::exception& e)
    {
        g_error = e.what();
        return nullptr;
    }
}

extern "C" void orc_destroy(void* handle)
{
    delete static_cast<ref_handle*>(handle);
}

extern "C" void orc_thresholds(void* handle, float* threshold, float* log10_threshold)
{
    const auto* h = static_cast<ref_handle*>(handle);
    /// same two expressions as the placer's constructor, place.cpp:87-88
    const auto thr = i2l::score_threshold(h->db->omega(), h->db->kmer_size());
    *threshold = thr;
    *log10_threshold = std::log10(thr);
}

extern "C" int orc_place(void* handle, const char* seq, const uint64_t* seq_off, uint32_t n,
                         uint32_t num_threads, uint32_t batch_size,
                         uint32_t* n_place, uint32_t* branch, float* score, double* lwr, uint64_t* count,
                         double* distal, double* pendant, uint32_t* group_rep, double* seconds)
{
    auto* h = static_cast<ref_handle*>(handle);
    try
    {
        if (num_threads > h->max_threads)
        {
            /// place.cpp:92-96,324-333: scratch is sized by the constructor's thread count
            throw std::runtime_error("num_threads exceeds the placer's max_threads");
        }
        if (batch_size == 0)
        {
            batch_size = n ? n : 1;
        }
        const uint64_t K = h->keep_at_most;
        double total = 0.0;
        for (uint32_t first = 0; first < n; first += batch_size)
        {
            const uint32_t last = std::min<uint64_t>(n, static_cast<uint64_t>(first) + batch_size);
            std::vector<i2l::seq_record> batch;
            batch.reserve(last - first);
            for (uint32_t i = first; i < last; ++i)
            {
                /// the header carries the record index so that results can be routed back
                batch.emplace_back(std::to_string(i), std::string(seq + seq_off[i], seq_off[i + 1] - seq_off[i]));
            }

            const auto t0 = std::chrono::steady_clock::now();
            const auto placed = h->placer->place(batch, num_threads);
            const auto t1 = std::chrono::steady_clock::now();
            total += std::chrono::duration<double>(t1 - t0).count();

            for (const auto& placed_seq : placed.placed_seqs)
            {
                const auto& headers = placed.sequence_map.at(placed_seq.sequence);
                const auto rep = static_cast<uint32_t>(std::stoul(std::string(headers.front())));
                for (const auto header : headers)
                {
                    const auto i = static_cast<uint64_t>(std::stoul(std::string(header)));
                    n_place[i] = static_cast<uint32_t>(placed_seq.placements.size());
                    if (group_rep) { group_rep[i] = rep; }
                    uint64_t j = i * K;
                    for (const auto& p : placed_seq.placements)
                    {
                        if (branch) { branch[j] = p.branch_id; }
                        if (score) { score[j] = p.score; }
                        if (lwr) { lwr[j] = p.weight_ratio; }
                        if (count) { count[j] = p.count; }
                        if (distal) { distal[j] = p.distal_length; }
                        if (pendant) { pendant[j] = p.pendant_length; }
                        ++j;
                    }
                }
            }
        }
        if (seconds) { *seconds = total; }
        return 0;
    }
    catch (const std::exception& e)
    {
        g_error = e.what();
        return 1;
    }
}

extern "C" uint64_t orc_kmers(void* handle, const char* seq, uint64_t len, uint64_t cap,
                              uint32_t* pos, uint32_t* key, uint32_t* group, int64_t* hit_len)
{
    const auto* h = static_cast<ref_handle*>(handle);
    const std::string_view view(seq, len);
    uint64_t n = 0;
    /// same iteration as query_kmers, place.cpp:294-313
    for (const auto& [kmer, keys] : i2l::to_kmers<i2l::one_ambiguity_policy>(view, h->db->kmer_size()))
    {
        for (const auto k : keys)
        {
            if (n < cap)
            {
                pos[n] = static_cast<uint32_t>(kmer.data() - seq);
                key[n] = k;
                group[n] = static_cast<uint32_t>(keys.size());
                const auto found = h->db->search(k);
                hit_len[n] = found ? static_cast<int64_t>((*found).size()) : -1;
            }
            ++n;
        }
    }
    return n;
}
