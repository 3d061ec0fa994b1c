// Host-side form of a phylo-k-mer database as flat arrays, and the `.epkdb` container it is stored in.
//
// The reference holds a database as i2l::phylo_kmer_db — a hash map key -> std::vector<pkdb_value>
// plus the tree index (call sites epik/src/epik/place.cpp:87,113-114,300; main.cpp:277-294) — read
// from an `.ipk` file by i2l::load. The i2l submodule (and with it the `.ipk` format) is absent from
// the reference tree, so this host layer carries the same content in CSR form: keys[i] owns
// entries[offsets[i] .. offsets[i+1]). `flatten()` converts any i2l-shaped database + tree into it.
#ifndef EPIK_B200_HOST_FLAT_DB_H
#define EPIK_B200_HOST_FLAT_DB_H

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <epik_b200.h>

namespace epik
{
    /// What i2l::phylo_kmer_db + i2l::phylo_tree hand to the placer, flattened.
    struct flat_db
    {
        uint32_t alphabet = EPK_DNA;              ///< EPK_DNA or EPK_AA (the reference picks it at link time)
        uint32_t kmer_size = 0;                   ///< db.kmer_size()
        float omega = 0.0f;                       ///< db.omega()
        std::vector<uint32_t> keys;               ///< unique k-mer codes
        std::vector<uint64_t> offsets;            ///< [n_keys + 1]
        std::vector<epk_entry> entries;           ///< (branch, score), pkdb_value order (place.cpp:358)
        std::vector<float> filter_value;          ///< empty or [n_keys]; larger = more informative
        std::vector<double> branch_length;        ///< by post-order id (place.cpp:103)
        std::vector<uint64_t> subtree_num_nodes;  ///< db.tree_index()[i] (place.cpp:113)
        std::vector<double> subtree_total_length; ///< db.tree_index()[i] (place.cpp:114)
        std::string newick;                       ///< the tree as jplace wants it: `label:length{edge}` (main.cpp:297)

        size_t num_nodes() const noexcept { return branch_length.size(); }
        size_t num_entries() const noexcept { return entries.size(); }

        epk_db_desc describe() const
        {
            if (offsets.size() != keys.size() + 1 || (!offsets.empty() && offsets.back() != entries.size())
                || subtree_num_nodes.size() != branch_length.size() || subtree_total_length.size() != branch_length.size()
                || (!filter_value.empty() && filter_value.size() != keys.size()))
            {
                throw std::runtime_error("Inconsistent flat database arrays");
            }
            epk_db_desc d{};
            d.alphabet = alphabet;
            d.kmer_size = kmer_size;
            d.omega = omega;
            d.n_keys = keys.size();
            d.keys = keys.data();
            d.offsets = offsets.data();
            d.entries = entries.data();
            d.filter_value = filter_value.empty() ? nullptr : filter_value.data();
            d.n_nodes = static_cast<uint32_t>(branch_length.size());
            d.branch_length = branch_length.data();
            d.subtree_num_nodes = subtree_num_nodes.data();
            d.subtree_total_length = subtree_total_length.data();
            return d;
        }
    };

    /// Flattens an i2l-shaped database and tree. Uses what EPIK's call sites prove (kmer_size(), omega(),
    /// tree_index(), get_node_count(), get_by_postorder_id(i)->get_branch_length()) plus iteration over
    /// (key, entries) pairs, which the placer itself never needs — [UNVERIFIED against upstream i2l].
    template<class Db, class Tree>
    flat_db flatten(const Db& db, const Tree& tree, uint32_t alphabet)
    {
        flat_db out;
        out.alphabet = alphabet;
        out.kmer_size = static_cast<uint32_t>(db.kmer_size());
        out.omega = static_cast<float>(db.omega());
        out.offsets.push_back(0);
        for (const auto& [key, list] : db)
        {
            out.keys.push_back(static_cast<uint32_t>(key));
            for (const auto& [branch, score] : list)
            {
                out.entries.push_back(epk_entry{ static_cast<uint32_t>(branch), static_cast<float>(score) });
            }
            out.offsets.push_back(out.entries.size());
        }
        const size_t n = tree.get_node_count();
        const auto& index = db.tree_index();
        for (size_t i = 0; i < n; ++i)
        {
            const auto node = tree.get_by_postorder_id(static_cast<uint32_t>(i));
            if (!node) { throw std::runtime_error("Could not find node by post-order id: " + std::to_string(i)); }
            out.branch_length.push_back(static_cast<double>((*node)->get_branch_length()));
            out.subtree_num_nodes.push_back(index[i].subtree_num_nodes);
            out.subtree_total_length.push_back(index[i].subtree_total_length);
        }
        return out;
    }

    namespace io
    {
        // `.epkdb`: "EPKDB1\0\0", then little-endian
        //   u32 alphabet, u32 kmer_size, f32 omega, u32 n_nodes, u64 n_keys, u64 n_entries, u64 has_filter, u64 newick_bytes,
        //   keys u32[n_keys], (pad to 8), offsets u64[n_keys+1], entries {u32,f32}[n_entries], filter f32[n_keys] (pad to 8),
        //   branch_length f64[n_nodes], subtree_num_nodes u64[n_nodes], subtree_total_length f64[n_nodes], newick bytes.
        constexpr char EPKDB_MAGIC[8] = { 'E', 'P', 'K', 'D', 'B', '1', 0, 0 };

        namespace detail
        {
            template<class T>
            void read_array(std::ifstream& in, std::vector<T>& v, size_t n, bool pad8)
            {
                v.resize(n);
                in.read(reinterpret_cast<char*>(v.data()), static_cast<std::streamsize>(n * sizeof(T)));
                if (pad8 && (n * sizeof(T)) % 8) { in.ignore(8 - (n * sizeof(T)) % 8); }
                if (!in) { throw std::runtime_error("Truncated database file"); }
            }
            template<class T>
            void write_array(std::ofstream& out, const std::vector<T>& v, bool pad8)
            {
                out.write(reinterpret_cast<const char*>(v.data()), static_cast<std::streamsize>(v.size() * sizeof(T)));
                static const char zeros[8] = { 0 };
                if (pad8 && (v.size() * sizeof(T)) % 8) { out.write(zeros, 8 - (v.size() * sizeof(T)) % 8); }
            }
        }

        inline flat_db load_epkdb(const std::string& filename)
        {
            std::ifstream in(filename, std::ios::binary);
            if (!in) { throw std::runtime_error("Could not open database file " + filename); }
            char magic[8];
            in.read(magic, 8);
            if (!in || std::memcmp(magic, EPKDB_MAGIC, 8) != 0) { throw std::runtime_error("Not an .epkdb database: " + filename); }
            flat_db db;
            uint32_t n_nodes = 0;
            uint64_t n_keys = 0, n_entries = 0, has_filter = 0, newick_bytes = 0;
            in.read(reinterpret_cast<char*>(&db.alphabet), 4);
            in.read(reinterpret_cast<char*>(&db.kmer_size), 4);
            in.read(reinterpret_cast<char*>(&db.omega), 4);
            in.read(reinterpret_cast<char*>(&n_nodes), 4);
            in.read(reinterpret_cast<char*>(&n_keys), 8);
            in.read(reinterpret_cast<char*>(&n_entries), 8);
            in.read(reinterpret_cast<char*>(&has_filter), 8);
            in.read(reinterpret_cast<char*>(&newick_bytes), 8);
            if (!in) { throw std::runtime_error("Truncated database file"); }
            // the header is not trusted: the counts must fit the file before anything is allocated for them
            {
                const auto here = in.tellg();
                in.seekg(0, std::ios::end);
                const uint64_t left = static_cast<uint64_t>(in.tellg() - here);
                in.seekg(here);
                if (n_keys > left / 12 || n_entries > left / 8 || n_nodes > left / 24 || newick_bytes > left)
                {
                    throw std::runtime_error("Corrupt database file: the header counts exceed the file size");
                }
            }
            detail::read_array(in, db.keys, n_keys, true);
            detail::read_array(in, db.offsets, n_keys + 1, false);
            // offsets describe the entries array: 0 first, non-decreasing, n_entries last (epk_db_create checks the
            // same again; here the message can name the file)
            if (db.offsets.front() != 0 || db.offsets.back() != n_entries || !std::is_sorted(db.offsets.begin(), db.offsets.end()))
            {
                throw std::runtime_error("Corrupt database file: posting-list offsets of " + filename + " are inconsistent");
            }
            detail::read_array(in, db.entries, n_entries, false);
            if (has_filter) { detail::read_array(in, db.filter_value, n_keys, true); }
            detail::read_array(in, db.branch_length, n_nodes, false);
            detail::read_array(in, db.subtree_num_nodes, n_nodes, false);
            detail::read_array(in, db.subtree_total_length, n_nodes, false);
            db.newick.resize(newick_bytes);
            in.read(db.newick.data(), static_cast<std::streamsize>(newick_bytes));
            if (!in) { throw std::runtime_error("Truncated database file"); }
            return db;
        }

        inline void save_epkdb(const flat_db& db, const std::string& filename)
        {
            std::ofstream out(filename, std::ios::binary);
            if (!out) { throw std::runtime_error("Could not create file " + filename); }
            const uint32_t n_nodes = static_cast<uint32_t>(db.num_nodes());
            const uint64_t n_keys = db.keys.size(), n_entries = db.entries.size();
            const uint64_t has_filter = db.filter_value.empty() ? 0 : 1, newick_bytes = db.newick.size();
            out.write(EPKDB_MAGIC, 8);
            out.write(reinterpret_cast<const char*>(&db.alphabet), 4);
            out.write(reinterpret_cast<const char*>(&db.kmer_size), 4);
            out.write(reinterpret_cast<const char*>(&db.omega), 4);
            out.write(reinterpret_cast<const char*>(&n_nodes), 4);
            out.write(reinterpret_cast<const char*>(&n_keys), 8);
            out.write(reinterpret_cast<const char*>(&n_entries), 8);
            out.write(reinterpret_cast<const char*>(&has_filter), 8);
            out.write(reinterpret_cast<const char*>(&newick_bytes), 8);
            detail::write_array(out, db.keys, true);
            detail::write_array(out, db.offsets, false);
            detail::write_array(out, db.entries, false);
            if (has_filter) { detail::write_array(out, db.filter_value, true); }
            detail::write_array(out, db.branch_length, false);
            detail::write_array(out, db.subtree_num_nodes, false);
            detail::write_array(out, db.subtree_total_length, false);
            out.write(db.newick.data(), static_cast<std::streamsize>(newick_bytes));
            if (!out) { throw std::runtime_error("Could not write " + filename); }
        }
    }
}

#endif
