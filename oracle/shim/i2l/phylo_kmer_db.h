// TEST INFRASTRUCTURE ONLY (oracle). i2l shim — phylo-k-mer database. PARITY UNPINNED (i2l absent).
//
// Shape required by EPIK's call sites:
//   place.cpp:87        db.omega(), db.kmer_size()
//   place.cpp:113-114   db.tree_index()[postorder_id].{subtree_num_nodes, subtree_total_length}
//   place.cpp:280,300   db.search(key) -> optional-like, copyable, `explicit bool`, operator* -> range of
//                       pkdb_value (.size(), operator[], iteration)
//   main.cpp:278-292    version(), sequence_type(), positions_loaded(), get_num_entries_loaded/total(), tree()
// Storage mirrors what upstream does as far as it is known (a hash map from key to a heap-allocated
// vector of (branch, score)); the map here is a plain linear-probing table so that the timed CPU
// baseline is not slowed down by std::unordered_map.
#ifndef I2L_SHIM_PHYLO_KMER_DB_H
#define I2L_SHIM_PHYLO_KMER_DB_H

#include <limits>
#include <stdexcept>
#include <i2l/phylo_kmer.h>

namespace i2l
{
    namespace protocol
    {
        /// main.cpp:278 — first serialization protocol version that carries the tree index.
        constexpr unsigned EARLIEST_INDEX = 4;
        constexpr unsigned CURRENT = 5;
    }

    struct tree_index_entry
    {
        size_t subtree_num_nodes;
        double subtree_total_length;
    };

    /// Non-owning optional reference (std::optional<T&> does not exist in C++17).
    template<class T>
    class optional_ref
    {
    public:
        optional_ref() = default;
        explicit optional_ref(const T& value) : _ptr(&value) {}
        explicit operator bool() const noexcept { return _ptr != nullptr; }
        const T& operator*() const noexcept { return *_ptr; }
        const T* operator->() const noexcept { return _ptr; }
    private:
        const T* _ptr = nullptr;
    };

    class phylo_kmer_db
    {
    public:
        using key_type = phylo_kmer::key_type;
        using value_type = std::vector<pkdb_value>;
        using search_result = optional_ref<value_type>;

        phylo_kmer_db(size_t kmer_size, phylo_kmer::score_type omega, std::string tree)
            : _kmer_size(kmer_size), _omega(omega), _tree(std::move(tree))
        {
            rehash(1024);
        }

        search_result search(key_type key) const noexcept
        {
            size_t i = bucket_of(key);
            while (true)
            {
                const auto& s = _slots[i];
                if (s.value == npos) { return search_result(); }
                if (s.key == key) { return search_result(_values[s.value]); }
                i = (i + 1) & _mask;
            }
        }

        /// Appends one (branch, score) pair to the list of `key`, creating the list if needed.
        void unsafe_insert(key_type key, const pkdb_value& value)
        {
            list_of(key).push_back(value);
            ++_num_entries_loaded;
        }

        /// Bulk variant used by the C API: the whole list of a key at once.
        void insert_list(key_type key, const pkdb_value* values, size_t n)
        {
            auto& list = list_of(key);
            list.insert(list.end(), values, values + n);
            _num_entries_loaded += n;
        }

        /// Iteration over (key, entries) pairs — what a consumer that flattens the database needs (the
        /// placer of the reference never iterates; upstream's iteration API is [UNVERIFIED]).
        class const_iterator
        {
        public:
            const_iterator(const phylo_kmer_db* db, size_t slot) : _db(db), _slot(slot) { skip(); }
            std::pair<key_type, const value_type&> operator*() const
            {
                return { _db->_slots[_slot].key, _db->_values[_db->_slots[_slot].value] };
            }
            const_iterator& operator++() { ++_slot; skip(); return *this; }
            bool operator!=(const const_iterator& o) const noexcept { return _slot != o._slot; }
        private:
            void skip() { while (_slot < _db->_slots.size() && _db->_slots[_slot].value == npos) { ++_slot; } }
            const phylo_kmer_db* _db;
            size_t _slot;
        };
        const_iterator begin() const { return const_iterator(this, 0); }
        const_iterator end() const { return const_iterator(this, _slots.size()); }

        size_t size() const noexcept { return _values.size(); }
        size_t kmer_size() const noexcept { return _kmer_size; }
        phylo_kmer::score_type omega() const noexcept { return _omega; }
        void set_omega(phylo_kmer::score_type omega) noexcept { _omega = omega; }
        std::string_view tree() const noexcept { return _tree; }
        const std::vector<tree_index_entry>& tree_index() const noexcept { return _tree_index; }
        std::vector<tree_index_entry>& tree_index() noexcept { return _tree_index; }
        unsigned version() const noexcept { return protocol::CURRENT; }
        std::string_view sequence_type() const noexcept { return sequence_type_name; }
        bool positions_loaded() const noexcept { return false; }
        size_t get_num_entries_loaded() const noexcept { return _num_entries_loaded; }
        size_t get_num_entries_total() const noexcept { return _num_entries_total; }
        void set_num_entries_total(size_t n) noexcept { _num_entries_total = n; }

    private:
        static constexpr size_t npos = std::numeric_limits<size_t>::max();
        struct slot { key_type key = 0; size_t value = npos; };

        size_t bucket_of(key_type key) const noexcept
        {
            uint64_t h = key;
            h *= 0x9E3779B97F4A7C15ull;
            return static_cast<size_t>(h >> 29) & _mask;
        }

        void rehash(size_t capacity)
        {
            std::vector<slot> old = std::move(_slots);
            _slots.assign(capacity, slot{});
            _mask = capacity - 1;
            for (const auto& s : old)
            {
                if (s.value == npos) { continue; }
                size_t i = bucket_of(s.key);
                while (_slots[i].value != npos) { i = (i + 1) & _mask; }
                _slots[i] = s;
            }
        }

        value_type& list_of(key_type key)
        {
            if ((_values.size() + 1) * 2 > _slots.size())
            {
                rehash(_slots.size() * 2);
            }
            size_t i = bucket_of(key);
            while (_slots[i].value != npos)
            {
                if (_slots[i].key == key) { return _values[_slots[i].value]; }
                i = (i + 1) & _mask;
            }
            _slots[i].key = key;
            _slots[i].value = _values.size();
            _values.emplace_back();
            return _values.back();
        }

        size_t _kmer_size;
        phylo_kmer::score_type _omega;
        std::string _tree;
        std::vector<tree_index_entry> _tree_index;
        std::vector<slot> _slots;
        size_t _mask = 0;
        std::vector<value_type> _values;
        size_t _num_entries_loaded = 0;
        size_t _num_entries_total = 0;
    };
}

#endif
