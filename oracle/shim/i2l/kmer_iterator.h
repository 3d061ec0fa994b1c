// TEST INFRASTRUCTURE ONLY (oracle). i2l shim — k-mer enumeration. PARITY UNPINNED (i2l absent).
//
// Contract from the only call site, place.cpp:294:
//     for (const auto& [kmer, keys] : i2l::to_kmers<i2l::one_ambiguity_policy>(seq, db.kmer_size()))
// keys.size()==1 -> exact k-mer; >1 -> exactly one ambiguous letter, one key per resolution.
// Windows with a gap/unknown letter or with two or more ambiguous letters yield nothing
// ([UNVERIFIED] vs upstream).
//
// Two statements of the same rule live here:
//   * encode_window(): the slow, obvious one (re-encodes each window letter by letter);
//   * to_kmers(): a lazy rolling iterator (what the timed CPU baseline runs, so that the baseline is
//     not handicapped by the shim). tests/test_oracle.py checks that they agree.
#ifndef I2L_SHIM_KMER_ITERATOR_H
#define I2L_SHIM_KMER_ITERATOR_H

#include <array>
#include <string_view>
#include <utility>
#include <vector>
#include <i2l/seq.h>

namespace i2l
{
    struct one_ambiguity_policy {};

    /// Exactly two members: place.cpp:294 destructures it as [kmer, keys].
    /// The window position is kmer.data() - seq.data().
    struct kmer_window
    {
        std::string_view kmer;
        std::vector<phylo_kmer::key_type> keys;
    };

    /// Encodes the window of `k` letters starting at `pos`. Returns false if the window is skipped.
    inline bool encode_window(std::string_view seq, size_t pos, size_t k, std::vector<phylo_kmer::key_type>& keys)
    {
        keys.clear();
        size_t ambiguous_at = k;
        std::vector<phylo_kmer::key_type> resolutions;
        phylo_kmer::key_type base = 0;
        for (size_t i = 0; i < k; ++i)
        {
            const auto l = seq::decode(seq[pos + i]);
            if (l.kind == seq::letter_kind::gap)
            {
                return false;
            }
            base <<= bits_per_letter;
            if (l.kind == seq::letter_kind::ambiguous)
            {
                if (ambiguous_at != k)
                {
                    return false;
                }
                ambiguous_at = i;
                resolutions = l.codes;
            }
            else
            {
                base |= l.codes[0];
            }
        }
        if (ambiguous_at == k)
        {
            keys.push_back(base);
        }
        else
        {
            const unsigned shift = bits_per_letter * static_cast<unsigned>(k - 1 - ambiguous_at);
            for (const auto code : resolutions)
            {
                keys.push_back(base | (code << shift));
            }
        }
        return true;
    }

    namespace impl
    {
        /// 256-entry view of seq::decode(), built once.
        struct letter_table
        {
            std::array<seq::letter, 256> letters;
            letter_table()
            {
                for (int c = 0; c < 256; ++c)
                {
                    letters[static_cast<size_t>(c)] = seq::decode(static_cast<char>(c));
                }
            }
        };
        inline const letter_table& table()
        {
            static const letter_table t;
            return t;
        }
    }

    /// Lazy range over the non-skipped windows of a sequence.
    class kmer_range
    {
    public:
        class iterator
        {
        public:
            iterator(std::string_view seq, size_t k, bool at_end)
                : _seq(seq), _k(k), _end(seq.size() < k || k == 0 ? 0 : seq.size())
                , _mask(k * bits_per_letter >= 32 ? ~0u : ((1u << (k * bits_per_letter)) - 1u))
            {
                if (at_end || _end == 0)
                {
                    _next = _end;
                    _done = true;
                    return;
                }
                advance();
            }

            const kmer_window& operator*() const { return _cur; }
            iterator& operator++() { advance(); return *this; }
            bool operator!=(const iterator& other) const { return _done != other._done; }

        private:
            /// Consumes letters until a window ending at _next-1 is emitted or the sequence ends.
            void advance()
            {
                const auto& letters = impl::table().letters;
                while (_next < _end)
                {
                    const size_t i = _next++;
                    const auto& l = letters[static_cast<unsigned char>(_seq[i])];
                    phylo_kmer::key_type code = 0;
                    if (l.kind == seq::letter_kind::gap)
                    {
                        _last_gap = static_cast<long>(i);
                    }
                    else if (l.kind == seq::letter_kind::ambiguous)
                    {
                        _prev_amb = _last_amb;
                        _last_amb = static_cast<long>(i);
                    }
                    else
                    {
                        code = l.codes[0];
                    }
                    _key = ((_key << bits_per_letter) & _mask) | code;

                    if (i + 1 < _k)
                    {
                        continue;
                    }
                    const long start = static_cast<long>(i + 1 - _k);
                    if (_last_gap >= start || _prev_amb >= start)
                    {
                        continue;
                    }
                    _cur.kmer = _seq.substr(static_cast<size_t>(start), _k);
                    _cur.keys.clear();
                    if (_last_amb >= start)
                    {
                        const unsigned shift = bits_per_letter * static_cast<unsigned>(static_cast<long>(i) - _last_amb);
                        for (const auto c : letters[static_cast<unsigned char>(_seq[static_cast<size_t>(_last_amb)])].codes)
                        {
                            _cur.keys.push_back(_key | (c << shift));
                        }
                    }
                    else
                    {
                        _cur.keys.push_back(_key);
                    }
                    return;
                }
                _done = true;
            }

            std::string_view _seq;
            size_t _k;
            size_t _end;
            phylo_kmer::key_type _mask;
            size_t _next = 0;
            phylo_kmer::key_type _key = 0;
            long _last_gap = -1;
            long _last_amb = -1;
            long _prev_amb = -1;
            bool _done = false;
            kmer_window _cur;
        };

        kmer_range(std::string_view seq, size_t k) : _seq(seq), _k(k) {}
        iterator begin() const { return iterator(_seq, _k, false); }
        iterator end() const { return iterator(_seq, _k, true); }

    private:
        std::string_view _seq;
        size_t _k;
    };

    template<class Policy>
    kmer_range to_kmers(std::string_view seq, size_t k)
    {
        return kmer_range(seq, k);
    }
}

#endif
