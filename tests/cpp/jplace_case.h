// Test data shared by the two jplace harnesses (the reference's writer and ours): a few batches of
// placements whose numbers exercise every branch of the number formatting (plain, leading zeros,
// exponents of one/two/three digits, integers, zeros of both signs, denormals, floats widened to double).
#pragma once
#include <string>
#include <string_view>
#include <vector>

namespace jplace_case
{
    struct storage
    {
        std::vector<std::string> sequences, headers;
    };

    inline const std::vector<double>& numbers()
    {
        static const std::vector<double> v = {
            0.31214008, 0.0966156451, 0.044845012, 1.0, 0.1, 100.0, 1e21, 1e22, 123456789012345678901234.0, 1e-5, 1.5e-6,
            1e-7, 2.5e-7, 3.2e-310, 4.9406564584124654e-324, 1.7976931348623157e308, 0.0, -0.0, -0.25, 0.15, 0.483333,
            2.0 / 3.0, 1.0 / 3.0, 5e-324, 12345.678, 0.000001, 0.0000001, 99999999999999999999.0, 0.5, 7.0, 1234567.0,
            static_cast<double>(-0.75930208f), static_cast<double>(-1.09263539f), static_cast<double>(-1.27790618f),
            static_cast<double>(0.0999774411f), static_cast<double>(-59.8765432f), 6.02214076e23, 1.602176634e-19 };
        return v;
    }

    // Collection = placed_collection, Placed = placed_sequence, of either implementation
    template<class Collection, class Placed>
    Collection make_batch(storage& st, int batch)
    {
        Collection c;
        const auto& num = numbers();
        const int n_seq = 5;
        // reserve so that the string_views stay valid
        st.sequences.reserve(64);
        st.headers.reserve(256);
        for (int i = 0; i < n_seq; ++i)
        {
            st.sequences.push_back("ACGT" + std::to_string(batch) + "_" + std::to_string(i));
            const std::string_view seq = st.sequences.back();
            const int n_names = 1 + (i % 3);
            for (int h = 0; h < n_names; ++h)
            {
                st.headers.push_back(h == 1 ? "read \"quoted\"\tb" + std::to_string(batch) : "read_" + std::to_string(batch) + "_" + std::to_string(i) + "_" + std::to_string(h));
                c.sequence_map[seq].push_back(st.headers.back());
            }
            Placed p;
            p.sequence = seq;
            const int n_place = (i == 4) ? 0 : 1 + i * 2;
            for (int j = 0; j < n_place; ++j)
            {
                const size_t at = static_cast<size_t>(batch * 17 + i * 7 + j * 3);
                p.placements.push_back({ static_cast<unsigned>(at % 4999), static_cast<float>(num[at % num.size()]),
                                         num[(at + 1) % num.size()], static_cast<size_t>(j),
                                         num[(at + 2) % num.size()], num[(at + 5) % num.size()] });
            }
            c.placed_seqs.push_back(std::move(p));
        }
        return c;
    }

    template<class Writer, class Collection, class Placed>
    void write_all(const std::string& filename)
    {
        storage st;
        const std::string tree = "((A:0.1{0},B:0.2{1}):0.05{2},C:0.3{3}):0.0{4};";
        Writer w(filename, "epik-dna -d DB.ipk -q \"q.fasta\" ", tree);
        w.start();
        for (int b = 0; b < 3; ++b)
        {
            const Collection c = make_batch<Collection, Placed>(st, b);
            w << c;
        }
        w.end();
    }
}
