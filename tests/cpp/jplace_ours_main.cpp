// TEST HARNESS: writes the shared test data with epik_b200's jplace writer (no GPU involved).
//   jp_ours <file>            the reference's call sequence: start(), operator<< per batch, end()
//   jp_ours <file> parallel   the driver's: batches formatted on threads (jplace_writer::format), the pieces
//                             of a "device pass" written with append_many (concurrent pwrites); two passes, the
//                             first of them with an empty batch in front
#include <string>
#include <vector>
#include <epik/place.h>
#include <epik/jplace.h>
#include "jplace_case.h"

int main(int argc, char** argv)
{
    if (argc < 2) { return 2; }
    using collection = epik::impl::placed_collection;
    using placed = epik::impl::placed_sequence;
    if (argc == 2)
    {
        jplace_case::write_all<epik::io::jplace_writer, collection, placed>(argv[1]);
        return 0;
    }
    jplace_case::storage st;
    const std::string tree = "((A:0.1{0},B:0.2{1}):0.05{2},C:0.3{3}):0.0{4};";
    epik::io::jplace_writer w(argv[1], "epik-dna -d DB.ipk -q \"q.fasta\" ", tree);
    w.start();
    std::vector<collection> batches;
    for (int b = 0; b < 3; ++b) { batches.push_back(jplace_case::make_batch<collection, placed>(st, b)); }
    const std::vector<std::vector<int>> passes = { { -1, 0, 1 }, { 2 } }; // -1: a batch without records
    for (const auto& pass : passes)
    {
        const collection empty;
        std::vector<std::string> text(pass.size());
        std::vector<size_t> n_records(pass.size());
        epik::impl::parallel_for(pass.size(), 3, [&](size_t i) {
            const collection& c = pass[i] < 0 ? empty : batches[static_cast<size_t>(pass[i])];
            text[i] = epik::io::jplace_writer::format(c);
            n_records[i] = c.placed_seqs.size();
        });
        w.append_many(text, n_records, 3);
    }
    w.end();
    return 0;
}
