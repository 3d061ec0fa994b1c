// TEST HARNESS: jplace_writer::format_arrays (the driver's path: text straight from the arrays of a device pass) must
// give the bytes of jplace_writer::format on the placed_collection built from the same arrays. Random placements with
// the doubles that exercise every branch of the number formatting, duplicate sequences, names that need escaping,
// records without placements. Prints "ok <bytes>" or the first difference.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>
#include <epik/fasta.h>
#include <epik/place.h>
#include <epik/jplace.h>

int main(int argc, char** argv)
{
    std::mt19937_64 rng(7);
    const size_t n = 3000, K = 7, batch = 700;
    const double specials[] = { 0.0, -0.0, 1.0, 100.0, 1e21, 1e22, 123456789012345678.0, 0.001, 1e-6, 1e-7, 9.999999778196308e21, 2.5e-7,
                                0.30000000000000004, 5e-324, 1.7976931348623157e308, -3.25, 1e-5, 12345.678, std::nan(""), INFINITY };
    std::vector<epik::io::seq_record> records;
    for (size_t i = 0; i < n; ++i)
    {
        std::string seq(20, 'A');
        const uint64_t pick = rng() % 400; // few distinct sequences: many duplicates inside a batch
        for (size_t j = 0; j < 20; ++j) { seq[j] = "ACGT"[(pick >> (2 * (j % 10))) & 3]; }
        std::string name = "read" + std::to_string(i);
        if (i % 97 == 0) { name += " \"quoted\"\\back\tTab\x01"; }
        records.emplace_back(name, seq);
    }
    std::vector<uint32_t> n_place(n), branch(n * K), count(n * K);
    std::vector<float> score(n * K);
    std::vector<double> lwr(n * K), distal(n * K), pendant(n * K);
    std::uniform_real_distribution<double> unit(0.0, 1.0);
    auto value = [&]() -> double {
        const uint64_t r = rng() % 10;
        if (r == 0) { return specials[rng() % (sizeof(specials) / sizeof(specials[0]))]; }
        if (r < 4) { return std::pow(10.0, unit(rng) * 60.0 - 40.0); }
        return unit(rng) * (r < 7 ? 1.0 : -300.0);
    };
    // the lengths are properties of the branch (the writer keeps their text per branch), now and then they are not
    // (the memo must check the value, not trust the key); a few branch ids beyond what the memo covers
    const size_t n_branches = 3000;
    std::vector<double> distal_of(n_branches), pendant_of(n_branches);
    for (size_t b = 0; b < n_branches; ++b) { distal_of[b] = value(); pendant_of[b] = value(); }
    for (size_t i = 0; i < n; ++i)
    {
        n_place[i] = static_cast<uint32_t>(rng() % (K + 1));
        for (size_t j = 0; j < K; ++j)
        {
            const uint32_t b = static_cast<uint32_t>(rng() % n_branches);
            branch[i * K + j] = rng() % 50 ? b : b + (1u << 22);
            score[i * K + j] = static_cast<float>(value());
            lwr[i * K + j] = value();
            distal[i * K + j] = rng() % 8 ? distal_of[b] : value();
            pendant[i * K + j] = rng() % 8 ? pendant_of[b] : value();
            count[i * K + j] = static_cast<uint32_t>(rng() % 140);
        }
    }
    epk_result rows{ n_place.data(), branch.data(), score.data(), lwr.data(), count.data(), distal.data(), pendant.data() };
    // grouping per batch as epk_group_batches leaves it: first record of the batch with the same letters
    std::vector<uint32_t> rep(n);
    for (size_t first = 0; first < n; first += batch)
    {
        const size_t last = std::min(n, first + batch);
        for (size_t i = first; i < last; ++i)
        {
            size_t j = first;
            while (records[j].sequence() != records[i].sequence()) { ++j; }
            rep[i] = static_cast<uint32_t>(j - first);
            // identical sequences share one placement (place.cpp:207): the arrays hold the same rows for them
            if (j != i)
            {
                n_place[i] = n_place[j];
            }
        }
    }
    size_t total = 0;
    for (size_t first = 0; first < n; first += batch)
    {
        const size_t last = std::min(n, first + batch);
        epik::impl::placed_collection c;
        for (size_t i = first; i < last; ++i)
        {
            const std::string_view seq(records[i].sequence());
            c.sequence_map[seq].push_back(records[i].header());
            if (rep[i] != i - first) { continue; }
            epik::impl::placed_sequence p;
            p.sequence = seq;
            for (size_t j = 0; j < n_place[i]; ++j)
            {
                const size_t at = i * K + j;
                p.placements.push_back({ branch[at], score[at], lwr[at], count[at], distal[at], pendant[at] });
            }
            c.placed_seqs.push_back(std::move(p));
        }
        const std::string want = epik::io::jplace_writer::format(c);
        size_t groups = 0;
        const std::string got = epik::io::jplace_writer::format_arrays(records, rows, K, rep.data(), first, last, &groups);
        if (got != want || groups != c.placed_seqs.size())
        {
            size_t at = 0;
            while (at < got.size() && at < want.size() && got[at] == want[at]) { ++at; }
            std::printf("batch at %zu differs at byte %zu: got ...%s\n want ...%s\n", first, at, got.substr(at > 40 ? at - 40 : 0, 120).c_str(),
                        want.substr(at > 40 ? at - 40 : 0, 120).c_str());
            return 1;
        }
        total += got.size();
    }
    // ---- the longest text there can be: names of control characters (six bytes each once escaped), every number at its
    // longest - the buffer is sized by a bound, it must hold ----
    {
        std::vector<epik::io::seq_record> worst;
        for (size_t i = 0; i < 3; ++i) { worst.emplace_back(std::string(300 + i, '\x01'), std::string(10 + i, 'A')); }
        const size_t wn = worst.size();
        std::vector<uint32_t> w_n(wn, static_cast<uint32_t>(K)), w_branch(wn * K, 4294967295u), w_count(wn * K, 1), w_rep{ 0, 1, 2 };
        std::vector<float> w_score(wn * K, -3.4028234663852886e38f);
        std::vector<double> w_lwr(wn * K, -1.7976931348623157e308), w_distal(wn * K, -2.2250738585072014e-308), w_pendant(wn * K, -1.2345678901234567e-100);
        epk_result w_rows{ w_n.data(), w_branch.data(), w_score.data(), w_lwr.data(), w_count.data(), w_distal.data(), w_pendant.data() };
        epik::impl::placed_collection c;
        for (size_t i = 0; i < wn; ++i)
        {
            c.sequence_map[worst[i].sequence()].push_back(worst[i].header());
            epik::impl::placed_sequence p;
            p.sequence = worst[i].sequence();
            for (size_t j = 0; j < K; ++j) { p.placements.push_back({ w_branch[j], w_score[j], w_lwr[j], w_count[j], w_distal[j], w_pendant[j] }); }
            c.placed_seqs.push_back(std::move(p));
        }
        const std::string want = epik::io::jplace_writer::format(c);
        std::string buffer;
        const size_t length = epik::io::jplace_writer::format_arrays_into(buffer, worst, w_rows, K, w_rep.data(), 0, wn);
        const size_t bound = epik::io::jplace_writer::arrays_text_bound(worst, K, 0, wn);
        if (length > bound || length > buffer.size() || buffer.compare(0, length, want) != 0)
        {
            std::printf("longest text: %zu bytes, bound %zu, buffer %zu, expected %zu\n", length, bound, buffer.size(), want.size());
            return 1;
        }
    }
    // ---- the whole file: format + append_many (objects, a pass at a time) against append_batches (arrays, text and
    // append in one stage), two passes each so that the second one starts in the middle of a page ----
    const std::string dir = argc > 1 ? argv[1] : "/tmp";
    auto slurp = [](const std::string& name) {
        std::FILE* f = std::fopen(name.c_str(), "rb");
        std::string all;
        char buf[65536];
        size_t got;
        while (f && (got = std::fread(buf, 1, sizeof(buf), f)) > 0) { all.append(buf, got); }
        if (f) { std::fclose(f); }
        return all;
    };
    const size_t half = 3 * batch; // records of the first pass
    const std::string want_name = dir + "/want.jplace";
    {
        epik::io::jplace_writer writer(want_name, "epik test", "(A:0.1,B:0.2);");
        writer.start();
        for (size_t pass_first : { size_t(0), half })
        {
            const size_t pass_last = pass_first == 0 ? half : n;
            std::vector<std::string> texts;
            std::vector<size_t> counts;
            for (size_t first = pass_first; first < pass_last; first += batch)
            {
                size_t groups = 0;
                texts.push_back(epik::io::jplace_writer::format_arrays(records, rows, K, rep.data(), first, std::min(pass_last, first + batch), &groups));
                counts.push_back(groups);
            }
            writer.append_many(texts, counts, 3);
        }
        writer.end();
    }
    const std::string want_file = slurp(want_name);
    for (size_t threads : { size_t(1), size_t(5), size_t(4) })
    {
        // (4 threads: without the mapping - what a nearly full file system gets, pwrite reports its errors)
        if (threads == 4) { setenv("EPK_JPLACE_NO_MMAP", "1", 1); }
        const std::string got_name = dir + "/got" + std::to_string(threads) + ".jplace";
        {
            epik::io::jplace_writer writer(got_name, "epik test", "(A:0.1,B:0.2);");
            writer.start();
            for (size_t pass_first : { size_t(0), half })
            {
                const size_t pass_last = pass_first == 0 ? half : n;
                const size_t n_batches = (pass_last - pass_first + batch - 1) / batch;
                const size_t bound = epik::io::jplace_writer::arrays_text_bound(records, K, pass_first, pass_last) + 64 * n_batches;
                writer.append_batches(n_batches, bound, threads, [&](size_t b, std::string& buffer, size_t* n_records) {
                    const size_t first = pass_first + b * batch;
                    return epik::io::jplace_writer::format_arrays_into(buffer, records, rows, K, rep.data(), first, std::min(pass_last, first + batch), n_records);
                });
            }
            writer.end();
        }
        const std::string got_file = slurp(got_name);
        if (got_file != want_file || got_file.size() < total)
        {
            size_t at = 0;
            while (at < got_file.size() && at < want_file.size() && got_file[at] == want_file[at]) { ++at; }
            std::printf("file written with %zu threads differs at byte %zu of %zu / %zu\n", threads, at, got_file.size(), want_file.size());
            return 1;
        }
    }
    // a bound that is too small is an error, and the file is left as it was (with and without the mapping)
    for (int mapped = 0; mapped < 2; ++mapped)
    {
        if (mapped) { unsetenv("EPK_JPLACE_NO_MMAP"); }
        const std::string name = dir + "/short.jplace";
        epik::io::jplace_writer writer(name, "epik test", "(A:0.1,B:0.2);");
        writer.start();
        const size_t before = slurp(name).size();
        bool refused = false;
        try
        {
            writer.append_batches(2, 1000, 2, [&](size_t b, std::string& buffer, size_t* n_records) {
                return epik::io::jplace_writer::format_arrays_into(buffer, records, rows, K, rep.data(), b * batch, (b + 1) * batch, n_records);
            });
        }
        catch (const std::runtime_error&) { refused = true; }
        if (!refused || slurp(name).size() != before)
        {
            std::printf("a text beyond its bound was not refused cleanly\n");
            return 1;
        }
    }
    // a batch whose text cannot be produced: the threads waiting for their turn behind it give up, the error comes
    // out, the file is left as it was
    {
        const std::string name = dir + "/thrown.jplace";
        epik::io::jplace_writer writer(name, "epik test", "(A:0.1,B:0.2);");
        writer.start();
        const size_t before = slurp(name).size();
        bool thrown = false;
        try
        {
            writer.append_batches(4, epik::io::jplace_writer::arrays_text_bound(records, K, 0, 4 * batch), 4,
                                  [&](size_t b, std::string& buffer, size_t* n_records) -> size_t {
                                      if (b == 1) { throw std::runtime_error("no text for this batch"); }
                                      return epik::io::jplace_writer::format_arrays_into(buffer, records, rows, K, rep.data(), b * batch, (b + 1) * batch, n_records);
                                  });
        }
        catch (const std::runtime_error&) { thrown = true; }
        if (!thrown || slurp(name).size() != before)
        {
            std::printf("an error inside a batch did not come out cleanly\n");
            return 1;
        }
        // and the writer is still usable
        writer.append_batches(1, epik::io::jplace_writer::arrays_text_bound(records, K, 0, batch), 2, [&](size_t, std::string& buffer, size_t* n_records) {
            return epik::io::jplace_writer::format_arrays_into(buffer, records, rows, K, rep.data(), 0, batch, n_records);
        });
        writer.end();
        if (slurp(name).size() <= before)
        {
            std::printf("the writer did not go on after an error\n");
            return 1;
        }
    }
    std::printf("ok %zu %zu\n", total, want_file.size());
    return 0;
}
