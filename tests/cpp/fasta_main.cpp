// TEST HARNESS: dumps what epik::io::batch_fasta reads: one line per record, "batch<TAB>header<TAB>sequence",
// then "bytes<TAB>N" with bytes_read() at the end.
#include <iostream>
#include <epik/fasta.h>

int main(int argc, char** argv)
{
    if (argc < 3) { return 2; }
    // optional: host threads, bytes per read block (tiny blocks put every kind of boundary inside a record)
    const size_t threads = argc > 3 ? std::stoul(argv[3]) : 1, block = argc > 4 ? std::stoul(argv[4]) : (32u << 20);
    epik::io::batch_fasta reader(argv[1], static_cast<size_t>(std::stoul(argv[2])), threads, block);
    if (threads > 1) { reader.set_records_per_thread(1); }
    size_t b = 0;
    while (true)
    {
        const auto batch = reader.next_batch();
        if (batch.empty()) { break; }
        for (const auto& r : batch) { std::cout << b << '\t' << r.header() << '\t' << r.sequence() << '\n'; }
        ++b;
    }
    std::cout << "bytes\t" << reader.bytes_read() << '\n';
    return 0;
}
