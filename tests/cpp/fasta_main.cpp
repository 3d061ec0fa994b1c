// TEST HARNESS: dumps what epik::io::batch_fasta reads: one line per record, "batch<TAB>header<TAB>sequence",
// then "bytes<TAB>N" with bytes_read() at the end.
#include <iostream>
#include <epik/fasta.h>

int main(int argc, char** argv)
{
    if (argc < 3) { return 2; }
    epik::io::batch_fasta reader(argv[1], static_cast<size_t>(std::stoul(argv[2])));
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
