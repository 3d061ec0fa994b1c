// TEST HARNESS: the reference's driver loop (main.cpp:294-366) written against epik_b200's headers with an
// i2l-SHAPED database and tree (the oracle's shim types stand in for the absent i2l): constructs
// epik::placer(db, tree, keep_at_most, keep_factor, threads) exactly as main.cpp:295 does, places batch by batch
// with place(batch, threads) and writes the jplace.   usage: facade_main DB.epkdb Q.fasta OUT.jplace BATCH_SIZE
#include <iostream>
#include <i2l/phylo_kmer_db.h>
#include <i2l/phylo_tree.h>
#include <epik/fasta.h>
#include <epik/jplace.h>
#include <epik/place.h>

int main(int argc, char** argv)
{
    if (argc < 5) { return 2; }
    try
    {
        const epik::flat_db flat = epik::io::load_epkdb(argv[1]);
        i2l::phylo_kmer_db db(flat.kmer_size, flat.omega, flat.newick);
        for (size_t i = 0; i < flat.keys.size(); ++i)
        {
            db.insert_list(flat.keys[i], reinterpret_cast<const i2l::pkdb_value*>(flat.entries.data() + flat.offsets[i]),
                           flat.offsets[i + 1] - flat.offsets[i]);
        }
        for (size_t i = 0; i < flat.num_nodes(); ++i)
        {
            db.tree_index().push_back({ static_cast<size_t>(flat.subtree_num_nodes[i]), flat.subtree_total_length[i] });
        }
        const i2l::phylo_tree tree(flat.branch_length);

        const size_t keep_at_most = 7, num_threads = 1;
        const double keep_factor = 0.01;
        auto placer = epik::placer(db, tree, keep_at_most, keep_factor, num_threads);   // main.cpp:295
        auto jplace = epik::io::jplace_writer(argv[3], "facade_main ", flat.newick);
        jplace.start();
        auto reader = epik::io::batch_fasta(argv[2], std::stoul(argv[4]));
        size_t placed_total = 0;
        while (true)
        {
            const auto batch = reader.next_batch();
            if (batch.empty()) { break; }
            const auto placed_batch = placer.place(batch, num_threads);                  // main.cpp:344
            jplace << placed_batch;                                                       // main.cpp:361
            placed_total += batch.size();
        }
        jplace.end();
        std::cout << "placed " << placed_total << std::endl;
    }
    catch (const std::runtime_error& error)
    {
        std::cerr << "Error: " << error.what() << std::endl;
        return -1;
    }
    return 0;
}
