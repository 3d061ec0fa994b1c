// epik-dna / epik-aa on B200: the driver of the reference (epik/src/epik/main.cpp:205-390) with the same
// flags, defaults, messages, exit codes, output file name and jplace layout, in front of the CUDA placer.
// Built twice like the reference (epik/CMakeLists.txt:60-76,109-128): -DSEQ_TYPE_AA selects the amino-acid
// binary. Differences that are not cosmetic:
//   * the database is an `.epkdb` file (epik/flat_db.h): the `.ipk` format lives in the absent i2l submodule;
//   * FASTA records are read `--batch-size` at a time as the reference does, but many such batches are placed
//     in one device pass (--device-batches, default 128); dedupe and "nm" grouping stay per --batch-size batch;
//   * -j/--jobs sizes the host side: grouping, result objects and jplace text of the --batch-size batches of a
//     device pass are produced on that many threads (the placement itself has no host threads);
//   * --device N picks the CUDA device; --devices N replicates the database on devices 0 .. N-1 ("all": every visible
//     one) and cuts every device pass into N slices (epk_place_many): one process, one FASTA, one jplace, N GPUs.
#include <chrono>
#include <future>
#include <memory>
#include <cmath>
#include <cstdlib>
#include <iomanip>
#include <iostream>
#include <limits>
#include <map>
#include <sstream>
#include <string>
#include <vector>

#include <epik/fasta.h>
#include <epik/flat_db.h>
#include <epik/jplace.h>
#include <epik/place.h>

namespace
{
#ifdef SEQ_TYPE_AA
    constexpr uint32_t BUILD_ALPHABET = EPK_AA;
#else
    constexpr uint32_t BUILD_ALPHABET = EPK_DNA;
#endif

    /// main.cpp:23-32
    std::string make_invocation(int argc, char** argv)
    {
        std::string invocation;
        for (int i = 0; i < argc; ++i)
        {
            invocation += argv[i];
            invocation += " ";
        }
        return invocation;
    }

    /// main.cpp:34-37: <output_dir>/placements_<basename(query)>.jplace
    std::string make_output_filename(const std::string& input_file, const std::string& output_dir)
    {
        const size_t slash = input_file.find_last_of('/');
        const std::string base = slash == std::string::npos ? input_file : input_file.substr(slash + 1);
        std::string dir = output_dir;
        if (!dir.empty() && dir.back() != '/') { dir += '/'; }
        return dir + "placements_" + base + ".jplace";
    }

    /// 1234567 -> "1.2M" (the reference prints counts and speeds this way, main.cpp:126-150)
    std::string to_human_readable(double value)
    {
        static const char* suffix[] = { "", "K", "M", "B", "T" };
        size_t i = 0;
        while (value >= 1000.0 && i < 4)
        {
            value /= 1000.0;
            ++i;
        }
        std::ostringstream oss;
        oss << std::fixed << std::setprecision(i == 0 ? 0 : 1) << value << suffix[i];
        return oss.str();
    }

    /// --max-ram: number, optional unit B/K/M/G by first letter, case-insensitive (main.cpp:156-194)
    size_t parse_human_readable(const std::string& max_ram)
    {
        double value = 0.0;
        char unit = 0;
        std::stringstream ss(max_ram);
        ss >> value;
        if (ss.fail()) { throw std::runtime_error("Could not parse --max-ram parameter: wrong numerical part"); }
        if (!ss.eof())
        {
            ss >> unit;
            if (ss.fail()) { throw std::runtime_error("Could not parse --max-ram parameter: wrong unit"); }
        }
        switch (std::toupper(unit))
        {
            case 0:
            case 'B': return static_cast<size_t>(value);
            case 'K': return static_cast<size_t>(value * 1024);
            case 'M': return static_cast<size_t>(value * 1024 * 1024);
            case 'G': return static_cast<size_t>(value * 1024 * 1024 * 1024);
            default: throw std::runtime_error("Unknown memory unit.");
        }
    }

    /// The flags of main.cpp:209-222; positionals and unknown trailing words are ignored as cxxopts does.
    struct options
    {
        std::map<std::string, std::string> values;
        bool help = false;

        static std::string canonical(const std::string& name)
        {
            static const std::map<std::string, std::string> shorts = {
                { "d", "database" }, { "q", "query" }, { "j", "jobs" }, { "o", "output-dir" }, { "h", "help" } };
            const auto it = shorts.find(name);
            return it == shorts.end() ? name : it->second;
        }

        void parse(int argc, char** argv)
        {
            for (int i = 1; i < argc; ++i)
            {
                std::string arg = argv[i];
                if (arg.size() < 2 || arg[0] != '-') { continue; }
                std::string name = arg.substr(arg[1] == '-' ? 2 : 1), value;
                bool has_value = false;
                const size_t eq = name.find('=');
                if (eq != std::string::npos)
                {
                    value = name.substr(eq + 1);
                    name = name.substr(0, eq);
                    has_value = true;
                }
                name = canonical(name);
                if (name == "help")
                {
                    help = true;
                    continue;
                }
                if (name == "pool-database" && !has_value) // (the one flag without an argument)
                {
                    values[name] = "1";
                    continue;
                }
                if (!has_value)
                {
                    if (i + 1 >= argc) { throw std::runtime_error("Option '" + name + "' is missing an argument"); }
                    value = argv[++i];
                }
                values[name] = value;
            }
        }
        bool count(const std::string& name) const { return values.count(name) != 0; }
        bool flag(const std::string& name) const { return count(name) && values.at(name) != "0"; }
        std::string get(const std::string& name, const std::string& fallback = "") const
        {
            const auto it = values.find(name);
            if (it != values.end()) { return it->second; }
            if (fallback.empty()) { throw std::runtime_error("Option '" + name + "' has no value"); }
            return fallback;
        }
    };

    const char* USAGE =
        "Evolutionary Placement with Informative K-mers\n"
        "Usage:\n"
        "  epik [OPTION...]\n\n"
        "  -d, --database arg      IPK database (.epkdb)\n"
        "  -q, --query arg         Input query file (.fasta)\n"
        "  -j, --jobs arg          Num threads (default: 1)\n"
        "      --batch-size arg    Batch size (default: 2000)\n"
        "      --omega arg         Determines the threshold value (default: 1.5)\n"
        "      --mu arg            Proportion of the database to load (default: 1.0)\n"
        "      --max-ram arg       Approximate database size to load, MB\n"
        "  -o, --output-dir arg    Output directory\n"
        "      --keep-at-most arg  Number of branches to report (default: 7)\n"
        "      --keep-factor arg   Minimum LWR to report (default: 0.01)\n"
        "      --device arg        CUDA device (default: 0)\n"
        "      --devices arg       Number of CUDA devices to place on, or 'all' (default: 1)\n"
        "      --pool-database     With --devices: the database is split over the devices' memory and\n"
        "                          read in place over NVLink instead of being replicated\n"
        "      --device-batches arg  Batches placed per device pass (default: 128)\n"
        "  -h, --help              Print usage\n";
}

int main(int argc, char** argv)
{
    std::ios::sync_with_stdio(false);
    options opts;
    if (argc == 1)
    {
        std::cout << USAGE << std::endl;
        return 0;
    }
    try { opts.parse(argc, argv); }
    catch (const std::runtime_error& error)
    {
        std::cerr << "Error: " << error.what() << std::endl;
        return -1;
    }
    if (opts.help)
    {
        std::cout << USAGE << std::endl;
        return 0;
    }

    try
    {
        const std::string db_file = opts.get("database");
        const std::string query_file = opts.get("query");
        const size_t batch_size = std::stoull(opts.get("batch-size", "2000"));
        const float user_omega = std::stof(opts.get("omega", "1.5"));
        const float user_mu = std::stof(opts.get("mu", "1.0"));
        const size_t keep_at_most = std::stoull(opts.get("keep-at-most", "7"));
        const double keep_factor = std::stod(opts.get("keep-factor", "0.01"));
        const std::string output_dir = opts.get("output-dir");
        const int device = std::stoi(opts.get("device", "0"));
        const size_t device_batches = std::max<size_t>(1, std::stoull(opts.get("device-batches", "128")));
        const size_t num_threads = std::max<size_t>(1, std::stoull(opts.get("jobs", "1")));

        if (user_mu < 0.0f || user_mu > 1.0f) { throw std::runtime_error("Mu has to a value in [0, 1]"); } // main.cpp:196-202

        size_t max_entries = std::numeric_limits<size_t>::max();
        if (opts.count("max-ram"))
        {
            const size_t max_ram = parse_human_readable(opts.get("max-ram"));
            max_entries = max_ram / sizeof(epk_entry); // sizeof(i2l::pkdb_value), main.cpp:257
            if (max_entries == 0) { throw std::runtime_error("Memory limit is too low"); }
            std::cout << "Max-RAM provided: will be loaded not more than " << to_human_readable(static_cast<double>(max_entries))
                      << " phylo-k-mers." << std::endl;
        }

        std::cout << "Loading database with mu=" << user_mu << " and omega=" << user_omega << "..." << std::endl;
        const epik::flat_db flat = epik::io::load_epkdb(db_file);
        if (flat.alphabet != BUILD_ALPHABET)
        {
            throw std::runtime_error(BUILD_ALPHABET == EPK_AA ? "This is epik-aa: the database holds nucleotide k-mers"
                                                              : "This is epik-dna: the database holds amino-acid k-mers");
        }
        const epik::load_filter filter{ user_mu, user_omega, max_entries };
        const std::string devices_arg = opts.get("devices", "1");
        const size_t n_devices = devices_arg == "all" ? static_cast<size_t>(std::max(1, epk_device_count()))
                                                      : std::max<size_t>(1, std::stoull(devices_arg));
        // the database is replicated: every GPU gets the filtered, flattened copy (uploaded concurrently)
        std::vector<std::unique_ptr<epik::device_database>> replicas(n_devices);
        if (opts.flag("pool-database") && n_devices > 1)
        {
            // a database larger than one GPU: an n-th of it in each HBM, every GPU reads all of it (epk_db_pool)
            std::vector<int> ids(n_devices);
            const int present = std::max(1, epk_device_count());
            for (size_t g = 0; g < n_devices; ++g) { ids[g] = static_cast<int>(g) % present; } // (more shards than GPUs: testing)
            replicas = epik::device_database::pooled(flat, &filter, ids);
        }
        else
        {
            epik::impl::parallel_for(n_devices, n_devices, [&](size_t g) {
                replicas[g] = std::make_unique<epik::device_database>(flat, &filter, n_devices == 1 ? device : static_cast<int>(g));
            });
        }
        const epik::device_database& db = *replicas[0];

        std::cout << "Database parameters:" << std::endl
                  << "\tSequence type: " << db.sequence_type() << std::endl
                  << "\tk: " << db.kmer_size() << std::endl
                  << "\tomega: " << db.omega() << std::endl
                  << "\tPositions loaded: false" << std::endl << std::endl;
        std::cout << "Loaded " << to_human_readable(static_cast<double>(db.get_num_entries_loaded())) << " of "
                  << to_human_readable(static_cast<double>(db.get_num_entries_total())) << " phylo-k-mers. " << std::endl << std::endl;

        std::vector<const epik::device_database*> replica_ptrs;
        for (const auto& r : replicas) { replica_ptrs.push_back(r.get()); }
        epik::placer placer(replica_ptrs, keep_at_most, keep_factor, num_threads);
        if (n_devices > 1) { std::cout << "Placing on " << n_devices << " CUDA devices." << std::endl; }
        const std::string jplace_filename = make_output_filename(query_file, output_dir);
        epik::io::jplace_writer jplace(jplace_filename, make_invocation(argc, argv), flat.newick);
        jplace.start();

        std::cout << "Instruction set: CUDA sm_100a" << std::endl;
        std::cout << "Placing " << query_file << "..." << std::endl;

        const auto begin = std::chrono::steady_clock::now();
        size_t num_seq_placed = 0, num_iterations = 0;
        double average_speed = 0.0;

        epik::io::batch_fasta reader(query_file, batch_size * device_batches, num_threads);
        // Two passes are in flight: while the jplace text of one device pass is produced (on the -j threads) and
        // appended, the next pass is read and placed. A pass owns its records: the results hold views into them.
        using records_t = decltype(reader.next_batch());
        struct pass_t
        {
            records_t records;
            epik::placer::placed_arrays placed{};
        };
        double stage_ms[4] = { 0, 0, 0, 0 }; // waited for the reader, placing, waited for the writer | inside the writer
        std::future<void> writing;
        std::future<records_t> reading = std::async(std::launch::async, [&reader] { return reader.next_batch(); });
        while (true)
        {
            auto pass = std::make_shared<pass_t>();
            const auto t_read0 = std::chrono::steady_clock::now();
            pass->records = reading.get();
            stage_ms[0] += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_read0).count();
            if (pass->records.empty()) { break; }
            reading = std::async(std::launch::async, [&reader] { return reader.next_batch(); }); // parsed while this pass is placed

            const auto begin_batch = std::chrono::steady_clock::now();
            pass->placed = placer.place_arrays(pass->records, batch_size);
            const auto end_batch = std::chrono::steady_clock::now();
            const size_t n_batches = (pass->records.size() + batch_size - 1) / batch_size;

            // the reference's figure, main.cpp:348-353: nominal batch size over the time spent in place()
            double ms = std::chrono::duration<double, std::milli>(end_batch - begin_batch).count() / static_cast<double>(n_batches);
            if (ms <= 0.0) { ms = 1e-3; }
            average_speed += 1000.0 * static_cast<double>(batch_size) / ms * static_cast<double>(n_batches);
            num_iterations += n_batches;
            num_seq_placed += pass->records.size();
            stage_ms[1] += std::chrono::duration<double, std::milli>(end_batch - begin_batch).count();

            // The text of the pass before this one is read from the placer's OTHER set of result arrays while this pass is
            // placed; the pass after this one overwrites that set, so that text must be out before going round again
            const auto t_write0 = std::chrono::steady_clock::now();
            if (writing.valid()) { writing.get(); } // passes are appended in order
            stage_ms[2] += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_write0).count();

            // the writer: every -j thread formats a batch straight from the arrays of the pass and copies the text to its
            // place in the file, known as soon as the batches before it are formatted (jplace_writer::append_batches)
            writing = std::async(std::launch::async, [pass, n_batches, batch_size, &jplace, num_threads, &stage_ms] {
                const auto t0 = std::chrono::steady_clock::now();
                const auto& placed = pass->placed;
                const size_t bound = epik::io::jplace_writer::arrays_text_bound(pass->records, placed.keep_at_most, 0, placed.n) + 64 * n_batches;
                jplace.append_batches(n_batches, bound, num_threads, [&](size_t b, std::string& buffer, size_t* n_records) {
                    return epik::io::jplace_writer::format_arrays_into(buffer, pass->records, placed.rows, placed.keep_at_most, placed.group_rep,
                                                                       b * batch_size, std::min(placed.n, (b + 1) * batch_size), n_records);
                });
                stage_ms[3] += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            });
        }
        if (writing.valid()) { writing.get(); }
        jplace.end();
        if (num_iterations) { average_speed /= static_cast<double>(num_iterations); }

        std::cout << std::endl << "Placed " << num_seq_placed << " sequences.\nAverage speed: "
                  << to_human_readable(average_speed) << " seq/s.\n";
        std::cout << "Output: " << jplace_filename << std::endl;
        const auto placement_time = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - begin).count();
        std::cout << "Placement time: " << placement_time / 1000 << "s (" << placement_time << " ms)" << std::endl;
        if (std::getenv("EPK_DRIVER_TIMING"))
        {
            std::cerr << "stages (ms): waited for the reader " << stage_ms[0] << ", placing " << stage_ms[1] << ", waited for the writer "
                      << stage_ms[2] << "; inside the writer (text and append) " << stage_ms[3] << std::endl;
        }
        std::cout << "Done." << '\n' << std::flush;
    }
    catch (const std::exception& error)
    {
        std::cerr << "Error: " << error.what() << std::endl;
        return -1;
    }
    return 0;
}
