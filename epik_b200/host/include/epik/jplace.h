// jplace (v3) output with the reference's layout: epik/src/epik/jplace.cpp:9-158, which drives a RapidJSON
// PrettyWriter (4-space indent) and flips kFormatSingleLineArray on and off to put every placement row
// and every [name, multiplicity] pair on one line. RapidJSON is not a dependency here: json_pretty_writer
// below restates the layout rules of that writer (where a value gets a newline, an indent, ", " or ": "),
// and the numbers are printed as RapidJSON prints them (shortest round-trip digits, plain notation for
// decimal exponents in (-6, 21], `d.ddde-7` otherwise, `0.0` / `-0.0` for zeros). Digits come from
// std::to_chars (shortest round trip); RapidJSON's Grisu2 is shortest for all but about 0.1 % of doubles,
// where it prints one more digit — those numbers parse to the same double but are not byte-identical.
#ifndef EPIK_B200_HOST_JPLACE_H
#define EPIK_B200_HOST_JPLACE_H

#include <atomic>
#include <charconv>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <string_view>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/statvfs.h>
#include <unistd.h>

#include <epik/place.h>

namespace epik::io
{
    /// The subset of a pretty-printing JSON writer jplace.cpp uses.
    class json_pretty_writer
    {
    public:
        explicit json_pretty_writer(std::string& out) : _out(out) {}

        void single_line_arrays(bool on) noexcept { _single_line = on; }

        void start_object() { prefix(); _levels.push_back({ 0, false }); _out.push_back('{'); }
        void end_object()
        {
            const bool empty = _levels.back().count == 0;
            _levels.pop_back();
            if (!empty) { _out.push_back('\n'); indent(); }
            _out.push_back('}');
        }
        void start_array() { prefix(); _levels.push_back({ 0, true }); _out.push_back('['); }
        void end_array()
        {
            const bool empty = _levels.back().count == 0;
            _levels.pop_back();
            if (!empty && !_single_line) { _out.push_back('\n'); indent(); }
            _out.push_back(']');
        }
        void key(std::string_view s) { string(s); }
        void string(std::string_view s)
        {
            prefix();
            static const char hex[] = "0123456789ABCDEF";
            _out.push_back('"');
            for (const char ch : s)
            {
                const unsigned char c = static_cast<unsigned char>(ch);
                switch (c)
                {
                    case '"': _out += "\\\""; break;
                    case '\\': _out += "\\\\"; break;
                    case '\b': _out += "\\b"; break;
                    case '\f': _out += "\\f"; break;
                    case '\n': _out += "\\n"; break;
                    case '\r': _out += "\\r"; break;
                    case '\t': _out += "\\t"; break;
                    default:
                        if (c < 0x20)
                        {
                            _out += "\\u00";
                            _out.push_back(hex[c >> 4]);
                            _out.push_back(hex[c & 15]);
                        }
                        else { _out.push_back(ch); }
                }
            }
            _out.push_back('"');
        }
        void uint(uint64_t v)
        {
            prefix();
            char buf[24];
            const auto r = std::to_chars(buf, buf + sizeof(buf), v);
            _out.append(buf, r.ptr);
        }
        void number(double v)
        {
            prefix();
            if (std::isnan(v) || std::isinf(v)) { return; } // RapidJSON refuses them after the prefix
            if (v == 0.0)
            {
                _out += std::signbit(v) ? "-0.0" : "0.0";
                return;
            }
            if (v < 0.0)
            {
                _out.push_back('-');
                v = -v;
            }
            // shortest round-trip digits and decimal exponent: d.ddddde[+-]XX
            char sci[40];
            const auto r = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);
            const char* e = static_cast<const char*>(std::memchr(sci, 'e', static_cast<size_t>(r.ptr - sci)));
            char digits[24];
            int length = 0;
            for (const char* p = sci; p < e; ++p)
            {
                if (*p != '.') { digits[length++] = *p; }
            }
            int exp10 = 0;
            std::from_chars(e[1] == '+' ? e + 2 : e + 1, r.ptr, exp10);
            const int kk = exp10 + 1;      // 10^(kk-1) <= v < 10^kk
            const int k = kk - length;     // v = digits * 10^k
            if (0 <= k && kk <= 21)
            {
                _out.append(digits, static_cast<size_t>(length));
                _out.append(static_cast<size_t>(k), '0');
                _out += ".0";
            }
            else if (0 < kk && kk <= 21)
            {
                _out.append(digits, static_cast<size_t>(kk));
                _out.push_back('.');
                _out.append(digits + kk, static_cast<size_t>(length - kk));
            }
            else if (-6 < kk && kk <= 0)
            {
                _out += "0.";
                _out.append(static_cast<size_t>(-kk), '0');
                _out.append(digits, static_cast<size_t>(length));
            }
            else
            {
                _out.push_back(digits[0]);
                if (length > 1)
                {
                    _out.push_back('.');
                    _out.append(digits + 1, static_cast<size_t>(length - 1));
                }
                _out.push_back('e');
                int x = kk - 1;
                if (x < 0)
                {
                    _out.push_back('-');
                    x = -x;
                }
                _out += std::to_string(x);
            }
        }

        /// A writer that formats further elements of an array `depth` levels down (root = 1) as if some had
        /// been written already: every element it gets starts with ",\n" + indent (the jplace writer formats
        /// batches of records on several threads and stitches the pieces).
        void prime_array_elements(size_t depth)
        {
            _levels.assign(depth, level{ 1, false });
            _levels.back() = level{ 1, true };
        }
        /// elements were appended to the innermost array behind this writer's back
        void add_elements(size_t n) { _levels.back().count += n; }

    private:
        struct level { size_t count; bool in_array; };

        void indent() { _out.append(_levels.size() * 4, ' '); }
        void prefix()
        {
            if (_levels.empty()) { return; }
            level& l = _levels.back();
            if (l.in_array)
            {
                if (l.count > 0)
                {
                    _out.push_back(',');
                    if (_single_line) { _out.push_back(' '); }
                }
                if (!_single_line) { _out.push_back('\n'); indent(); }
            }
            else
            {
                if (l.count > 0)
                {
                    if (l.count % 2 == 0) { _out += ",\n"; }
                    else { _out += ": "; }
                }
                else { _out.push_back('\n'); }
                if (l.count % 2 == 0) { indent(); }
            }
            ++l.count;
        }

        std::string& _out;
        std::vector<level> _levels;
        bool _single_line = false;
    };

    /// \brief Writes a collection of placements to a .jplace formatted file (jplace.h:19-53): same
    /// constructor, start(), operator<<, end(); the buffer is appended to the file per batch (jplace.cpp:33-36).
    class jplace_writer
    {
    public:
        jplace_writer(const std::string& filename, const std::string& invocation, std::string_view newick_tree)
            : _filename(filename), _out(filename), _writer(_buffer), _invocation(invocation), _tree(newick_tree)
        {
            if (!_out) { throw std::runtime_error("Could not create file " + filename); }
        }
        jplace_writer(const jplace_writer&) = delete;
        jplace_writer& operator=(const jplace_writer&) = delete;

        void start()
        {
            _writer.single_line_arrays(true);
            _writer.start_object();
            _writer.key("metadata");
            _writer.start_object();
            _writer.key("invocation");
            _writer.string(_invocation);
            _writer.end_object();
            _writer.key("tree");
            _writer.string(_tree);
            _writer.key("version");
            _writer.uint(3);
            _writer.key("fields");
            _writer.start_array();
            for (const char* f : { "edge_num", "likelihood", "like_weight_ratio", "distal_length", "pendant_length" })
            {
                _writer.string(f);
            }
            _writer.end_array();
            _writer.single_line_arrays(false);
            _writer.key("placements");
            _writer.start_array();
            flush();
        }

        jplace_writer& operator<<(const impl::placed_collection& placed)
        {
            append(format(placed), placed.placed_seqs.size());
            return *this;
        }

        /// The text operator<< appends for one batch, as further elements of "placements" (every record starts
        /// with ",\n" + indent). Touches no writer state: batches can be formatted concurrently and appended in
        /// order (the 35 shortest-round-trip doubles of a record are where a 30 M reads/s placer spends its time
        /// when it has to write jplace).
        static std::string format(const impl::placed_collection& placed)
        {
            std::string text;
            text.reserve(placed.placed_seqs.size() * 900);
            json_pretty_writer w(text);
            w.prime_array_elements(2);
            for (const auto& placed_seq : placed.placed_seqs)
            {
                w.start_object();
                w.key("p");
                w.start_array();
                for (const auto& p : placed_seq.placements)
                {
                    w.single_line_arrays(false);
                    w.start_array();
                    w.single_line_arrays(true);
                    w.uint(p.branch_id);
                    w.number(static_cast<double>(p.score)); // float widened to double (jplace.cpp:131)
                    w.number(p.weight_ratio);
                    w.number(p.distal_length);
                    w.number(p.pendant_length);
                    w.end_array();
                }
                w.single_line_arrays(false);
                w.end_array();
                w.key("nm");
                w.start_array();
                for (const auto header : placed.sequence_map.at(placed_seq.sequence))
                {
                    w.single_line_arrays(false);
                    w.start_array();
                    w.single_line_arrays(true);
                    w.string(header);
                    w.uint(1);
                    w.end_array();
                }
                w.single_line_arrays(false);
                w.end_array();
                w.end_object();
            }
            return text;
        }

        /// Appends what format() returned for a batch of n_records records (jplace.cpp:33-36: per batch).
        /// The same text as format(), straight from the arrays of a device pass (placer::place_arrays) for the records
        /// [first, last) of one --batch-size batch, without building a placed_collection: one record per group of
        /// identical sequences, in order of first occurrence, its names in record order. The layout of a record is
        /// fixed, so this writes through a raw cursor instead of the generic writer (2.3 GB/s of text from 16 threads
        /// was what bounded the driver). Byte-identical to format() on the same placements (host test).
        template<class Record>
        static std::string format_arrays(const std::vector<Record>& records, const epk_result& rows, size_t K, const uint32_t* group_rep,
                                         size_t first, size_t last, size_t* n_records_out = nullptr)
        {
            std::string text;
            text.resize(format_arrays_into(text, records, rows, K, group_rep, first, last, n_records_out));
            return text;
        }

        /// An upper bound of the text of the records [first, last) with up to K placements each: 48 bytes of frame and up
        /// to 26 per number, escapes can at most multiply a name by 6
        template<class Record>
        static size_t arrays_text_bound(const std::vector<Record>& records, size_t K, size_t first, size_t last)
        {
            size_t name_bytes = 0;
            for (size_t i = first; i < last; ++i) { name_bytes += std::string_view(records[i].header()).size(); }
            return (last - first) * (144 + K * 160) + name_bytes * 6 + 64;
        }

        /// format_arrays into the front of a buffer that only grows (a thread formats batch after batch into the same
        /// one: a fresh 2 MB string per batch is 2 MB of page faults and zeroes); returns the length of the text.
        template<class Record>
        static size_t format_arrays_into(std::string& text, const std::vector<Record>& records, const epk_result& rows, size_t K,
                                         const uint32_t* group_rep, size_t first, size_t last, size_t* n_records_out = nullptr)
        {
            const size_t n = last - first;
            // members of a group, in record order, as a chain through `next`
            std::vector<uint32_t> next(n, 0xFFFFFFFFu), tail(n);
            size_t n_groups = 0, name_bytes = 0;
            for (size_t j = 0; j < n; ++j)
            {
                const uint32_t r = group_rep[first + j];
                if (r == j) { ++n_groups; }
                else { next[tail[r]] = static_cast<uint32_t>(j); }
                tail[r] = static_cast<uint32_t>(j);
                name_bytes += std::string_view(records[first + j].header()).size();
            }
            const size_t bound = n_groups * (96 + K * 160) + n * 48 + name_bytes * 6 + 64; // (<= arrays_text_bound)
            if (text.size() < bound) { text.resize(bound); }
            char* const begin = text.data();
            char* at = begin;
            // the two lengths of a placement are properties of its branch: their text is kept per branch (and thread)
            thread_local std::vector<memo_entry> distal_memo, pendant_memo;
            auto put = [&at](const char* lit, size_t len) { std::memcpy(at, lit, len); at += len; };
#define EPK_PUT(LIT) put(LIT, sizeof(LIT) - 1)
            for (size_t j = 0; j < n; ++j)
            {
                if (group_rep[first + j] != j) { continue; }
                const size_t i = first + j;
                EPK_PUT(",\n        {\n            \"p\": [");
                const size_t m = rows.n_place[i];
                for (size_t p = 0; p < m; ++p)
                {
                    const size_t row = i * K + p;
                    if (p) { *at++ = ','; }
                    EPK_PUT("\n                [");
                    at = std::to_chars(at, at + 24, static_cast<uint64_t>(rows.branch[row])).ptr;
                    EPK_PUT(", ");
                    at = raw_number(at, static_cast<double>(rows.score[row])); // float widened to double (jplace.cpp:131)
                    EPK_PUT(", ");
                    at = raw_number(at, rows.lwr[row]);
                    EPK_PUT(", ");
                    at = memo_number(distal_memo, rows.branch[row], rows.distal[row], at);
                    EPK_PUT(", ");
                    at = memo_number(pendant_memo, rows.branch[row], rows.pendant[row], at);
                    *at++ = ']';
                }
                if (m) { EPK_PUT("\n            ]"); }
                else { *at++ = ']'; }
                EPK_PUT(",\n            \"nm\": [");
                bool first_name = true;
                for (uint32_t member = static_cast<uint32_t>(j); member != 0xFFFFFFFFu; member = next[member])
                {
                    if (!first_name) { *at++ = ','; }
                    first_name = false;
                    EPK_PUT("\n                [");
                    at = raw_string(at, std::string_view(records[first + member].header()));
                    EPK_PUT(", 1]");
                }
                EPK_PUT("\n            ]\n        }");
            }
#undef EPK_PUT
            if (n_records_out) { *n_records_out = n_groups; }
            return static_cast<size_t>(at - begin);
        }

        void append(const std::string& text, size_t n_records)
        {
            if (n_records)
            {
                // the very first record of the file has no comma in front of it
                _buffer.append(text, _n_records == 0 ? 1 : 0, std::string::npos);
                _n_records += n_records;
                _writer.add_elements(n_records);
            }
            flush();
        }

        /// Whether the file's new tail may be written through a shared mapping: only with room for it on the file system
        /// (a store into a mapped page that cannot be allocated is a SIGBUS, a pwrite there is an error that can be
        /// reported) and unless EPK_JPLACE_NO_MMAP is set (tests: the pwrite path).
        static bool may_map(int fd, size_t bytes)
        {
            if (std::getenv("EPK_JPLACE_NO_MMAP")) { return false; }
            struct statvfs fs;
            if (::fstatvfs(fd, &fs) != 0) { return false; }
            const unsigned long long room = static_cast<unsigned long long>(fs.f_bavail) * fs.f_frsize;
            return room >= static_cast<unsigned long long>(bytes) + (64ull << 20);
        }

        /// append() for the batches of one device pass at once: the pieces go to their places in the file with
        /// concurrent pwrite()s (one thread moves ~1.2 GB/s into the page cache; the text of a pass is ~230 MB).
        void append_many(std::vector<std::string>& texts, const std::vector<size_t>& n_records, size_t threads)
        {
            flush(); // what is buffered (the file's opening) goes first
            std::vector<size_t> offset(texts.size() + 1, 0);
            size_t total_records = 0;
            for (size_t b = 0; b < texts.size(); ++b)
            {
                if (n_records[b] && _n_records + total_records == 0) { texts[b].erase(0, 1); } // no comma before the first record
                total_records += n_records[b];
                offset[b + 1] = offset[b] + texts[b].size();
            }
            const int fd = ::open(_filename.c_str(), O_RDWR);
            if (fd < 0) { throw std::runtime_error("Could not open file " + _filename); }
            const size_t base = _file_bytes;
            // Through a shared mapping of the file's new tail where the file system allows it: concurrent pwrites to ONE
            // file take turns on its inode lock (3 GB/s with 16 threads on tmpfs - a third of a device pass's time went
            // there), page faults of a mapping do not.
            bool mapped = false;
            if (offset.back() != 0 && may_map(fd, offset.back()) && ::ftruncate(fd, static_cast<off_t>(base + offset.back())) == 0)
            {
                const size_t page = static_cast<size_t>(::sysconf(_SC_PAGESIZE));
                const size_t map_from = base / page * page, lead = base - map_from;
                void* m = ::mmap(nullptr, lead + offset.back(), PROT_READ | PROT_WRITE, MAP_SHARED, fd, static_cast<off_t>(map_from));
                if (m != MAP_FAILED)
                {
                    char* to = static_cast<char*>(m) + lead;
                    impl::parallel_for(texts.size(), threads, [&](size_t b) { std::memcpy(to + offset[b], texts[b].data(), texts[b].size()); });
                    ::munmap(m, lead + offset.back());
                    mapped = true;
                }
            }
            try
            {
                impl::parallel_for(mapped ? 0 : texts.size(), threads, [&](size_t b) {
                    size_t done = 0;
                    while (done < texts[b].size())
                    {
                        const ssize_t w = ::pwrite(fd, texts[b].data() + done, texts[b].size() - done,
                                                   static_cast<off_t>(base + offset[b] + done));
                        if (w <= 0) { throw std::runtime_error("Could not write to " + _filename); }
                        done += static_cast<size_t>(w);
                    }
                });
            }
            catch (...)
            {
                ::close(fd);
                throw;
            }
            ::close(fd);
            _file_bytes += offset.back();
            _n_records += total_records;
            if (total_records) { _writer.add_elements(total_records); }
        }

        /// Text and append in one stage (the driver's writer): `format(b, buffer, &n_records)` leaves the text of batch b
        /// at the front of `buffer` - one per thread, it only grows - and returns its length; the batch's place in the
        /// file is known as soon as the batches before it are formatted (they are handed out in order), and the thread
        /// copies its text there while it is still in its cache: no text of a pass is ever kept, nothing is read back
        /// from memory to be appended. `max_bytes` bounds the text of all batches together (arrays_text_bound): the file
        /// is extended and mapped that far, and cut back to what was written.
        template<class Format>
        void append_batches(size_t n_batches, size_t max_bytes, size_t threads, Format&& format)
        {
            flush(); // what is buffered (the file's opening) goes first
            if (n_batches == 0) { return; }
            const int fd = ::open(_filename.c_str(), O_RDWR);
            if (fd < 0) { throw std::runtime_error("Could not open file " + _filename); }
            const size_t base = _file_bytes;
            void* m = MAP_FAILED;
            size_t lead = 0;
            if (max_bytes != 0 && may_map(fd, max_bytes) && ::ftruncate(fd, static_cast<off_t>(base + max_bytes)) == 0)
            {
                const size_t page = static_cast<size_t>(::sysconf(_SC_PAGESIZE));
                const size_t map_from = base / page * page;
                lead = base - map_from;
                m = ::mmap(nullptr, lead + max_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, static_cast<off_t>(map_from));
            }
            char* const to = m == MAP_FAILED ? nullptr : static_cast<char*>(m) + lead; // (nullptr: pwrite, as append_many)
            std::vector<size_t> start(n_batches + 1, 0);
            std::atomic<size_t> published{ 0 }; // start[0 .. published] are final
            std::atomic<bool> failed{ false };
            size_t records_so_far = _n_records; // (written by the thread whose turn it is to publish)
            auto release = [&](size_t keep_bytes) {
                if (m != MAP_FAILED) { ::munmap(m, lead + max_bytes); }
                const bool cut = ::ftruncate(fd, static_cast<off_t>(base + keep_bytes)) == 0;
                ::close(fd);
                return cut;
            };
            try
            {
                impl::parallel_for(n_batches, threads, [&](size_t b) {
                    thread_local std::string buffer;
                    try
                    {
                        size_t n_records = 0;
                        size_t bytes = format(b, buffer, &n_records);
                        while (published.load(std::memory_order_acquire) != b)
                        {
                            if (failed.load(std::memory_order_relaxed)) { return; }
                            // (asleep, not spinning: the thread waited for may be waiting for a core, the host threads
                            // are shared with the reader and the packing of the next pass; a batch takes milliseconds)
                            std::this_thread::sleep_for(std::chrono::microseconds(50));
                        }
                        const char* from = buffer.data();
                        if (n_records && records_so_far == 0) // no comma before the first record of the file
                        {
                            ++from;
                            --bytes;
                        }
                        records_so_far += n_records;
                        start[b + 1] = start[b] + bytes;
                        if (start[b + 1] > max_bytes) { throw std::runtime_error("jplace text beyond its bound"); }
                        published.store(b + 1, std::memory_order_release);
                        if (to) { std::memcpy(to + start[b], from, bytes); }
                        else
                        {
                            size_t done = 0;
                            while (done < bytes)
                            {
                                const ssize_t w = ::pwrite(fd, from + done, bytes - done, static_cast<off_t>(base + start[b] + done));
                                if (w <= 0) { throw std::runtime_error("Could not write to " + _filename); }
                                done += static_cast<size_t>(w);
                            }
                        }
                    }
                    catch (...)
                    {
                        failed.store(true);
                        throw;
                    }
                });
            }
            catch (...)
            {
                release(0);
                throw;
            }
            if (!release(start[n_batches])) { throw std::runtime_error("Could not write to " + _filename); }
            _file_bytes += start[n_batches];
            if (records_so_far != _n_records) { _writer.add_elements(records_so_far - _n_records); }
            _n_records = records_so_far;
        }

        void end()
        {
            _writer.end_array();
            _writer.end_object();
            flush();
        }

    private:
        /// json_pretty_writer::string / ::number through a raw cursor (same bytes)
        static char* raw_string(char* at, std::string_view s)
        {
            static const char hex[] = "0123456789ABCDEF";
            *at++ = '"';
            for (const char ch : s)
            {
                const unsigned char c = static_cast<unsigned char>(ch);
                if (c >= 0x20 && c != '"' && c != '\\') { *at++ = ch; continue; }
                *at++ = '\\';
                switch (c)
                {
                    case '"': *at++ = '"'; break;
                    case '\\': *at++ = '\\'; break;
                    case '\b': *at++ = 'b'; break;
                    case '\f': *at++ = 'f'; break;
                    case '\n': *at++ = 'n'; break;
                    case '\r': *at++ = 'r'; break;
                    case '\t': *at++ = 't'; break;
                    default: *at++ = 'u'; *at++ = '0'; *at++ = '0'; *at++ = hex[c >> 4]; *at++ = hex[c & 15];
                }
            }
            *at++ = '"';
            return at;
        }
        /// The text of a number that is the same whenever `key` is: a memo checked against the bits of the value, so it
        /// is right whatever the caller's idea of "the same" (a half of the 35 doubles of a record are the distal and
        /// pendant lengths of its branches, and a shortest-round-trip double is the expensive part of a record).
        struct memo_entry
        {
            uint64_t bits = 0;
            uint8_t length = 0xFF; // 0xFF: empty
            char text[47];
        };
        static char* memo_number(std::vector<memo_entry>& memo, size_t key, double v, char* at)
        {
            if (key >= (size_t(1) << 22)) { return raw_number(at, v); } // (bounded: 168 MB per thread at the very most)
            if (key >= memo.size()) { memo.resize(std::max<size_t>(key + 1, 2 * memo.size())); }
            memo_entry& entry = memo[key];
            uint64_t bits;
            std::memcpy(&bits, &v, sizeof(bits));
            if (entry.length == 0xFF || entry.bits != bits)
            {
                entry.length = static_cast<uint8_t>(raw_number(entry.text, v) - entry.text);
                entry.bits = bits;
            }
            std::memcpy(at, entry.text, entry.length);
            return at + entry.length;
        }
        static char* raw_number(char* at, double v)
        {
            if (std::isnan(v) || std::isinf(v)) { return at; } // RapidJSON refuses them after the prefix
            if (v == 0.0)
            {
                const char* zero = std::signbit(v) ? "-0.0" : "0.0";
                const size_t len = std::strlen(zero);
                std::memcpy(at, zero, len);
                return at + len;
            }
            if (v < 0.0)
            {
                *at++ = '-';
                v = -v;
            }
            if (v >= 1e-5 && v < 1e15)
            {
                // the usual scores, ratios and lengths: the same shortest digits, laid out by the library (the test
                // compares with the generic writer over this range too)
                char* const end = std::to_chars(at, at + 40, v, std::chars_format::fixed).ptr;
                if (std::memchr(at, '.', static_cast<size_t>(end - at))) { return end; }
                end[0] = '.'; end[1] = '0';
                return end + 2;
            }
            char sci[40];
            const auto r = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);
            const char* e = static_cast<const char*>(std::memchr(sci, 'e', static_cast<size_t>(r.ptr - sci)));
            char digits[24];
            int length = 0;
            for (const char* p = sci; p < e; ++p)
            {
                if (*p != '.') { digits[length++] = *p; }
            }
            int exp10 = 0;
            std::from_chars(e[1] == '+' ? e + 2 : e + 1, r.ptr, exp10);
            const int kk = exp10 + 1;
            const int k = kk - length;
            if (0 <= k && kk <= 21)
            {
                std::memcpy(at, digits, static_cast<size_t>(length)); at += length;
                std::memset(at, '0', static_cast<size_t>(k)); at += k;
                *at++ = '.'; *at++ = '0';
            }
            else if (0 < kk && kk <= 21)
            {
                std::memcpy(at, digits, static_cast<size_t>(kk)); at += kk;
                *at++ = '.';
                std::memcpy(at, digits + kk, static_cast<size_t>(length - kk)); at += length - kk;
            }
            else if (-6 < kk && kk <= 0)
            {
                *at++ = '0'; *at++ = '.';
                std::memset(at, '0', static_cast<size_t>(-kk)); at += -kk;
                std::memcpy(at, digits, static_cast<size_t>(length)); at += length;
            }
            else
            {
                *at++ = digits[0];
                if (length > 1)
                {
                    *at++ = '.';
                    std::memcpy(at, digits + 1, static_cast<size_t>(length - 1)); at += length - 1;
                }
                *at++ = 'e';
                int x = kk - 1;
                if (x < 0)
                {
                    *at++ = '-';
                    x = -x;
                }
                at = std::to_chars(at, at + 8, x).ptr;
            }
            return at;
        }

        void flush()
        {
            if (!_out.is_open()) { _out.open(_filename, std::ios_base::app); }
            _out << _buffer;
            _file_bytes += _buffer.size();
            _buffer.clear();
            _out.close();
        }

        std::string _filename;
        std::ofstream _out;
        std::string _buffer;
        json_pretty_writer _writer;
        std::string _invocation;
        std::string _tree;
        size_t _n_records = 0, _file_bytes = 0;
    };
}

#endif
