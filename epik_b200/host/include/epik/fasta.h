// Batch FASTA reader: the host-side stand-in for i2l::io::batch_fasta / i2l::seq_record as EPIK uses them
// (epik/src/epik/main.cpp:332-337,358: `batch_fasta(file, batch_size)`, `next_batch()` -> vector of
// records, empty at end of file, `bytes_read()` for the progress bar; place.cpp:73-79: `sequence()`,
// `header()`). The i2l implementation is absent from the reference tree; what is [UNVERIFIED] about it:
// the header is taken as the whole line after '>' (trailing CR/blank removed), sequence lines are
// concatenated with white space removed, letters are kept as they are (case, gaps and ambiguity codes
// are the k-mer codec's business).
#ifndef EPIK_B200_HOST_FASTA_H
#define EPIK_B200_HOST_FASTA_H

#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <algorithm>
#include <string_view>
#include <thread>
#include <vector>

namespace epik::io
{
    class seq_record
    {
    public:
        seq_record() = default;
        seq_record(std::string header, std::string sequence) : _header(std::move(header)), _sequence(std::move(sequence)) {}
        std::string_view header() const noexcept { return _header; }     ///< NUL-terminated (jplace.cpp:152 relies on it)
        std::string_view sequence() const noexcept { return _sequence; }
    private:
        std::string _header;
        std::string _sequence;
    };

    /// Records are parsed on `threads` host threads: the file is read in large blocks, record starts ('>' at the
    /// beginning of a line) are found with memchr, and the records of a batch are cut into contiguous runs, one per
    /// thread (one thread parses 2.3 M 150-bp records a second - less than a single GPU places).
    class batch_fasta
    {
    public:
        batch_fasta(const std::string& filename, size_t batch_size, size_t threads = 1, size_t block_bytes = 32u << 20)
            : _file(std::fopen(filename.c_str(), "rb")), _batch_size(std::max<size_t>(1, batch_size)),
              _threads(std::max<size_t>(1, std::min<size_t>(threads, 64))), _block(std::max<size_t>(1, block_bytes))
        {
            if (!_file) { throw std::runtime_error("Cannot open file: " + filename); }
        }
        batch_fasta(const batch_fasta&) = delete;
        batch_fasta& operator=(const batch_fasta&) = delete;
        ~batch_fasta() { if (_file) { std::fclose(_file); } }

        /// Up to batch_size records; empty at end of file.
        std::vector<seq_record> next_batch()
        {
            // ---- enough bytes for batch_size whole records (a record is whole once the next one starts, or at the end) ----
            while (!_eof && _starts.size() <= _batch_size) { read_more(); }
            const size_t whole = _eof ? _starts.size() : (_starts.empty() ? 0 : _starts.size() - 1);
            const size_t n = std::min(_batch_size, whole);
            std::vector<seq_record> batch(n);
            if (n == 0)
            {
                if (_eof) { _consumed += _end - _begin; _begin = _end; }
                return batch;
            }
            const size_t stop = n < _starts.size() ? _starts[n] : _end; // first byte after the batch
            const char* data = _buffer.data();
            const size_t threads = std::min(_threads, std::max<size_t>(1, n / _records_per_thread));
            auto parse_run = [&](size_t t) {
                for (size_t r = n * t / threads; r < n * (t + 1) / threads; ++r)
                {
                    batch[r] = parse_record(data + _starts[r], data + (r + 1 < _starts.size() ? _starts[r + 1] : _end));
                }
            };
            if (threads == 1) { parse_run(0); }
            else
            {
                std::vector<std::thread> pool;
                for (size_t t = 0; t < threads; ++t) { pool.emplace_back(parse_run, t); }
                for (auto& th : pool) { th.join(); }
            }
            _consumed += stop - _begin;
            _begin = stop;
            _starts.erase(_starts.begin(), _starts.begin() + static_cast<std::ptrdiff_t>(n));
            return batch;
        }

        size_t bytes_read() const noexcept { return _consumed; }

    private:
        /// One record: data[0] == '>', the header line, then sequence lines up to `end`.
        static seq_record parse_record(const char* p, const char* end)
        {
            const char* nl = static_cast<const char*>(std::memchr(p, '\n', static_cast<size_t>(end - p)));
            const char* line_end = nl ? nl : end;
            std::string_view header(p + 1, static_cast<size_t>(line_end - p - 1));
            while (!header.empty() && (header.back() == '\r' || header.back() == ' ' || header.back() == '\t')) { header.remove_suffix(1); }
            std::string sequence;
            sequence.reserve(static_cast<size_t>(end - line_end));
            p = nl ? nl + 1 : end;
            while (p < end)
            {
                nl = static_cast<const char*>(std::memchr(p, '\n', static_cast<size_t>(end - p)));
                line_end = nl ? nl : end;
                std::string_view line(p, static_cast<size_t>(line_end - p));
                while (!line.empty() && (line.back() == '\r' || line.back() == ' ' || line.back() == '\t')) { line.remove_suffix(1); }
                if (std::memchr(line.data(), ' ', line.size()) == nullptr && std::memchr(line.data(), '\t', line.size()) == nullptr)
                {
                    sequence.append(line.data(), line.size()); // the usual line: letters only
                }
                else
                {
                    for (const char c : line)
                    {
                        if (c != ' ' && c != '\t') { sequence.push_back(c); }
                    }
                }
                p = nl ? nl + 1 : end;
            }
            return seq_record(std::string(header), std::move(sequence));
        }

        /// Appends a block of the file to the buffer (dropping what has been consumed) and notes the record starts in it.
        void read_more()
        {
            const size_t BLOCK = _block;
            if (_begin > 0 && _begin >= (_end - _begin))
            {
                // compact: the unconsumed tail moves to the front, the record starts with it
                std::memmove(_buffer.data(), _buffer.data() + _begin, _end - _begin);
                for (auto& st : _starts) { st -= _begin; }
                _scanned -= _begin;
                _end -= _begin;
                _begin = 0;
            }
            if (_buffer.size() < _end + BLOCK) { _buffer.resize(std::max(_buffer.size() * 2, _end + BLOCK)); }
            const size_t got = std::fread(_buffer.data() + _end, 1, BLOCK, _file);
            if (got == 0) { _eof = true; }
            _end += got;
            // record starts: '>' as the first byte of a line (what comes before the first one - comments - is skipped)
            const char* data = _buffer.data();
            while (_scanned < _end)
            {
                const char* hit = static_cast<const char*>(std::memchr(data + _scanned, '>', _end - _scanned));
                if (!hit) { _scanned = _end; break; }
                const size_t at = static_cast<size_t>(hit - data);
                const bool line_start = at > 0 ? data[at - 1] == '\n' : _consumed + _begin == 0; // (byte 0 of the file)
                if (line_start) { _starts.push_back(at); }
                _scanned = at + 1;
            }
        }

        std::FILE* _file;
        size_t _batch_size, _threads, _block;
        std::vector<char> _buffer;
        std::vector<size_t> _starts;   // offsets in _buffer of the '>' of every record found in [_begin, _end)
        size_t _begin = 0, _end = 0, _scanned = 0, _consumed = 0;
        size_t _records_per_thread = 256; // fewer records than this are not worth a thread
    public:
        void set_records_per_thread(size_t n) noexcept { _records_per_thread = std::max<size_t>(1, n); } ///< (tests)
    private:
        bool _eof = false;
    };
}

#endif
