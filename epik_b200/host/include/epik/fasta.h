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
#include <string_view>
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

    class batch_fasta
    {
    public:
        batch_fasta(const std::string& filename, size_t batch_size)
            : _file(std::fopen(filename.c_str(), "rb")), _batch_size(batch_size), _buffer(1u << 22)
        {
            if (!_file) { throw std::runtime_error("Cannot open file: " + filename); }
        }
        batch_fasta(const batch_fasta&) = delete;
        batch_fasta& operator=(const batch_fasta&) = delete;
        ~batch_fasta() { if (_file) { std::fclose(_file); } }

        /// Up to batch_size records; empty at end of file.
        std::vector<seq_record> next_batch()
        {
            std::vector<seq_record> batch;
            batch.reserve(_batch_size);
            std::string_view line;
            while (batch.size() < _batch_size)
            {
                if (!_have_header)
                {
                    // skip to the next header line
                    bool found = false;
                    while (read_line(line))
                    {
                        if (!line.empty() && line[0] == '>') { found = true; break; }
                    }
                    if (!found) { break; }
                    _header.assign(line.data() + 1, line.size() - 1);
                    _have_header = true;
                }
                std::string sequence;
                bool next_header = false;
                while (read_line(line))
                {
                    if (!line.empty() && line[0] == '>') { next_header = true; break; }
                    if (std::memchr(line.data(), ' ', line.size()) == nullptr && std::memchr(line.data(), '\t', line.size()) == nullptr)
                    {
                        sequence.append(line.data(), line.size()); // the usual line: letters only
                        continue;
                    }
                    for (const char c : line)
                    {
                        if (c != ' ' && c != '\t') { sequence.push_back(c); }
                    }
                }
                batch.emplace_back(std::move(_header), std::move(sequence));
                _header.clear();
                _have_header = next_header;
                if (next_header) { _header.assign(line.data() + 1, line.size() - 1); }
                else { break; } // end of file
            }
            return batch;
        }

        size_t bytes_read() const noexcept { return _bytes_consumed; }

    private:
        /// One line without its terminator (LF or CRLF) and without trailing blanks; false at end of file. The view
        /// points into the read buffer (a line that straddles two buffer fills is assembled in _straddle) and is
        /// valid until the next call.
        bool read_line(std::string_view& line)
        {
            _straddle.clear();
            bool any = false, whole = false;
            const char* start = nullptr;
            size_t length = 0;
            while (true)
            {
                if (_pos == _len)
                {
                    _len = std::fread(_buffer.data(), 1, _buffer.size(), _file);
                    _pos = 0;
                    if (_len == 0) { break; }
                }
                any = true;
                start = _buffer.data() + _pos;
                const char* nl = static_cast<const char*>(std::memchr(start, '\n', _len - _pos));
                if (nl)
                {
                    length = static_cast<size_t>(nl - start);
                    _bytes_consumed += length + 1;
                    _pos += length + 1;
                    whole = _straddle.empty();
                    if (!whole) { _straddle.append(start, length); }
                    break;
                }
                _straddle.append(start, _len - _pos); // no terminator in what is left of the buffer
                _bytes_consumed += _len - _pos;
                _pos = _len;
            }
            line = whole ? std::string_view(start, length) : std::string_view(_straddle);
            while (!line.empty() && (line.back() == '\r' || line.back() == ' ' || line.back() == '\t')) { line.remove_suffix(1); }
            return any;
        }

        std::FILE* _file;
        size_t _batch_size;
        std::vector<char> _buffer;
        size_t _pos = 0, _len = 0, _bytes_consumed = 0;
        std::string _header, _straddle;
        bool _have_header = false;
    };
}

#endif
