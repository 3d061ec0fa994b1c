// TEST INFRASTRUCTURE ONLY (oracle). i2l shim — sequence record. PARITY UNPINNED (i2l absent).
// place.cpp:73-79 needs sequence() and header() convertible to std::string_view with storage that
// outlives the batch; jplace.cpp:152 needs header().data() to be NUL-terminated.
#ifndef I2L_SHIM_SEQ_RECORD_H
#define I2L_SHIM_SEQ_RECORD_H

#include <string>
#include <string_view>

namespace i2l
{
    class seq_record
    {
    public:
        seq_record(std::string header, std::string sequence)
            : _header(std::move(header)), _sequence(std::move(sequence)) {}

        std::string_view header() const noexcept { return _header; }
        std::string_view sequence() const noexcept { return _sequence; }

    private:
        std::string _header;
        std::string _sequence;
    };
}

#endif
