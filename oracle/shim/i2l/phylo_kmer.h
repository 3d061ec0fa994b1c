// TEST INFRASTRUCTURE ONLY (oracle). Never included by the product under epik_b200/.
//
// i2l shim: the `phylo42/i2l` submodule (pinned at 43805ff20710cf8c253dc4703d1ec72de7757129 in the
// reference's .SUBMODULES.json:9-15) is absent from /root/reference. This header restates the part of
// its contract that EPIK's placer relies on, as proven by the call sites in
// epik/src/epik/place.cpp and epik/include/epik/place.h. PARITY UNPINNED for
// everything in this directory: upstream i2l could not be read, compiled or tested here.
#ifndef I2L_SHIM_PHYLO_KMER_H
#define I2L_SHIM_PHYLO_KMER_H

#include <algorithm>   // place.cpp:153,196 use std::partial_sort / std::copy_if without including it
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <string>
#include <string_view>
#include <vector>

namespace i2l
{
    /// place.h:50-51 use phylo_kmer::{branch_type, score_type}; main.cpp:325 uses key_type.
    struct phylo_kmer
    {
        using key_type = uint32_t;
        using branch_type = uint32_t;
        using score_type = float;
    };

    /// place.cpp:358 destructures it as [postorder_node_id, score]; intrinsic.h:19,23 name the
    /// members .branch and .score; main.cpp:257 uses sizeof() as the RAM accounting unit (8 bytes).
    struct pkdb_value
    {
        phylo_kmer::branch_type branch;
        phylo_kmer::score_type score;
    };
    static_assert(sizeof(pkdb_value) == 8, "RAM accounting unit of --max-ram (main.cpp:257)");

    /// Number of letters of the alphabet this build of the shim was compiled for.
    /// The reference links i2l::dna or i2l::aa (epik/CMakeLists.txt:72,124); the shim follows with
    /// -DSEQ_TYPE_AA (default: DNA).
#ifdef SEQ_TYPE_AA
    constexpr unsigned alphabet_size = 20;
    constexpr unsigned bits_per_letter = 5;
    constexpr const char* sequence_type_name = "Proteins";
#else
    constexpr unsigned alphabet_size = 4;
    constexpr unsigned bits_per_letter = 2;
    constexpr const char* sequence_type_name = "DNA";
#endif

    /// place.cpp:87: `_threshold{ i2l::score_threshold(db.omega(), db.kmer_size()) }`.
    /// RAPPAS definition (omega / |alphabet|)^k, evaluated in float. [UNVERIFIED vs upstream i2l]
    inline phylo_kmer::score_type score_threshold(phylo_kmer::score_type omega, size_t kmer_size)
    {
        return std::pow(omega / static_cast<phylo_kmer::score_type>(alphabet_size),
                        static_cast<phylo_kmer::score_type>(kmer_size));
    }
}

#endif
