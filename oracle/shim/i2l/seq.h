// TEST INFRASTRUCTURE ONLY (oracle). i2l shim — alphabet tables. PARITY UNPINNED (i2l absent).
//
// Restates the published 2-bit (DNA) / 5-bit (amino-acid) letter codes of phylo42/i2l @ 43805ff as
// recalled in SURVEY.md Appendix B. Written naively on purpose (switch statements, std::vector),
// so that it is an independent statement of the codec the product implements with LUTs.
#ifndef I2L_SHIM_SEQ_H
#define I2L_SHIM_SEQ_H

#include <cctype>
#include <vector>
#include <i2l/phylo_kmer.h>

namespace i2l::seq
{
    enum class letter_kind { plain, ambiguous, gap };

    struct letter
    {
        letter_kind kind;
        /// one code for plain letters, all resolutions (in this order) for ambiguous ones
        std::vector<phylo_kmer::key_type> codes;
    };

#ifdef SEQ_TYPE_AA
    /// RAPPAS order R,H,K,D,E,S,T,N,Q,C,G,P,A,I,L,M,F,W,Y,V -> 0..19; U->C, O->L;
    /// B->{D,N}, Z->{E,Q}, J->{I,L}, X->all; '-', '.', '*' and anything else are gaps.
    inline letter decode(char raw)
    {
        const char c = static_cast<char>(std::toupper(static_cast<unsigned char>(raw)));
        switch (c)
        {
            case 'R': return { letter_kind::plain, { 0 } };
            case 'H': return { letter_kind::plain, { 1 } };
            case 'K': return { letter_kind::plain, { 2 } };
            case 'D': return { letter_kind::plain, { 3 } };
            case 'E': return { letter_kind::plain, { 4 } };
            case 'S': return { letter_kind::plain, { 5 } };
            case 'T': return { letter_kind::plain, { 6 } };
            case 'N': return { letter_kind::plain, { 7 } };
            case 'Q': return { letter_kind::plain, { 8 } };
            case 'C': case 'U': return { letter_kind::plain, { 9 } };
            case 'G': return { letter_kind::plain, { 10 } };
            case 'P': return { letter_kind::plain, { 11 } };
            case 'A': return { letter_kind::plain, { 12 } };
            case 'I': return { letter_kind::plain, { 13 } };
            case 'L': case 'O': return { letter_kind::plain, { 14 } };
            case 'M': return { letter_kind::plain, { 15 } };
            case 'F': return { letter_kind::plain, { 16 } };
            case 'W': return { letter_kind::plain, { 17 } };
            case 'Y': return { letter_kind::plain, { 18 } };
            case 'V': return { letter_kind::plain, { 19 } };
            case 'B': return { letter_kind::ambiguous, { 3, 7 } };
            case 'Z': return { letter_kind::ambiguous, { 4, 8 } };
            case 'J': return { letter_kind::ambiguous, { 13, 14 } };
            case 'X': return { letter_kind::ambiguous,
                               { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19 } };
            default: return { letter_kind::gap, {} };
        }
    }
#else
    /// A=0 C=1 G=2 T=3 (U as T), case-insensitive; IUPAC codes expand to their base sets in
    /// code order; '-', '.' and anything else are gaps.
    inline letter decode(char raw)
    {
        const char c = static_cast<char>(std::toupper(static_cast<unsigned char>(raw)));
        switch (c)
        {
            case 'A': return { letter_kind::plain, { 0 } };
            case 'C': return { letter_kind::plain, { 1 } };
            case 'G': return { letter_kind::plain, { 2 } };
            case 'T': case 'U': return { letter_kind::plain, { 3 } };
            case 'R': return { letter_kind::ambiguous, { 0, 2 } };
            case 'Y': return { letter_kind::ambiguous, { 1, 3 } };
            case 'S': return { letter_kind::ambiguous, { 1, 2 } };
            case 'W': return { letter_kind::ambiguous, { 0, 3 } };
            case 'K': return { letter_kind::ambiguous, { 2, 3 } };
            case 'M': return { letter_kind::ambiguous, { 0, 1 } };
            case 'B': return { letter_kind::ambiguous, { 1, 2, 3 } };
            case 'D': return { letter_kind::ambiguous, { 0, 2, 3 } };
            case 'H': return { letter_kind::ambiguous, { 0, 1, 3 } };
            case 'V': return { letter_kind::ambiguous, { 0, 1, 2 } };
            case 'N': return { letter_kind::ambiguous, { 0, 1, 2, 3 } };
            default: return { letter_kind::gap, {} };
        }
    }
#endif
}

#endif
