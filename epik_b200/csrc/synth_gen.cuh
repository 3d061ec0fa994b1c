// On-device generator of synthetic phylo-k-mer databases ("gen2" of epik_b200/synth.py).
//
// BASELINE config 5 (k = 12, ~20 k branches, ~4 G postings = 32 GB) cannot be built from host arrays in a bench run:
// numpy needs minutes and > 100 GB for it. The database is therefore a pure function of (seed, key, position in the
// list), evaluated here one key range at a time straight into the build's chunk buffer, and restated in numpy for
// any subset of keys (synth.gen2_lists) so that the reference can be fed exactly the lists a set of reads looks up.
// Every step is integer arithmetic or a single correctly rounded IEEE operation: the two sides agree bit for bit.
//   present(key)   = stream(3, key) < presence_threshold
//   length(key)    = length_table[stream(4, key) >> (64 - table_bits)]          (quantiles of the length distribution)
//   branch(key, j) = (stream(7, key) % N + sum_{i <= j} (stream(6, key << 24 ^ i) % 7 + 1)) % N   (distinct: length <= N / 8)
//   score(key, j)  = min(float(score_lo + (-0.3 - score_lo) * unit(stream(8, key << 24 ^ j))), -0.3f)
// with stream(tag, x) = mix64(x * 0xD6E8FEB86659FD93 + mix64(seed << 8 ^ tag)), unit(h) = ((h >> 11) + 0.5) * 2^-53.
// Test / bench infrastructure (the reference has no generator: it ships no fixtures at all).
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace epk
{
struct synth_params
{
    uint64_t base_present, base_length, base_stride, base_start, base_score; // mix64(seed << 8 ^ tag), tags 3 4 6 7 8
    uint64_t presence_threshold;
    const uint32_t* length_table; // device
    uint32_t length_shift;        // 64 - table_bits
    uint32_t n_nodes;
    uint32_t alphabet, k;
    double score_lo, score_span;  // score = score_lo + score_span * unit
};

__host__ __device__ inline uint64_t mix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ inline uint64_t stream_at(uint64_t base, uint64_t idx) { return mix64(idx * 0xD6E8FEB86659FD93ull + base); }

// candidate index (0 .. |alphabet|^k) -> k-mer code (amino acids: 5 bits per letter, letter i of the index in base 20)
__host__ __device__ inline uint32_t synth_key_of(uint32_t alphabet, uint32_t k, uint64_t candidate)
{
    if (alphabet == 0) { return static_cast<uint32_t>(candidate); }
    uint32_t key = 0;
    for (uint32_t i = 0; i < k; ++i)
    {
        key |= static_cast<uint32_t>(candidate % 20u) << (5u * i);
        candidate /= 20u;
    }
    return key;
}

// present[c] = 1 / 0, length[c] = list length (0 when absent), per candidate
__global__ void synth_lengths_kernel(const synth_params sp, uint64_t n_candidates, uint64_t* present, uint64_t* length)
{
    const uint64_t c = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (c >= n_candidates) { return; }
    const uint64_t key = synth_key_of(sp.alphabet, sp.k, c);
    const bool here = stream_at(sp.base_present, key) < sp.presence_threshold;
    present[c] = here ? 1u : 0u;
    length[c] = here ? sp.length_table[stream_at(sp.base_length, key) >> sp.length_shift] : 0u;
}

// keys[at[c]] = key of candidate c, offsets[at[c]] = entries before it (for the present candidates)
__global__ void synth_scatter_kernel(const synth_params sp, uint64_t n_candidates, const uint64_t* present, const uint64_t* at,
                                     const uint64_t* before, uint32_t* keys, uint64_t* offsets)
{
    const uint64_t c = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (c >= n_candidates || !present[c]) { return; }
    keys[at[c]] = synth_key_of(sp.alphabet, sp.k, c);
    offsets[at[c]] = before[c];
}

// One warp per list: the entries of keys [0, n_keys) of the arrays given (pointers already moved to the key range),
// entry j of key i at dst[offsets[i] + j] (dst biased by the caller).
__global__ void synth_entries_kernel(const synth_params sp, const uint32_t* keys, const uint64_t* offsets, uint64_t n_keys, uint2* dst)
{
    const uint64_t i = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (i >= n_keys) { return; }
    const uint64_t key = keys[i];
    const uint64_t b = offsets[i], n = offsets[i + 1] - b;
    uint64_t run = stream_at(sp.base_start, key) % sp.n_nodes; // start + strides so far
    for (uint64_t j0 = 0; j0 < n; j0 += 32)
    {
        const uint64_t j = j0 + lane;
        const uint64_t eid = (key << 24) ^ j;
        uint32_t stride = j < n ? static_cast<uint32_t>(stream_at(sp.base_stride, eid) % 7u) + 1u : 0u;
        uint32_t incl = stride;
        for (int d = 1; d < 32; d <<= 1)
        {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= static_cast<uint32_t>(d)) { incl += o; }
        }
        if (j < n)
        {
            const uint32_t branch = static_cast<uint32_t>((run + incl) % sp.n_nodes);
            const uint64_t h = stream_at(sp.base_score, eid);
            const double unit = __dmul_rn(__dadd_rn(static_cast<double>(h >> 11), 0.5), 1.0 / 9007199254740992.0);
            float score = __double2float_rn(__dadd_rn(sp.score_lo, __dmul_rn(sp.score_span, unit)));
            score = fminf(score, -0.3f);
            dst[b + j] = make_uint2(branch, __float_as_uint(score));
        }
        run += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
}
} // namespace epk
