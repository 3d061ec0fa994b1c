// epk_gather_microbench: the random-gather ceiling of the roofline, measured by a kernel that shares nothing
// with the placement kernel but the database ("a measured random-gather ceiling using the
// same list-length distribution"). Every probe is a uniformly random k-mer code looked up in the database's own
// key table; a hit gathers the key's whole posting list. There are no accumulators, no shared-memory budget
// and no sweep: the launch runs at full occupancy, so its byte rate is what the memory system gives this access
// pattern (8 / 16-byte table probes + runs of 8-byte postings a few hundred bytes long), not what one design
// manages to ask of it.
//   method 0: 128-bit loads into registers, 4 lists in flight per warp, 32 warps per SM
//   method 1: the same with 8 lists in flight per warp
//   method 2: bulk asynchronous copies (cp.async.bulk, completion on an mbarrier) of whole lists into a ring of
//             shared-memory stages per warp; the warp reads the stage back from shared memory
#include <algorithm>
#include "common.cuh"
#include "db.h"

namespace epk
{
namespace
{
    constexpr uint32_t FULL_MASK = 0xFFFFFFFFu;

    __device__ __forceinline__ uint64_t mix64(uint64_t x)
    {
        x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
        x ^= x >> 27; x *= 0x94D049BB133111EBull;
        return x ^ (x >> 31);
    }

    // a uniformly random k-mer of the database's alphabet (every letter a valid, unambiguous code)
    __device__ __forceinline__ uint32_t random_key(const device_db& db, uint64_t seed, uint64_t index)
    {
        uint64_t r = mix64(seed + 0x9E3779B97F4A7C15ull * (index + 1));
        if (db.bits == 2) { return static_cast<uint32_t>(r) & db.key_mask; }
        uint32_t key = 0;
        for (uint32_t j = 0; j < db.k; ++j)
        {
            key = (key << db.bits) | static_cast<uint32_t>(r % 20u);
            r = r / 20u + (r << 37); // k <= 6 letters: plenty of bits left
        }
        return key;
    }

    template<uint32_t TABLE>
    __device__ __forceinline__ uint64_t probe(const device_db& db, uint32_t key) { return db_search<TABLE>(db, key); }

    struct tally { unsigned long long hits, entries, checksum; };

    __device__ __forceinline__ void fold(tally* out, uint64_t hits, uint64_t entries, uint32_t sink, uint32_t lane)
    {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
        {
            hits += __shfl_xor_sync(FULL_MASK, hits, d);
            entries += __shfl_xor_sync(FULL_MASK, entries, d);
            sink ^= __shfl_xor_sync(FULL_MASK, sink, d);
        }
        if (lane == 0)
        {
            atomicAdd(&out->hits, static_cast<unsigned long long>(hits));
            atomicAdd(&out->entries, static_cast<unsigned long long>(entries));
            atomicXor(&out->checksum, static_cast<unsigned long long>(sink));
        }
    }

    // methods 0 / 1: lane l of a warp probes key 32 * batch + l; the warp then walks the hits U at a time, lane l
    // fetching postings 2l, 2l + 1 of each with one 128-bit load (longer lists: further rounds of 64 postings)
    template<uint32_t TABLE, int U>
    __global__ void __launch_bounds__(512, 2) gather_ldg_kernel(const device_db db, uint64_t n_probes, uint64_t seed, tally* out)
    {
        const uint32_t lane = threadIdx.x & 31u;
        const uint64_t warp = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) >> 5;
        const uint64_t n_warps = (gridDim.x * static_cast<uint64_t>(blockDim.x)) >> 5;
        uint64_t hits = 0, entries = 0;
        uint32_t sink = 0;
        for (uint64_t first = warp * 32; first < n_probes; first += n_warps * 32)
        {
            uint64_t slot = 0;
            if (first + lane < n_probes) { slot = probe<TABLE>(db, random_key(db, seed, first + lane)); }
            const uint32_t len = static_cast<uint32_t>(slot & LEN_MASK);
            hits += len != 0;
            entries += len;
            uint32_t todo = __ballot_sync(FULL_MASK, len != 0);
            while (todo)
            {
                const uint4* list[U];
                uint32_t n[U];
#pragma unroll
                for (int u = 0; u < U; ++u)
                {
                    const int src = todo ? __ffs(todo) - 1 : 0;
                    const uint64_t s = __shfl_sync(FULL_MASK, slot, src);
                    n[u] = todo ? static_cast<uint32_t>(s & LEN_MASK) : 0u;
                    list[u] = reinterpret_cast<const uint4*>(db.postings + (s >> LEN_BITS));
                    todo &= todo - 1;
                }
                uint32_t longest = 0;
#pragma unroll
                for (int u = 0; u < U; ++u) { longest = max(longest, n[u]); }
                for (uint32_t pos = 0; pos < longest; pos += 64)
                {
                    uint4 v[U];
#pragma unroll
                    for (int u = 0; u < U; ++u)
                    {
                        v[u] = make_uint4(0, 0, 0, 0);
                        if (pos + 2 * lane < n[u]) { v[u] = __ldg(list[u] + pos / 2 + lane); }
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) { sink ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w; }
                }
            }
        }
        fold(out, hits, entries, sink, lane);
    }

    // method 2: whole lists (pieces of at most STAGE_BYTES) land in shared memory through bulk asynchronous copies
    constexpr uint32_t STAGES = 4;
    constexpr uint32_t STAGE_BYTES = 2048;
    constexpr uint32_t BULK_WARPS = 8;

    __device__ __forceinline__ void bulk_issue(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
    }
    __device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity)
    {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "WAIT_%=:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
            "@!p bra WAIT_%=;\n\t}"
            ::"r"(bar), "r"(parity) : "memory");
    }

    template<uint32_t TABLE>
    __global__ void __launch_bounds__(BULK_WARPS * 32) gather_bulk_kernel(const device_db db, uint64_t n_probes, uint64_t seed, tally* out)
    {
        extern __shared__ __align__(128) unsigned char smem[];
        const uint32_t lane = threadIdx.x & 31u;
        const uint32_t w = threadIdx.x >> 5;
        unsigned char* stage0 = smem + w * STAGES * STAGE_BYTES;
        uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BULK_WARPS * STAGES * STAGE_BYTES) + w * STAGES;
        uint32_t* meta = reinterpret_cast<uint32_t*>(smem + BULK_WARPS * STAGES * (STAGE_BYTES + 8)) + w * STAGES;
        const uint32_t stage_sh = static_cast<uint32_t>(__cvta_generic_to_shared(stage0));
        const uint32_t bar_sh = static_cast<uint32_t>(__cvta_generic_to_shared(bars));
        if (lane < STAGES) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_sh + 8u * lane) : "memory"); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        __syncwarp();

        const uint64_t warp = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) >> 5;
        const uint64_t n_warps = (gridDim.x * static_cast<uint64_t>(blockDim.x)) >> 5;
        uint64_t hits = 0, entries = 0;
        uint32_t sink = 0, issued = 0, consumed = 0;
        auto consume_one = [&] {
            const uint32_t s = consumed % STAGES;
            bar_wait(bar_sh + 8u * s, (consumed / STAGES) & 1u);
            const uint32_t bytes = meta[s];
            for (uint32_t o = 16u * lane; o < bytes; o += 512u)
            {
                const uint4 v = *reinterpret_cast<const uint4*>(stage0 + s * STAGE_BYTES + o);
                sink ^= v.x ^ v.y ^ v.z ^ v.w;
            }
            __syncwarp();
            ++consumed;
        };
        for (uint64_t first = warp * 32; first < n_probes; first += n_warps * 32)
        {
            uint64_t slot = 0;
            if (first + lane < n_probes) { slot = probe<TABLE>(db, random_key(db, seed, first + lane)); }
            const uint32_t len = static_cast<uint32_t>(slot & LEN_MASK);
            hits += len != 0;
            entries += len;
            uint32_t todo = __ballot_sync(FULL_MASK, len != 0);
            while (todo)
            {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                const uint64_t s = __shfl_sync(FULL_MASK, slot, src);
                const uint32_t bytes_all = (static_cast<uint32_t>(s & LEN_MASK) + 1u) / 2u * 16u; // lists are padded to 16 bytes
                const unsigned char* from = reinterpret_cast<const unsigned char*>(db.postings + (s >> LEN_BITS));
                for (uint32_t o = 0; o < bytes_all; o += STAGE_BYTES)
                {
                    if (issued - consumed == STAGES) { consume_one(); }
                    const uint32_t st = issued % STAGES;
                    const uint32_t bytes = min(bytes_all - o, STAGE_BYTES);
                    if (lane == 0)
                    {
                        meta[st] = bytes;
                        bulk_issue(stage_sh + st * STAGE_BYTES, from + o, bytes, bar_sh + 8u * st);
                    }
                    __syncwarp();
                    ++issued;
                }
            }
        }
        while (consumed < issued) { consume_one(); }
        fold(out, hits, entries, sink, lane);
    }
} // namespace
} // namespace epk

using namespace epk;

extern "C" epk_status epk_gather_microbench(const epk_db* db, uint64_t n_probes, uint64_t seed, uint32_t method,
                                            uint32_t reps, epk_microbench* out)
{
    return guarded([&] {
        if (!db || !out || reps == 0) { throw error(EPK_E_INVALID, "null argument"); }
        if (method > 2) { throw error(EPK_E_INVALID, "method must be 0, 1 or 2"); }
        EPK_CUDA(cudaSetDevice(db->device));
        int n_sms = 0;
        EPK_CUDA(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, db->device));
        device_buffer<tally> d_tally(1);
        cudaEvent_t e0, e1;
        EPK_CUDA(cudaEventCreate(&e0));
        EPK_CUDA(cudaEventCreate(&e1));
        const bool direct = db->table_kind == TABLE_DIRECT;
        const size_t bulk_smem = BULK_WARPS * STAGES * (STAGE_BYTES + 8 + 4);
        if (method == 2)
        {
            EPK_CUDA(cudaFuncSetAttribute(gather_bulk_kernel<TABLE_DIRECT>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bulk_smem)));
            EPK_CUDA(cudaFuncSetAttribute(gather_bulk_kernel<TABLE_HASH>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bulk_smem)));
        }
        auto run = [&] {
            EPK_CUDA(cudaMemsetAsync(d_tally.get(), 0, sizeof(tally)));
            if (method == 2)
            {
                const unsigned grid = 3u * n_sms;
                if (direct) { gather_bulk_kernel<TABLE_DIRECT><<<grid, BULK_WARPS * 32, bulk_smem>>>(db->d, n_probes, seed, d_tally.get()); }
                else { gather_bulk_kernel<TABLE_HASH><<<grid, BULK_WARPS * 32, bulk_smem>>>(db->d, n_probes, seed, d_tally.get()); }
            }
            else
            {
                const unsigned grid = 2u * n_sms;
                if (method == 0)
                {
                    if (direct) { gather_ldg_kernel<TABLE_DIRECT, 4><<<grid, 512>>>(db->d, n_probes, seed, d_tally.get()); }
                    else { gather_ldg_kernel<TABLE_HASH, 4><<<grid, 512>>>(db->d, n_probes, seed, d_tally.get()); }
                }
                else
                {
                    if (direct) { gather_ldg_kernel<TABLE_DIRECT, 8><<<grid, 512>>>(db->d, n_probes, seed, d_tally.get()); }
                    else { gather_ldg_kernel<TABLE_HASH, 8><<<grid, 512>>>(db->d, n_probes, seed, d_tally.get()); }
                }
            }
            EPK_CUDA(cudaGetLastError());
        };
        run(); // warm-up
        EPK_CUDA(cudaDeviceSynchronize());
        EPK_CUDA(cudaEventRecord(e0));
        for (uint32_t r = 0; r < reps; ++r) { run(); }
        EPK_CUDA(cudaEventRecord(e1));
        EPK_CUDA(cudaEventSynchronize(e1));
        float ms = 0.0f;
        EPK_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        tally t{};
        EPK_CUDA(cudaMemcpy(&t, d_tally.get(), sizeof(t), cudaMemcpyDeviceToHost));
        out->ms_per_launch = ms / reps;
        out->n_probes = n_probes;
        out->n_hits = t.hits;
        out->n_entries = t.entries;
        out->bytes = 16ull * n_probes + 8ull * t.entries; // the accounting of SURVEY.md 8d: 16 B per look-up, 8 B per posting
        out->checksum = t.checksum;
    });
}
