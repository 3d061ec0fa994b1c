// epk_db_pool: one database spread over the HBM of several GPUs, read in place over NVLink.
//
// Database-sharded mode as round 1 built it moves per-query ROWS between GPUs (place_kernel.cuh, "partial rows"):
// every GPU encodes and probes every query, and the owner adds G rows. It never reached the rate of one GPU. This file
// turns the exchange round: the QUERIES stay where they are - each GPU places its own slice with the fused kernel,
// unchanged - and the DATABASE is what is shared. The shards of the key space (epk_db_create_shard /
// epk_db_create_synthetic, one per GPU) are mapped one after the other into ONE range of virtual addresses that every
// GPU of the process can read (CUDA virtual memory management: cuMemCreate on the shard's GPU, cuMemMap into the
// common range, cuMemSetAccess for all the GPUs), and every GPU gets a key table of its own - 8 bytes per possible
// code, 128 MB at k = 12 - whose slots point into that range. A posting list is then gathered from whichever HBM holds
// it: local loads for 1 / G of the lists, NVLink loads (P2P through NVSwitch) for the rest, issued by the same
// ld.global.nc.v4 of the same kernel. Per query 8 bytes per posting cross the link, not 6 bytes per branch and shard.
// Nothing is re-associated: the additions of a branch run in window order as on one GPU, so the results are the fused
// kernel's to the bit - ambiguous windows, load filter and range teams included - and no launch form is missing.
//
// One process drives the GPUs (as the reference is one process, main.cpp:332-366; epk_place_many cuts a pass over the
// placers). The driver entry points are taken from the runtime (cudaGetDriverEntryPoint): the library does not link
// libcuda and still loads on a machine without a driver.
#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <cuda.h>

#include "common.cuh"
#include "db.h"

using namespace epk;

namespace
{
    struct driver_api
    {
        CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
        CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
        CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
        CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
        CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
        CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
        CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
        CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
        CUresult (*GetErrorString)(CUresult, const char**) = nullptr;

        template<class F>
        static void resolve(const char* name, F& fn)
        {
            void* p = nullptr;
            cudaDriverEntryPointQueryResult status = cudaDriverEntryPointSymbolNotFound;
            EPK_CUDA(cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &status));
            if (status != cudaDriverEntryPointSuccess || !p) { throw error(EPK_E_CUDA, std::string("driver entry point not found: ") + name); }
            fn = reinterpret_cast<F>(p);
        }
        driver_api()
        {
            resolve("cuMemGetAllocationGranularity", MemGetAllocationGranularity);
            resolve("cuMemCreate", MemCreate);
            resolve("cuMemAddressReserve", MemAddressReserve);
            resolve("cuMemMap", MemMap);
            resolve("cuMemSetAccess", MemSetAccess);
            resolve("cuMemUnmap", MemUnmap);
            resolve("cuMemRelease", MemRelease);
            resolve("cuMemAddressFree", MemAddressFree);
            resolve("cuGetErrorString", GetErrorString);
        }
        void check(CUresult r, const char* what) const
        {
            if (r == CUDA_SUCCESS) { return; }
            const char* text = nullptr;
            GetErrorString(r, &text);
            throw error(r == CUDA_ERROR_OUT_OF_MEMORY ? EPK_E_NOMEM : EPK_E_CUDA, std::string(what) + ": " + (text ? text : "unknown driver error"));
        }
    };

    // The common range and the physical memory mapped into it; lives as long as one pooled database does.
    struct pool_memory
    {
        driver_api api;
        CUdeviceptr base = 0;
        size_t reserved = 0;
        std::vector<std::pair<size_t, size_t>> mapped; // (offset, bytes) of every mapping
        std::vector<CUmemGenericAllocationHandle> handles;
        int any_device = 0;
        ~pool_memory()
        {
            cudaSetDevice(any_device);
            cudaDeviceSynchronize();
            for (const auto& m : mapped) { api.MemUnmap(base + m.first, m.second); }
            for (const auto h : handles) { api.MemRelease(h); }
            if (base) { api.MemAddressFree(base, reserved); }
        }
    };

    // merged[i] = slot of key i in the pool: the shard's slot with its offset moved by the shard's place in the range
    __global__ void merge_table_kernel(const uint64_t* __restrict__ shard_table, uint64_t n_slots, uint64_t bias_entries,
                                       unsigned long long* merged, uint32_t* clash)
    {
        const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
        if (i >= n_slots) { return; }
        const uint64_t v = shard_table[i];
        if ((v & LEN_MASK) == 0) { return; }
        if (merged[i] != 0ull) { atomicOr(clash, 1u); return; } // two shards hold the same key: not shards of one database
        merged[i] = v + (bias_entries << LEN_BITS);
    }
}

extern "C" epk_status epk_db_pool(epk_db* const* shards, uint32_t n_shards, epk_db** out)
{
    return guarded([&] {
        if (!shards || !out || n_shards == 0) { throw error(EPK_E_INVALID, "null argument"); }
        for (uint32_t g = 0; g < n_shards; ++g) { out[g] = nullptr; }
        if (n_shards > 64) { throw error(EPK_E_INVALID, "at most 64 shards"); }
        const epk_db& first = *shards[0];
        uint64_t total_entries = 0;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            const epk_db* s = shards[g];
            if (!s) { throw error(EPK_E_INVALID, "null shard"); }
            if (s->shard_count != n_shards || s->shard_index != g) { throw error(EPK_E_INVALID, "shards[g] must be shard g of n_shards (epk_db_create_shard)"); }
            if (s->table_kind != TABLE_DIRECT)
            {
                throw error(EPK_E_UNSUPPORTED, "pooled databases need the direct-address key table (alphabet bits x k <= 26)");
            }
            if (s->pool) { throw error(EPK_E_INVALID, "a pooled database cannot be pooled again"); }
            if (s->d.k != first.d.k || s->d.alphabet != first.d.alphabet || s->d.n_nodes != first.d.n_nodes || s->omega != first.omega
                || s->d.range_headers != first.d.range_headers || s->table_slots != first.table_slots
                || s->n_entries_total != first.n_entries_total)
            {
                throw error(EPK_E_INVALID, "the shards do not belong to one database");
            }
            total_entries += s->postings.size();
        }
        (void)total_entries;

        auto pool = std::make_shared<pool_memory>();
        pool->any_device = first.device;
        const driver_api& api = pool->api;
        // ---- the common range: every shard's postings at a multiple of the allocation granularity ----
        std::vector<size_t> offset(n_shards), bytes(n_shards);
        size_t total = 0;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            EPK_CUDA(cudaSetDevice(shards[g]->device));
            EPK_CUDA(cudaFree(nullptr)); // (the device's context is current for the driver calls below)
            CUmemAllocationProp prop{};
            prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
            prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            prop.location.id = shards[g]->device;
            size_t granularity = 0;
            api.check(api.MemGetAllocationGranularity(&granularity, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
            granularity = std::max<size_t>(granularity, 2u << 20);
            offset[g] = (total + granularity - 1) / granularity * granularity;
            bytes[g] = (std::max<size_t>(shards[g]->postings.bytes(), 16) + granularity - 1) / granularity * granularity;
            total = offset[g] + bytes[g];
        }
        // (a segment descriptor names its first posting by a 32-bit index of posting PAIRS: place_kernel.cuh)
        if (total / 8 >= (1ull << 33)) { throw error(EPK_E_UNSUPPORTED, "a pooled database holds at most 2^33 postings (64 GB)"); }
        api.check(api.MemAddressReserve(&pool->base, total, 0, 0, 0), "cuMemAddressReserve");
        pool->reserved = total;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            EPK_CUDA(cudaSetDevice(shards[g]->device));
            CUmemAllocationProp prop{};
            prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
            prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            prop.location.id = shards[g]->device;
            CUmemGenericAllocationHandle handle = 0;
            api.check(api.MemCreate(&handle, bytes[g], &prop, 0), "cuMemCreate");
            pool->handles.push_back(handle);
            api.check(api.MemMap(pool->base + offset[g], bytes[g], 0, handle, 0), "cuMemMap");
            pool->mapped.emplace_back(offset[g], bytes[g]);
        }
        // every GPU that holds a shard reads (and, for the copy below, writes) the whole range
        std::vector<CUmemAccessDesc> access;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            bool seen = false;
            for (const auto& a : access) { seen = seen || a.location.id == shards[g]->device; }
            if (seen) { continue; }
            CUmemAccessDesc a{};
            a.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            a.location.id = shards[g]->device;
            a.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
            access.push_back(a);
        }
        api.check(api.MemSetAccess(pool->base, total, access.data(), access.size()), "cuMemSetAccess (are the GPUs peers of each other?)");
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            EPK_CUDA(cudaSetDevice(shards[g]->device));
            EPK_CUDA(cudaMemcpy(reinterpret_cast<void*>(pool->base + offset[g]), shards[g]->postings.get(), shards[g]->postings.bytes(),
                                cudaMemcpyDeviceToDevice));
        }

        // ---- per GPU: a key table whose slots point into the range ----
        std::vector<std::unique_ptr<epk_db>> made(n_shards);
        float score_lo = 0.0f, score_hi = 0.0f;
        bool any_scores = false, fast_ok = true;
        uint64_t entries_loaded = 0, keys_loaded = 0;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            const epk_db& s = *shards[g];
            if (s.n_entries_loaded)
            {
                score_lo = any_scores ? std::min(score_lo, s.d.score_lo) : s.d.score_lo;
                score_hi = any_scores ? std::max(score_hi, s.d.score_hi) : s.d.score_hi;
                any_scores = true;
            }
            fast_ok = fast_ok && s.d.fast_ok != 0;
            entries_loaded += s.n_entries_loaded;
            keys_loaded += s.n_keys_loaded;
        }
        for (uint32_t d = 0; d < n_shards; ++d)
        {
            const epk_db& mine = *shards[d];
            EPK_CUDA(cudaSetDevice(mine.device));
            auto db = std::make_unique<epk_db>();
            db->device = mine.device;
            db->table_kind = TABLE_DIRECT;
            db->table_slots = mine.table_slots;
            db->direct.resize(mine.table_slots);
            EPK_CUDA(cudaMemset(db->direct.get(), 0, db->direct.bytes()));
            device_buffer<uint64_t> peer_table(mine.table_slots);
            device_buffer<uint32_t> clash(1);
            EPK_CUDA(cudaMemset(clash.get(), 0, 4));
            for (uint32_t g = 0; g < n_shards; ++g)
            {
                const uint64_t* table = shards[g]->direct.get();
                if (shards[g]->device != mine.device)
                {
                    EPK_CUDA(cudaMemcpyPeer(peer_table.get(), mine.device, table, shards[g]->device, mine.table_slots * 8));
                    table = peer_table.get();
                }
                merge_table_kernel<<<blocks_for(mine.table_slots, 256), 256>>>(table, mine.table_slots, offset[g] / 8,
                                                                                 reinterpret_cast<unsigned long long*>(db->direct.get()), clash.get());
                EPK_CUDA(cudaGetLastError());
                EPK_CUDA(cudaDeviceSynchronize()); // (peer_table is refilled next)
            }
            uint32_t h_clash = 0;
            EPK_CUDA(cudaMemcpy(&h_clash, clash.get(), 4, cudaMemcpyDeviceToHost));
            if (h_clash) { throw error(EPK_E_INVALID, "two shards hold the same key: not the shards of one database"); }
            db->letters.resize(256);
            EPK_CUDA(cudaMemcpy(db->letters.get(), mine.letters.get(), 256 * 4, cudaMemcpyDeviceToDevice));
            db->d = mine.d;
            db->d.direct = db->direct.get();
            db->d.hashed = nullptr;
            db->d.postings = reinterpret_cast<const uint2*>(pool->base);
            db->d.score_lo = score_lo;
            db->d.score_hi = score_hi;
            db->d.fast_ok = fast_ok ? 1u : 0u;
            db->distal = mine.distal;
            db->pendant = mine.pendant;
            db->omega = mine.omega;
            db->n_entries_total = mine.n_entries_total;
            db->n_entries_loaded = entries_loaded;
            db->n_keys_loaded = keys_loaded;
            db->shard_index = 0;
            db->shard_count = 1; // every key is reachable: to the placer this is a whole database
            db->pool = pool;
            db->pool_local_bytes = bytes[d];
            made[d] = std::move(db);
        }
        for (uint32_t g = 0; g < n_shards; ++g) { out[g] = made[g].release(); }
    });
}
