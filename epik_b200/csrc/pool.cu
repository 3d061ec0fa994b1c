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
#include <unistd.h>

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
        CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
        CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;

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
            resolve("cuMemExportToShareableHandle", MemExportToShareableHandle);
            resolve("cuMemImportFromShareableHandle", MemImportFromShareableHandle);
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


// ---- the same pool with one PROCESS per GPU --------------------------------------------------------------------
//
// Under torchrun / MPI every GPU has its own process and its own address space: the physical memory of a shard then
// travels as a POSIX file descriptor (cuMemExportToShareableHandle; the caller passes it to the peers over a unix
// socket, SCM_RIGHTS) and its key table as a CUDA IPC handle. Every process builds the same address range for itself.
//   1. epk_db_pool_share(shard)        my postings move into exportable memory; I get {fd, bytes, table handle} to send
//   2. (the caller exchanges those)
//   3. epk_db_pool_join(shard, parts)  maps all the parts - mine and the imported ones - into one range, merges the key
//                                      tables into one of mine, returns the pooled database on my device
namespace
{
    struct exported_part
    {
        driver_api api;
        CUmemGenericAllocationHandle handle = 0;
        CUdeviceptr at = 0;
        size_t bytes = 0;
        int device = 0;
        ~exported_part()
        {
            cudaSetDevice(device);
            cudaDeviceSynchronize();
            if (at) { api.MemUnmap(at, bytes); api.MemAddressFree(at, bytes); }
            if (handle) { api.MemRelease(handle); }
        }
    };
}

extern "C" epk_status epk_db_pool_share(epk_db* shard, int* fd_out, uint64_t* bytes_out, unsigned char* table_handle_out)
{
    return guarded([&] {
        if (!shard || !fd_out || !bytes_out || !table_handle_out) { throw error(EPK_E_INVALID, "null argument"); }
        if (shard->table_kind != TABLE_DIRECT) { throw error(EPK_E_UNSUPPORTED, "pooled databases need the direct-address key table (alphabet bits x k <= 26)"); }
        if (shard->pool || shard->exported) { throw error(EPK_E_INVALID, "this database is pooled or shared already"); }
        EPK_CUDA(cudaSetDevice(shard->device));
        EPK_CUDA(cudaFree(nullptr));
        auto part = std::make_shared<exported_part>();
        part->device = shard->device;
        const driver_api& api = part->api;
        CUmemAllocationProp prop{};
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = shard->device;
        prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t granularity = 0;
        api.check(api.MemGetAllocationGranularity(&granularity, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
        granularity = std::max<size_t>(granularity, 2u << 20);
        part->bytes = (std::max<size_t>(shard->postings.bytes(), 16) + granularity - 1) / granularity * granularity;
        api.check(api.MemCreate(&part->handle, part->bytes, &prop, 0), "cuMemCreate (exportable)");
        api.check(api.MemAddressReserve(&part->at, part->bytes, 0, 0, 0), "cuMemAddressReserve");
        api.check(api.MemMap(part->at, part->bytes, 0, part->handle, 0), "cuMemMap");
        CUmemAccessDesc access{};
        access.location = prop.location;
        access.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        api.check(api.MemSetAccess(part->at, part->bytes, &access, 1), "cuMemSetAccess");
        EPK_CUDA(cudaMemcpy(reinterpret_cast<void*>(part->at), shard->postings.get(), shard->postings.bytes(), cudaMemcpyDeviceToDevice));
        int fd = -1;
        api.check(api.MemExportToShareableHandle(&fd, part->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
        cudaIpcMemHandle_t table{};
        EPK_CUDA(cudaIpcGetMemHandle(&table, shard->direct.get()));
        static_assert(sizeof(table) == EPK_IPC_HANDLE_BYTES, "IPC handle size");
        std::memcpy(table_handle_out, &table, sizeof(table));
        // the shard keeps answering (its own postings stay until it is destroyed); the exported copy lives with it and
        // with every process that maps it
        shard->exported = part;
        *fd_out = fd;
        *bytes_out = part->bytes;
    });
}

extern "C" epk_status epk_db_pool_join(epk_db* shard, uint32_t n_shards, const int* fds, const uint64_t* bytes,
                                       const unsigned char* table_handles, epk_db** out)
{
    return guarded([&] {
        if (!shard || !fds || !bytes || !table_handles || !out) { throw error(EPK_E_INVALID, "null argument"); }
        *out = nullptr;
        if (n_shards == 0 || n_shards > 64 || shard->shard_count != n_shards) { throw error(EPK_E_INVALID, "the shard is not one of n_shards"); }
        if (!shard->exported) { throw error(EPK_E_INVALID, "epk_db_pool_share comes first"); }
        const uint32_t me = shard->shard_index;
        const auto mine = std::static_pointer_cast<exported_part>(shard->exported);
        EPK_CUDA(cudaSetDevice(shard->device));
        EPK_CUDA(cudaFree(nullptr));
        auto pool = std::make_shared<pool_memory>();
        pool->any_device = shard->device;
        const driver_api& api = pool->api;
        std::vector<size_t> offset(n_shards);
        size_t total = 0;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            if (bytes[g] == 0 || bytes[g] % (2u << 20) != 0) { throw error(EPK_E_INVALID, "part sizes are multiples of the allocation granularity"); }
            offset[g] = total;
            total += bytes[g];
        }
        if (bytes[me] != mine->bytes) { throw error(EPK_E_INVALID, "my own part has another size than I shared"); }
        if (total / 8 >= (1ull << 33)) { throw error(EPK_E_UNSUPPORTED, "a pooled database holds at most 2^33 postings (64 GB)"); }
        api.check(api.MemAddressReserve(&pool->base, total, 0, 0, 0), "cuMemAddressReserve");
        pool->reserved = total;
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            CUmemGenericAllocationHandle handle = 0;
            if (g == me)
            {
                // (a second reference of my own allocation: exported again and imported, so that the pool owns one)
                handle = 0;
                int fd = -1;
                api.check(api.MemExportToShareableHandle(&fd, mine->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
                const CUresult r = api.MemImportFromShareableHandle(&handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
                ::close(fd);
                api.check(r, "cuMemImportFromShareableHandle (own part)");
            }
            else
            {
                api.check(api.MemImportFromShareableHandle(&handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fds[g])), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
                          "cuMemImportFromShareableHandle (is the descriptor this process's own copy, received over a unix socket?)");
            }
            pool->handles.push_back(handle);
            api.check(api.MemMap(pool->base + offset[g], bytes[g], 0, handle, 0), "cuMemMap");
            pool->mapped.emplace_back(offset[g], static_cast<size_t>(bytes[g]));
        }
        CUmemAccessDesc access{};
        access.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        access.location.id = shard->device;
        access.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        api.check(api.MemSetAccess(pool->base, total, &access, 1), "cuMemSetAccess (are the GPUs peers of each other?)");

        auto db = std::make_unique<epk_db>();
        db->device = shard->device;
        db->table_kind = TABLE_DIRECT;
        db->table_slots = shard->table_slots;
        db->direct.resize(shard->table_slots);
        EPK_CUDA(cudaMemset(db->direct.get(), 0, db->direct.bytes()));
        device_buffer<uint32_t> clash(1);
        EPK_CUDA(cudaMemset(clash.get(), 0, 4));
        for (uint32_t g = 0; g < n_shards; ++g)
        {
            const uint64_t* table = shard->direct.get();
            void* opened = nullptr;
            if (g != me)
            {
                cudaIpcMemHandle_t h{};
                std::memcpy(&h, table_handles + static_cast<size_t>(g) * EPK_IPC_HANDLE_BYTES, sizeof(h));
                EPK_CUDA(cudaIpcOpenMemHandle(&opened, h, cudaIpcMemLazyEnablePeerAccess));
                table = static_cast<const uint64_t*>(opened);
            }
            merge_table_kernel<<<blocks_for(shard->table_slots, 256), 256>>>(table, shard->table_slots, offset[g] / 8,
                                                                              reinterpret_cast<unsigned long long*>(db->direct.get()), clash.get());
            const cudaError_t launched = cudaGetLastError();
            const cudaError_t done = cudaDeviceSynchronize();
            if (opened) { cudaIpcCloseMemHandle(opened); }
            EPK_CUDA(launched);
            EPK_CUDA(done);
        }
        uint32_t h_clash = 0;
        EPK_CUDA(cudaMemcpy(&h_clash, clash.get(), 4, cudaMemcpyDeviceToHost));
        if (h_clash) { throw error(EPK_E_INVALID, "two shards hold the same key: not the shards of one database"); }
        db->letters.resize(256);
        EPK_CUDA(cudaMemcpy(db->letters.get(), shard->letters.get(), 256 * 4, cudaMemcpyDeviceToDevice));
        db->d = shard->d;
        db->d.direct = db->direct.get();
        db->d.hashed = nullptr;
        db->d.postings = reinterpret_cast<const uint2*>(pool->base);
        // (the score range of the OTHER shards is not known here: the sweep's short arithmetic forms are only used when
        // the caller says all shards allow them - see epk_db_pool_set_score_range)
        db->distal = shard->distal;
        db->pendant = shard->pendant;
        db->omega = shard->omega;
        db->n_entries_total = shard->n_entries_total;
        db->n_entries_loaded = shard->n_entries_loaded; // (of my shard; the caller sums)
        db->n_keys_loaded = shard->n_keys_loaded;
        db->shard_index = 0;
        db->shard_count = 1;
        db->pool = pool;
        db->pool_local_bytes = bytes[me];
        *out = db.release();
    });
}

// The range of the scores over ALL the shards (each process knows its own: epk_db_get_score_range) decides which
// arithmetic the sweep may use; the processes agree on it (min of lo, max of hi, AND of fast_ok) and set it.
extern "C" epk_status epk_db_get_score_range(const epk_db* db, float* lo, float* hi, uint32_t* fast_ok, uint64_t* entries_loaded, uint64_t* keys_loaded)
{
    return guarded([&] {
        if (!db || !lo || !hi || !fast_ok || !entries_loaded || !keys_loaded) { throw error(EPK_E_INVALID, "null argument"); }
        *lo = db->d.score_lo;
        *hi = db->d.score_hi;
        *fast_ok = db->d.fast_ok;
        *entries_loaded = db->n_entries_loaded;
        *keys_loaded = db->n_keys_loaded;
    });
}

extern "C" epk_status epk_db_pool_set_score_range(epk_db* pooled, float lo, float hi, uint32_t fast_ok, uint64_t entries_loaded, uint64_t keys_loaded)
{
    return guarded([&] {
        if (!pooled || !pooled->pool) { throw error(EPK_E_INVALID, "not a pooled database"); }
        pooled->d.score_lo = lo;
        pooled->d.score_hi = hi;
        pooled->d.fast_ok = fast_ok ? 1u : 0u;
        pooled->n_entries_loaded = entries_loaded;
        pooled->n_keys_loaded = keys_loaded;
    });
}
