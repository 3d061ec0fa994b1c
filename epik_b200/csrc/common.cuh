// Host-side helpers shared by the C-ABI translation units: error plumbing (no exception ever
// crosses the ABI — reference errors surface as std::runtime_error caught in main.cpp:384-388,
// ours as epk_status + epk_last_error) and RAII wrappers around CUDA allocations.
#pragma once

#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <utility>
#include <cuda_runtime.h>
#include "../../include/epik_b200.h"

namespace epk
{
struct error : std::runtime_error
{
    epk_status status;
    error(epk_status s, const std::string& what) : std::runtime_error(what), status(s) {}
};

void set_last_error(const std::string& message);

#define EPK_CUDA(call)                                                                                  \
    do                                                                                                  \
    {                                                                                                   \
        const cudaError_t epk_cuda_err = (call);                                                        \
        if (epk_cuda_err != cudaSuccess)                                                                \
        {                                                                                               \
            throw ::epk::error(epk_cuda_err == cudaErrorMemoryAllocation ? EPK_E_NOMEM : EPK_E_CUDA,     \
                               std::string(#call) + ": " + cudaGetErrorString(epk_cuda_err));           \
        }                                                                                               \
    } while (0)

// Runs body(), maps exceptions to a status code and records the message.
template<class F>
epk_status guarded(F&& body) noexcept
{
    try
    {
        body();
        return EPK_OK;
    }
    catch (const error& e)
    {
        set_last_error(e.what());
        return e.status;
    }
    catch (const std::bad_alloc&)
    {
        set_last_error("out of host memory");
        return EPK_E_NOMEM;
    }
    catch (const std::exception& e)
    {
        set_last_error(e.what());
        return EPK_E_INVALID;
    }
}

template<class T>
class device_buffer
{
public:
    device_buffer() = default;
    explicit device_buffer(size_t count) { resize(count); }
    device_buffer(const device_buffer&) = delete;
    device_buffer& operator=(const device_buffer&) = delete;
    device_buffer(device_buffer&& o) noexcept : _ptr(o._ptr), _count(o._count) { o._ptr = nullptr; o._count = 0; }
    device_buffer& operator=(device_buffer&& o) noexcept
    {
        if (this != &o) { release(); _ptr = o._ptr; _count = o._count; o._ptr = nullptr; o._count = 0; }
        return *this;
    }
    ~device_buffer() { release(); }

    void resize(size_t count)
    {
        release();
        if (count)
        {
            EPK_CUDA(cudaMalloc(reinterpret_cast<void**>(&_ptr), count * sizeof(T)));
            _count = count;
        }
    }
    void release() noexcept
    {
        if (_ptr) { cudaFree(_ptr); }
        _ptr = nullptr;
        _count = 0;
    }
    T* get() const noexcept { return _ptr; }
    size_t size() const noexcept { return _count; }
    size_t bytes() const noexcept { return _count * sizeof(T); }

private:
    T* _ptr = nullptr;
    size_t _count = 0;
};

template<class T>
class pinned_buffer
{
public:
    pinned_buffer() = default;
    pinned_buffer(const pinned_buffer&) = delete;
    pinned_buffer& operator=(const pinned_buffer&) = delete;
    ~pinned_buffer() { release(); }
    void resize(size_t count)
    {
        release();
        if (count)
        {
            EPK_CUDA(cudaMallocHost(reinterpret_cast<void**>(&_ptr), count * sizeof(T)));
            _count = count;
        }
    }
    void release() noexcept
    {
        if (_ptr) { cudaFreeHost(_ptr); }
        _ptr = nullptr;
        _count = 0;
    }
    T* get() const noexcept { return _ptr; }
    size_t size() const noexcept { return _count; }

private:
    T* _ptr = nullptr;
    size_t _count = 0;
};

inline unsigned blocks_for(uint64_t n, unsigned threads)
{
    return static_cast<unsigned>((n + threads - 1) / threads);
}
} // namespace epk
