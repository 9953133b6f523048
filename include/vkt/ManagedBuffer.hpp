// Migratable byte buffer (reference include/cpp/vkt/ManagedBuffer.hpp:20-276).
//
// The buffer remembers the policy it was last allocated under; migrate() compares that policy's device with the
// calling thread's current policy and, when they differ, allocates on the new device, copies, and frees the old
// allocation (reference :168-198).  Like the reference this is header-inline: the real ABI of migration is the
// four Memory.hpp functions plus Get/SetThreadExecutionPolicy.
#pragma once

#include <cstddef>

#include "ExecutionPolicy.hpp"
#include "ManagedResource.hpp"
#include "Memory.hpp"

namespace vkt
{
    template <typename T>
    class ManagedBuffer
    {
    public:
        typedef T value_type;

        ManagedBuffer(std::size_t size = 0) : size_(size)
        {
            allocate(size);
            resourceHandle_ = RegisterManagedResource(this);
        }

        ManagedBuffer(ManagedBuffer& rhs) : size_(rhs.size_), lastAllocationPolicy_(rhs.lastAllocationPolicy_)
        {
            rhs.migrate();
            allocate(rhs.size_);
            copy(rhs);
            resourceHandle_ = RegisterManagedResource(this);
        }

        ManagedBuffer(ManagedBuffer&& rhs) : size_(rhs.size_), lastAllocationPolicy_(rhs.lastAllocationPolicy_)
        {
            rhs.migrate();
            // both buffers now live on the current policy's device: take the allocation over
            data_ = rhs.data_;
            lastAllocationPolicy_ = rhs.lastAllocationPolicy_;
            rhs.data_ = nullptr;
            rhs.size_ = 0;
            resourceHandle_ = RegisterManagedResource(this);
        }

        virtual ~ManagedBuffer()
        {
            UnregisterManagedResource(resourceHandle_);
            deallocate();
        }

        ManagedBuffer& operator=(ManagedBuffer& rhs)
        {
            if (&rhs != this) {
                rhs.migrate();
                deallocate();
                size_ = rhs.size_;
                allocate(rhs.size_);
                copy(rhs);
            }
            return *this;
        }

        ManagedBuffer& operator=(ManagedBuffer&& rhs)
        {
            if (&rhs != this) {
                rhs.migrate();
                deallocate();
                data_ = rhs.data_;
                size_ = rhs.size_;
                lastAllocationPolicy_ = rhs.lastAllocationPolicy_;
                rhs.data_ = nullptr;
                rhs.size_ = 0;
            }
            return *this;
        }

        ResourceHandle getResourceHandle() const { return resourceHandle_; }

        // Moves the data to the device named by the calling thread's policy (no-op when it is already there).
        void migrate()
        {
            ExecutionPolicy ep = GetThreadExecutionPolicy();
            if (ep.device == lastAllocationPolicy_.device)
                return;
            T* moved = nullptr;
            Allocate((void**)&moved, size_ * sizeof(T));                       // on the new device
            if (!moved)
                return;   // out of memory there: the bytes stay where they are (host::describe reports it to the caller)
            Memcpy(moved, data_, size_ * sizeof(T),
                   ep.device == ExecutionPolicy::Device::GPU ? CopyKind::HostToDevice : CopyKind::DeviceToHost);
            SetThreadExecutionPolicy(lastAllocationPolicy_);                   // free with the backend that allocated
            Free(data_);
            SetThreadExecutionPolicy(ep);
            lastAllocationPolicy_ = ep;
            data_ = moved;
        }

        // volkit-b200 extension: migrate() for a buffer whose whole content is about to be overwritten (InputStream::read
        // under the GPU policy): allocate on the new device, free the old allocation, copy nothing
        void migrateDiscard()
        {
            ExecutionPolicy ep = GetThreadExecutionPolicy();
            if (ep.device == lastAllocationPolicy_.device)
                return;
            T* fresh = nullptr;
            Allocate((void**)&fresh, size_ * sizeof(T));
            if (!fresh)
                return;
            SetThreadExecutionPolicy(lastAllocationPolicy_);
            Free(data_);
            SetThreadExecutionPolicy(ep);
            lastAllocationPolicy_ = ep;
            data_ = fresh;
        }

        // volkit-b200 extension: where the bytes live right now, and the pointer WITHOUT migrating
        bool residesOnHost() const { return lastAllocationPolicy_.device == ExecutionPolicy::Device::CPU; }
        T* dataInPlace() { return data_; }

    protected:
        void allocate(std::size_t size)
        {
            lastAllocationPolicy_ = GetThreadExecutionPolicy();
            Allocate((void**)&data_, size * sizeof(T));
        }

        void deallocate()
        {
            if (!data_) return;
            ExecutionPolicy cur = GetThreadExecutionPolicy();
            SetThreadExecutionPolicy(lastAllocationPolicy_);
            Free(data_);
            SetThreadExecutionPolicy(cur);
            data_ = nullptr;
        }

        void resize(std::size_t size)
        {
            migrate();
            T* grown = nullptr;
            Allocate((void**)&grown, size * sizeof(T));
            const bool gpu = lastAllocationPolicy_.device == ExecutionPolicy::Device::GPU;
            const std::size_t keep = (size < size_ ? size : size_) * sizeof(T);
            if (data_ && keep) Memcpy(grown, data_, keep, gpu ? CopyKind::DeviceToDevice : CopyKind::HostToHost);
            Free(data_);
            data_ = grown;
            size_ = size;
        }

        void fill(T const& value)
        {
            migrate();
            MemsetRange(data_, &value, size_ * sizeof(T), sizeof(T));
        }
        void fill(T& value) { fill(static_cast<T const&>(value)); }

        // rhs has been migrated to the current policy's device by the caller
        void copy(ManagedBuffer& rhs)
        {
            const bool gpu = lastAllocationPolicy_.device == ExecutionPolicy::Device::GPU;
            const bool rgpu = rhs.lastAllocationPolicy_.device == ExecutionPolicy::Device::GPU;
            CopyKind ck = gpu ? (rgpu ? CopyKind::DeviceToDevice : CopyKind::HostToDevice)
                              : (rgpu ? CopyKind::DeviceToHost : CopyKind::HostToHost);
            Memcpy(data_, rhs.data_, size_ * sizeof(T), ck);
        }

        T* data_ = nullptr;
        std::size_t size_ = 0;

    private:
        ExecutionPolicy lastAllocationPolicy_ = {};
        ResourceHandle resourceHandle_ = ResourceHandle(-1);
    };
} // vkt
