// 3-D array of objects, x-fastest (reference include/cpp/vkt/Array3D.hpp:16-136): dims(), operator[](Vec3i), begin/end,
// resize, fill, data, numElements, empty.
//
// One deliberate difference.  The reference derives Array3D<T> from ManagedBuffer<T>, i.e. the elements are raw,
// never-constructed bytes that follow the execution policy.  For T = StructuredVolume (the only instantiation the
// hot path uses, BrickDecompose) that lets BrickDecomposeResize move-assign onto garbage (src/vkt/Decompose.cpp:127:
// whether it frees wild pointers depends on the heap's history) and would ship C++ objects to the device under the GPU
// policy.  Here the elements are ordinary host objects, constructed and destroyed with the array; it is each brick's
// voxel buffer that migrates, on its own, like any other volume.
#pragma once

#include <cstddef>
#include <utility>

#include "linalg.hpp"

namespace vkt
{
    template <typename T>
    class Array3D
    {
    public:
        typedef T value_type;
        typedef T* iterator;
        typedef T const* const_iterator;

        Array3D() = default;
        explicit Array3D(Vec3i const& dims) : data_(count(dims) ? new T[count(dims)]() : nullptr), dims_(dims) {}

        // T may only be copyable from a non-const reference (StructuredVolume, like every ManagedBuffer)
        Array3D(Array3D& rhs) : Array3D(rhs.dims_)
        {
            for (std::size_t i = 0; i < numElements(); ++i) data_[i] = rhs.data_[i];
        }
        Array3D(Array3D&& rhs) : data_(rhs.data_), dims_(rhs.dims_)
        {
            rhs.data_ = nullptr;
            rhs.dims_ = Vec3i{ 0, 0, 0 };
        }
        ~Array3D() { delete[] data_; }

        Array3D& operator=(Array3D& rhs)
        {
            if (&rhs != this) {
                Array3D copy(rhs);
                swap(copy);
            }
            return *this;
        }
        Array3D& operator=(Array3D&& rhs)
        {
            if (&rhs != this) {
                Array3D taken(std::move(rhs));
                swap(taken);
            }
            return *this;
        }

        // keeps the leading elements in linear order, as the reference's buffer resize keeps the leading bytes
        void resize(Vec3i const& dims)
        {
            Array3D grown(dims);
            std::size_t keep = numElements() < grown.numElements() ? numElements() : grown.numElements();
            for (std::size_t i = 0; i < keep; ++i) grown.data_[i] = std::move(data_[i]);
            swap(grown);
        }

        void fill(T& value) { for (std::size_t i = 0; i < numElements(); ++i) data_[i] = value; }
        void fill(T const& value) { fill(const_cast<T&>(value)); }

        iterator begin() { return data_; }
        const_iterator begin() const { return data_; }
        const_iterator cbegin() { return data_; }
        iterator end() { return data_ + numElements(); }
        const_iterator end() const { return data_ + numElements(); }
        const_iterator cend() { return data_ + numElements(); }

        T& operator[](Vec3i const& index) { return data_[linear(index)]; }
        T const& operator[](Vec3i const& index) const { return data_[linear(index)]; }

        bool empty() const { return numElements() == 0; }
        T* data() { return data_; }
        T const* data() const { return data_; }
        Vec3i dims() const { return dims_; }
        std::size_t numElements() const { return count(dims_); }

    private:
        static std::size_t count(Vec3i const& d) { return d.x * std::size_t(d.y) * d.z; }
        std::size_t linear(Vec3i const& i) const { return i.z * dims_.x * std::size_t(dims_.y) + i.y * std::size_t(dims_.x) + i.x; }
        void swap(Array3D& o) { std::swap(data_, o.data_); std::swap(dims_, o.dims_); }

        T*    data_ = nullptr;
        Vec3i dims_ = { 0, 0, 0 };
    };
} // vkt
