// scalix-b200: shape value types.  Same public surface and semantics as
// /root/reference scalix/shape.cuh:57-130 (first index is the fastest varying one;
// elements() is the product of all extents; the initializer-list constructor throws
// std::invalid_argument on a size mismatch in host code).
#pragma once

#include "constexpr_assign_array.cuh"
#include "throw_exception.hpp"
#include <cstddef>
#include <initializer_list>
#include <ostream>

namespace sclx {

using uint = unsigned int;

template<uint Rank>
class shape_like_t {
  public:
    constexpr shape_like_t() = default;

    __host__ __device__ constexpr shape_like_t(std::initializer_list<size_t> list) {
#ifndef __CUDA_ARCH__
        if (list.size() != Rank) {
            throw_exception<std::invalid_argument>(
                "Shape initializer list size does not match rank",
                "sclx::shape_like_t::"
            );
        }
#endif
        const size_t* src = list.begin();
        for (uint i = 0; i < Rank; ++i) extents_[i] = src[i];
    }

    constexpr shape_like_t(const shape_like_t&)            = default;
    constexpr shape_like_t& operator=(const shape_like_t&) = default;

    __host__ __device__ constexpr shape_like_t(const size_t (&extents)[Rank]) {
        for (uint i = 0; i < Rank; ++i) extents_[i] = extents[i];
    }

    __host__ __device__ constexpr const size_t& operator[](size_t i) const {
        return extents_[i];
    }
    __host__ __device__ constexpr size_t& operator[](size_t i) { return extents_[i]; }

    [[nodiscard]] __host__ __device__ constexpr uint rank() const { return Rank; }

    __host__ __device__ constexpr bool operator==(const shape_like_t& other) const {
        for (uint i = 0; i < Rank; ++i)
            if (extents_[i] != other.extents_[i]) return false;
        return true;
    }
    __host__ __device__ constexpr bool operator!=(const shape_like_t& other) const {
        return !(*this == other);
    }

    // "(a, b, c)"
    __host__ friend std::ostream& operator<<(std::ostream& os, const shape_like_t& s) {
        os << "(";
        for (uint i = 0; i < Rank; ++i) os << (i ? ", " : "") << s.extents_[i];
        return os << ")";
    }

  protected:
    size_t extents_[Rank]{};
};

template<uint Rank>
class shape_t : public shape_like_t<Rank> {
  public:
    using shape_like_t<Rank>::shape_like_t;
    constexpr shape_t() = default;
    __host__ __device__ constexpr shape_t(const shape_like_t<Rank>& other)
        : shape_like_t<Rank>(other) {}

    __host__ __device__ constexpr size_t elements() const {
        size_t count = 1;
        for (uint i = 0; i < Rank; ++i) count *= this->extents_[i];
        return count;
    }
};

}  // namespace sclx
