// scalix-b200: multi-dimensional index.  Semantics of
// /root/reference scalix/index.cuh:42-109: column-major linearisation (index 0 fastest,
// last dimension slowest) and its inverse.
#pragma once

#include "shape.cuh"

namespace sclx {

using index_t = size_t;

template<uint Rank = 1>
class md_index_t : public shape_like_t<Rank> {
  public:
    __host__ __device__ constexpr md_index_t() : shape_like_t<Rank>() {}
    __host__ __device__ constexpr md_index_t(std::initializer_list<size_t> list)
        : shape_like_t<Rank>(list) {}
    __host__ __device__ constexpr md_index_t(const shape_like_t<Rank>& other)
        : shape_like_t<Rank>(other) {}

    // sum_i idx[i] * prod_{j<i} shape[j]
    __host__ __device__ constexpr index_t as_linear(const shape_t<Rank>& shape) const {
        index_t linear = (*this)[Rank - 1];
        for (uint i = Rank - 1; i-- > 0;) linear = linear * shape[i] + (*this)[i];
        return linear;
    }

    __host__ __device__ constexpr md_index_t operator+(const md_index_t& other) const {
        md_index_t r(*this);
        r += other;
        return r;
    }
    __host__ __device__ constexpr md_index_t operator-(const md_index_t& other) const {
        md_index_t r(*this);
        r -= other;
        return r;
    }
    __host__ __device__ constexpr md_index_t& operator+=(const md_index_t& other) {
        for (uint i = 0; i < Rank; ++i) (*this)[i] += other[i];
        return *this;
    }
    __host__ __device__ constexpr md_index_t& operator-=(const md_index_t& other) {
        for (uint i = 0; i < Rank; ++i) (*this)[i] -= other[i];
        return *this;
    }

    // inverse of as_linear; the last coordinate absorbs any overflow
    template<uint OtherRank = Rank>
    static __host__ __device__ constexpr md_index_t<OtherRank>
    create_from_linear(index_t linear, const shape_t<OtherRank>& shape) {
        md_index_t<OtherRank> idx;
        for (uint i = 0; i + 1 < OtherRank; ++i) {
            const index_t q = linear / shape[i];
            idx[i]          = linear - q * shape[i];
            linear          = q;
        }
        idx[OtherRank - 1] = linear;
        return idx;
    }
};

}  // namespace sclx
