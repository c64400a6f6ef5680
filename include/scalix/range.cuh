// scalix-b200: iteration range (API of /root/reference scalix/range.cuh:40-54).
#pragma once

#include "shape.cuh"

namespace sclx {

template<uint Rank>
class md_range_t : public shape_t<Rank> {
  public:
    constexpr md_range_t() = default;
    __host__ __device__ constexpr md_range_t(std::initializer_list<size_t> list)
        : shape_t<Rank>(list) {}
    __host__ __device__ constexpr explicit md_range_t(const shape_t<Rank>& shape)
        : shape_t<Rank>(shape) {}
    constexpr md_range_t(const md_range_t&)            = default;
    constexpr md_range_t& operator=(const md_range_t&) = default;
};

}  // namespace sclx
