// scalix-b200: algorithm::reduce_last_dim (surface of /root/reference
// scalix/algorithm/reduce_last_dim.cuh:141-269, four overloads):
//     result[s] = identity (f) fold_j arr[s + stride*j],   stride = product of the leading dims.
//
//   reference : per device a SERIAL host loop of `stride` thrust::reduce calls over a strided
//               iterator (one 4-byte word used per 32-byte sector, temp alloc + 2 kernels + D2H
//               + sync per call), then a mutex-ordered host fold through Unified Memory
//               (ibid. :184-211).
//   here      : ONE kernel per partition that reads every byte once, coalesced, and produces all
//               `stride` partial results (reduce_last_dim_kernel); partial rows land in
//               peer-visible scratch and a tiny kernel folds them in FIXED partition order
//               (fold_rows_contig_kernel) — float results are bit-stable run to run.
// `f` must be associative and commutative (as thrust::reduce requires of the reference's
// callers) and `identity` neutral for it.
#pragma once
#include "../detail/b200/host_launch.cuh"
#include "../execute_kernel.cuh"
#include "../fill.cuh"
#include "functional.cuh"

#include <sstream>

namespace sclx::algorithm {

// kept for source compatibility with code that names the reference's helper iterator
// (scalix/algorithm/reduce_last_dim.cuh:45-139); the kernels here do not use it
template<class T>
class strided_iterator {
  public:
    using value_type        = T;
    using difference_type   = std::ptrdiff_t;
    using pointer           = T*;
    using reference         = T&;
    using iterator_category = std::random_access_iterator_tag;

    __host__ __device__ strided_iterator(T* ptr, size_t stride) : ptr_(ptr), stride_(stride) {}
    __host__ __device__ reference operator*() const { return *ptr_; }
    __host__ __device__ reference operator[](difference_type n) const { return ptr_[n * difference_type(stride_)]; }
    __host__ __device__ strided_iterator& operator++() { ptr_ += stride_; return *this; }
    __host__ __device__ strided_iterator operator++(int) { auto t = *this; ptr_ += stride_; return t; }
    __host__ __device__ strided_iterator& operator--() { ptr_ -= stride_; return *this; }
    __host__ __device__ strided_iterator& operator+=(difference_type n) { ptr_ += n * difference_type(stride_); return *this; }
    __host__ __device__ strided_iterator& operator-=(difference_type n) { ptr_ -= n * difference_type(stride_); return *this; }
    __host__ __device__ strided_iterator operator+(difference_type n) const { auto t = *this; t += n; return t; }
    __host__ __device__ strided_iterator operator-(difference_type n) const { auto t = *this; t -= n; return t; }
    __host__ __device__ difference_type operator-(const strided_iterator& o) const { return (ptr_ - o.ptr_) / difference_type(stride_); }
    __host__ __device__ bool operator==(const strided_iterator& o) const { return ptr_ == o.ptr_; }
    __host__ __device__ bool operator!=(const strided_iterator& o) const { return ptr_ != o.ptr_; }
    __host__ __device__ bool operator<(const strided_iterator& o) const { return ptr_ < o.ptr_; }

  private:
    T* ptr_;
    size_t stride_;
};

template<class T, class ResultT, uint Rank, class F>
__host__ void reduce_last_dim(
    const array<const T, Rank>& arr,
    array<ResultT, Rank - 1>& result,
    const T& identity,
    F&& f
) {
    static_assert(Rank >= 2, "reduce_last_dim needs at least two dimensions");
    sclx::shape_t<Rank - 1> result_shape;
    for (uint i = 0; i < Rank - 1; ++i) result_shape[i] = arr.shape()[i];
    if (result_shape != result.shape()) {
        std::stringstream error_stream;
        error_stream << "Result array has wrong shape: "
                     << "expected: " << result_shape << ", got: " << result.shape();
        throw_exception<std::invalid_argument>(error_stream.str(), "sclx::algorithm::");
    }
    using functor_t = std::decay_t<F>;
    static_assert(std::is_trivially_copyable_v<T>, "reduce_last_dim: trivially copyable elements only");

    const int caller_device = cuda::traits::current_device();
    const auto device_info  = get_device_split_info(arr);
    const int nparts        = int(device_info.size());
    const size_t rows       = result_shape.elements();
    const T* in             = arr.data().get();
    functor_t op(f);

    if constexpr (std::is_same_v<T, ResultT>) {
        T* out = result.data().get();
        if (nparts == 1) {
            b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
                b200::launch_reduce_last_dim<T, functor_t>(
                    in + p.offset, rows, rows ? p.elements / rows : 0, identity, out, op, p
                );
            });
            b200::sync_devices(b200::devices_of(device_info));
        } else {
            T* partial_rows = b200::scratch<T>(size_t(nparts) * rows);
            b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
                b200::launch_reduce_last_dim<T, functor_t>(
                    in + p.offset, rows, rows ? p.elements / rows : 0, identity,
                    partial_rows + size_t(p.ordinal) * rows, op, p
                );
            });
            // fixed-order fold on the first partition's device, ordered after all partials
            const sclx::detail::device_info_t folder{device_info[0]};
            sclx::detail::run_launch_group(
                folder,
                [&](int, int device_id, size_t, size_t, cudaStream_t stream) {
                    unsigned grid = unsigned((rows + 255) / 256);
                    const unsigned cap = unsigned(b200::sm_count_of(device_id)) * 8u;
                    if (grid > cap) grid = cap;
                    if (grid == 0) return;
                    b200::fold_rows_contig_kernel<T, functor_t>
                        <<<grid, 256, 0, stream>>>(partial_rows, nparts, rows, identity, out, op);
                }
            );
            auto devices = b200::devices_of(device_info);
            b200::sync_devices(devices);
        }
    } else {
        array<T, Rank - 1> tmp(result_shape);
        reduce_last_dim(arr, tmp, identity, f);
        assign_array(tmp, result).get();
    }
    cuda::set_device(caller_device);
}

template<class T, class ResultT, uint Rank, class F>
__host__ void reduce_last_dim(
    const array<T, Rank>& arr,
    array<ResultT, Rank - 1>& result,
    const T& identity,
    F&& f
) {
    return reduce_last_dim(static_cast<const array<const T, Rank>&>(arr), result, identity, f);
}

template<class T, uint Rank, class F>
__host__ array<T, Rank - 1>
reduce_last_dim(const array<const T, Rank>& arr, const T& identity, F&& f) {
    sclx::shape_t<Rank - 1> result_shape;
    for (uint i = 0; i < Rank - 1; ++i) result_shape[i] = arr.shape()[i];
    sclx::array<T, Rank - 1> result(result_shape);
    reduce_last_dim(arr, result, identity, f);
    return result;
}

template<class T, uint Rank, class F>
__host__ array<T, Rank - 1>
reduce_last_dim(const array<T, Rank>& arr, const T& identity, F&& f) {
    return reduce_last_dim(static_cast<const array<const T, Rank>&>(arr), identity, f);
}

}  // namespace sclx::algorithm
