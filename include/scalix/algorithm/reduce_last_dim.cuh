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
//               peer-visible scratch (plain NVLink stores) and the partition that finishes LAST
//               folds them in FIXED partition order inside the same kernel (row_exchange_t) —
//               one launch group, float results bit-stable run to run.
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
        b200::call_workspaces ws(device_info);  // leased for this call only
        if (nparts == 1) {
            b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
                b200::launch_reduce_last_dim<T, functor_t>(
                    in + p.offset, rows, rows ? p.elements / rows : 0, identity, out, op, p, ws
                );
            });
            b200::sync_devices(b200::devices_of(device_info));
        } else {
            // ONE launch group: every partition's kernel stores its row in peer-visible
            // scratch and the last one to arrive folds all rows in fixed partition order
            // straight into `result` (row_exchange_t, reduce.cuh) — no second group, no
            // cross-device event waits, no host fold
            const size_t row_bytes = (size_t(nparts) * rows * sizeof(T) + 255) / 256 * 256;
            char* block            = ws.scratch<char>(row_bytes + 256);
            b200::row_exchange_t xchg;
            xchg.rows      = block;
            xchg.arrivals  = reinterpret_cast<unsigned*>(block + row_bytes);
            xchg.final_out = out;
            xchg.nparts    = nparts;
            // the partition that folds last also tells the host: one polled word of pinned host
            // memory instead of N stream synchronisations in a row
            struct flag_lease {
                unsigned* flag = nullptr;
                ~flag_lease() { sclx_b200_rt_host_flag_release(flag); }
            } done;
            unsigned ticket = 0;
            cuda::detail::check_rt(sclx_b200_rt_host_flag_acquire(&done.flag, &ticket));
            xchg.done_flag   = done.flag;
            xchg.done_ticket = ticket;
            b200::for_each_partition(arr, device_info, [&](const b200::partition_t& p) {
                b200::row_exchange_t mine = xchg;
                mine.ordinal              = p.ordinal;
                b200::launch_reduce_last_dim<T, functor_t>(
                    in + p.offset, rows, rows ? p.elements / rows : 0, identity, nullptr, op, p, ws, mine
                );
            });
            auto devices = b200::devices_of(device_info);
            if (rows > 0 && arr.elements() > 0)
                cuda::detail::check_rt(
                    sclx_b200_rt_host_flag_wait(done.flag, ticket, int(devices.size()), devices.data())
                );
            else b200::sync_devices(devices);
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
