// scalix-b200: algorithm::elementwise_reduce (surface of /root/reference
// scalix/algorithm/elementwise_reduce.cuh:39-121):
//     result[i] = op(...op(op(a[i], b[i]), c[i])..., z[i])        asynchronous.
// This is the vectorisable form of an axpy-style update: every thread gathers 8 results from
// independent coalesced loads before storing (generate_kernel), which the opaque
// per-element lambda of kernel_handler::launch cannot do.
#pragma once
#include "../detail/b200/host_launch.cuh"
#include "../execute_kernel.cuh"
#include "../throw_exception.hpp"
#include "functional.cuh"

#include <thrust/tuple.h>

namespace sclx::algorithm {

// true when all arrays have the same shape (pairwise, left to right)
template<class T, class U, uint Rank, class... Ts>
bool check_shapes(
    const sclx::array<T, Rank>& a,
    const sclx::array<U, Rank>& b,
    const sclx::array<Ts, Rank>&... args
) {
    if (a.shape() != b.shape()) return false;
    if constexpr (sizeof...(args) > 0) return check_shapes(b, args...);
    return true;
}

// left fold of op over one index of several arrays: compute(i, op) = op(...op(op(a[i], b[i]),
// c[i])..., z[i]), evaluated by compile-time recursion over the pack so that the operand
// order (and therefore non-commutative or rounding-sensitive ops) matches the reference.
// Holds raw pointers, not array handles: the functor travels into generate_kernel by value
// and the arrays themselves are kept alive by the caller until the returned future is ready.
template<class R, class T, class U, uint Rank, class... Ts>
struct reduce_along_index {
    const T* a_;
    const U* b_;
    thrust::tuple<const Ts*...> args_;

    constexpr static uint num_args = sizeof...(Ts);

    __host__ reduce_along_index(
        const sclx::array<T, Rank>& a,
        const sclx::array<U, Rank>& b,
        const sclx::array<Ts, Rank>&... args
    )
        : a_(a.data().get()),
          b_(b.data().get()),
          args_(args.data().get()...) {}

    template<class BinaryOp, uint I>
    __host__ __device__ R
    compute_impl(const index_t& idx, BinaryOp&& op, const thrust::integral_constant<uint, I>&) const {
        if constexpr (I > 0) {
            return op(
                compute_impl(idx, op, thrust::integral_constant<uint, I - 1>()),
                thrust::get<I - 1>(args_)[idx]
            );
        } else {
            return op(a_[idx], b_[idx]);
        }
    }

    template<class BinaryOp>
    __host__ __device__ R compute(const index_t& idx, BinaryOp&& op) const {
        return compute_impl(idx, op, thrust::integral_constant<uint, num_args>());
    }
};

REGISTER_SCALIX_KERNEL_TAG(elementwise_reduce_kernel);  // kept for source compatibility

template<class BinaryOp, class R, class T, class U, uint Rank, class... Ts>
std::future<void> elementwise_reduce(
    BinaryOp&& op,
    sclx::array<R, Rank> result,
    sclx::array<T, Rank> a,
    sclx::array<U, Rank> b,
    sclx::array<Ts, Rank>... args
) {
    if (!check_shapes(a, b, args...) || result.shape() != a.shape()) {
        throw_exception<std::invalid_argument>("input shapes must match", "sclx::algorithm::");
    }
    reduce_along_index<R, T, U, Rank, Ts...> functor(a, b, args...);
    return sclx::execute_kernel([=](const kernel_handler&) {
        R* dst = result.data().get();
        std::decay_t<BinaryOp> fn(op);
        b200::for_each_partition(result, [&](const b200::partition_t& p) {
            b200::launch_generate(
                dst + p.offset, p.elements, p.offset,
                [=] __device__(size_t i) { return functor.compute(i, fn); }, p
            );
        });
    });
}

}  // namespace sclx::algorithm
