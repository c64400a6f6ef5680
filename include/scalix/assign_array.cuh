// scalix-b200: assign_array (surface of /root/reference scalix/assign_array.cuh:38-57):
// result[i] = static_cast<U>(source[i]) over the flattened storage, asynchronous.
// Known access pattern => generate_kernel (independent coalesced loads before any store)
// instead of the per-element lambda launcher.
#pragma once
#include "detail/b200/host_launch.cuh"
#include "execute_kernel.cuh"

namespace sclx {

template<class T, class U, uint Rank>
__host__ std::future<void> assign_array(array<T, Rank> source, array<U, Rank> result) {
    if (source.elements() != result.elements()) {
        throw_exception<std::runtime_error>(
            "Cannot assign array of different size", "sclx::array::"
        );
    }
    return execute_kernel([=](kernel_handler&) {
        const T* src = source.data().get();
        U* dst       = result.data().get();
        b200::for_each_partition(result, [&](const b200::partition_t& p) {
            b200::launch_generate(
                dst + p.offset, p.elements, p.offset,
                [=] __device__(size_t i) { return static_cast<U>(src[i]); }, p
            );
        });
    });
}

}  // namespace sclx
