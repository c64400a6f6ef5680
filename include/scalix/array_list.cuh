// scalix-b200: fixed-size list of same-type arrays, indexable on host and device
// (surface of /root/reference scalix/array_list.cuh:36-153).  Every single-array launch is
// wrapped into array_list<T, Rank, 1> by kernel_handler::launch.
#pragma once
#include "array.cuh"
#include "constexpr_assign_array.cuh"

#include <array>
#include <initializer_list>
#include <vector>

namespace sclx {

// N handles (not N copies of the data): an sclx::array is a reference-counted view, so the
// list is N pointers + shapes, cheap to pass by value into a kernel.  Inside a device lambda
// `list[k][idx]` addresses element idx of the k-th array; on the host the launcher iterates the
// list to validate last extents and partitioning before it queues anything.
template<class T, uint Rank, uint N>
class array_list {
  public:
    array_list()                                 = default;
    array_list(const array_list&)                = default;
    array_list& operator=(const array_list&)     = default;
    array_list(array_list&&) noexcept            = default;
    array_list& operator=(array_list&&) noexcept = default;

    // array_list<T> -> array_list<const T> (read-only views of the same storage)
    template<class T_ = const T>
    __host__ operator array_list<T_, Rank, N>() const {  // NOLINT
        static_assert(std::is_same_v<T_, const T>, "Can only cast to const");
        std::vector<array<T_, Rank>> const_arrays(arrays_, arrays_ + N);
        return array_list<T_, Rank, N>(const_arrays);
    }

    // the four ways the reference allows a list to be built: braces, std::vector, raw pointer
    // to N arrays, std::array; the first two check the count against N (std::invalid_argument)
    __host__ array_list(std::initializer_list<array<T, Rank>> arrays) {
        check_count(arrays.size());
        constexpr_assign_array<N>(arrays_, arrays.begin());
    }

    __host__ explicit array_list(const std::vector<array<T, Rank>>& arrays) {
        check_count(arrays.size());
        constexpr_assign_array<N>(arrays_, arrays.data());
    }

    __host__ explicit array_list(const array<T, Rank>* arrays) {
        constexpr_assign_array<N>(arrays_, arrays);
    }

    __host__ explicit array_list(const std::array<array<T, Rank>, N>& arrays) {
        constexpr_assign_array<N>(arrays_, arrays.data());
    }

    // k-th array (no bounds check, as in the reference)
    __host__ __device__ array<T, Rank>& operator[](size_t i) { return arrays_[i]; }
    __host__ __device__ const array<T, Rank>& operator[](size_t i) const { return arrays_[i]; }

    __host__ __device__ constexpr size_t size() const { return N; }

    __host__ __device__ array<T, Rank>* begin() { return arrays_; }
    __host__ __device__ array<T, Rank>* end() { return arrays_ + N; }
    __host__ __device__ const array<T, Rank>* begin() const { return arrays_; }
    __host__ __device__ const array<T, Rank>* end() const { return arrays_ + N; }

  private:
    static void check_count(size_t count) {
        if (count != N) {
            throw_exception<std::invalid_argument>(
                "array_list must be initialized with exactly N arrays",
                "sclx::array_list::"
            );
        }
    }

    array<T, Rank> arrays_[N]{};
};

// Run-time sized list.  The handles live in pinned, device-mapped host memory so the list
// can be indexed from kernels.  (The reference's version cannot be instantiated — its
// storage pointer has the wrong type, scalix/array_list.cuh:114-118,150 — so only the
// declared interface is reproduced.)
template<class T, uint Rank>
class dynamic_array_list {
  public:
    dynamic_array_list()                                         = default;
    dynamic_array_list(const dynamic_array_list&)                = default;
    dynamic_array_list(dynamic_array_list&&) noexcept            = default;
    dynamic_array_list& operator=(const dynamic_array_list&)     = default;
    dynamic_array_list& operator=(dynamic_array_list&&) noexcept = default;

    __host__ dynamic_array_list(const std::vector<array<T, Rank>>& arrays) {  // NOLINT
        assign(arrays.data(), arrays.size());
    }

    __host__ dynamic_array_list(const std::initializer_list<array<T, Rank>>& arrays) {
        assign(arrays.begin(), arrays.size());
    }

    __host__ __device__ array<T, Rank>& operator[](size_t i) { return data_.get()[i]; }
    __host__ __device__ const array<T, Rank>& operator[](size_t i) const { return data_.get()[i]; }

    __host__ __device__ const size_t& size() const { return size_; }

    __host__ __device__ array<T, Rank>* begin() { return data_.get(); }
    __host__ __device__ array<T, Rank>* end() { return data_.get() + size_; }
    __host__ __device__ const array<T, Rank>* begin() const { return data_.get(); }
    __host__ __device__ const array<T, Rank>* end() const { return data_.get() + size_; }

  private:
    __host__ void assign(const array<T, Rank>* src, size_t count) {
        using elem_t = array<T, Rank>;
        void* raw    = nullptr;
        cuda::cuda_exception::raise_if_not_success(
            cudaHostAlloc(&raw, (count ? count : 1) * sizeof(elem_t),
                          cudaHostAllocPortable | cudaHostAllocMapped),
            std::experimental::source_location::current(),
            "sclx::dynamic_array_list::"
        );
        auto* elems = static_cast<elem_t*>(raw);
        for (size_t i = 0; i < count; ++i) new (elems + i) elem_t(src[i]);
        data_ = detail::unified_ptr<elem_t>(elems, [count](elem_t* p) {
            for (size_t i = 0; i < count; ++i) p[i].~elem_t();
            cudaFreeHost(p);
        });
        size_ = count;
    }

    detail::unified_ptr<array<T, Rank>> data_{};
    size_t size_{};
};

}  // namespace sclx
