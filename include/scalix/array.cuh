// scalix-b200: sclx::array<T, Rank> — the distributed N-d array of the hot path.
//
// Public surface of /root/reference scalix/array.cuh:63-1051 (unified_ptr, array,
// array_memory_info_t, get_device_split_info, is_same_device_split, reshape, flatten) with a
// different memory manager underneath:
//
//   reference : ONE cudaMallocManaged allocation + cudaMemAdvise(PreferredLocation /
//               AccessedBy / ReadMostly) + cudaMemPrefetchAsync per slice
//               (scalix/array.cuh:241-248,517-615); data moves by page faults.
//   here      : ONE contiguous virtual range (CUDA VMM) whose pages are explicit physical
//               allocations on the device that owns the partition, mapped read/write into
//               EVERY device (sclx_b200_array_create).  The same pointer value is valid on
//               all GPUs: local accesses hit HBM, cross-partition accesses are plain
//               NVLink/NVSwitch peer loads, stores and atomics — nothing faults or migrates.
//               Host code may still dereference elements (the reference's examples do):
//               a software mirror at the same addresses is filled on first touch (in spans
//               that grow with sequential access) and written back before the next device
//               access (runtime.cpp).  User-space loads and stores fault pages in; a SYSCALL
//               (read(2), fread, recv ...) cannot, so bulk I/O into array memory goes after
//               prefetch_async({cpu_device_id}), which makes the range host-resident,
//               readable and writable, and pinned for PCIe-rate copies when it is large.
//
// The partition arithmetic is the reference's: the LAST dimension is split into contiguous
// blocks of ceil(last/ndev) columns, surplus devices are dropped (ibid. :471-508).
// Device ids are the runtime's logical devices (see cuda.hpp).
#pragma once

#include "cuda.hpp"
#include "index.cuh"
#include "range.cuh"
#include "shape.cuh"
#include "throw_exception.hpp"

#include <algorithm>
#include <future>
#include <iostream>
#include <memory>
#include <tuple>
#include <type_traits>
#include <vector>
// serialization helpers need cereal only where the consumer has it on the include path
// (the reference vendors it as extern/cereal and includes it unconditionally, array.cuh:59-61)
#if defined(__has_include)
#if __has_include(<cereal/cereal.hpp>)
#define SCLX_B200_HAS_CEREAL 1
#ifndef __CUDA_ARCH__
#include <cereal/archives/binary.hpp>
#include <cereal/cereal.hpp>
#endif
#endif
#endif

namespace sclx {

namespace detail {

template<class T>
using dummy_shared_ptr_t = char[sizeof(std::shared_ptr<T>)];

// frees whatever a unified_ptr may own: a runtime array, else plain CUDA memory
inline void release_device_memory(const void* ptr) {
    if (ptr == nullptr) return;
    void* p = const_cast<void*>(ptr);
    if (sclx_b200_array_owner(p) >= 0) {
        sclx_b200_array_free(p);
        return;
    }
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost) {
        cudaFreeHost(p);
        return;
    }
    cudaGetLastError();
    cudaFree(p);
}

template<class T>
class unified_ptr;

template<class T, class U>
__host__ __device__ unified_ptr<T> static_pointer_cast(const unified_ptr<U>& ptr);
template<class T, class U>
__host__ unified_ptr<T> dynamic_pointer_cast(const unified_ptr<U>& ptr);
template<class T, class U>
__host__ __device__ unified_ptr<T> const_pointer_cast(const unified_ptr<U>& ptr);
template<class T, class U>
__host__ __device__ unified_ptr<T> reinterpret_pointer_cast(const unified_ptr<U>& ptr);

// Host: reference-counted owner + raw pointer.  Device: raw pointer + same-size padding, so
// an object holding a unified_ptr has ONE layout and can be copied into a kernel's
// parameter space byte for byte.
template<class T>
class unified_ptr {
  public:
    template<class T_, class U_>
    friend __host__ __device__ unified_ptr<T_> static_pointer_cast(const unified_ptr<U_>&);
    template<class T_, class U_>
    friend __host__ unified_ptr<T_> dynamic_pointer_cast(const unified_ptr<U_>&);
    template<class T_, class U_>
    friend __host__ __device__ unified_ptr<T_> const_pointer_cast(const unified_ptr<U_>&);
    template<class T_, class U_>
    friend __host__ __device__ unified_ptr<T_> reinterpret_pointer_cast(const unified_ptr<U_>&);
    friend class unified_ptr<const T>;
    friend class unified_ptr<std::remove_const_t<T>>;

    unified_ptr() = default;

    // implicit conversion to pointer-to-const
    template<class T_ = const T>
    __host__ __device__ operator unified_ptr<T_>() const {  // NOLINT
        static_assert(std::is_same_v<const T, T_>, "Invalid cast");
        unified_ptr<T_> out;
        out.raw_ptr_ = raw_ptr_;
#ifndef __CUDA_ARCH__
        out.ptr_ = ptr_;
#endif
        return out;
    }

    // adopts `ptr`: the last owner destroys the object and releases the memory
    __host__ __device__ explicit unified_ptr(T* ptr) : raw_ptr_(ptr) {
#ifndef __CUDA_ARCH__
        ptr_ = std::shared_ptr<T>(ptr, [](T* p) {
            if (p) p->~T();
            release_device_memory(p);
        });
#endif
    }

    template<class Deleter>
    __host__ __device__ explicit unified_ptr(T* ptr, Deleter deleter) : raw_ptr_(ptr) {
#ifndef __CUDA_ARCH__
        ptr_ = std::shared_ptr<T>(ptr, deleter);
#endif
    }

    __host__ __device__ unified_ptr(const unified_ptr& other) : raw_ptr_(other.raw_ptr_) {
#ifndef __CUDA_ARCH__
        ptr_ = other.ptr_;
#endif
    }

    __host__ __device__ unified_ptr(unified_ptr&& other) noexcept : raw_ptr_(other.raw_ptr_) {
#ifndef __CUDA_ARCH__
        ptr_ = std::move(other.ptr_);
#endif
        other.raw_ptr_ = nullptr;
    }

    __host__ __device__ unified_ptr& operator=(const unified_ptr& other) {
        if (this == &other) return *this;
#ifndef __CUDA_ARCH__
        ptr_ = other.ptr_;
#endif
        raw_ptr_ = other.raw_ptr_;
        return *this;
    }

    __host__ __device__ unified_ptr& operator=(unified_ptr&& other) noexcept {
        if (this == &other) return *this;
#ifndef __CUDA_ARCH__
        ptr_ = std::move(other.ptr_);
#endif
        raw_ptr_       = other.raw_ptr_;
        other.raw_ptr_ = nullptr;
        return *this;
    }

    __host__ __device__ T& operator*() const { return *raw_ptr_; }

    __host__ __device__ auto operator->() const {
#ifndef __CUDA_ARCH__
        return ptr_;
#else
        return raw_ptr_;
#endif
    }

    __host__ __device__ T* get() const { return raw_ptr_; }

    __host__ __device__ explicit operator bool() const { return raw_ptr_ != nullptr; }

    __host__ auto use_count() const -> decltype(std::shared_ptr<T>{}.use_count()) {
#ifndef __CUDA_ARCH__
        return ptr_.use_count();
#else
        return 0;
#endif
    }

    // for views: the parent keeps the memory alive
    static constexpr auto no_delete = [](T*) {};

    // scalix-b200 extension: a pointer into `owner`'s allocation that SHARES its ownership
    // (views built by array::get_slice / get_range keep the allocation alive)
    __host__ __device__ static unified_ptr aliasing(const unified_ptr& owner, T* ptr) {
        unified_ptr out;
        out.raw_ptr_ = ptr;
#ifndef __CUDA_ARCH__
        out.ptr_ = std::shared_ptr<T>(owner.ptr_, ptr);
#else
        (void)owner;
#endif
        return out;
    }

    class safe_delete_array {
      public:
        __host__ safe_delete_array(size_t n) : n_(n) {}  // NOLINT
        __host__ void operator()(T* ptr) {
            if constexpr (!std::is_trivially_destructible_v<T>) {
                if (ptr)
                    for (size_t i = 0; i < n_; ++i) ptr[i].~T();
            }
            release_device_memory(ptr);
        }

      private:
        size_t n_;
    };

  private:
    T* raw_ptr_{};
#ifndef __CUDA_ARCH__
    std::shared_ptr<T> ptr_{};
#else
    dummy_shared_ptr_t<T> ptr_{};
#endif
};

// One zero-filled allocation of `size` elements on the CURRENT device, visible to all
// devices and (through the mirror) to the host.  Same name and role as the reference's
// allocate_cuda_usm (scalix/array.cuh:241-248); not Unified Memory any more.
template<class T>
__host__ unified_ptr<T> allocate_cuda_usm(size_t size) {
    using raw_t = std::remove_const_t<T>;
    cuda::detail::ensure_runtime();
    const int device = cuda::traits::current_device();
    const size_t zero = 0;
    void* base        = nullptr;
    cuda::detail::check_rt(
        sclx_b200_array_create(size * sizeof(raw_t), 1, &device, &zero, &base)
    );
    return unified_ptr<T>{
        static_cast<T*>(base), typename unified_ptr<T>::safe_delete_array(size)};
}

// A single object reachable from host and device: pinned, mapped host memory.
template<class T>
__host__ unified_ptr<std::decay_t<T>> make_unified_ptr(T&& value = T{}) {
    using value_type = std::decay_t<T>;
    void* raw        = nullptr;
    cuda::cuda_exception::raise_if_not_success(
        cudaHostAlloc(&raw, sizeof(value_type), cudaHostAllocPortable | cudaHostAllocMapped),
        std::experimental::source_location::current(),
        "sclx::detail::"
    );
    auto* ptr = new (raw) value_type(std::forward<T>(value));
    return unified_ptr<value_type>{ptr};
}

template<class T, class U>
__host__ __device__ unified_ptr<T> static_pointer_cast(const unified_ptr<U>& ptr) {
    unified_ptr<T> out;
    out.raw_ptr_ = static_cast<T*>(ptr.raw_ptr_);
#ifndef __CUDA_ARCH__
    out.ptr_ = std::static_pointer_cast<T>(ptr.ptr_);
#endif
    return out;
}

template<class T, class U>
__host__ unified_ptr<T> dynamic_pointer_cast(const unified_ptr<U>& ptr) {
    unified_ptr<T> out;
    out.raw_ptr_ = dynamic_cast<T*>(ptr.raw_ptr_);
#ifndef __CUDA_ARCH__
    out.ptr_ = std::dynamic_pointer_cast<T>(ptr.ptr_);
#endif
    return out;
}

template<class T, class U>
__host__ __device__ unified_ptr<T> reinterpret_pointer_cast(const unified_ptr<U>& ptr) {
    unified_ptr<T> out;
    out.raw_ptr_ = reinterpret_cast<T*>(ptr.raw_ptr_);
#ifndef __CUDA_ARCH__
    out.ptr_ = std::reinterpret_pointer_cast<T>(ptr.ptr_);
#endif
    return out;
}

template<class T, class U>
__host__ __device__ unified_ptr<T> const_pointer_cast(const unified_ptr<U>& ptr) {
    unified_ptr<T> out;
    out.raw_ptr_ = const_cast<T*>(ptr.raw_ptr_);
#ifndef __CUDA_ARCH__
    out.ptr_ = std::const_pointer_cast<T>(ptr.ptr_);
#endif
    return out;
}

}  // namespace detail

enum class data_capture_mode { copy, capture };

enum copy_policy {
    devicedevice = cudaMemcpyDeviceToDevice,
    devicehost   = cudaMemcpyDeviceToHost,
    hostdevice   = cudaMemcpyHostToDevice,
    hosthost     = cudaMemcpyHostToHost
};

// where the partitions of an array live; host-side bookkeeping
template<class T>
struct array_memory_info_t {
    detail::unified_ptr<const T> data_start{};
    detail::unified_ptr<size_t> elements_per_device{};
    detail::unified_ptr<int> devices{};
    uint num_devices{};

    static_assert(!std::is_const_v<T>, "T must not be const");
};

enum class exec_topology { distributed, replicated };

template<class T, uint Rank>
class array;

template<class T, uint Rank>
__host__ std::vector<std::tuple<int, size_t, size_t>>
get_device_split_info(const array<T, Rank>& arr);

template<class T, uint Rank>
class array {
  public:
    using mem_info_t = array_memory_info_t<std::remove_const_t<T>>;
    using split_t    = std::vector<std::tuple<int, size_t, size_t>>;  // (device, first col, cols)

    array() = default;

    __host__ array(std::initializer_list<size_t> shape, bool call_prefetch = true)
        : shape_(shape) {
        allocate_partitioned();
        (void)call_prefetch;  // pages are placed at allocation; nothing to prefetch
    }

    __host__ explicit array(const shape_t<Rank>& shape, bool call_prefetch = true)
        : shape_(shape) {
        allocate_partitioned();
        (void)call_prefetch;
    }

    // non-owning / shared view (also the device-side constructor)
    __host__ __device__ array(const shape_t<Rank>& shape, detail::unified_ptr<T> data)
        : data_(std::move(data)),
          shape_(shape) {}

    // copy `data` into a new partitioned allocation, or adopt a device-accessible pointer
    __host__ array(
        const shape_t<Rank>& shape,
        T* data,
        bool call_prefetch     = true,
        data_capture_mode mode = data_capture_mode::copy,
        copy_policy policy     = copy_policy::hostdevice
    )
        : shape_(shape) {
        (void)call_prefetch;
        if (mode == data_capture_mode::copy) {
            allocate_partitioned();
            cuda::detail::check_rt(sclx_b200_rt_before_device_access());
            cuda::detail::check_rt(sclx_b200_rt_quiesce());
            cuda::cuda_exception::raise_if_not_success(
                cudaMemcpy(
                    const_cast<std::remove_const_t<T>*>(data_.get()),
                    data,
                    elements() * sizeof(T),
                    static_cast<cudaMemcpyKind>(policy)
                ),
                std::experimental::source_location::current(),
                "sclx::array::"
            );
        } else {
            data_       = detail::unified_ptr<T>{data};
            alloc_base_ = data;
            set_primary_devices(false);
        }
    }

    template<class U = T, class = std::enable_if_t<!std::is_const_v<U>>>
    __host__ __device__ operator array<const U, Rank>() const {  // NOLINT
        array<const U, Rank> out;
        out.shape_       = shape_;
        out.data_        = data_;
        out.memory_info_ = memory_info_;
        out.alloc_base_  = alloc_base_;
        return out;
    }

    __host__ void copy_range_from(
        const md_index_t<Rank>& start,
        const md_index_t<Rank>& end,
        const T* data,
        copy_policy policy = copy_policy::hostdevice
    ) {
        const auto flat_start = start.as_linear(shape_);
        const auto flat_end   = end.as_linear(shape_);
        cuda::detail::check_rt(sclx_b200_rt_before_device_access());
        cuda::detail::check_rt(sclx_b200_rt_quiesce());
        cuda::cuda_exception::raise_if_not_success(
            cudaMemcpy(
                data_.get() + flat_start,
                data,
                (flat_end - flat_start) * sizeof(T),
                static_cast<cudaMemcpyKind>(policy)
            ),
            std::experimental::source_location::current(),
            "sclx::array::"
        );
    }

    // element access; on the host this goes through the software mirror
    __host__ __device__ T& operator[](const md_index_t<Rank>& index) const {
        return data_.get()[index.as_linear(shape_)];
    }

    __host__ __device__ T& operator[](index_t index) const { return data_.get()[index]; }

    template<class... Args>
    __host__ __device__ T& operator()(const Args&... args) const {
        return this->operator[](md_index_t<Rank>{static_cast<size_t>(args)...});
    }

    __host__ __device__ const shape_t<Rank>& shape() const { return shape_; }

    __host__ __device__ const size_t& trailing_dimension() const { return shape_[Rank - 1]; }

    __host__ __device__ detail::unified_ptr<T> data() const { return data_; }

    __host__ __device__ T* begin() const { return data_.get(); }

    __host__ __device__ T* end() const { return data_.get() + elements(); }

    __host__ __device__ size_t elements() const { return shape_.elements(); }

    [[nodiscard]] __host__ __device__ constexpr uint rank() const { return Rank; }

    // ---- partitioning ---------------------------------------------------------------
    // all logical devices, in order
    __host__ array& set_primary_devices(bool call_prefetch = true) {
        const int device_count = cuda::traits::device_count();
        std::vector<int> devices(device_count);
        for (int i = 0; i < device_count; ++i) devices[i] = i;
        return set_primary_devices(default_split(devices, /*warn=*/false), call_prefetch);
    }

    __host__ array&
    set_primary_devices(const std::vector<int>& devices, bool call_prefetch = true) {
        if (devices.empty())
            throw_exception<std::invalid_argument>("Invalid device list; it is empty.", "sclx::array::");
        const int max_dev_id   = *std::max_element(devices.begin(), devices.end());
        const int min_dev_id   = *std::min_element(devices.begin(), devices.end());
        const int device_count = cuda::traits::device_count();
        if (max_dev_id >= device_count || min_dev_id < 0) {
            throw_exception<std::invalid_argument>(
                "Invalid device list; device ids must be less than the number "
                "of devices, which is "
                    + std::to_string(device_count) + ".",
                "sclx::array::"
            );
        }
        return set_primary_devices(default_split(devices, /*warn=*/true), call_prefetch);
    }

    // split info = (device_id, first column, columns) along the last dimension.  Moves the
    // physical pages of the allocation to the new owners; pointer and contents are kept.
    __host__ array&
    set_primary_devices(const split_t& device_split_info, bool call_prefetch = true) {
        (void)call_prefetch;
        if (is_slice()) {
            throw_exception<std::runtime_error>(
                "Array is a get_slice and cannot have its primary devices changed.",
                "sclx::array::"
            );
        }
        for (const auto& [device_id, slice_idx, slice_range] : device_split_info) {
            (void)device_id;
            (void)slice_idx;
            if (slice_range == 0 && elements() != 0) {
                throw_exception<std::invalid_argument>(
                    "Invalid device split info; slice length must be greater than 0.",
                    "sclx::array::"
                );
            }
        }
        std::vector<int> devices;
        std::vector<size_t> byte_starts;
        partition_bytes(device_split_info, devices, byte_starts);
        void* base = const_cast<std::remove_const_t<T>*>(data_.get());
        if (base != nullptr && sclx_b200_array_owner(base) >= 0) {
            cuda::detail::check_rt(sclx_b200_array_repartition(
                base, int(devices.size()), devices.data(), byte_starts.data()
            ));
        }
        record_split(device_split_info);
        return *this;
    }

    // read-mostly hints have no meaning without Unified Memory
    __host__ array& set_read_mostly() { return *this; }
    __host__ array& unset_read_mostly() { return *this; }

    // "prefetch to the CPU" makes the range HOST-RESIDENT in one bulk copy (per owner, in
    // parallel): readable and writable without further faults — syscalls into it work, as on
    // the reference's managed memory — and written back as a whole before the next device
    // access.  A prefetch to devices only makes pending host stores visible (pages never move
    // on access).
    __host__ std::future<void> prefetch_async(
        const std::vector<int>& device_ids,
        exec_topology topology = exec_topology::replicated
    ) {
        if (topology != exec_topology::distributed && topology != exec_topology::replicated) {
            throw_exception<std::invalid_argument>("Invalid exec_topology specified.", "sclx::array::");
        }
        bool to_host = false;
        for (int id : device_ids) to_host |= (id == cuda::traits::cpu_device_id);
        return make_coherent(to_host);
    }

    __host__ std::future<void> prefetch_async(exec_topology topology = exec_topology::replicated) {
        if (topology != exec_topology::distributed && topology != exec_topology::replicated) {
            throw_exception<std::invalid_argument>("Invalid execution topology.", "sclx::array::");
        }
        return make_coherent(false);
    }

    __host__ const mem_info_t& memory_info() const { return *memory_info_.get(); }

    __host__ bool has_memory_info() const { return memory_info_.get() != nullptr; }

    __host__ __device__ bool is_slice() const {
        return alloc_base_ != nullptr && alloc_base_ != static_cast<const void*>(data_.get());
    }

    // ---- views ------------------------------------------------------------------------
    // drop trailing dimensions: slice_idx[0] indexes the LAST dimension, slice_idx[1] the one
    // before it, ... (same convention as the reference, scalix/array.cuh:720-778)
    template<uint IndexRank>
    __host__ __device__ array<T, Rank - IndexRank> get_slice(md_index_t<IndexRank> slice_idx) const {
        static_assert(IndexRank < Rank, "slice index must leave at least one dimension");
        size_t stride = elements();
        size_t offset = 0;
        for (uint i = 0; i < IndexRank; ++i) {
            stride /= shape_[Rank - 1 - i];
            offset += slice_idx[i] * stride;
        }
        shape_t<Rank - IndexRank> shape;
        for (uint i = 0; i < Rank - IndexRank; ++i) shape[i] = shape_[i];
        return make_view<Rank - IndexRank>(shape, offset);
    }

    __host__ __device__ auto get_slice(md_index_t<1> slice_idx) const {
        if constexpr (Rank == 1) {
            return this->operator[](slice_idx);
        } else {
            shape_t<Rank - 1> shape;
            for (uint i = 0; i < Rank - 1; ++i) shape[i] = shape_[i];
            return make_view<Rank - 1>(shape, slice_idx[0] * (elements() / shape_[Rank - 1]));
        }
    }

    // columns [start, end) of the last dimension
    __host__ __device__ array get_range(md_index_t<1> start, md_index_t<1> end) const {
        shape_t<Rank> shape(shape_);
        shape[Rank - 1]   = end[0] - start[0];
        const size_t lead = shape_[Rank - 1] ? elements() / shape_[Rank - 1] : 0;
        return make_view<Rank>(shape, start[0] * lead);
    }

    template<class T_, uint Rank_>
    friend class array;

    friend void swap(array& lhs, array& rhs) noexcept {
        using std::swap;
        swap(lhs.data_, rhs.data_);
        swap(lhs.memory_info_, rhs.memory_info_);
        swap(lhs.shape_, rhs.shape_);
        swap(lhs.alloc_base_, rhs.alloc_base_);
    }

  private:
    detail::unified_ptr<T> data_{};
    shape_t<Rank> shape_{};
    detail::unified_ptr<mem_info_t> memory_info_{};  // host heap, shared with all views
    const void* alloc_base_{};                       // first byte of the owning allocation

    template<uint NewRank>
    __host__ __device__ array<T, NewRank>
    make_view(const shape_t<NewRank>& shape, size_t element_offset) const {
        array<T, NewRank> view{
            shape, detail::unified_ptr<T>::aliasing(data_, data_.get() + element_offset)};
#ifndef __CUDA_ARCH__
        view.memory_info_ = memory_info_;
#endif
        view.alloc_base_ = alloc_base_ ? alloc_base_ : static_cast<const void*>(data_.get());
        return view;
    }

    __host__ size_t leading_elements() const {
        return shape_[Rank - 1] ? elements() / shape_[Rank - 1] : 0;
    }

    // ceil(last/ndev) columns per device, surplus devices dropped
    __host__ split_t default_split(const std::vector<int>& devices, bool warn) const {
        const size_t last  = shape_[Rank - 1];
        const size_t ndev  = devices.size();
        const size_t split = (last + ndev - 1) / ndev;
        split_t info;
        size_t used = split ? (last + split - 1) / split : 0;
        if (used == 0) used = 1;  // empty array: one empty partition on the first device
        if (warn && ndev > used) {
            std::cerr << "Warning: too many devices specified for array. "
                      << "Using " << used << " devices instead." << std::endl;
        }
        for (size_t d = 0; d < used && d < ndev; ++d) {
            const size_t first = d * split;
            info.emplace_back(devices[d], first, std::min(last - first, split));
        }
        return info;
    }

    __host__ void partition_bytes(
        const split_t& info,
        std::vector<int>& devices,
        std::vector<size_t>& byte_starts
    ) const {
        const size_t lead = leading_elements();
        for (const auto& [device_id, slice_idx, slice_range] : info) {
            (void)slice_range;
            devices.push_back(device_id);
            byte_starts.push_back(devices.size() == 1 ? 0 : slice_idx * lead * sizeof(T));
        }
    }

    __host__ void record_split(const split_t& info) {
        using raw_t = std::remove_const_t<T>;
        auto* mi    = new mem_info_t;
        const size_t n = info.size();
        auto* per_dev  = new size_t[n];
        auto* devs     = new int[n];
        const size_t lead = leading_elements();
        for (size_t d = 0; d < n; ++d) {
            per_dev[d] = std::get<2>(info[d]) * lead;
            devs[d]    = std::get<0>(info[d]);
        }
        mi->data_start = detail::unified_ptr<const raw_t>(
            const_cast<const raw_t*>(data_.get()), detail::unified_ptr<const raw_t>::no_delete
        );
        mi->elements_per_device = detail::unified_ptr<size_t>(per_dev, [](size_t* p) { delete[] p; });
        mi->devices             = detail::unified_ptr<int>(devs, [](int* p) { delete[] p; });
        mi->num_devices         = uint(n);
        if (memory_info_.get() != nullptr) {
            // views share the record: update it in place
            *memory_info_.get() = *mi;
            delete mi;
        } else {
            memory_info_ = detail::unified_ptr<mem_info_t>(mi, [](mem_info_t* p) { delete p; });
        }
    }

    __host__ void allocate_partitioned() {
        using raw_t = std::remove_const_t<T>;
        static_assert(!std::is_const_v<T>, "cannot allocate an array of const elements");
        cuda::detail::ensure_runtime();
        const int device_count = cuda::traits::device_count();
        std::vector<int> all(device_count);
        for (int i = 0; i < device_count; ++i) all[i] = i;
        const split_t info = default_split(all, /*warn=*/false);
        std::vector<int> devices;
        std::vector<size_t> byte_starts;
        partition_bytes(info, devices, byte_starts);
        void* base = nullptr;
        cuda::detail::check_rt(sclx_b200_array_create(
            elements() * sizeof(raw_t), int(devices.size()), devices.data(), byte_starts.data(), &base
        ));
        data_ = detail::unified_ptr<T>(
            static_cast<T*>(base), typename detail::unified_ptr<T>::safe_delete_array(elements())
        );
        alloc_base_ = base;
        record_split(info);
    }

    __host__ std::future<void> make_coherent(bool to_host) {
        if (to_host) {
            cuda::detail::check_rt(
                sclx_b200_rt_prefetch_to_host(data_.get(), elements() * sizeof(T))
            );
        } else {
            cuda::detail::check_rt(sclx_b200_rt_before_device_access());
        }
        std::promise<void> done;
        done.set_value();
        return done.get_future();
    }
};

// Host: a COPY with the new shape (device-to-device, freshly partitioned); device: an
// aliasing view (same semantics as /root/reference scalix/array.cuh:882-900).
template<class T, uint RankOld, uint RankNew>
__host__ __device__ array<T, RankNew>
reshape(const array<T, RankOld>& arr, const shape_t<RankNew>& shape) {
#ifdef __CUDA_ARCH__
    return array<T, RankNew>{shape, arr.data()};
#else
    if (arr.shape().elements() != shape.elements()) {
        throw_exception<std::invalid_argument>(
            "Cannot reshape array with different number of elements"
        );
    }
    return array<T, RankNew>{
        shape, arr.data().get(), true, data_capture_mode::copy, copy_policy::devicedevice};
#endif
}

template<class T, uint Rank>
__host__ __device__ array<T, 1> flatten(const array<T, Rank>& arr) {
    return reshape(arr, shape_t<1>{arr.elements()});
}

// (device id, first column, columns) for every partition that holds part of `arr`, along
// arr's LAST dimension and relative to arr (slice-aware).  A column that straddles a
// partition boundary belongs to the partition holding its first element.  Contract of
// /root/reference scalix/array.cuh:973-1028, exact for any slice (the reference is exact
// only for device-aligned ones).
template<class T, uint Rank>
__host__ std::vector<std::tuple<int, size_t, size_t>>
get_device_split_info(const array<T, Rank>& arr) {
    std::vector<std::tuple<int, size_t, size_t>> splits;
    const size_t last = arr.shape()[Rank - 1];
    if (!arr.has_memory_info()) {
        splits.emplace_back(cuda::traits::current_device(), size_t(0), last);
        return splits;
    }
    const auto& info    = arr.memory_info();
    const size_t total  = arr.elements();
    const size_t lead   = last ? total / last : 0;
    const size_t offset = size_t(arr.data().get() - info.data_start.get());
    size_t part_begin   = 0;
    for (uint d = 0; d < info.num_devices; ++d) {
        const size_t part_end = part_begin + info.elements_per_device.get()[d];
        const size_t lo       = std::max(part_begin, offset);
        const size_t hi       = std::min(part_end, offset + total);
        if (hi > lo && lead > 0) {
            const size_t first_col = (lo - offset + lead - 1) / lead;
            const size_t end_col   = std::min(last, (hi - offset + lead - 1) / lead);
            if (end_col > first_col)
                splits.emplace_back(info.devices.get()[d], first_col, end_col - first_col);
        }
        part_begin = part_end;
    }
    if (splits.empty()) {
        const int dev = info.num_devices ? info.devices.get()[0] : cuda::traits::current_device();
        splits.emplace_back(dev, size_t(0), last);
    }
    return splits;
}

#ifdef SCLX_B200_HAS_CEREAL
// On-disk format of /root/reference scalix/array.cuh:907-950: Rank extents, then the payload
// (one raw block for cereal::BinaryInputArchive, element by element for every other archive
// — including BinaryOutputArchive, so a binary round trip is element-wise out, raw in; both
// are the same bytes).  Host element access goes through the array's host mirror.
template<class Archive, class T, uint Rank>
__host__ void serialize_array(Archive& archive, const sclx::array<T, Rank>& array) {
#ifndef __CUDA_ARCH__
    for (uint d = 0; d < Rank; ++d) archive(array.shape()[d]);
    if constexpr (std::is_same_v<Archive, cereal::BinaryInputArchive>) {
        // loading into an existing array: staged, because a stream may hand the destination
        // pointer straight to read(2), and a syscall cannot fault the host mirror in
        using raw_t = std::remove_const_t<T>;
        std::vector<raw_t> staged(array.elements());
        archive(cereal::binary_data(staged.data(), staged.size() * sizeof(T)));
        std::copy(staged.begin(), staged.end(), const_cast<raw_t*>(array.begin()));
    } else {
        std::for_each(array.begin(), array.end(), [&](auto& element) { archive(element); });
    }
#endif
}

template<class Archive, class T, uint Rank>
__host__ void deserialize_array(Archive& archive, sclx::array<T, Rank>& array) {
#ifndef __CUDA_ARCH__
    sclx::shape_t<Rank> shape;
    for (uint d = 0; d < Rank; ++d) archive(shape[d]);
    array = sclx::array<T, Rank>{shape};
    if constexpr (std::is_same_v<Archive, cereal::BinaryInputArchive>) {
        // staged: a stream may read() straight into the destination, and a syscall cannot
        // fault the host mirror in the way a user-space store does
        std::vector<std::remove_const_t<T>> staged(array.elements());
        archive(cereal::binary_data(staged.data(), staged.size() * sizeof(T)));
        std::copy(staged.begin(), staged.end(), array.begin());
    } else {
        std::for_each(array.begin(), array.end(), [&](auto& element) { archive(element); });
    }
#endif
}
#endif  // SCLX_B200_HAS_CEREAL

__host__ inline bool is_same_device_split(
    const std::vector<std::tuple<int, size_t, size_t>>& lhs,
    const std::vector<std::tuple<int, size_t, size_t>>& rhs
) {
    return lhs == rhs;
}

}  // namespace sclx
