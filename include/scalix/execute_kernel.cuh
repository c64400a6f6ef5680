// scalix-b200: sclx::execute_kernel, kernel_handler and local_array.
//
// Public surface of /root/reference scalix/execute_kernel.cuh:45-515.  What changes underneath
// (BASELINE.json north_star, subsystem 2):
//
//   reference : execute_kernel spawns a std::async thread per call; launch creates one stream
//               per device, toggles cudaMemAdvise(ReadMostly) on every result array twice
//               and blocks until all streams are idle (ibid. :286,300-302,506-513).
//   here      : the host lambda runs at once on the calling thread and only QUEUES kernels on
//               the persistent per-device streams; ordering between consecutive launches is
//               kept on the device with events.  The returned std::future<void> behaves like
//               the reference's (never deferred, blocks when dropped while work is running,
//               delivers errors through get()); see execute_kernel below.
#pragma once

#include "array.cuh"
#include "array_list.cuh"
#include "array_tuple.cuh"
#include "cuda.hpp"
#include "detail/execute_kernel.cuh"
#include "kernel_info.cuh"
#include "range.cuh"
#include "shape.cuh"

#include <chrono>
#include <future>

namespace sclx {

class kernel_handler {
  public:
    // defaults of every launch overload below (values of sclx::cuda::traits::kernel): 1-D blocks
    // of 256 threads; "no cap" for the grid, which the dispatcher then sizes to the machine
    static constexpr uint tb_rank_default = cuda::traits::kernel::default_block_shape.rank();
    static constexpr shape_t<1> block_default() { return cuda::traits::kernel::default_block_shape; }
    static constexpr size_t grid_default() { return cuda::traits::kernel::default_grid_size; }

    // ---- range launches ---------------------------------------------------------------
    // launch(range, array_list, f): the general range form (reference :49-72).  All result
    // arrays must share their last extent (std::invalid_argument otherwise) and are brought to
    // the first one's partitioning (with the reference's warning) if they differ; partition p
    // runs the range indices whose last coordinate falls in its slice, on p's device, on that
    // device's persistent stream.  `block_shape` only shapes local_thread_id();
    // `grid_size` caps the grid (default: sized to the machine by launch_config.cuh).
    template<
        class KernelTag = detail::default_kernel_tag,
        uint RangeRank,
        class T,
        uint ResultRank,
        uint NResult,
        class F,
        uint ThreadBlockRank = tb_rank_default>
    __host__ void launch(
        md_range_t<RangeRank> range,
        array_list<T, ResultRank, NResult> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch_array_list_impl<KernelTag>(range, result, block_shape, grid_size, std::forward<F>(f));
    }

    // same for a run-time sized list (reference :74-96)
    template<
        class KernelTag = detail::default_kernel_tag,
        uint RangeRank,
        class T,
        uint ResultRank,
        class F,
        uint ThreadBlockRank = tb_rank_default>
    __host__ void launch(
        md_range_t<RangeRank> range,
        dynamic_array_list<T, ResultRank> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch_array_list_impl<KernelTag>(range, result, block_shape, grid_size, std::forward<F>(f));
    }

    // single result array (reference :98-120): wrapped in a one-element array_list
    template<
        class KernelTag = detail::default_kernel_tag,
        uint RangeRank,
        class T,
        uint ResultRank,
        class F,
        uint ThreadBlockRank = tb_rank_default>
    __host__ void launch(
        md_range_t<RangeRank> range,
        array<T, ResultRank> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch<KernelTag>(
            range, array_list<T, ResultRank, 1>({result}), std::forward<F>(f), block_shape, grid_size
        );
    }

    // heterogeneous results (reference :195-216): array_tuple of arrays with equal last extents
    template<
        class KernelTag = detail::default_kernel_tag,
        uint RangeRank,
        class F,
        uint ThreadBlockRank = tb_rank_default,
        class... ArrayTypes>
    __host__ void launch(
        md_range_t<RangeRank> range,
        array_tuple<ArrayTypes...> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch_tuple_impl<KernelTag>(range, result, block_shape, grid_size, std::forward<F>(f));
    }

    // ---- index-generator launches -------------------------------------------------------
    // launch(index_generator, results, f) (reference :122-146): EVERY partition walks the whole
    // generator range, maps each range index through the generator and calls f only for the
    // indices whose last coordinate it owns — writes stay local, reads may be anywhere.  The
    // generator is copied to the device by value; its index_range()'s last extent must equal
    // the results' last extent.
    template<
        class KernelTag = detail::default_kernel_tag,
        class IndexGenerator,
        class T,
        uint ResultRank,
        uint NResult,
        class F,
        uint ThreadBlockRank = tb_rank_default>
    __host__ void launch(
        IndexGenerator&& index_generator,
        array_list<T, ResultRank, NResult> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch_array_list_impl<KernelTag>(
            std::forward<IndexGenerator>(index_generator), result, block_shape, grid_size,
            std::forward<F>(f)
        );
    }

    // index generator, run-time sized list (reference :148-170)
    template<
        class KernelTag = detail::default_kernel_tag,
        class IndexGenerator,
        class T,
        uint ResultRank,
        class F,
        uint ThreadBlockRank = tb_rank_default>
    __host__ void launch(
        IndexGenerator&& index_generator,
        dynamic_array_list<T, ResultRank> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch_array_list_impl<KernelTag>(
            std::forward<IndexGenerator>(index_generator), result, block_shape, grid_size,
            std::forward<F>(f)
        );
    }

    // index generator, single result array (reference :172-193)
    template<
        class KernelTag = detail::default_kernel_tag,
        class IndexGenerator,
        class T,
        uint ResultRank,
        class F,
        uint ThreadBlockRank = tb_rank_default>
    __host__ void launch(
        IndexGenerator&& index_generator,
        array<T, ResultRank> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch<KernelTag>(
            std::forward<IndexGenerator>(index_generator),
            array_list<T, ResultRank, 1>({result}),
            std::forward<F>(f),
            block_shape,
            grid_size
        );
    }

    // index generator, heterogeneous results (reference :218-239)
    template<
        class KernelTag = detail::default_kernel_tag,
        class IndexGenerator,
        class F,
        uint ThreadBlockRank = tb_rank_default,
        class... ArrayTypes>
    __host__ void launch(
        IndexGenerator&& index_generator,
        array_tuple<ArrayTypes...> result,
        F&& f,
        shape_t<ThreadBlockRank> block_shape = block_default(),
        size_t grid_size                     = grid_default()
    ) const {
        launch_tuple_impl<KernelTag>(
            std::forward<IndexGenerator>(index_generator), result, block_shape, grid_size,
            std::forward<F>(f)
        );
    }

    // block-level barrier and device-wide memory fence for device lambdas that captured the
    // handler by value (a lambda filling a local_array cooperatively needs the first one
    // before any thread reads what its neighbours wrote)
    __device__ void syncthreads() const { __syncthreads(); }
    __device__ void threadfence() const { __threadfence(); }

    template<class T, uint Rank>
    friend class local_array;

  private:
    __device__ static char* get_local_mem(const size_t& offset) {
        extern __shared__ char local_mem[];
        return local_mem + offset;
    }

    // Every result array must be partitioned like the first one: the launcher issues one
    // kernel per partition of result[0] and a thread may only write what its device owns.
    // A differing array is re-partitioned on the spot (its pages move, pointer and contents
    // stay) and the reference's warning is printed.
    template<class ArrayListType>
    __host__ static auto unify_split(ArrayListType& result) {
        auto device_split_first = get_device_split_info(result[0]);
        for (auto& array : result) {
            if (!is_same_device_split(device_split_first, get_device_split_info(array))) {
                array.set_primary_devices(device_split_first);
                std::cerr << "Warning: Not every array in the list has the same "
                             "memory split, setting primary devices for all "
                             "arrays to the same as the first array in the list"
                          << std::endl;
            }
        }
        return get_device_split_info(result[0]);
    }

    template<class KernelTag, uint RangeRank, uint ThreadBlockRank, class ArrayListType, class F>
    __host__ void launch_array_list_impl(
        md_range_t<RangeRank> range,
        ArrayListType&& result,
        shape_t<ThreadBlockRank> block_shape,
        size_t grid_size,
        F&& f
    ) const {
        const size_t last_dim = result[0].shape()[result[0].rank() - 1];
        for (auto& array : result) {
            if (array.shape()[array.rank() - 1] != last_dim) {
                throw_exception<std::invalid_argument>(
                    "All arrays in the result must have the same last dimension"
                );
            }
        }
        auto device_info = unify_split(result);
        KernelTag::template execute<RangeRank, ThreadBlockRank>(
            device_info, range, block_shape, grid_size, local_mem_size_, std::forward<F>(f)
        );
    }

    template<
        class KernelTag,
        class IndexGenerator,
        uint ThreadBlockRank,
        class ArrayListType,
        class F,
        class = std::void_t<decltype(std::decay_t<IndexGenerator>::index_rank)>>
    __host__ void launch_array_list_impl(
        IndexGenerator&& index_generator,
        ArrayListType&& result,
        shape_t<ThreadBlockRank> block_shape,
        size_t grid_size,
        F&& f
    ) const {
        using generator_t = std::remove_reference_t<IndexGenerator>;
        for (auto& array : result) {
            if (index_generator.index_range()[generator_t::index_rank - 1]
                != array.shape()[array.rank() - 1]) {
                throw std::invalid_argument("Index generator indices and "
                                            "result array must have the same "
                                            "last dimension");
            }
        }
        auto device_info = unify_split(result);
        KernelTag::template execute<IndexGenerator, ThreadBlockRank>(
            device_info, index_generator, block_shape, grid_size, local_mem_size_,
            std::forward<F>(f)
        );
    }

    template<class KernelTag, uint RangeRank, uint ThreadBlockRank, class F, class... ArrayTypes>
    __host__ void launch_tuple_impl(
        md_range_t<RangeRank> range,
        array_tuple<ArrayTypes...> result,
        shape_t<ThreadBlockRank> block_shape,
        size_t grid_size,
        F&& f
    ) const {
        if (!result.last_dims_all_equal()) {
            throw_exception<std::invalid_argument>(
                "All arrays in the result must have the same last dimension"
            );
        }
        auto device_info = get_device_split_info(
            thrust::get<0>(static_cast<thrust::tuple<ArrayTypes...>&>(result))
        );
        KernelTag::template execute<RangeRank, ThreadBlockRank>(
            device_info, range, block_shape, grid_size, local_mem_size_, std::forward<F>(f)
        );
    }

    template<
        class KernelTag,
        class IndexGenerator,
        uint ThreadBlockRank,
        class F,
        class... ArrayTypes,
        class = std::void_t<decltype(std::decay_t<IndexGenerator>::index_rank)>>
    __host__ void launch_tuple_impl(
        IndexGenerator&& index_generator,
        array_tuple<ArrayTypes...> result,
        shape_t<ThreadBlockRank> block_shape,
        size_t grid_size,
        F&& f
    ) const {
        using generator_t = std::remove_reference_t<IndexGenerator>;
        auto& first = thrust::get<0>(static_cast<thrust::tuple<ArrayTypes...>&>(result));
        if (index_generator.index_range()[generator_t::index_rank - 1]
            != first.trailing_dimension()) {
            throw_exception<std::invalid_argument>(
                "Index generator indices and result array must have the same last dimension"
            );
        }
        if (!result.last_dims_all_equal()) {
            throw_exception<std::invalid_argument>(
                "All arrays in the result must have the same last dimension"
            );
        }
        auto device_info = get_device_split_info(first);
        KernelTag::template execute<IndexGenerator, ThreadBlockRank>(
            device_info, index_generator, block_shape, grid_size, local_mem_size_,
            std::forward<F>(f)
        );
    }

    size_t local_mem_size_ = 0;  // dynamic shared memory requested by local_arrays
};

// A view of the launch's dynamic shared memory.  Construct it on the host inside the
// execute_kernel lambda (this reserves the bytes), capture it BY VALUE into a `mutable`
// device lambda.
template<class T, uint Rank>
class local_array {
  public:
    local_array() = default;

    // host side: reserve elements()*sizeof(T) bytes of the launch's dynamic shared memory,
    // aligned for T, after whatever earlier local_arrays of the same handler took
    __host__ local_array(kernel_handler& handler, const shape_t<Rank>& shape) : shape_(shape) {
        const size_t misalign = handler.local_mem_size_ % alignof(T);
        if (misalign != 0) handler.local_mem_size_ += alignof(T) - misalign;
        offset_ = handler.local_mem_size_;
        handler.local_mem_size_ += elements() * sizeof(T);
    }

    // device side: element access, column-major like sclx::array (index 0 fastest)
    __device__ T& operator[](const md_index_t<Rank>& idx) { return data()[idx.as_linear(shape_)]; }

    __device__ T& operator[](const index_t& idx) { return data()[idx]; }

    template<class... Args>
    __device__ T& operator()(const Args&... args) {
        return operator[](md_index_t<Rank>{static_cast<const size_t&>(args)...});
    }

    __host__ __device__ size_t elements() const { return shape_.elements(); }

    __host__ __device__ const shape_t<Rank>& shape() const { return shape_; }

    // first use inside a kernel turns the host-computed byte offset into a pointer into this
    // block's dynamic shared memory; the lambda must therefore be `mutable`
    __device__ T* data() {
        if (!data_) data_ = reinterpret_cast<T*>(kernel_handler::get_local_mem(offset_));
        return data_;
    }

    __device__ T* begin() { return data(); }

    __device__ T* end() { return begin() + elements(); }

  private:
    size_t offset_{};
    shape_t<Rank> shape_;
    T* data_{};  // resolved on first use inside the kernel
};

namespace detail {

// how long execute_kernel polls for its kernels before it hands the wait to a thread
#ifndef SCALIX_B200_FUTURE_POLL_US
#define SCALIX_B200_FUTURE_POLL_US 25
#endif

inline std::future<void> ready_future(std::exception_ptr error = nullptr) {
    std::promise<void> done;
    if (error) done.set_exception(error);
    else done.set_value();
    return done.get_future();
}

}  // namespace detail

// Runs `f(handler)` on the calling thread — it queues kernels on the persistent per-device
// streams — and returns the future of /root/reference scalix/execute_kernel.cuh:506-513 with
// the same observable behaviour: wait_for() reports ready/timeout (never deferred), get()
// delivers exceptions thrown by `f` and CUDA errors of the queued work, and a future that is
// DROPPED leaves nothing running (the reference's std::async future blocks in its destructor,
// relied on at examples/index_generator.cu:78-86).
//   * Short work (done within SCALIX_B200_FUTURE_POLL_US of polling, i.e. less than a thread
//     spawn costs): the future is returned already completed.
//   * Longer work: std::async(std::launch::async) waits for the streams, exactly the
//     reference's future type — the caller overlaps host work with the kernels, and dropping
//     the future blocks until they have finished.  The spawn overlaps the kernels themselves.
template<class F>
__host__ std::future<void> execute_kernel(F&& f) {
    cuda::detail::ensure_runtime();
    auto& used = detail::launch_record::devices();
    const std::vector<int> outer(std::move(used));  // nested execute_kernel calls
    used.clear();
    std::exception_ptr error;
    try {
        kernel_handler handler;
        f(handler);
    } catch (...) {
        error = std::current_exception();
    }
    std::vector<int> devices = used;
    used                     = outer;
    for (int d : devices) detail::launch_record::note(d);
    if (devices.empty()) return detail::ready_future(error);
    auto wait_for_devices = [devices]() {
        cuda::detail::check_rt(sclx_b200_rt_sync_devices(int(devices.size()), devices.data()));
    };
    if (error) {
        // kernels queued before `f` threw still finish before the future completes
        try {
            wait_for_devices();
        } catch (...) {
        }
        return detail::ready_future(error);
    }
    const auto t0 = std::chrono::steady_clock::now();
    do {
        if (sclx_b200_rt_devices_idle(int(devices.size()), devices.data())) {
            try {
                wait_for_devices();  // nothing left to wait for: collects errors, notes the sync
            } catch (...) {
                return detail::ready_future(std::current_exception());
            }
            return detail::ready_future();
        }
    } while (std::chrono::steady_clock::now() - t0
             < std::chrono::microseconds(SCALIX_B200_FUTURE_POLL_US));
    return std::async(std::launch::async, wait_for_devices);
}

}  // namespace sclx
