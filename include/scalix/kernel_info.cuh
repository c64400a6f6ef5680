// scalix-b200: per-launch metadata handed (by value) to every device lambda as its second
// argument.  Public surface of /root/reference scalix/kernel_info.cuh:39-114 with the
// thread-id arithmetic done in 64 bits (the reference multiplies blockIdx.x*blockDim.x in
// 32 bits, ibid. :86-101, and cannot address >= 2^32 elements per device).
#pragma once

#include "index.cuh"
#include "range.cuh"
#include "shape.cuh"

namespace sclx {

template<uint ProblemRank = 1, uint ThreadBlockRank = 1>
class kernel_info {
  public:
    kernel_info(const kernel_info&)                = default;
    kernel_info(kernel_info&&) noexcept            = default;
    kernel_info& operator=(const kernel_info&)     = default;
    kernel_info& operator=(kernel_info&&) noexcept = default;

    // built by the launcher once per partition: `device_range` is `global_range` with the
    // last extent cut to the partition's slice, `start_index` is zero except for the slice's
    // first column, `device_id` is the partition's ordinal
    __host__ kernel_info(
        const md_range_t<ProblemRank>& global_range,
        const md_range_t<ProblemRank>& device_range,
        const shape_t<ThreadBlockRank>& thread_block_shape,
        const md_index_t<ProblemRank>& start_index,
        const int& device_id
    )
        : thread_block_shape_(thread_block_shape),
          global_range_(global_range),
          device_range_(device_range),
          start_index_(start_index),
          start_linear_(start_index.as_linear(global_range)),
          device_id_(device_id) {}

    // ---- what was launched --------------------------------------------------------------
    __host__ __device__ const shape_t<ThreadBlockRank>& thread_block_shape() const {
        return thread_block_shape_;
    }
    __host__ __device__ const md_range_t<ProblemRank>& global_range() const {
        return global_range_;
    }
    __host__ __device__ const md_range_t<ProblemRank>& device_range() const {
        return device_range_;
    }
    __host__ __device__ const md_index_t<ProblemRank>& start_index() const {
        return start_index_;
    }

    // ---- who am I -----------------------------------------------------------------------
    // position inside the thread block, decomposed against thread_block_shape(); the usual
    // index into a local_array (examples/scaling_performance.cu:66-72)
    __device__ md_index_t<ThreadBlockRank> local_thread_id() const {
        return md_index_t<>::create_from_linear(threadIdx.x, thread_block_shape_);
    }
    __device__ index_t local_thread_linear_id() const { return threadIdx.x; }

    // Linear index (in global_range) of the element this thread is working on.  The
    // reference launches one thread per element, so its blockIdx*blockDim+threadIdx+start
    // IS the element index and its own algorithms use it as such
    // (/root/reference scalix/algorithm/transform.cuh:58-59); with the persistent
    // grid-stride launch used here the completed trips are added to keep that meaning.
    __device__ index_t global_thread_linear_id() const {
        return index_t(blockIdx.x) * blockDim.x + threadIdx.x + start_linear_
             + index_t(stride_count_) * (index_t(gridDim.x) * blockDim.x);
    }
    __device__ md_index_t<ProblemRank> global_thread_id() const {
        return md_index_t<>::create_from_linear(global_thread_linear_id(), global_range_);
    }

    // ordinal of this partition in the array's split (NOT the CUDA device id), as in
    // /root/reference scalix/detail/execute_kernel.cuh:147-153
    __host__ __device__ const int& device_id() const { return device_id_; }

    // grid-stride bookkeeping: the dispatcher bumps stride_count() after every trip, so
    // global_thread_linear_id() keeps naming the element in flight
    __device__ uint grid_stride() const { return gridDim.x * blockDim.x; }
    __device__ const uint& stride_count() const { return stride_count_; }
    __device__ void increment_stride_count() { ++stride_count_; }

    // scalix-b200 extension: linear offset of start_index() in global_range()
    __host__ __device__ index_t start_linear() const { return start_linear_; }

  private:
    shape_t<ThreadBlockRank> thread_block_shape_;
    md_range_t<ProblemRank> global_range_;
    md_range_t<ProblemRank> device_range_;
    md_index_t<ProblemRank> start_index_;
    index_t start_linear_{};
    uint stride_count_{};
    int device_id_;
};

}  // namespace sclx
