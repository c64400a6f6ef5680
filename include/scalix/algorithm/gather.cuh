// scalix-b200: algorithm::gather — out[i] = src[index[i]] over partitioned arrays.
//
// EXTENSION: the reference has no gather entry point; its users write the gather as a lambda
// over the range launcher (/root/reference examples/distributed_indexing.cu:146-154,
// examples/scaling_performance.cu:59-78) and Unified Memory faults the remote pages in.  That
// idiom keeps working here unchanged (cross-partition reads are NVLink peer loads on the
// array's peer-mapped range), but one load per index is capped by NVLink's request rate across
// GPUs and by HBM's random-access rate on one (detail/b200/gather_exchange.cuh).  This entry
// point is the framework-owned form of the same operation: `index` holds linear (storage-order)
// indices into `src`; `out` and `index` have the same length; result bits are those of the
// lambda form.
//
//   small problems, sources that fit L2, shapes the binning does not cover -> direct kernel
//   otherwise -> binned exchange (K1 partition / K2 serve from L2-resident regions / K3 place)
//
// Blocking, like the reference's algorithms.
#pragma once
#include "../detail/b200/gather_exchange.cuh"
#include "../detail/b200/host_launch.cuh"
#include "../execute_kernel.cuh"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <type_traits>
#include <vector>

namespace sclx::algorithm {

// `exchange` = the binned form (also on one device, where the "exchange" is between L2-sized
// regions of the same partition)
enum class gather_mode { automatic, direct, exchange };

namespace detail {

template<size_t Bytes>
struct gather_carrier;
template<>
struct gather_carrier<4> {
    using type = unsigned;
};
template<>
struct gather_carrier<8> {
    using type = unsigned long long;
};

inline size_t gx_align(size_t v) { return (v + 255) & ~size_t(255); }

inline unsigned gx_capacity(size_t n_requester, unsigned buckets) {
    // 1.25 x the uniform share + one tile (never more than the requester's whole partition),
    // plus the expected padding of every tile's run to a multiple of kGxRunAlign entries
    const size_t share = (n_requester + buckets - 1) / buckets;
    size_t cap         = share + share / 4 + b200::kGxTile;
    if (cap > n_requester) cap = n_requester;
    const size_t tiles = (n_requester + b200::kGxTile - 1) / b200::kGxTile;
    cap += tiles * b200::kGxRunAlign;
    cap = (cap + 31) & ~size_t(31);
    return unsigned(cap);
}

// bytes of one source region: two of them (the one being served and the one being prefetched)
// plus the streaming queues have to live in the 126 MB L2 together
inline size_t gx_region_bytes() {
    if (const char* e = std::getenv("SCLX_B200_GATHER_REGION_KB")) {
        const long kb = std::atol(e);
        if (kb > 0) return size_t(kb) << 10;
    }
    return size_t(32) << 20;
}

}  // namespace detail

template<class T, uint Rank, class I>
__host__ void gather(
    const array<const T, Rank>& src,
    const array<I, 1>& index,
    array<T, 1>& out,
    gather_mode mode = gather_mode::automatic
) {
    static_assert(std::is_trivially_copyable_v<T>, "gather: element type must be trivially copyable");
    static_assert(std::is_integral_v<std::remove_const_t<I>>, "gather: index type must be integral");
    if (index.elements() != out.elements())
        throw_exception<std::invalid_argument>(
            "index and result arrays must have the same length", "sclx::algorithm::"
        );
    const size_t n = out.elements();
    if (n == 0) return;
    const int caller_device = cuda::traits::current_device();
    const auto out_info     = get_device_split_info(out);
    const auto src_info     = get_device_split_info(src);
    const int nreq          = int(out_info.size());
    const int nown          = int(src_info.size());
    const size_t src_last   = src.shape()[Rank - 1];
    const size_t src_lead   = src_last ? src.elements() / src_last : 0;
    using index_t_          = std::remove_const_t<I>;
    const index_t_* idx_ptr = index.data().get();

    bool exchange_ok = (sizeof(T) == 4 || sizeof(T) == 8) && nown >= 1 && nown <= b200::kGxMaxParts
                       && nreq <= b200::kGxMaxParts && n < (size_t(1) << 31) * size_t(nreq);
    // owners must be uniform blocks of `per` elements starting at p*per (true for every
    // array partitioned by set_primary_devices; not for arbitrary slices), distinct devices
    size_t per = 0;
    if (exchange_ok) {
        per = std::get<2>(src_info[0]) * src_lead;
        for (int p = 0; p < nown && exchange_ok; ++p) {
            if (std::get<1>(src_info[p]) * src_lead != size_t(p) * per) exchange_ok = false;
            for (int o = 0; o < p; ++o)
                if (std::get<0>(src_info[o]) == std::get<0>(src_info[p])) exchange_ok = false;
        }
        for (int r = 0; r < nreq && exchange_ok; ++r)
            for (int o = 0; o < r; ++o)
                if (std::get<0>(out_info[o]) == std::get<0>(out_info[r])) exchange_ok = false;
        if (per == 0 || per > (size_t(1) << 32)) exchange_ok = false;
    }
    if (mode == gather_mode::exchange && !exchange_ok)
        throw_exception<std::invalid_argument>(
            "binned gather needs 4- or 8-byte elements and a source partitioned in uniform blocks over "
            "1..16 distinct devices",
            "sclx::algorithm::"
        );
    // bucket geometry: regions of 2^shift elements inside every owner's partition
    unsigned region_shift = 0, regions = 1;
    if (exchange_ok) {
        size_t region_elems = detail::gx_region_bytes() / sizeof(T);
        while ((size_t(1) << (region_shift + 1)) <= region_elems) ++region_shift;
        regions = unsigned((per + (size_t(1) << region_shift) - 1) >> region_shift);
        while (size_t(regions) * size_t(nown) > size_t(b200::kGxMaxBuckets) / 2) {  // <= 256 buckets: runs of >= 16
            ++region_shift;
            regions = unsigned((per + (size_t(1) << region_shift) - 1) >> region_shift);
        }
    }
    // worth it when requests cross devices, or when the source is too large for L2 to absorb
    // the random reads by itself
    const bool pays = nown > 1 || src.elements() * sizeof(T) > (size_t(96) << 20);
    const bool use_exchange
        = exchange_ok
       && (mode == gather_mode::exchange || (mode == gather_mode::automatic && pays && n >= (size_t(1) << 18)));

    if constexpr (sizeof(T) != 4 && sizeof(T) != 8) {
        // direct kernel on raw element type
        b200::for_each_partition(out, out_info, [&](const b200::partition_t& p) {
            if (p.elements == 0) return;
            const size_t work = (p.elements + 255) / 256;
            b200::gather_direct_kernel<T, index_t_, 256, 4>
                <<<b200::persistent_grid(work, 8, b200::sm_count_of(p.device)), 256, 0, p.stream>>>(
                    idx_ptr + p.offset, p.elements, src.data().get(), out.data().get() + p.offset
                );
        });
    } else {
        using U           = typename detail::gather_carrier<sizeof(T)>::type;
        const U* src_base = reinterpret_cast<const U*>(src.data().get());
        U* out_base       = reinterpret_cast<U*>(out.data().get());
        if (!use_exchange) {
            b200::for_each_partition(out, out_info, [&](const b200::partition_t& p) {
                if (p.elements == 0) return;
                // one 1024-index chunk per CTA: random-access kernels run faster de-synchronised
                const size_t work = std::min<size_t>((p.elements + 1023) / 1024, size_t(0x7fffffff));
                b200::gather_direct_kernel<U, index_t_, 256, 4>
                    <<<unsigned(work), 256, 0, p.stream>>>(
                        idx_ptr + p.offset, p.elements, src_base, out_base + p.offset
                    );
            });
        } else {
            // the per-device exchange workspace is shared: one binned gather at a time
            static std::mutex workspace_mutex;
            std::lock_guard<std::mutex> workspace_lock(workspace_mutex);
            const unsigned buckets = regions * unsigned(nown);
            // ---- workspace layout, per logical device ------------------------------------
            std::vector<size_t> n_r(nreq), tiles_r(nreq);
            std::vector<unsigned> cap(nreq);
            for (int r = 0; r < nreq; ++r) {
                n_r[r]     = std::get<2>(out_info[r]);
                tiles_r[r] = (n_r[r] + b200::kGxTile - 1) / b200::kGxTile;
                cap[r]     = n_r[r] ? detail::gx_capacity(n_r[r], buckets) : 0;
            }
            struct dev_layout {
                int device;
                size_t bytes = 0;
                char* base   = nullptr;
            };
            std::vector<dev_layout> devs;
            auto dev_slot = [&](int device) -> dev_layout& {
                for (auto& d : devs)
                    if (d.device == device) return d;
                devs.push_back(dev_layout{device});
                return devs.back();
            };
            auto reserve = [&](int device, size_t bytes) {
                dev_layout& d    = dev_slot(device);
                const size_t off = d.bytes;
                d.bytes += detail::gx_align(bytes);
                return off;
            };
            // offsets (filled in below once the bases are known)
            std::vector<std::vector<size_t>> queue_off(nreq, std::vector<size_t>(nown)),
                resp_off(nreq, std::vector<size_t>(nown));
            std::vector<size_t> counters_off(nreq), pos_off(nreq), tbase_off(nreq), tcnt_off(nreq), done_off(nreq);
            std::vector<std::vector<size_t>> pub_off(nreq, std::vector<size_t>(nown));
            for (int r = 0; r < nreq; ++r) {
                const int rdev  = std::get<0>(out_info[r]);
                counters_off[r] = reserve(rdev, sizeof(unsigned) * buckets);
                pos_off[r]      = reserve(rdev, tiles_r[r] * b200::kGxTile * sizeof(unsigned short));
                tbase_off[r]    = reserve(rdev, tiles_r[r] * buckets * sizeof(unsigned));
                tcnt_off[r]     = reserve(rdev, tiles_r[r] * buckets * sizeof(unsigned short));
                // Queues AND response buffers live in the requester's memory: K1 writes its
                // queues locally, the owner's K2 pulls them in 4 KiB chunks and pushes the
                // responses back in 4 KiB chunks (full NVLink lines both ways), K3 is local.
                // Measured on 2 B200 (profiles/r02_gather_n2_push_vs_pull.txt): pushing the
                // queue RUNS (~128 B each) into the owner's memory made K1 2.5x slower, pulling
                // the response runs made K3 1.8x slower.
                for (int q = 0; q < nown; ++q) {
                    resp_off[r][q]  = reserve(rdev, size_t(regions) * cap[r] * sizeof(U));
                    queue_off[r][q] = reserve(rdev, size_t(regions) * cap[r] * sizeof(unsigned));
                    pub_off[r][q]   = reserve(std::get<0>(src_info[q]), size_t(regions) * sizeof(unsigned));
                }
                done_off[r] = reserve(rdev, 256);
            }
            for (auto& d : devs) {
                void* ptr = nullptr;
                cuda::detail::check_rt(sclx_b200_rt_exchange_workspace(d.device, d.bytes, &ptr));
                d.base = static_cast<char*>(ptr);
            }
            auto at = [&](int device, size_t off) { return dev_slot(device).base + off; };

            b200::gx_geometry_t geo;
            geo.div          = b200::gx_make_div(per);
            geo.magic32      = per < (size_t(1) << 32) ? unsigned((size_t(1) << 32) / per) : 0u;
            geo.region_shift = region_shift;
            geo.regions      = regions;
            geo.buckets      = buckets;
            std::vector<b200::gx_requester_t<U>> rq(nreq);
            for (int r = 0; r < nreq; ++r) {
                const int rdev = std::get<0>(out_info[r]);
                auto& x        = rq[r];
                x.out          = out_base + std::get<1>(out_info[r]);
                x.src_base     = src_base;
                x.n            = n_r[r];
                x.nowners      = nown;
                x.geo          = geo;
                x.cap          = cap[r];
                x.counters     = reinterpret_cast<unsigned*>(at(rdev, counters_off[r]));
                x.pos          = reinterpret_cast<unsigned short*>(at(rdev, pos_off[r]));
                x.tile_base    = reinterpret_cast<unsigned*>(at(rdev, tbase_off[r]));
                x.tile_cnt     = reinterpret_cast<unsigned short*>(at(rdev, tcnt_off[r]));
                x.ctas_done    = reinterpret_cast<unsigned*>(at(rdev, done_off[r]));
                for (int q = 0; q < nown; ++q) {
                    x.queue[q]        = reinterpret_cast<unsigned*>(at(rdev, queue_off[r][q]));
                    x.resp[q]         = reinterpret_cast<U*>(at(rdev, resp_off[r][q]));
                    x.counters_pub[q] = reinterpret_cast<unsigned*>(at(std::get<0>(src_info[q]), pub_off[r][q]));
                }
            }
            std::vector<b200::gx_owner_t<U>> ow(nown);
            for (int q = 0; q < nown; ++q) {
                auto& y        = ow[q];
                y.src_local    = src_base + size_t(q) * per;
                y.part_elems   = std::get<2>(src_info[q]) * src_lead;
                y.nreq         = nreq;
                y.regions      = regions;
                y.region_shift = region_shift;
                for (int r = 0; r < nreq; ++r) {
                    y.queue[r]   = rq[r].queue[q];
                    y.counter[r] = rq[r].counters_pub[q];
                    y.cap[r]     = cap[r];
                    y.resp[r]    = rq[r].resp[q];
                }
            }
            static const int prefetch_mode = [] {
                const char* e = std::getenv("SCLX_B200_GATHER_PREFETCH");
                return e ? std::atoi(e) : 1;
            }();
            // SCLX_B200_GATHER_TRACE=1: wait after every phase and print its wall time (tuning aid)
            static const bool trace = std::getenv("SCLX_B200_GATHER_TRACE") != nullptr;
            std::vector<int> all_devices = b200::devices_of(out_info);
            for (int d : b200::devices_of(src_info))
                if (std::find(all_devices.begin(), all_devices.end(), d) == all_devices.end()) all_devices.push_back(d);
            auto phase_clock = std::chrono::steady_clock::now();
            auto phase_done  = [&](const char* what) {
                if (!trace) return;
                b200::sync_devices(all_devices);
                const auto now = std::chrono::steady_clock::now();
                std::fprintf(stderr, "gather %-9s %8.3f ms\n", what,
                             std::chrono::duration<double, std::milli>(now - phase_clock).count());
                phase_clock = now;
            };
            // ---- three launch groups --------------------------------------------------------
            b200::for_each_partition(out, out_info, [&](const b200::partition_t& p) {
                const auto& x = rq[p.ordinal];
                if (x.n == 0) return;
                cuda::cuda_exception::raise_if_not_success(
                    cudaMemsetAsync(x.counters, 0, sizeof(unsigned) * buckets, p.stream),
                    std::experimental::source_location::current(),
                    "sclx::algorithm::"
                );
                cuda::cuda_exception::raise_if_not_success(
                    cudaMemsetAsync(x.ctas_done, 0, sizeof(unsigned), p.stream),
                    std::experimental::source_location::current(),
                    "sclx::algorithm::"
                );
                // one tile per CTA (staggered CTA starts de-synchronise the load / bucket / store
                // phases of the chip; a single resident wave runs them in lock step)
                const unsigned grid = unsigned(std::min<size_t>(tiles_r[p.ordinal], size_t(1) << 20));
                auto launch_k1 = [&](auto kernel) {
                    cuda::cuda_exception::raise_if_not_success(
                        cudaFuncSetAttribute(
                            kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(b200::kGxPartitionSmem)
                        ),
                        std::experimental::source_location::current(),
                        "sclx::algorithm::"
                    );
                    kernel<<<grid, b200::kGxThreads, b200::kGxPartitionSmem, p.stream>>>(x, idx_ptr + p.offset);
                };
                if (nown == 1) launch_k1(b200::gx_partition_kernel<U, index_t_, b200::kGxOneOwner>);
                else if (src.elements() <= (size_t(1) << 32) && per < (size_t(1) << 32))
                    launch_k1(b200::gx_partition_kernel<U, index_t_, b200::kGxOwner32>);
                else launch_k1(b200::gx_partition_kernel<U, index_t_, b200::kGxOwner64>);
            });
            phase_done("partition");
            b200::for_each_partition(src, src_info, [&](const b200::partition_t& p) {
                unsigned max_cap = 0;
                for (int r = 0; r < nreq; ++r) max_cap = std::max(max_cap, cap[r]);
                if (max_cap == 0) return;
                const dim3 grid(
                    (max_cap + b200::kGxServeChunk - 1) / b200::kGxServeChunk, unsigned(nreq), regions
                );
                b200::gx_serve_kernel<U>
                    <<<grid, b200::kGxServeThreads, 0, p.stream>>>(ow[p.ordinal], prefetch_mode);
            });
            phase_done("serve");
            b200::for_each_partition(out, out_info, [&](const b200::partition_t& p) {
                const auto& x = rq[p.ordinal];
                if (x.n == 0) return;
                const unsigned grid = unsigned(std::min<size_t>(tiles_r[p.ordinal], size_t(1) << 20));
                b200::gx_place_kernel<U><<<grid, b200::kGxThreads, 0, p.stream>>>(x);
            });
            phase_done("place");
        }
    }
    std::vector<int> devices = b200::devices_of(out_info);
    for (int d : b200::devices_of(src_info))
        if (std::find(devices.begin(), devices.end(), d) == devices.end()) devices.push_back(d);
    b200::sync_devices(devices);
    cuda::set_device(caller_device);
}

template<class T, uint Rank, class I>
__host__ void
gather(const array<T, Rank>& src, const array<I, 1>& index, array<T, 1>& out, gather_mode mode = gather_mode::automatic) {
    gather(static_cast<const array<const T, Rank>&>(src), index, out, mode);
}

}  // namespace sclx::algorithm
