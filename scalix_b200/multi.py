"""Host side of the one-process-per-GPU (torchrun) form of the partitioned algorithms.

The reference is one process driving N devices (scalix/array.cuh:467-513 partitions the last
dimension; scalix/algorithm/inclusive_scan.cuh:105-157 and reduce_last_dim.cuh:208-211 combine
the per-device results on the HOST through Unified Memory).  Under torchrun every rank owns
one partition of the same global array; the only cross-rank steps are the two tiny exchanges
below.  The per-partition compute is injected (`local_*` callables): bench.py / the GPU tests
pass the sm_100a kernels of libscalix_b200.so, the CPU (gloo) tests pass the checker — this
module itself never touches oracle/ and has no CPU compute of its own.

    scan          : total_r = local_total()            # pass A, one read of the partition
                    totals  = all_gather(total_r)      # world x sizeof(T) bytes
                    local_scan(lower_totals=totals[:rank])   # carry folded inside the kernel
    reduce_last_dim: row_r  = local_rows()             # [rows] partial per rank
                    rows    = all_gather(row_r)        # world x rows x sizeof(T)
                    out     = fold(rows) in FIXED rank order (bit-stable floats, unlike the
                              mutex-ordered host fold of reduce_last_dim.cuh:208-211)
"""
import torch
import torch.distributed as dist


def rank_partition(last_dim, world, rank):
    """(start_col, n_cols) of `rank`'s partition, or (last_dim, 0) when the split uses fewer
    devices than ranks — same arithmetic as array::set_primary_devices
    (scalix/array.cuh:471-508): ceil(last/ndev) columns per device, trailing devices dropped."""
    from . import partition

    parts = partition(last_dim, world)
    if rank < len(parts):
        return parts[rank]
    return (last_dim, 0)


def world_size(group=None):
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def all_gather_flat(t, group=None):
    """all ranks' copies of the 1-D tensor `t`, concatenated in rank order (same device/dtype)"""
    w = world_size(group)
    if w == 1:
        return t.clone()
    out = torch.empty(w * t.numel(), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    return out


def exchange_scan_totals(total, group=None):
    """`total`: 1-element tensor holding this partition's fold.  Returns (all totals in rank
    order, number of lower partitions): exactly the `lower_totals` / `n_lower` arguments of
    sclx_b200_inclusive_scan, which folds them left to right as the reference's host chain does
    (inclusive_scan.cuh:122: last(d) = f(last(d-1), last(d)))."""
    rank = dist.get_rank(group) if world_size(group) > 1 else 0
    return all_gather_flat(total.reshape(1), group), rank


def distributed_inclusive_scan(local_total, local_scan, total_like, group=None):
    """local_total() -> 1-element tensor; local_scan(totals_tensor, n_lower) runs the carry-seeded
    scan of this rank's partition.  Rank 0 needs no pass A result but still contributes its
    total (later ranks need it)."""
    w = world_size(group)
    if w == 1:
        return local_scan(total_like, 0)
    totals, n_lower = exchange_scan_totals(local_total(), group)
    return local_scan(totals, n_lower)


def distributed_reduce_last_dim(local_rows, fold, group=None):
    """local_rows() -> [rows] tensor of this partition's partial rows; fold(parts[w*rows], w)
    -> [rows] folds them in rank order.  Every rank ends with the full result (the reference
    leaves it in managed memory visible to all devices)."""
    row = local_rows()
    w = world_size(group)
    if w == 1:
        return fold(row, 1)
    return fold(all_gather_flat(row, group), w)
