"""N>1 host logic (scalix_b200/multi.py) on CPU: world_size 2 and 3 over gloo.

Each rank owns one last-dimension partition of the same index-keyed synthetic array; the
per-partition compute is the CHECKER (oracle/, injected — multi.py has no compute of its own),
so what is under test is the product's host side: partition arithmetic, the two exchanges and
the operand order of the carry / fold.  The assembled result must equal the serial oracle and
the reference's scan + host chain + fix-up restatement."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, ret):
    sys.path.insert(0, REPO)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from scalix_b200 import multi

        if case == "scan":
            leading, last = 4, 10_007
            start, ncols = multi.rank_partition(last, world, rank)
            local = oracle.masked_i64(leading * ncols, 0x5CA5, 0xFFFF, first=leading * start)
            box = {}

            def local_total():
                return torch.tensor([oracle.reduce(local, oracle.I64, oracle.PLUS, 0)], dtype=torch.int64)

            def local_scan(totals, n_lower):
                carry = np.int64(0)
                for t in totals[:n_lower].tolist():  # left to right, like the kernel and the host chain
                    carry = np.int64(carry + np.int64(t))
                box["out"] = (oracle.inclusive_scan(local, oracle.I64, oracle.PLUS) + carry) if local.size else local
                return box["out"]

            out = multi.distributed_inclusive_scan(local_total, local_scan, torch.zeros(1, dtype=torch.int64))
            pieces = [None] * world
            dist.all_gather_object(pieces, (start, np.asarray(out)))
            if rank == 0:
                got = np.concatenate([p for _, p in sorted(pieces, key=lambda x: x[0])])
                full = oracle.masked_i64(leading * last, 0x5CA5, 0xFFFF)
                parts = oracle.partition(last, world)
                assert np.array_equal(got, oracle.inclusive_scan(full, oracle.I64, oracle.PLUS))
                assert np.array_equal(got, oracle.inclusive_scan_partitioned(
                    full, [leading * ln for _, ln in parts], oracle.I64, oracle.PLUS))
        elif case == "rld":
            rows, last = 64, 5_003
            start, ncols = multi.rank_partition(last, world, rank)
            local = oracle.uniform_f32(rows * ncols, 0x5CA11, first=rows * start)

            def local_rows():
                return torch.from_numpy(oracle.reduce_last_dim(local, rows, [ncols], oracle.F32, oracle.PLUS, 0.0))

            def fold(parts, w):
                acc = np.zeros(rows, np.float32)
                p = parts.numpy().reshape(w, rows)
                for d in range(w):  # fixed rank order
                    acc = (acc + p[d]).astype(np.float32)
                return torch.from_numpy(acc)

            out = multi.distributed_reduce_last_dim(local_rows, fold).numpy()
            every = [None] * world
            dist.all_gather_object(every, out)
            assert all(np.array_equal(e, out) for e in every), "ranks disagree on the folded rows"
            full = oracle.uniform_f32(rows * last, 0x5CA11)
            want = oracle.reduce_last_dim_f64acc(full, rows)
            assert np.all(np.abs(out.astype(np.float64) - want) <= 1e-5 * np.abs(want))
        elif case == "partition":
            # more ranks than columns: trailing ranks own nothing (array.cuh:471-489 drops devices)
            got = multi.rank_partition(1, world, rank)
            assert got == ((0, 1) if rank == 0 else (1, 0))
        ret[rank] = "ok"
    except BaseException as exc:  # surface in the parent
        ret[rank] = f"{type(exc).__name__}: {exc}"
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", ["scan", "rld", "partition"])
def test_multi_rank_host_logic_gloo(sclx, case, world):
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, case, ret), nprocs=world, join=True)
        assert [ret.get(r) for r in range(world)] == ["ok"] * world
