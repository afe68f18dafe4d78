"""The N > 1 path on CPU: two gloo ranks shard a record stream, run the same host logic, and reduce their step times."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_records, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    import ftf_loader
    ftf = ftf_loader.load()
    from flink_tensorflow_b200 import sharding
    import oracle
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        xs = np.random.default_rng(5).uniform(-100, 100, n_records).astype(np.float32)     # the same stream on every rank
        mine = list(sharding.shard_indices(n_records, rank, world))
        # host side of the per-record path: build the Example, the STRING tensor, its wire image; parse it back
        ys = []
        lib = ftf._abi.load_library()
        for i in mine:
            tv = ftf.messages_to_tensor([ftf.example(x=float(xs[i]))])
            back = ftf.TensorValue.read(tv.write())
            out = np.empty(1, np.float32)
            assert lib.ftf_parse_examples_f32(back.buffer.ctypes.data, back.buffer.nbytes, 1, b"x", 0, 0.0, out.ctypes.data) == 0
            ys.append(float(oracle.affine_f32(out, 0.5, 2.0)[0]))        # the device step is covered by the -m gpu tests
        dist.barrier()
        ms = sharding.max_over_ranks(10.0 + 5.0 * rank)                   # rank 1 is "slower"
        thr, ms2 = sharding.aggregate_throughput(len(mine), 10.0 + 5.0 * rank)
        gathered = [None] * world
        dist.all_gather_object(gathered, ys)
        if rank == 0:
            q.put((ms, thr, ms2, sharding.merge_sharded(gathered, n_records), xs.tolist()))
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_records_and_reduce_times():
    world, n = 2, 37
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    [p.start() for p in procs]
    ms, thr, ms2, merged, xs = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert ms == 15.0 and ms2 == 15.0                                      # MAX over ranks
    assert thr == pytest.approx((n // 2 + n % 2) * 2 / 0.015)              # rank 0's units x world / slowest time
    want = (np.float32(0.5) * np.asarray(xs, np.float32) + np.float32(2.0)).tolist()
    assert merged == want                                                  # every record exactly once, in order


def test_shard_indices_cover_everything_once():
    sys.path.insert(0, ROOT)
    import ftf_loader
    ftf_loader.load()
    from flink_tensorflow_b200 import sharding
    for n in (0, 1, 7, 64, 1000):
        for world in (1, 2, 4, 8):
            seen = sorted(i for r in range(world) for i in sharding.shard_indices(n, r, world))
            assert seen == list(range(n))
            sizes = [len(sharding.shard_indices(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    assert [sharding.device_for_subtask(i, 8) for i in (0, 7, 8, 13)] == [0, 7, 0, 5]
    with pytest.raises(ValueError):
        sharding.shard_indices(4, 2, 2)
