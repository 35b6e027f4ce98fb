"""Multi-GPU host logic on CPU: two gloo ranks shard a batch by read, compress
their shards (with the oracle -- this is a test), all-gather the per-read byte
positions and must arrive at the single-process container index."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_reads(nreads, seed):
    rng = np.random.default_rng(seed)
    lens = rng.integers(1, 9000, nreads)
    return [(470 + np.cumsum(rng.integers(-30, 31, int(n)))).astype(np.int16) for n in lens]


def _worker(rank, world, port, nreads, method, q):
    import torch
    import torch.distributed as dist
    from honours_b200 import shard
    from oracle import cpu_impl
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc = cpu_impl.load("oracle")
        reads = _make_reads(nreads, 123)
        r0, r1 = shard.equal_count_ranges(nreads, world)[rank]
        per = -(-nreads // world)
        streams = [orc.press(method, s)[1] for s in reads[r0:r1]]
        sizes = np.zeros(per, dtype=np.int64)           # padded to the common count
        sizes[: r1 - r0] = [len(s) for s in streams]
        local = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        gathered = shard.allgather_offsets(torch.from_numpy(local), world).numpy().reshape(world, per + 1)
        index = shard.global_index(gathered.astype(np.uint64))
        # this rank's shard lands in one contiguous sub-range of the container
        mine = index[rank * per: rank * per + (r1 - r0)]
        q.put((rank, index.tolist(), int(mine[0]), b"".join(streams)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("method", ["svb16_zd", "vbe21_zd"])
def test_two_rank_container_index_matches_single_process(oracle, method):
    import torch.multiprocessing as mp
    world, nreads = 2, 37
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nreads, method, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    reads = _make_reads(nreads, 123)
    expect_streams = [oracle.press(method, s)[1] for s in reads]
    expect_off = np.concatenate([[0], np.cumsum([len(s) for s in expect_streams])])
    per = -(-nreads // world)
    assert res[0][1] == res[1][1]                        # every rank derives the same index
    index = np.array(res[0][1], dtype=np.uint64)
    assert int(index[-1]) == int(expect_off[-1])
    container = bytearray(int(index[-1]))
    for rank, _, base, blob in res:
        container[base: base + len(blob)] = blob
        r0 = rank * per
        n = min(nreads, r0 + per) - r0
        assert index[rank * per: rank * per + n].tolist() == expect_off[r0: r0 + n].tolist()
    assert bytes(container) == b"".join(expect_streams)


def test_partition_is_balanced_and_covers():
    from honours_b200 import shard
    rng = np.random.default_rng(1)
    lens = rng.integers(2000, 200000, 5000)
    for nr in (1, 2, 4, 8):
        parts = shard.partition_reads(lens, nr)
        assert parts[0][0] == 0 and parts[-1][1] == lens.size
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        tot = [int(lens[a:b].sum()) for a, b in parts]
        assert max(tot) - min(tot) <= 2 * int(lens.max())


def test_global_index_matches_definition():
    from honours_b200 import shard
    g = np.array([[0, 5, 9], [0, 2, 2], [0, 1, 7]], dtype=np.uint64)
    assert shard.global_index(g).tolist() == [0, 5, 9, 11, 11, 12, 18]
