"""CPU, world_size 2 over gloo: the multi-GPU plan without GPUs.

Each rank takes the contiguous block range the C ABI assigns it (pim_snappy_shard_range, the split of
snappy_compress.c:495), produces that range's block stream with the checker (the oracle stands in for the
device here -- this test is about the partition and the stitch, not the kernels), and rank 0 stitches
header + shards in rank order.  The stitched file must equal the one-piece file byte for byte, the ranges
must tile the blocks, and the timing reduction bench.py uses (MAX over ranks) must agree on every rank."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, bs, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import pim_compression_b200 as P
    from oracle.oracle import Oracle

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle = Oracle()
        data = P.synth_host(P.SYNTH_MIX, 0x5EED, n)            # every rank regenerates the same seeded input
        nb = int(P.lib().pim_snappy_num_blocks(n, bs))
        lo, hi = P.shard_range(nb, world, rank)
        piece = data[lo * bs:min(n, hi * bs)]
        shard = oracle.compress(piece, bs)
        hdr, _, _ = oracle.read_header(shard)
        body = shard[hdr:] if piece.size else np.zeros(0, np.uint8)  # a shard is a block stream without the file header
        gathered = [None] * world
        dist.all_gather_object(gathered, (lo, hi, body.tobytes()))
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)  # bench.py's max-over-ranks on a fake per-rank time
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok_max = float(t.item()) == float(world)
        if rank == 0:
            whole = oracle.compress(data, bs)
            whdr, _, _ = oracle.read_header(whole)
            stitched = whole[:whdr].tobytes() + b"".join(g[2] for g in gathered)
            tiles = gathered[0][0] == 0 and gathered[-1][1] == nb and all(gathered[i][1] == gathered[i + 1][0] for i in range(world - 1))
            q.put((ok_max, tiles, stitched == whole.tobytes(), [(g[0], g[1]) for g in gathered]))
        else:
            q.put((ok_max, True, True, None))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,bs", [((3 << 20) + 777, 32768), (100000, 4096), (70000, 65536)])
def test_two_ranks_partition_and_stitch(P, oracle, n, bs):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, bs, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[0] for r in res), "MAX over ranks disagrees"
    assert all(r[1] for r in res), "shard ranges do not tile the blocks"
    assert all(r[2] for r in res), "stitched shards differ from the one-piece file"


def test_shard_ranges_match_reference_split(P):
    # ceil(nb / n) per shard, contiguous, trailing shards may be empty (snappy_compress.c:495)
    for nb in (0, 1, 2, 7, 8, 9, 312, 2048, 32768):
        for ng in (1, 2, 4, 8):
            per = (nb + ng - 1) // ng
            got = [P.shard_range(nb, ng, g) for g in range(ng)]
            assert got == [(min(nb, per * g), min(nb, per * g + per)) for g in range(ng)]
            assert got[0][0] == 0 and got[-1][1] == nb
