"""Multi-GPU host logic on CPU: image sharding (no data-path collective) exercised with
torch.distributed gloo, world_size 2 — each rank runs the ORACLE on its shard (standing in for the
device), the gathered per-image checksums must equal the single-process result."""
import os
import socket
import sys

import numpy as np
import pytest

from libultrahdr_dev_b200 import shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n,world", [(0, 1), (1, 2), (7, 2), (8, 8), (1024, 8), (5, 8), (13, 4)])
def test_shards_partition_the_batch(n, world):
    for fn in (shard.shard_round_robin, shard.shard_contiguous):
        parts = [fn(n, r, world).indices for r in range(world)]
        assert sorted(i for p in parts for i in p) == list(range(n))
        sizes = [len(p) for p in parts]
        assert max(sizes) - min(sizes) <= 1
    for i in range(n):
        assert i in shard.shard_round_robin(n, shard.owner_round_robin(i, world), world).indices
    with pytest.raises(ValueError):
        shard.shard_round_robin(4, 2, 2)


def test_merge_checksums_detects_gaps_and_duplicates():
    assert shard.merge_checksums([[(0, 5), (2, 7)], [(1, 6)]]) == [5, 6, 7]
    with pytest.raises(ValueError):
        shard.merge_checksums([[(0, 5)], [(0, 6)]])
    with pytest.raises(ValueError):
        shard.merge_checksums([[(0, 5)], [(2, 6)]])
    assert shard.whole_job_throughput([2, 2], 100, [1.0, 2.0]) == 200.0


def _image_checksum(orc, index, w=64, h=32):
    from libultrahdr_dev_b200 import abi, synth

    sdr_b, hdr_b = synth.natural_pair(w, h, index)
    sdr, hdr = abi.host_yuv420(sdr_b, w, h, 0), abi.host_p010(hdr_b, w, h, 2)
    st, mp, md, _ = orc.generate_gainmap(sdr, hdr, abi.TF_HLG)
    st2, out, _ = orc.apply_gainmap(sdr, abi.host_map(mp, w // 4, h // 4), md, abi.OUTPUT_HDR_HLG, 3.4e38)
    assert st == 0 and st2 == 0
    return int(mp.astype(np.int64).sum()) * 1000003 + int(out.view(np.uint32).astype(np.int64).sum())


def _worker(rank, world, port, n_images, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from oracle.pyoracle import CpuImpl

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = CpuImpl.oracle()
    mine = shard.shard_round_robin(n_images, rank, world)
    part = [(i, _image_checksum(orc, i)) for i in mine.indices]
    # control plane only: gather checksums, reduce a timing with MAX (what bench.py does with NCCL)
    gathered = [None] * world
    dist.all_gather_object(gathered, part)
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.barrier()
    if rank == 0:
        q.put((shard.merge_checksums(gathered), float(t.item())))
    dist.destroy_process_group()


def test_world_size_2_gloo_sharded_batch_equals_single_process(orc):
    import torch.multiprocessing as mp

    n_images, world = 7, 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_images, q)) for r in range(world)]
    for p in procs:
        p.start()
    merged, tmax = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert tmax == float(world)
    assert merged == [_image_checksum(orc, i) for i in range(n_images)]
