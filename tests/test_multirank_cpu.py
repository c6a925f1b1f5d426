"""CPU tier, world_size 2 over gloo: the N > 1 host logic -- row-block sharding, per-rank rendering of only the
owned rows, gather of the shards on rank 0 -- gives the same frame as one rank.  The per-rank renderer here is the
host emulation of the device code (tests/emul); on the GPU the same sharding rule runs inside the CUDA library
(tests/test_gpu_parity.py::test_sharded_render_equals_whole_frame)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, W, H, block_rows, tmp):
    for p in (ROOT, os.path.join(ROOT, "opencl-raytracer_b200"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    import emul
    from pyrt import scenes as S
    from pyrt import shard, types as T

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    objs, lights = S.synthetic_scene(1200, 150)
    z = T.camera_zdir(H)
    rows = shard.shard_rows(H, block_rows, rank, world)
    rays = T.camera_rays(W, H, z, rows=rows)
    out, _, _, c = emul.render(objs, lights, rays, 4, 2, 2, 1, nthreads=1)
    # gather: shards have different sizes (ragged last block) -> pad to the largest
    n_max = torch.tensor([len(out)])
    dist.all_reduce(n_max, op=dist.ReduceOp.MAX)
    pad = torch.zeros((int(n_max), 4), dtype=torch.float32)
    pad[: len(out)] = torch.from_numpy(out)
    counts = torch.tensor([c["ref_primary"], c["traced_shadow"], c["traced_reflect"]], dtype=torch.int64)
    dist.all_reduce(counts)
    bufs = [torch.zeros_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, bufs, dst=0)
    if rank == 0:
        frame = np.full((W * H, 4), -1.0, dtype=np.float32)
        for r in range(world):
            rr = shard.shard_rows(H, block_rows, r, world)
            shard.scatter_rows(frame, W, rr, bufs[r].numpy()[: len(rr) * W])
        np.save(os.path.join(tmp, "frame.npy"), frame)
        np.save(os.path.join(tmp, "counts.npy"), counts.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("block_rows", [4, 16])
def test_two_ranks_equal_one(tmp_path, block_rows):
    import torch.multiprocessing as mp

    import emul
    from helpers import frame_rays
    from pyrt import scenes as S

    W, H = 64, 37  # 37 rows: ragged last block, unequal shards
    emul.lib()  # build the emulation library once, before the workers race for it
    port = 29600 + os.getpid() % 300 + block_rows
    mp.spawn(_worker, args=(2, port, W, H, block_rows, str(tmp_path)), nprocs=2, join=True)
    frame = np.load(tmp_path / "frame.npy")
    counts = np.load(tmp_path / "counts.npy")
    objs, lights = S.synthetic_scene(1200, 150)
    _, rays = frame_rays(W, H)
    whole, _, _, c = emul.render(objs, lights, rays, 4, 2, 2, 1)
    assert np.array_equal(frame, whole)
    assert counts.tolist() == [c["ref_primary"], c["traced_shadow"], c["traced_reflect"]]


def test_shard_rows_partition():
    from pyrt import shard

    for H, br, n in ((4320, 16, 8), (37, 4, 2), (100, 32, 3), (5, 8, 4)):
        allrows = np.concatenate([shard.shard_rows(H, br, s, n) for s in range(n)])
        assert sorted(allrows.tolist()) == list(range(H))
    with pytest.raises(ValueError):
        shard.shard_rows(10, 0, 0, 1)


def test_every_ranks_launch_sequences_cover_the_frame_once(rt):
    """The readback plan of a sharded rt_render (rt_plan_sequences, host logic of the C-ABI library): over all ranks and all of a
    rank's launch sequences every row of the frame is rendered and copied exactly once, whatever the copy : kernel ratio says."""
    from pyrt import shard

    W = 7680
    for H, br, n in ((4320, 16, 8), (4320, 16, 4), (2160, 16, 2), (37, 4, 2), (100, 32, 3)):
        for ratio in (0.5, 1.25, 3.0):
            seen = np.zeros(H, dtype=np.int32)
            for s in range(n):
                rows = shard.shard_rows(H, br, s, n)
                b = rt.plan_sequences(W, H, br, s, n, 0, 0.0, ratio).astype(np.int64)
                blocks = list(range(s, -(-H // br), n))
                assert b[-1] == len(blocks)
                mine = []
                for k in range(len(b) - 1):
                    for u in range(b[k], b[k + 1]):
                        r0 = blocks[u] * br
                        mine.extend(range(r0, min(r0 + br, H)))
                assert mine == rows.tolist()
                seen[mine] += 1
            assert np.all(seen == 1)
