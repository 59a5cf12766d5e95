"""Multi-GPU frame sharding on CPU: world_size-2 gloo processes each render their
block of an orbit with a stand-in scene and the host gather reproduces the single-rank result."""
import os
import socket

import numpy as np
import pytest


def test_shard_range_partitions_exactly(vf):
    for n in (0, 1, 7, 16, 1024):
        for world in (1, 2, 3, 4, 8):
            blocks = [vf.shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[k][1] == blocks[k + 1][0] for k in range(world - 1))
            assert max(b - a for a, b in blocks) - min(b - a for a, b in blocks) <= 1
    with pytest.raises(ValueError):
        vf.shard_range(4, 2, 2)


def test_orbit_cameras_shape(vf):
    cams = vf.orbit_cameras(8)
    assert len(cams) == 8 and len(cams[0]) == 6
    eye = np.array([c[0] for c in cams])
    assert np.allclose(np.hypot(eye[:, 0], eye[:, 2]), 3 * np.sqrt(2)) and np.allclose(eye[:, 1], 2.0)
    assert vf.orbit_cameras(8, start=2, stop=5) == cams[2:5]


class _FakeScene:
    """Deterministic stand-in renderer: frame = f(camera eye); lets the sharding logic run on CPU."""
    width, height = 8, 6

    def render_batch_rgba(self, cameras):
        out = np.zeros((len(cameras), self.height, self.width, 4), np.uint8)
        for k, c in enumerate(cameras):
            out[k] = int((c[0][0] * 10 + c[0][2] * 7) % 251)
        return out


def _worker(rank, world, port, n_frames, q):
    import torch.distributed as dist
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "vulkan-forge_b200", "python"))
    import vulkan_forge as vf
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, frames = vf.render_orbit(_FakeScene(), n_frames)            # rank/world from the env
    dist.barrier()                                                   # the only collective: timing fence
    gathered = [None] * world
    dist.all_gather_object(gathered, (lo, frames.tobytes(), frames.shape))   # test-side host gather
    dist.destroy_process_group()
    if rank == 0:
        q.put(gathered)


def test_two_rank_gloo_orbit_equals_single_rank(vf):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    n_frames, world = 7, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    gathered = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    full = np.zeros((n_frames, 6, 8, 4), np.uint8)
    for lo, buf, shape in gathered:
        full[lo:lo + shape[0]] = np.frombuffer(buf, np.uint8).reshape(shape)
    _, single = vf.render_orbit(_FakeScene(), n_frames, rank=0, world=1)
    assert np.array_equal(full, single)
