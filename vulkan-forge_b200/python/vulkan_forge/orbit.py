"""Frame-batch sharding across GPUs (additive).

The path shards trivially: frames of a camera orbit are independent, so rank p of P renders the
contiguous block [floor(pN/P), floor((p+1)N/P)) on its own GPU and the frames are gathered on
the host.  No collective is on the data path; one process per GPU (torchrun) is the launcher.
"""
from __future__ import annotations

import math
import os

import numpy as np


def shard_range(n_frames: int, rank: int, world: int):
    """Contiguous block of frame indices owned by `rank` (SURVEY 8e partitioning)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("need 0 <= rank < world")
    return (rank * n_frames) // world, ((rank + 1) * n_frames) // world


def bind_host_to_gpu(local_index: int = 0) -> dict:
    """Pins the calling process to the CPUs next to GPU `local_index` (as CUDA numbers it, i.e. after
    CUDA_VISIBLE_DEVICES), so that pinned host buffers allocated afterwards land on the GPU's NUMA
    node and the per-GPU DMA streams of a one-process-per-GPU job do not share one socket's memory.
    Call it before the first CUDA allocation.  Returns what was done; never raises (a container may
    hide the topology or restrict the CPU set)."""
    info = {"bound": False}
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = [v.strip() for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip()]
        if local_index < len(vis) and not vis[local_index].isdigit():
            h = pynvml.nvmlDeviceGetHandleByUUID(vis[local_index])
        else:
            h = pynvml.nvmlDeviceGetHandleByIndex(int(vis[local_index]) if local_index < len(vis) else local_index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        near = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (int(wd) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        info.update(near=len(near), allowed=len(allowed))
        both = near & allowed
        if both and both != allowed:
            os.sched_setaffinity(0, both)
            info.update(bound=True, cpus=len(both))
    except Exception as e:  # noqa: BLE001 - topology is a hint, never a requirement
        info["error"] = f"{type(e).__name__}: {e}"
    return info


def orbit_cameras(n_frames: int, radius: float = 3.0 * math.sqrt(2.0), height: float = 2.0,
                  fovy_deg: float = 45.0, znear: float = 0.1, zfar: float = 100.0, start: int = 0,
                  stop: int | None = None):
    """Camera tuples (eye, target, up, fovy, znear, zfar) for frames [start, stop) of an
    n_frames orbit: eye = (R cos t_k, height, R sin t_k), t_k = 2 pi k / n."""
    stop = n_frames if stop is None else stop
    cams = []
    for k in range(start, stop):
        t = 2.0 * math.pi * k / n_frames
        cams.append(((radius * math.cos(t), height, radius * math.sin(t)), (0.0, 0.0, 0.0),
                     (0.0, 1.0, 0.0), fovy_deg, znear, zfar))
    return cams


def render_orbit(scene, n_frames: int, rank: int | None = None, world: int | None = None,
                 gather=None, **orbit_kw):
    """Render this rank's block of an orbit with `scene` and return (first_index, frames).

    rank/world default to torchrun's RANK/WORLD_SIZE.  `gather`, if given, is called as
    gather(first_index, frames) -- e.g. a writer into a shared (N,H,W,4) host array; there is
    deliberately no device-side exchange.
    """
    rank = int(os.environ.get("RANK", 0)) if rank is None else rank
    world = int(os.environ.get("WORLD_SIZE", 1)) if world is None else world
    lo, hi = shard_range(n_frames, rank, world)
    if hi == lo:
        frames = np.empty((0, scene.height, scene.width, 4), np.uint8)
    else:
        frames = scene.render_batch_rgba(orbit_cameras(n_frames, start=lo, stop=hi, **orbit_kw))
    if gather is not None:
        gather(lo, frames)
    return lo, frames
