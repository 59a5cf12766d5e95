#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native vulkan-forge render path.

Workload (BASELINE.json configs[2], the one `metric` is quoted on): Scene 1920x1080, synthetic
4096x4096 DEM (seed 0), grid 4096 (every texel a vertex), terrain LUT, NORTHSTAR shading (sun
35/120, exposure 1.1, Reinhard, 1-99 % percentile range).  A step = FRAMES (default 32) orbit frames per GPU over one DEM.

  value : frames/s, whole job, DEM resident in HBM, frames left in HBM (CUDA events on the launch
          stream, L2 flushed between steps, max over ranks)
  e2e   : frames/s through the public Python API with HOST buffers: every step uploads the DEM from
          pinned host memory, renders, and gathers the frames into pinned host memory
  roofline / cpu_baseline : see DESIGN.md "Measurement"
`--impl reference` times the reference's CPU path instead: wgpu on a software adapter is
unavailable on this box and the Rust crate cannot be built, so it is the C oracle (oracle/)
on all host threads -- the one place besides cpu_baseline where bench.py executes oracle/.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "vulkan-forge_b200", "python")


def _load_by_path(name, rel):
    """Pure-Python helper modules of the package (synthetic DEM, orbit cameras) loaded by file path, WITHOUT importing
    the package: the reference arm must not load the product's native libraries."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(name, os.path.join(PKG, "vulkan_forge", rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


W, H, G, DEM_N, DEM_SEED, CMAP = 1920, 1080, 4096, 4096, 0, "terrain"
WORKLOAD = "c3: Scene 1920x1080, DEM 4096^2 seed 0, grid 4096, terrain LUT, NORTHSTAR (sun 35/120, exposure 1.1, auto 1-99%)"
ALGO_BYTES_FRAME = 4 * min(G * G, DEM_N * DEM_N) + 4 * W * H + 1024 + 176


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.p = None
        vis = [v.strip() for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip()]
        if index < len(vis):                               # nvidia-smi counts physical devices (index or UUID both accepted)
            index = vis[index]
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            pass

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out = self.p.communicate(timeout=5)[0]
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ncu_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (None if absent)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json"))).get(kernel)
    except Exception:
        return None


def orbit_uniform_blocks(vf, scene, n, frames):
    return scene.camera_blocks(vf.orbit_cameras(n)[:frames])


def host_threads():
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1, which must not
    throttle the CPU arm)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def workload_config(frames):
    """The same object in both arms (the driver compares them)."""
    return {"workload": WORKLOAD, "frames_per_step_per_gpu": frames, "l2": "flushed between steps (256 MiB write)",
            "sharding": "independent frames per GPU, no collective"}


def cpu_reference_fps(frames_budget_s, dem, uniforms, hmin, hmax, min_frames=2, max_frames=64):
    """Reference CPU path stand-in (oracle, all host threads) on a bounded sample of the workload."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import vf_oracle as oracle
    lut = np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{CMAP}.rgba"), np.uint8)
    t0 = time.perf_counter()
    n = 0
    while n < max_frames and (n < min_frames or time.perf_counter() - t0 < frames_budget_s):
        oracle.render_terrain(W, H, G, dem, uniforms[n % len(uniforms)], lut, mode=oracle.MODE_NORTHSTAR, depth=False,
                              h_min=hmin, h_max=hmax, threads=host_threads())
        n += 1
    dt = time.perf_counter() - t0
    return n / dt, n, host_threads(), dt


def run_reference(args):
    """--impl reference: CPU only; rank 0 alone works.  Nothing of the product is imported here: cameras and the DEM
    come from the package's pure-Python helpers loaded by path, uniforms and frames from oracle/."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import vf_oracle as oracle
    synthetic_dem = _load_by_path("vf_synthetic", "_synthetic.py").synthetic_dem
    orbit_cameras = _load_by_path("vf_orbit", "orbit.py").orbit_cameras
    dem = synthetic_dem(DEM_SEED, DEM_N)
    hmin, hmax = (float(x) for x in oracle.h_range(dem))
    cams = orbit_cameras(max(args.frames, 1) * 4)
    us = [oracle.look_at_uniforms(W, H, c[0], c[1], c[2], c[3], c[4], c[5], sun=oracle.sun_dir(35.0, 120.0), exposure=1.1,
                                  h_min=hmin, h_max=hmax, normalize_sun=False) for c in cams]
    lut = np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{CMAP}.rgba"), np.uint8)
    per_step = args.ref_frames                        # bounded sample: the first `per_step` frames of every step

    def step(k):
        for j in range(per_step):
            oracle.render_terrain(W, H, G, dem, us[(k * args.frames + j) % len(us)], lut, mode=oracle.MODE_NORTHSTAR,
                                  h_min=hmin, h_max=hmax, threads=host_threads())
    for k in range(args.warmup):
        step(k)
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(k)
    dt = time.perf_counter() - t0
    fps = args.steps * per_step / dt
    line = {
        "impl": "reference", "metric": "terrain frames/s (1080p, 4096^2 DEM)", "value": fps, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": args.frames / fps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.frames),
        "mpix_per_s": fps * W * H / 1e6,
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": host_threads(), "kind": "port",
                         "sample": f"{per_step} of the {args.frames} frames of each of {args.steps} steps ({dt:.1f} s); oracle/ (C + OpenMP, all host threads); "
                                   "ms_per_step is the sample's rate scaled to a whole step; the reference's wgpu software path is unavailable on this box"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=32, help="frames per step per GPU")
    ap.add_argument("--ref-frames", type=int, default=2, help="frames per step of the CPU reference arm")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    sys.path.insert(0, PKG)
    import numpy as np
    import torch
    import vulkan_forge as vf
    from vulkan_forge._synthetic import synthetic_dem
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    numa = vf.bind_host_to_gpu(local) if world > 1 and not os.environ.get("VF_NO_BIND") else {"bound": False}
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ext = vf._vulkan_forge

    dem = torch.from_numpy(synthetic_dem(DEM_SEED, DEM_N)).pin_memory().numpy()     # pinned host input
    sc = ext.Scene(W, H, grid=G, colormap=CMAP)
    stream = torch.cuda.Stream()
    sc.set_stream(stream.cuda_stream)
    sc.set_height_from_r32f(dem)
    sc.set_shading("northstar", depth=False)
    sc.set_sun(35.0, 120.0)
    sc.set_exposure(1.1)
    hmin, hmax = sc.set_auto_height_range()
    F = args.frames
    n_orbit = F * max(world, 1) * 4
    lo = (rank * F) % n_orbit
    blocks = sc.camera_blocks(vf.orbit_cameras(n_orbit)[lo:lo + F])                  # each rank its own cameras
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")                 # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: device-resident ----
    for _ in range(args.warmup):
        sc.render_batch_device(blocks)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with torch.cuda.stream(stream):
        for k in range(args.steps):
            flush.zero_()                                                            # L2 flush, outside the events
            ev[k][0].record(stream)
            sc.render_batch_device(blocks)
            ev[k][1].record(stream)
    barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    clocks = sampler.stop() if sampler else None

    # per-kernel times of one frame (CUDA events inside the library, same stream), L2 flushed
    sc.enable_timing(True)
    kt = {"ms_geom": [], "ms_bin": [], "ms_raster": [], "ms_total": []}
    for k in range(max(5, args.warmup)):
        with torch.cuda.stream(stream):
            flush.zero_()
        sc.render_batch_device(blocks[k % F:k % F + 1])
        m = sc.render_metrics()
        for key in kt:
            kt[key].append(m[key])
    sc.enable_timing(False)
    kmed = {k: statistics.median(v) for k, v in kt.items()}

    # ---- e2e: public API, host buffers, H2D + D2H inside the timed region ----
    # Every step uploads the DEM from pinned host memory (64 MiB) and gathers its F frames into pinned
    # host memory; the steps are pipelined (upload k+1 and batch k overlap: set_height_from_r32f_async /
    # render_batch_rgba_async / collect), as a tile- or orbit-batch job would run them.
    def e2e_run(steps):
        acc, prev = 0, None
        for _ in range(steps):
            sc.set_height_from_r32f_async(dem)             # 64 MiB H2D from pinned memory
            tk = sc.render_batch_rgba_async(blocks)        # F frames, D2H into pinned host memory
            if prev is not None:
                acc += int(sc.collect(prev)[:, ::270, ::480, :3].sum())     # device->host result actually read
            prev = tk
        acc += int(sc.collect(prev)[:, ::270, ::480, :3].sum())
        return acc
    e2e_run(args.warmup)
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()

    # ---- e2e_png: the same steps, frames delivered as .png files in pinned host memory (encoded on the device: only the
    # compressed bytes cross the bus) ----
    # Pipelined like the raw leg: the DEM of step k+1 uploads (64 MiB H2D per step, inside the timed region) while step k
    # renders, encodes and drains (render_batch_png_raw(..., next_height=) = vf_scene_render_batch_png_next).
    def png_run(steps):
        total = 0
        sc.set_height_from_r32f_async(dem)
        for k in range(steps):
            buf, sizes = sc.render_batch_png_raw(blocks, next_height=dem if k + 1 < steps else None)
            total += int(sizes.sum()) + int(buf[0, 1])          # device->host result actually read
        return total
    png_steps = max(4, args.steps // 2)
    png_run(4)                                                   # the host slots of the files settle (pinned allocations) in the first three
    barrier()
    t0 = time.perf_counter()
    png_bytes = png_run(png_steps)
    torch.cuda.synchronize()
    png_s = time.perf_counter() - t0
    barrier()

    # ---- parity gate of this very run (outside the timed regions, rank 0): frame 0 of the timed batch against the oracle ----
    parity = None
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import vf_oracle as oracle
        lut = np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{CMAP}.rgba"), np.uint8)
        u0 = np.asarray(blocks[0])
        got = np.asarray(sc.render_batch_rgba(blocks[0:1])[0])
        sc.set_uniforms_raw(u0)
        gvis = sc.debug_visibility()
        ref, rvis = oracle.render_terrain(W, H, G, dem, u0, lut, mode=oracle.MODE_NORTHSTAR, depth=False, h_min=float(hmin),
                                          h_max=float(hmax), threads=host_threads(), want_prim=True)
        d = np.abs(got.astype(np.int16) - ref.astype(np.int16)).max(-1)
        olo, ohi = oracle.h_range(dem)
        parity = {"frames_checked": 1, "frame": "frame 0 of the timed batch", "within_1lsb": float((d <= 1).mean()), "bit_identical": bool((d == 0).all()),
                  "coverage_mismatch": int((gvis != rvis).sum()), "percentile_range_bit_exact": bool(np.float32(hmin) == olo and np.float32(hmax) == ohi),
                  "against": "oracle/ (CPU restatement; pixel parity with the reference's wgpu renderer is unpinned: it cannot run on this box)"}

    t = torch.tensor([dev_ms, e2e_s, png_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_s, png_s = float(t[0]), float(t[1]), float(t[2])
    png_fps = png_steps * F * world / png_s
    total_frames = args.steps * F * world
    fps = total_frames / (dev_ms / 1e3)
    e2e_fps = total_frames / e2e_s

    if rank == 0:
        peak, peak_kind = peaks()
        dom = max(("ms_geom", "ms_bin", "ms_raster"), key=lambda k: kmed[k])
        algo = {"ms_geom": 4 * G * G + 16 * G + 176, "ms_bin": 0, "ms_raster": 4 * W * H + 1024 + 176}
        dom_bytes = algo[dom] if algo[dom] else ALGO_BYTES_FRAME
        achieved = dom_bytes / (kmed[dom] * 1e-3) / 1e9
        # whole frame: from the timed steps themselves (the per-kernel events of kmed add a few microseconds per frame and keep
        # the next frame's geometry pass from starting under the shading tail)
        frame_ms = dev_ms / (args.steps * F)
        frame_achieved = ALGO_BYTES_FRAME / (frame_ms * 1e-3) / 1e9
        # Second reading of the same kernel: it is issue-bound, not HBM-bound.  Warp
        # instructions per launch come from the committed ncu capture, the time is the live one above.
        issue, winst = None, ncu_traffic({"ms_geom": "k_geom", "ms_raster": "k_raster"}.get(dom, "") + "_warp_inst")
        if winst and clocks and clocks.get("sm_mhz"):
            slots = 148 * 4 * clocks["sm_mhz"] * 1e6 * kmed[dom] * 1e-3          # one warp instruction per SM sub-partition per clock
            issue = {"warp_inst_per_launch": winst, "issue_slots": slots, "frac": winst / slots, "source": "warp instructions: committed ncu capture (profiles/dram_traffic.json); time and clock: this run"}
        cpu = None
        if world == 1:                                     # reported at N=1 only
            us = [np.asarray(blocks[i]) for i in range(min(F, 4))]
            cfps, cn, cores, cdt = cpu_reference_fps(args.cpu_seconds, dem, us, float(hmin), float(hmax))
            cpu = {"value": cfps, "unit": "frames/s", "cores": cores, "kind": "port",
                   "sample": f"{cn} frames of the same workload in {cdt:.1f} s; oracle/ (C + OpenMP, all host threads)"}
        line = {
            "metric": "terrain frames/s (1080p, 4096^2 DEM)", "value": fps, "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(F), "host_binding": numa,
            "mpix_per_s": fps * W * H / 1e6,
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": int(dem.nbytes + F * 176), "d2h_bytes_per_step": int(F * W * H * 4),
                    "mpix_per_s": e2e_fps * W * H / 1e6},
            "e2e_png": {"value": png_fps, "unit": "frames/s", "steps": png_steps, "h2d_bytes_per_step": int(dem.nbytes + F * 176),
                        "d2h_bytes_per_step": int(png_bytes // png_steps),
                        "note": "same steps, frames delivered as .png files in pinned host memory, encoded on the device; the DEM of step k+1 uploads while step k renders (Scene.render_batch_png_raw(cameras, next_height=...))"},
            "gpu_launches": args.steps * F * 3,   # k_geom_march + k_resolve_lean + k_shade_list per frame (value leg); k_binsort / k_raster only for frames that queue large or clipped triangles
            "parity": parity,
            "roofline": {"bound": "hbm", "limited_by": "instruction issue (see issue)", "kernel": {"ms_geom": "k_geom_march", "ms_bin": "k_binsort", "ms_raster": "k_resolve_lean + k_shade_list"}[dom],
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic({"ms_geom": "k_geom", "ms_bin": "k_bin", "ms_raster": "k_raster"}[dom]),
                         "traffic_source": "committed ncu --set full capture (profiles/dram_traffic.json), not measured in this run",
                         "peak_source": peak_kind, "algorithmic_bytes_per_launch": dom_bytes, "kernel_ms": kmed[dom],
                         "frame": {"algorithmic_bytes": ALGO_BYTES_FRAME, "ms": frame_ms, "achieved": frame_achieved, "frac": frame_achieved / peak, "source": "timed steps (incl. the L2 flush between them)"},
                         "kernels_ms": kmed, "issue": issue},
            "cpu_baseline": cpu,
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
