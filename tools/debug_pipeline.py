"""Replays 'pipeline' cases the API fuzzer logged (tools/fuzz_parity.py api) under several submit patterns and says which
frames differ from the oracle.  usage: python tools/debug_pipeline.py LOG [max_cases]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vulkan-forge_b200", "python")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import vf_oracle as o
import vulkan_forge as vf
from vulkan_forge._synthetic import synthetic_dem
ext = vf._vulkan_forge
lut = np.fromfile(os.path.join(ROOT, "tests", "golden", "lut_viridis.rgba"), np.uint8)
cases = [json.loads(l[len("API MISMATCH "):]) for l in open(sys.argv[1]) if l.startswith("API MISMATCH")][:int(sys.argv[2]) if len(sys.argv) > 2 else 4]
for d in cases:
    W, H, G, depth, mode = d["W"], d["H"], d["G"], d["depth"], d["mode"]
    cams = [tuple(tuple(x) if isinstance(x, list) else x for x in c) for c in d["cams"]]
    dems = [synthetic_dem(sd, G) * np.float32(sc_) for sd, sc_ in d["dems"]]
    def scene():
        sc = ext.Scene(W, H, grid=G, colormap="viridis")
        sc.set_shading(mode, depth=depth); sc.set_height_range(-0.4, 0.5)
        return sc
    ref = scene()
    blocks = ref.camera_blocks(cams)
    orc = [np.stack([o.render_terrain(W, H, G, dm, blocks[i], lut, mode=o.MODE_NORTHSTAR if mode == "northstar" else o.MODE_SHIPPED, depth=depth,
                                      h_min=-0.4, h_max=0.5) for i in range(len(cams))]) for dm in dems]
    def diff(got, which):
        if mode == "northstar":
            return [int((got[i] != orc[which][i]).any(-1).sum()) for i in range(len(cams))]
        return [int((np.abs(got[i].astype(int) - orc[which][i].astype(int)).max(-1) > 1).sum()) for i in range(len(cams))]
    print("case", {k: d[k] for k in ("W", "H", "G", "depth", "mode", "dems")}, "cams", len(cams), "logged", d.get("failed"))
    ref.set_height_from_r32f(dems[0])
    fr = []
    for c in cams:
        ref.set_camera_look_at(*c); fr.append(ref.render_rgba().copy())
    print("  sync single frames vs oracle      ", diff(np.stack(fr), 0), "big", ref.render_metrics()["big_prims"])
    print("  sync batch vs oracle              ", diff(ref.render_batch_rgba(cams), 0))
    sc = scene(); sc.set_height_from_r32f(dems[0])
    print("  sync upload + async batch, collect", diff(sc.collect(sc.render_batch_rgba_async(cams)), 0))
    print("  same again                        ", diff(sc.collect(sc.render_batch_rgba_async(cams)), 0))
    sc = scene(); sc.set_height_from_r32f_async(dems[0])
    print("  async-api upload + async batch    ", diff(sc.collect(sc.render_batch_rgba_async(cams)), 0))
    sc.set_height_from_r32f_async(dems[1])
    print("  second image, async               ", diff(sc.collect(sc.render_batch_rgba_async(cams)), 1))
    sc.set_height_from_r32f_async(dems[0])
    print("  first image again, async          ", diff(sc.collect(sc.render_batch_rgba_async(cams)), 0))
    import torch
    # V1: batch 0 in flight, then ONLY the next upload (no second batch), then collect
    sc = scene(); sc.set_height_from_r32f_async(dems[0]); tk = sc.render_batch_rgba_async(cams); sc.set_height_from_r32f_async(dems[1])
    got = sc.collect(tk); print("  V1 batch 0 + upload of image 1 only  ", diff(got, 0), "vs image 1:", diff(got, 1))
    # V2: batch 0 in flight, then a second batch on the SAME image (no upload)
    sc = scene(); sc.set_height_from_r32f_async(dems[0]); tk = sc.render_batch_rgba_async(cams); tk2 = sc.render_batch_rgba_async(cams)
    print("  V2 two batches in flight, no upload  ", diff(sc.collect(tk), 0), diff(sc.collect(tk2), 0))
    # V3: as V1 with pinned host arrays
    pdems = [torch.from_numpy(dm).pin_memory().numpy() for dm in dems]
    sc = scene(); sc.set_height_from_r32f_async(pdems[0]); tk = sc.render_batch_rgba_async(cams); sc.set_height_from_r32f_async(pdems[1])
    print("  V3 as V1, pinned arrays              ", diff(sc.collect(tk), 0))
    # V4: as V1 but the GPU is idle before the upload is issued
    sc = scene(); sc.set_height_from_r32f_async(dems[0]); tk = sc.render_batch_rgba_async(cams); torch.cuda.synchronize(); sc.set_height_from_r32f_async(dems[1])
    print("  V4 as V1, device idle before upload  ", diff(sc.collect(tk), 0))
    sc = scene(); pend = []
    for k in range(4):
        sc.set_height_from_r32f_async(dems[k & 1]); pend.append((k & 1, sc.render_batch_rgba_async(cams)))
    print("  4 in flight (fuzzer pattern w/o early collect)", [diff(sc.collect(tk), w) for w, tk in pend])
