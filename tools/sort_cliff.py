#!/usr/bin/env python
"""Where is the cliff of the per-tile sort?  (missing #5.)  Times gsb_isect_emit through ops.isect_tiles for
maps with k keys per tile, k = 128 ... 16384 (640x480: 1200 tiles), with the per-tile sort (method "tile": registers up to
1024 keys, shared memory up to 4096, in place in global memory beyond) and with the global stable radix sort (method
"radix": cub::DeviceRadixSort, the reference's shape), and checks that both give the same order."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from gsplatloc_b200 import ops

dev = torch.device("cuda", 0)
W, H, ts = 640, 480, 16
tw, th = W // ts, H // ts
g = torch.Generator(device=dev).manual_seed(1)
for k in (128, 256, 512, 1024, 2048, 4096, 8192, 16384):
    N = k * tw * th
    # centres strictly inside a tile, radius 1: every Gaussian touches exactly one tile
    tx = torch.randint(0, tw, (N,), device=dev, generator=g); ty = torch.randint(0, th, (N,), device=dev, generator=g)
    off = 3.0 + 10.0 * torch.rand((N, 2), device=dev, generator=g)
    m2 = (torch.stack([tx, ty], 1).float() * ts + off)[None].contiguous()
    radii = torch.ones((1, N), dtype=torch.int32, device=dev)
    depths = (1.0 + 9.0 * torch.rand((1, N), device=dev, generator=g)).contiguous()
    out = {}
    res = {}
    for method in ("tile", "radix"):
        for _ in range(2):
            r = ops.isect_tiles(m2, radii, depths, ts, tw, th, method=method)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            r = ops.isect_tiles(m2, radii, depths, ts, tw, th, method=method)
        e1.record(); torch.cuda.synchronize()
        out[method] = e0.elapsed_time(e1) / 5
        res[method] = r
    same = bool(torch.equal(res["tile"][2], res["radix"][2]) and torch.equal(res["tile"][1], res["radix"][1]))
    cnt = torch.bincount((ty * tw + tx), minlength=tw * th)
    print(json.dumps({"keys_per_tile_mean": k, "keys_per_tile_max": int(cnt.max()), "M": N, "tile_ms": round(out["tile"], 4),
                      "radix_ms": round(out["radix"], 4), "identical": same}), flush=True)
