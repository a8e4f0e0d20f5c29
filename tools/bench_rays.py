#!/usr/bin/env python
"""Ray-query throughput through the C ABI (host buffers in and out): n rays against a settled C2 scene."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from bradot_b200 import server  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="c2")
ap.add_argument("--warmup", type=int, default=200)
ap.add_argument("--rays", type=int, default=1000000)
a = ap.parse_args()
sc, label = bench.build_scene(a.scene, 1.0)
ps = server.PhysicsServerCUDA(0)
ps.load_scene(sc)
for _ in range(a.warmup):
    ps.step(sc.dt)
ps.sync()
rng = np.random.default_rng(1)
frm = rng.uniform(-14, 14, (a.rays, 3))
frm[:, 1] = rng.uniform(0.1, 25, a.rays)
to = frm + rng.normal(size=(a.rays, 3)) * 10
rays = np.concatenate([frm, to], 1).astype(np.float32)
ps.intersect_rays(rays[:1000])
best = 1e9
for _ in range(3):
    t0 = time.perf_counter()
    hit, body, pos, nrm = ps.intersect_rays(rays)
    best = min(best, time.perf_counter() - t0)
print("%s: %d rays in %.2f ms = %.1f M rays/s (host in/out, tree rebuild included), %.1f %% hit" %
      (label, a.rays, best * 1e3, a.rays / best / 1e6, 100.0 * hit.mean()))
ps.finish()
