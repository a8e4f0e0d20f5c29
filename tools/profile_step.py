#!/usr/bin/env python
"""Minimal driver for ncu: load a scene, run W untimed steps, then K steps (one context, no second arm)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from bradot_b200 import server  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="c2")
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--warmup", type=int, default=60)
ap.add_argument("--steps", type=int, default=3)
a = ap.parse_args()
sc, label = bench.build_scene(a.scene, a.scale)
ps = server.PhysicsServerCUDA(0)
ps.load_scene(sc)
for _ in range(a.warmup + a.steps):
    ps.step(sc.dt)
st = ps.stats()
print(label, st)
ps.finish()
