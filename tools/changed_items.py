"""How many stored broadphase boxes change per step in the settled regimes bench.py times?  (decides whether an incremental
pair update pays)"""
import sys
import numpy as np
sys.path.insert(0, ".")
import bench
from bradot_b200 import server
for name, scale in (("c5", 0.25), ("c2", 1.0), ("c3", 1.0)):
    sc, what = bench.build_scene(name, scale)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(bench.SETTLE[name]):
        ps.step(sc.dt)
    prev = ps.bodies_read(server.FIELD_ITEM_BOX)[:, 6:12].copy()
    fr = []
    for _ in range(10):
        ps.step(sc.dt)
        cur = ps.bodies_read(server.FIELD_ITEM_BOX)[:, 6:12]
        fr.append(float(np.any(cur != prev, axis=1).mean()))
        prev = cur.copy()
    st = ps.stats()
    print(name, what, "changed fraction per step:", ["%.4f" % f for f in fr], "active", st["active_objects"], "of", len(sc.bodies), flush=True)
    ps.finish()
