"""What fraction of the broadphase pairs would a tight-AABB pre-test of the narrowphase reject?  (settled regimes)"""
import sys
import numpy as np
sys.path.insert(0, ".")
import bench
from bradot_b200 import server
for name, scale in (("c5", 0.05), ("c3", 0.05), ("c2", 0.3)):
    sc, what = bench.build_scene(name, scale)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(bench.SETTLE[name]):
        ps.step(sc.dt)
    ids, pf, cf = ps.pairs_read()
    box = ps.bodies_read(server.FIELD_ITEM_BOX)
    cmin, cmax = box[:, 0:3], box[:, 0:3] + box[:, 3:6]
    smin, smax = box[:, 6:9], box[:, 9:12]
    a, b = ids[:, 0], ids[:, 1]
    finite = np.isfinite(cmin[a]).all(1) & np.isfinite(cmin[b]).all(1) & (np.abs(cmin[a]) < 1e6).all(1) & (np.abs(cmin[b]) < 1e6).all(1)
    disjoint = ((cmin[a] > cmax[b]) | (cmin[b] > cmax[a])).any(1) & finite
    col = (ids[:, 3] & 1) > 0
    has = ids[:, 2] > 0
    print(name, what, "pairs", len(ids), "colliding %.3f" % col.mean(), "with contacts %.3f" % has.mean(), "cache boxes disjoint %.3f" % disjoint.mean(),
          "disjoint among idle (no contacts, not colliding) %.3f" % disjoint[~col & ~has].mean(), flush=True)
    ps.finish()
