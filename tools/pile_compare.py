"""Free-running statistics of a tall sphere pile (16 x 40 x 16 = 10 240 spheres, 700 steps): this backend against the reference
(oracle/_ref/oracle_step), every 100 steps: lowest / mean height, largest and rms speed.  Takes ~4 min of host CPU."""
import sys, time, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import parity
from bradot_b200 import scenes, server
sc = scenes.spheres_container(16, 40, 16, half=4.6)
steps = 700
t0 = time.time()
_, dump = parity.run_oracle(sc, steps, dump_from=100, dump_every=100)
print("oracle %.1f s" % (time.time() - t0), flush=True)
ps = server.PhysicsServerCUDA(0); ps.load_scene(sc)
dyn = sc.bodies["mode"] >= scenes.MODE_RIGID
for s in range(1, steps + 1):
    ps.step(sc.dt)
    if s % 100 == 0:
        xf = ps.bodies_read(0); v = ps.bodies_read(1)
        r = dump[("bodies", s)]
        print(s, "gpu min y %.3f mean y %.3f max|v| %.3f rms v %.4f | ref min y %.3f mean y %.3f max|v| %.3f rms v %.4f" % (
            xf[dyn, 10].min(), xf[dyn, 10].mean(), np.abs(v[dyn]).max(), np.sqrt((v[dyn] ** 2).mean()),
            r[dyn, 10].min(), r[dyn, 10].mean(), np.abs(r[dyn, 12:15]).max(), np.sqrt((r[dyn, 12:15] ** 2).mean())), flush=True)
