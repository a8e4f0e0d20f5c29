#!/usr/bin/env python
"""Regenerate the golden dumps from the reference's own implementation (oracle/_ref/oracle_step).

The reference ships no golden vectors for the physics step, so these fixtures ARE outputs of the
reference itself, produced by the compiled oracle on small seeded scenes.  Run from the repo root:
    python tests/golden/make_golden.py
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity  # noqa: E402
from bradot_b200 import scenes  # noqa: E402

# name -> (scene factory, total steps, first dumped step): every step from `first` on is dumped
FIXTURES = {
    "pyramid21": (lambda: scenes.pyramid(21, base=6), 24, 18),
    "mixed27": (lambda: scenes.mixed_pile(3, half=2.5), 64, 58),
    "chains12": (lambda: scenes.convex_chains(n_chains=4, chain_len=3, n_hulls=3), 84, 79),
    "freefall16": (lambda: scenes.free_fall(16), 6, 0),
    # round 2: the joint types and body / shape parameters added then (each scene a handful of bodies)
    "sliders4": (lambda: scenes.slider_rigs(4), 30, 27),
    "sixdof4": (lambda: scenes.sixdof_rigs(4), 30, 27),
    "cones2": (lambda: scenes.hanging_chains(2, 4, cone_twist=True), 30, 27),
    "level2": (lambda: scenes.level_geometry(2), 40, 37),
}


def scene(name):
    return FIXTURES[name][0]()


if __name__ == "__main__":
    for name, (make, steps, first) in FIXTURES.items():
        out = os.path.join(HERE, name + ".bdmp")
        parity.run_oracle(make(), steps, dump_from=first, dump_every=1, keep=out)
        print(name, os.path.getsize(out), "bytes")
