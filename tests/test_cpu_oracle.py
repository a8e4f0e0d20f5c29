"""CPU-side oracle tests (no GPU): the golden dumps load, the compiled reference reproduces them bit for bit
(which pins the oracle build), and a numpy restatement of the two simplest stages agrees with them."""
import os
import sys

import numpy as np
import pytest

import parity
from bradot_b200 import scenes

sys.path.insert(0, os.path.join(parity.ROOT, "tests", "golden"))
sys.path.insert(0, os.path.join(parity.ROOT, "oracle"))
import make_golden  # noqa: E402
import restate  # noqa: E402


def _golden(name):
    return scenes.read_dump(os.path.join(parity.GOLDEN, name + ".bdmp"))


@pytest.mark.parametrize("name", sorted(make_golden.FIXTURES))
def test_golden_fixture_is_complete(name):
    d = _golden(name)
    make, steps, first = make_golden.FIXTURES[name]
    sc = make()
    assert d[("mass_props", -1)].shape == (len(sc.bodies), 16)
    for s in range(max(first, 1), steps + 1):
        assert d[("bodies", s)].shape == (len(sc.bodies), 20)
        assert len(d[("pair_ids", s)]) == len(d[("contacts", s)])
        assert np.all(np.isfinite(d[("bodies", s)]))


@pytest.mark.parametrize("name", sorted(make_golden.FIXTURES))
def test_compiled_reference_reproduces_golden(name):
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")
    make, steps, first = make_golden.FIXTURES[name]
    _, d = parity.run_oracle(make(), steps, dump_from=first, dump_every=1)
    g = _golden(name)
    assert set(d) == set(g)
    for k in g:
        assert np.array_equal(d[k], g[k]), k


def test_numpy_free_flight_restatement_matches_golden():
    d = _golden("freefall16")
    sc = make_golden.scene("freefall16")
    mp = d[("mass_props", -1)]
    b = d[("bodies", 0)]
    for i in range(len(sc.bodies)):
        xf, lin, ang = b[i, 0:12], b[i, 12:15], b[i, 15:18]
        for s in range(1, 7):
            xf, lin, ang = restate.free_flight_step(xf, lin, ang, mp[i], sc.dt)
            ref = d[("bodies", s)][i]
            assert np.max(np.abs(xf - ref[0:12])) < 2e-5, (i, s)
            assert np.max(np.abs(lin - ref[12:15])) < 2e-5 and np.max(np.abs(ang - ref[15:18])) < 2e-5, (i, s)


@pytest.mark.parametrize("name", ["pyramid21", "mixed27", "chains12"])
def test_pair_set_is_the_overlap_set_of_stored_boxes(name):
    """The claim the GPU broadphase rests on: after update(), pairs == overlapping stored boxes of the previous step."""
    d = _golden(name)
    make, steps, first = make_golden.FIXTURES[name]
    sc = make()
    static = sc.bodies["mode"] == scenes.MODE_STATIC
    for s in range(first + 1, steps + 1):
        it = d[("items", s - 1)]
        want = restate.overlap_pairs(it[:, 6:9], it[:, 9:12], static, sc.bodies["layer"], sc.bodies["mask"])
        ids = d[("pair_ids", s)]
        got = {(int(min(a, b)), int(max(a, b))) for a, b in ids[:, :2]}
        assert got == want, (name, s, len(got), len(want))
