"""GPU parity tests: the CUDA step against the reference's own implementation (the oracle).

Tolerances are BASELINE.json's: pair sets bit-exact, contacts 1e-5 relative, single-step body
state 1e-4 (teacher-forced: the GPU state is overwritten with the oracle's state of step k, one
step is taken with the oracle's sweep order imported, and the result is compared with step k+1).
"""
import numpy as np
import pytest

import parity
from bradot_b200 import scenes, server

pytestmark = pytest.mark.gpu

CONTACT_TOL = 1e-5   # relative, contact points and normals
STATE_TOL = 1e-4     # single-step transforms and velocities


def _need_oracle():
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")


def test_free_fall_integration_matches_reference():
    _need_oracle()
    sc = scenes.free_fall(96)
    _, dump = parity.run_oracle(sc, 30, dump_from=0, dump_every=10)
    ps = parity.new_server(sc, dump)
    for s in range(1, 31):
        ps.step(sc.dt)
        if s % 10 == 0:
            ref = dump[("bodies", s)]
            xf = ps.bodies_read(server.FIELD_TRANSFORM)
            lv = ps.bodies_read(server.FIELD_LINEAR_VELOCITY)
            av = ps.bodies_read(server.FIELD_ANGULAR_VELOCITY)
            assert parity.rel_err(xf, ref[:, 0:12], 1.0) < STATE_TOL
            assert parity.rel_err(lv, ref[:, 12:15], 1.0) < STATE_TOL
            assert parity.rel_err(av, ref[:, 15:18], 1.0) < STATE_TOL
            box = ps.bodies_read(server.FIELD_ITEM_BOX)
            assert parity.rel_err(box[:, 0:6], dump[("items", s)][:, 0:6], 1.0) < 1e-5
    assert ps.stats()["collision_pairs"] == 0


def _teacher_forced(sc, steps, check_steps):
    # step 0 -> 1 is not checked: the reference flushes its pending shape updates at the start of the
    # first step (bradot_physics_server_3d.cpp:1728-1733), which re-stores every broadphase box before pairing
    _, dump = parity.run_oracle(sc, steps, dump_from=0, dump_every=1)
    ps = parity.new_server(sc, dump)
    worst = {"contact": 0.0, "state": 0.0}
    for k in check_steps:
        parity.load_state(ps, dump, k)
        ps.set_solve_order(parity.solve_order(dump, k + 1))
        ps.step(sc.dt)
        st = ps.stats()
        assert st["overflow"] == 0, "colliding pairs missing from the imported sweep order at step %d" % (k + 1)
        ids, pf, cf = ps.pairs_read()
        rids, rpf, rcf = parity.oracle_pairs(dump, k + 1)
        # --- broadphase pair set: bit-exact
        gk, rk = parity.pair_keys(ids), parity.pair_keys(rids)
        order = np.argsort(rk)
        assert np.array_equal(gk, rk[order]), "pair set differs at step %d: gpu %d ref %d" % (k + 1, len(gk), len(rk))
        rids, rpf, rcf = rids[order], rpf[order], rcf[order]
        # --- colliding-pair list: bit-exact (restricted to awake islands: the reference only visits those)
        awake = dump[("bodies", k)][:, 18] > 0
        touched = awake[rids[:, 0]] | awake[rids[:, 1]]
        assert np.array_equal((ids[:, 3] & 1)[touched], (rids[:, 3] & 1)[touched]), "colliding-pair list differs at step %d" % (k + 1)
        assert np.array_equal(ids[touched, 2], rids[touched, 2]), "contact counts differ at step %d" % (k + 1)
        # --- contacts (points in both local frames, normals)
        for q in range(4):
            m = touched & (rids[:, 2] > q)
            worst["contact"] = max(worst["contact"], parity.rel_err(cf[m, q, 0:9], rcf[m, q, 0:9], 1.0))
        # --- body state after the step
        ref = dump[("bodies", k + 1)]
        xf = ps.bodies_read(server.FIELD_TRANSFORM)
        lv = ps.bodies_read(server.FIELD_LINEAR_VELOCITY)
        av = ps.bodies_read(server.FIELD_ANGULAR_VELOCITY)
        e = max(parity.rel_err(xf, ref[:, 0:12], 1.0), parity.rel_err(lv, ref[:, 12:15], 1.0), parity.rel_err(av, ref[:, 15:18], 1.0))
        worst["state"] = max(worst["state"], e)
    ps.finish()
    return worst


def test_teacher_forced_pyramid():
    _need_oracle()
    w = _teacher_forced(scenes.pyramid(55, base=10), 40, [1, 2, 5, 10, 20, 39])
    assert w["contact"] < CONTACT_TOL, w
    assert w["state"] < STATE_TOL, w


def test_teacher_forced_spheres():
    _need_oracle()
    w = _teacher_forced(scenes.spheres_container(6, 6, 6, half=2.0), 80, [1, 10, 30, 50, 60, 79])
    assert w["contact"] < CONTACT_TOL, w
    assert w["state"] < STATE_TOL, w


def test_teacher_forced_mixed_pile():
    _need_oracle()
    w = _teacher_forced(scenes.mixed_pile(5, half=3.5), 90, [1, 20, 40, 60, 75, 89])
    assert w["contact"] < CONTACT_TOL, w
    assert w["state"] < STATE_TOL, w
