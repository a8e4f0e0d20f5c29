"""GPU parity tests of the smaller rows: joint parameters / hinge limits / motors / solver priorities (a15, a16), axis locks
and body damp modes (a2, a7), per-space solver iterations (a16), body_set_state and free() at run time (boundary),
inactive spaces, and the reference's own helper vectors (tests/core/math/test_geometry_3d.h).  The checker is the
reference's own code (oracle/_ref/oracle_step)."""
import os

import numpy as np
import pytest

import parity
from bradot_b200 import scenes, server

pytestmark = pytest.mark.gpu

STATE_TOL = 1e-4
CONTACT_TOL = 1e-5


def _need_oracle():
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")


def test_reference_segment_vectors_on_the_device():
    """tests/core/math/test_geometry_3d.h:39-48 and :133 through the device helpers the capsule routines call."""
    ps = server.PhysicsServerCUDA(0)
    out = ps.test_segment_math(0, [[1, -1, 1, 1, 1, -1, -1, -2, -1, -1, 1, 1]])
    assert np.allclose(out[0, 0:3], (1, -0.2, 0.2), atol=1e-5) and np.allclose(out[0, 3:6], (-1, -0.2, 0.2), atol=1e-5)   # is_equal_approx
    # "Closest Distance Between Segments" == 2.0f exactly: the distance of the closest points
    out = ps.test_segment_math(0, [[1, -2, 0, 1, 2, 0, -1, 2, 0, -1, -2, 0]])
    assert float(np.linalg.norm(out[0, 0:3] - out[0, 3:6])) == 2.0
    out = ps.test_segment_math(1, [[2, 1, 4, 1, 1, 1, 5, 5, 5]])
    assert np.allclose(out[0], (2.33333, 2.33333, 2.33333), atol=1e-5)
    # degenerate inputs: a point against a zero-length segment, parallel segments
    out = ps.test_segment_math(1, [[3, 3, 3, 1, 1, 1, 1, 1, 1]])
    assert np.array_equal(out[0], np.array([1, 1, 1], np.float32))
    out = ps.test_segment_math(0, [[0, 0, 0, 1, 0, 0, 0, 1, 0, 1, 1, 0]])
    assert np.all(np.isfinite(out)) and abs(float(np.linalg.norm(out[0, 0:3] - out[0, 3:6])) - 1.0) < 1e-6
    ps.finish()


def test_joint_parameters_limits_motors_and_priorities():
    """Hinge limit (accumulated, clamped impulse) and motor, pin bias / damping / impulse clamp, extra sweeps for solver
    priority 2: teacher-forced with the reference's sweep order, body state after the step within 1e-4 (measured: bit-equal)."""
    _need_oracle()
    sc = scenes.hanging_chains()
    _, dump = parity.run_oracle(sc, 80)
    # the limits do engage: without them a hinge with +-2 rad/s would leave [-0.35, 0.5] within the first second
    w = parity.teacher_forced(sc, 80, [1, 5, 15, 30, 50, 79], dump=dump)
    assert w["state"] < STATE_TOL, w
    # and the parameters matter: the same scene with default joints departs from the reference
    plain = scenes.hanging_chains()
    plain.joints["flags"] &= scenes.JOINT_NO_COLLIDE
    plain.joints["priority"] = 0
    ps = parity.new_server(plain, dump)
    parity.load_state(ps, dump, 30)
    ps.set_solve_order(parity.solve_order(dump, 31))
    ps.step(sc.dt)
    av = ps.bodies_read(server.FIELD_ANGULAR_VELOCITY)
    assert parity.rel_err(av, dump[("bodies", 31)][:, 15:18], 1.0) > 1e-2
    ps.finish()


def test_joint_parameters_can_change_and_joints_can_go_at_run_time():
    _need_oracle()
    sc = scenes.hanging_chains(3, 4)
    plain = scenes.hanging_chains(3, 4)
    plain.joints["flags"] &= scenes.JOINT_NO_COLLIDE
    plain.joints["priority"] = 0
    _, dump = parity.run_oracle(sc, 20)
    ps = parity.new_server(plain, dump)
    ps.joints_write(0, sc.joints)              # hinge_joint_set_param / set_flag / joint_set_solver_priority after creation
    parity.load_state(ps, dump, 10)
    ps.set_solve_order(parity.solve_order(dump, 11))
    ps.step(sc.dt)
    lv = ps.bodies_read(server.FIELD_LINEAR_VELOCITY)
    assert parity.rel_err(lv, dump[("bodies", 11)][:, 12:15], 1.0) < STATE_TOL
    # a destroyed joint stops acting: the lowest link of chain 0 falls freely
    ps.set_solve_order(np.zeros((0, 3), np.int32))
    last = 3                                    # joint 3 links bodies 3 and 4 of chain 0 (anchor = body 0)
    ps.joints_destroy(last)
    y0 = ps.bodies_read(server.FIELD_TRANSFORM)[4, 10]
    for _ in range(30):
        ps.step(sc.dt)
    assert ps.bodies_read(server.FIELD_TRANSFORM)[4, 10] < y0 - 0.5
    ps.finish()


def test_axis_locks_and_damp_modes_in_free_flight():
    """body_set_axis_lock (bradot_body_3d.cpp:685-699) and BODY_DAMP_MODE_REPLACE (:586-606): 40 free-running steps."""
    _need_oracle()
    sc = scenes.locks_and_damping(64)
    _, dump = parity.run_oracle(sc, 40, dump_from=0, dump_every=10)
    ps = parity.new_server(sc, dump)
    for s in range(1, 41):
        ps.step(sc.dt)
        if s % 10 == 0:
            ref = dump[("bodies", s)]
            st = ps.bodies_read_state()
            assert parity.rel_err(st[:, 0:12], ref[:, 0:12], 1.0) < STATE_TOL
            assert parity.rel_err(st[:, 12:18], ref[:, 12:18], 1.0) < STATE_TOL
    # the locks did something: a locked linear axis kept its velocity component at zero
    locks = (sc.bodies["flags"] >> scenes.BODY_LOCK_SHIFT) & 63
    lv = ps.bodies_read(server.FIELD_LINEAR_VELOCITY)
    for ax in range(3):
        sel = (locks & (1 << ax)) != 0
        assert sel.any() and np.all(lv[sel, ax] == 0.0)
    ps.finish()


def test_axis_locks_and_damp_modes_with_contacts():
    _need_oracle()
    sc = scenes.locks_and_damping(48, on_plane=True)
    w = parity.teacher_forced(sc, 60, [1, 15, 30, 45, 59])
    assert w["colliding"] > 0 and w["contact"] < CONTACT_TOL and w["state"] < STATE_TOL, w


def test_each_space_sweeps_its_own_number_of_iterations():
    """SPACE_PARAM_SOLVER_ITERATIONS is per space (bradot_space_3d.h solver_iterations): 4 spaces with 4, 9, 14, 19 sweeps
    in the same launches, sweep order imported space by space."""
    _need_oracle()
    sc = scenes.batched_worlds(4, lattice=(3, 3, 3))
    sc.iterations, sc.iter_step = 4, 5
    w = parity.teacher_forced(sc, 70, [30, 45, 55, 69])
    assert w["colliding"] > 0 and w["contact"] < CONTACT_TOL and w["state"] < STATE_TOL, w
    # the counts matter: with 19 sweeps everywhere space 0 ends a step with other velocities
    _, dump = parity.run_oracle(sc, 56, dump_from=55)
    same = scenes.batched_worlds(4, lattice=(3, 3, 3))
    same.iterations = 19
    ps = parity.new_server(same, dump)
    parity.load_state(ps, dump, 55)
    ps.set_solve_order(parity.solve_order(dump, 56))
    ps.step(sc.dt)
    lv = ps.bodies_read(server.FIELD_LINEAR_VELOCITY)
    sp0 = sc.bodies["space"] == 0
    assert parity.rel_err(lv[sp0], dump[("bodies", 56)][sp0, 12:15], 1.0) > 1e-4
    ps.finish()


def _apply_events(ps, events, step, freed):
    for (s, body, kind, v) in events:
        if s != step or body in freed:
            continue
        if kind == 0:
            ps.body_set_state([body], server.BODY_STATE_TRANSFORM, v[:12])
        elif kind == 1:
            ps.body_set_state([body], server.BODY_STATE_LINEAR_VELOCITY, v[:3])
        elif kind == 2:
            ps.body_set_state([body], server.BODY_STATE_ANGULAR_VELOCITY, v[:3])
        elif kind == 3:
            ps.bodies_destroy([body])
            freed.add(body)
        elif kind == 4:
            ps.body_set_state([body], server.BODY_STATE_SLEEPING, v[:1])
        elif kind == 5:
            ps.body_set_state([body], server.BODY_STATE_CAN_SLEEP, v[:1])


def test_disabled_shapes_follow_the_reference():
    """body_set_shape_disabled (bradot_collision_object_3d.cpp:69-96; update_mass_properties skips the shape, bradot_body_3d.cpp:63-65):
    every third body of a settling pile has its shape disabled -- it falls through the others and through the floor, pairs
    with nothing, is seen by no ray.  Free-running: neither pair list holds such a body, positions of the disabled bodies (pure free fall) agree to 1e-4, the pile's to 2 mm early on."""
    _need_oracle()
    sc = scenes.mixed_pile(4, half=3.0)
    dyn = np.nonzero(sc.bodies["mode"] == scenes.MODE_RIGID)[0]
    off = dyn[::3]
    sc.bodies["flags"][off] |= scenes.BODY_SHAPE_DISABLED
    steps = 60
    rays = np.zeros((len(off), 6), np.float32)
    _, dump = parity.run_oracle(sc, steps, dump_from=0, dump_every=1)
    ps = parity.new_server(sc, dump)
    for k in range(1, steps + 1):
        ps.step(sc.dt)
        st = ps.bodies_read_state()
        ref = dump[("bodies", k)]
        assert np.max(np.abs(st[off, 0:18] - ref[off, 0:18])) < 1e-4, k
        if k <= 25:
            assert np.max(np.abs(st[:, 9:12] - ref[:, 9:12])) < 2e-3, k
        if k in (10, 25):       # (not the first step: the reference re-stores every box before its first pairing, DESIGN.md known deviations)
            ids, _, _ = ps.pairs_read()
            rids, _, _ = parity.oracle_pairs(dump, k)
            assert not np.isin(ids[:, 0:2], off).any() and not np.isin(rids[:, 0:2], off).any()
            # (the lists themselves are not compared in a free run: speculative box-only pairs may appear a step apart)
            assert abs(len(ids) - len(rids)) <= len(rids) // 10
    assert np.all(st[off, 10] < -0.4) and np.mean(st[np.setdiff1d(dyn, off), 10] > 0.0) > 0.9      # through the floor / on it
    rays[:, 0:3] = st[off, 9:12] + np.array([0, 1.0, 0], np.float32)
    rays[:, 3:6] = st[off, 9:12] - np.array([0, 1.0, 0], np.float32)
    hit, body, _, _ = ps.intersect_rays(rays)
    assert not np.isin(body[hit], off).any()
    mp = ps.bodies_read(server.FIELD_MASS_PROPS)
    assert np.allclose(mp[off, 1:4], 1.0) and np.allclose(mp[off, 4:7], 0.0)          # a body without shapes: unit inertia, centre of mass at the origin
    ps.finish()


def test_custom_inertia_and_centre_of_mass_follow_the_reference():
    """body_set_param(BODY_PARAM_INERTIA / BODY_PARAM_CENTER_OF_MASS) and body_reset_mass_properties (bradot_body_3d.cpp:168-172,
    209-229, update_mass_properties :56-150): tumbling bodies dropped on a plane get, at step 2, a full custom inertia, a
    partial one with a custom centre of mass, or a centre of mass alone (each followed by a mass write, which is when the
    reference recomputes the tensor about the new centre); at step 45 some are reset.  Free-running, every step: active flags
    equal, positions to 2 mm; and the same scene without the script ends up elsewhere."""
    _need_oracle()
    sc = scenes.locks_and_damping(24, on_plane=True)
    sc.bodies["flags"] &= ~np.uint32(63 << scenes.BODY_LOCK_SHIFT)        # no axis locks here: the bodies must be free to turn
    dyn = np.nonzero(sc.bodies["mode"] == scenes.MODE_RIGID)[0]
    events = []

    def ev(step, body, kind, v=()):
        row = np.zeros(12, np.float32)
        row[:len(v)] = v
        events.append((step, int(body), kind, row))
    for k, b in enumerate(dyn):
        if k % 4 == 0:
            ev(2, b, 8, (0.02 + 0.01 * (k % 3), 0.05, 0.03))
        elif k % 4 == 1:
            ev(2, b, 8, (0.06, 0.0, 0.0))
            ev(2, b, 9, (0.05, -0.1, 0.0))
        elif k % 4 == 2:
            ev(2, b, 9, (0.0, -0.15, 0.05))
        if k % 4 != 3:
            ev(2, b, 11, (float(sc.bodies["mass"][b]),))
        if k % 4 in (0, 1) and k % 8 < 4:
            ev(45, b, 10)
    steps = 90
    _, dump = parity.run_oracle(sc, steps, events=events)
    _, plain = parity.run_oracle(sc, steps, dump_steps=[steps])
    assert np.max(np.abs(dump[("bodies", steps)][:, 0:12] - plain[("bodies", steps)][:, 0:12])) > 0.05
    ps = parity.new_server(sc, dump)
    for s in range(1, steps + 1):
        touched = []
        for (st_, body, kind, v) in events:
            if st_ != s:
                continue
            if kind == 8:
                ps.bodies_set_mass_override([body], inertia=v[:3])
            elif kind == 9:
                ps.bodies_set_mass_override([body], center_of_mass=v[:3])
            elif kind == 10:
                ps.bodies_set_mass_override([body])
            touched.append(body)
        if touched:
            touched = sorted(set(touched))
            ps.bodies_update(touched, sc.bodies[touched])
        ps.step(sc.dt)
        ref = dump[("bodies", s)]
        st = ps.bodies_read_state()
        assert np.array_equal(st[:, 18], ref[:, 18]), "active flags differ at step %d" % s
        assert np.max(np.abs(st[:, 9:12] - ref[:, 9:12])) < 2e-3, "positions differ at step %d: %g" % (s, np.max(np.abs(st[:, 9:12] - ref[:, 9:12])))
    mp = ps.bodies_read(server.FIELD_MASS_PROPS)
    b0, b2 = int(dyn[4]), int(dyn[2])        # k = 4: full custom inertia, never reset; k = 2: centre of mass alone
    assert np.allclose(1.0 / mp[b0, 1:4], (0.03, 0.05, 0.03), rtol=1e-6) and np.allclose(mp[b0, 7:16], np.eye(3).reshape(-1))
    assert np.allclose(mp[b2, 4:7], (0.0, -0.15, 0.05))
    ps.finish()


def test_body_set_state_and_free_follow_the_reference():
    """BradotBody3D::set_state (bradot_body_3d.cpp:352-417) and free(RID) (bradot_physics_server_3d.cpp:1572-1590) at run
    time: a teleported sleeping body wakes, is orthonormalised and gets a new broadphase box; a moved static body wakes what
    rests on it; velocities wake; SLEEPING clears velocities; CAN_SLEEP false wakes; freed bodies leave.  Free-running from
    the same start, the active flag of every body equals the reference's after every step, positions agree to 2 mm."""
    _need_oracle()
    sc = scenes.scripted_events()
    events = scenes.scripted_events_script()
    steps = 160
    _, dump = parity.run_oracle(sc, steps, events=events)
    ps = parity.new_server(sc, dump)
    freed = set()
    for s in range(1, steps + 1):
        _apply_events(ps, events, s, freed)
        ps.step(sc.dt)
        ref = dump[("bodies", s)]
        st = ps.bodies_read_state()
        alive = np.array([i not in freed for i in range(len(sc.bodies))])
        assert np.array_equal(st[alive, 18], ref[alive, 18]), "active flags differ at step %d: %r vs %r" % (s, st[:, 18], ref[:, 18])
        assert np.max(np.abs(st[alive, 9:12] - ref[alive, 9:12])) < 2e-3, "positions differ at step %d" % s
        # (the pair COUNT is not compared in a free run: speculative box-only pairs may appear a step or two apart, because the
        # reference re-stores a moving item's box also when it pokes out of its BVH leaf's bound -- DESIGN.md, known deviations)
    # the script did what it says
    assert dump[("bodies", 70)][3, 18] == 1 and dump[("bodies", 69)][3, 18] == 0        # teleport woke body 3
    assert dump[("bodies", 80)][8, 18] == 1 and dump[("bodies", 79)][8, 18] == 0        # the moved platform woke its boxes
    ps.finish()


def test_freed_slots_are_reused_and_stay_consistent():
    sc = scenes.scripted_events()
    ps = parity.new_server(sc)
    for _ in range(30):
        ps.step(sc.dt)
    n0 = ps.body_count()
    ps.bodies_destroy([4, 6])
    ps.step(sc.dt)
    ids, _, _ = ps.pairs_read()
    assert not np.isin(ids[:, :2], [4, 6]).any()                  # their pairs are gone with the next step
    rec = sc.bodies[[4, 6]].copy()
    rec["xf"][:, 10] = 3.0
    slots = ps.bodies_create_slots(rec)
    assert sorted(slots.tolist()) == [4, 6] and ps.body_count() == n0
    extra = ps.bodies_create_slots(rec[:1])
    assert extra[0] == n0 and ps.body_count() == n0 + 1          # no free slot left: appended
    for _ in range(90):
        ps.step(sc.dt)
    y = ps.bodies_read(server.FIELD_TRANSFORM)[:, 10]
    assert abs(y[4] - 0.25) < 0.02 and abs(y[6] - 0.25) < 0.02   # fell and came to rest on the ground
    ps.finish()


def test_inactive_space_is_not_stepped_and_resumes_unchanged():
    sc = scenes.batched_worlds(3, lattice=(3, 3, 3))
    ps = parity.new_server(sc)
    twin = parity.new_server(sc)
    for _ in range(50):
        ps.step(sc.dt)
        twin.step(sc.dt)
    ps.space_set_active(1, False)
    frozen = ps.bodies_read_state()
    sp1 = sc.bodies["space"] == 1
    for _ in range(20):
        ps.step(sc.dt)
    now = ps.bodies_read_state()
    assert np.array_equal(now[sp1], frozen[sp1])                  # space 1 did not move
    assert not np.array_equal(now[~sp1], frozen[~sp1])            # the others did
    assert ps.stats()["active_objects"] <= int((~sp1 & (sc.bodies["mode"] >= 2)).sum())
    ps.space_set_active(1, True)
    ps.step(sc.dt)
    twin.step(sc.dt)
    # space 1 takes its 51st step 20 steps later than its twin, from the same state with the same cached contacts; only the
    # sweep order may differ (it depends on the pair counts of the other spaces)
    a = ps.bodies_read_state()[sp1]
    b = twin.bodies_read_state()[sp1]
    assert not np.array_equal(a, frozen[sp1])
    dpos, dvel = float(np.max(np.abs(a[:, 0:12] - b[:, 0:12]))), float(np.max(np.abs(a[:, 12:18] - b[:, 12:18])))
    assert dpos < 3e-3 and dvel < 0.1, (dpos, dvel, np.argmax(np.max(np.abs(a[:, 0:18] - b[:, 0:18]), axis=1)), a[:, 18].sum(), b[:, 18].sum())
    ps.finish()
    twin.finish()


def test_body_update_after_upload_changes_mode_layers_and_parameters():
    sc = scenes.scripted_events()
    ps = parity.new_server(sc)
    for _ in range(40):
        ps.step(sc.dt)
    # body_set_mode(STATIC) on a resting box, body_set_collision_mask(0) on another: it falls through the ground
    rec = sc.bodies[[2, 3]].copy()
    rec["mode"][0] = scenes.MODE_STATIC
    rec["mask"][1] = 0
    rec["layer"][1] = 0
    ps.bodies_update([2, 3], rec)
    ps.body_set_state([3], server.BODY_STATE_SLEEPING, [0.0])
    y3 = ps.bodies_read(server.FIELD_TRANSFORM)[3, 10]
    for _ in range(30):
        ps.step(sc.dt)
    st = ps.bodies_read_state()
    assert st[2, 18] == 0 and abs(st[2, 10] - 0.25) < 0.02
    assert st[3, 10] < y3 - 0.5
    # body_set_param(MASS) keeps the body where it is and changes its inverse mass
    rec = sc.bodies[[5]].copy()
    rec["mass"] = 4.0
    ps.bodies_update([5], rec)
    mp = ps.bodies_read(server.FIELD_MASS_PROPS, 5, 1)
    assert abs(mp[0, 0] - 0.25) < 1e-7
    ps.finish()


def test_contact_reporting_matches_the_reference():
    """body_set_max_contacts_reported + the contacts BradotBody3D::add_contact keeps (bradot_body_pair_3d.cpp:417-428,
    bradot_body_3d.h:353-395): teacher-forced, the SET of contacts reported to every body equals the reference's -- also
    where a body touches more contacts than it asked for (the N deepest stay) -- with positions, normals and depths to 1e-5,
    velocities and impulses to 1e-4.  A STATIC body never clears its list in the reference (integrate_forces, which resets
    contact_count, returns at its first line for it: bradot_body_3d.cpp:478-480,671): it keeps the N deepest contacts it has
    EVER had, and a sleeping body keeps those of its last waking step (it is not in the active list integrate_forces runs
    over).  Both need the history, so they are checked free-running from step 0 instead."""
    _need_oracle()
    sc = scenes.pyramid(36, base=8)
    flags = sc.bodies["flags"].copy()
    flags[0] |= 5 << scenes.BODY_REPORT_SHIFT                      # the ground plane: 8 boxes rest on it, it keeps 5 contacts
    flags[1:] |= np.where(np.arange(len(flags) - 1) % 2 == 0, 6, 0).astype(np.uint32) << scenes.BODY_REPORT_SHIFT
    flags[10] = (flags[10] & 0xFF00FFFF) | (2 << scenes.BODY_REPORT_SHIFT)   # a box in the wall that keeps only 2 of its contacts
    sc.bodies["flags"] = flags
    steps, checks = 60, [5, 20, 40, 59]
    _, dump = parity.run_oracle(sc, steps, dump_steps=checks + [k + 1 for k in checks] + [2])
    ps = parity.new_server(sc, dump)
    seen = 0
    for k in checks:
        parity.load_state(ps, dump, k)
        ps.set_solve_order(parity.solve_order(dump, k + 1))
        ps.step(sc.dt)
        counts, offsets, rec = ps.contacts_read()
        ref = dump.get(("reports", k + 1), np.zeros((0, 21), np.float32))
        for body in np.nonzero((sc.bodies["flags"] >> scenes.BODY_REPORT_SHIFT) & 0xFF)[0]:
            if sc.bodies["mode"][body] == scenes.MODE_STATIC or dump[("bodies", k)][body, 18] == 0:
                continue                           # static or asleep: keeps older contacts (checked free-running below)
            mine = rec[offsets[body]:offsets[body] + counts[body]]
            theirs = ref[ref[:, 0] == body]
            assert len(mine) == len(theirs), "body %d reports %d contacts, the reference %d (step %d)" % (body, len(mine), len(theirs), k + 1)
            if not len(mine):
                continue
            cap = int((sc.bodies["flags"][body] >> scenes.BODY_REPORT_SHIFT) & 0xFF)
            assert len(mine) <= cap
            # same set: order both by (collider, position)
            o1 = np.lexsort((mine["local_pos"][:, 2], mine["local_pos"][:, 0], mine["collider"]))
            o2 = np.lexsort((theirs[:, 3], theirs[:, 1], theirs[:, 14]))
            m, t = mine[o1], theirs[o2]
            assert np.array_equal(m["collider"], t[:, 14].astype(np.int32)), (body, m["collider"], t[:, 14])
            assert parity.rel_err(m["local_pos"], t[:, 1:4], 1.0) < CONTACT_TOL and parity.rel_err(m["local_normal"], t[:, 4:7], 1.0) < CONTACT_TOL
            assert parity.rel_err(m["depth"], t[:, 7], 1.0) < CONTACT_TOL and parity.rel_err(m["collider_pos"], t[:, 11:14], 1.0) < CONTACT_TOL
            assert parity.rel_err(m["local_velocity"], t[:, 8:11], 1.0) < 1e-4 and parity.rel_err(m["collider_velocity"], t[:, 15:18], 1.0) < 1e-4
            assert parity.rel_err(m["impulse"], t[:, 18:21], 1.0) < 1e-4
            seen += len(mine)
    assert seen > 50
    # one body through the single-body call
    c = ps.body_contacts(1, 16)
    assert len(c) == counts[1] > 0 and np.array_equal(np.sort(c["depth"]), np.sort(rec[offsets[1]:offsets[1] + counts[1]]["depth"]))
    ps.finish()
    # the ground, free-running from the start: the 5 deepest contacts it has had since step 1
    ps = parity.new_server(sc)
    for k in range(1, 41):
        ps.step(sc.dt)
        if k in (2, 21, 40):
            ref = dump[("reports", k)]
            theirs = ref[ref[:, 0] == 0]
            mine = ps.body_contacts(0, 16)
            assert len(mine) == len(theirs) == 5
            # (the eight base boxes land alike to the last bits, so WHICH of them holds a place is a tie; the depths are not)
            # free-running, this pyramid is 1e-5 from the reference at step 3 and 2 mm at step 20 (sweep order)
            assert np.max(np.abs(np.sort(mine["depth"]) - np.sort(theirs[:, 7]))) < (1e-5 if k == 2 else 1e-3)
            assert set(mine["collider"]) <= set(range(1, 9)) and np.allclose(mine["local_normal"], theirs[:, 4:7], atol=1e-3)
    # by now the pile sleeps in both; box 1 still reports the contacts it had when it fell asleep
    for k in range(41, 60):
        ps.step(sc.dt)
    ref, st = dump[("reports", 59)], ps.bodies_read_state()
    assert st[1, 18] == 0 and dump[("bodies", 59)][1, 18] == 0
    theirs, mine = ref[ref[:, 0] == 1], ps.body_contacts(1, 16)
    assert len(mine) == len(theirs) > 0 and np.array_equal(np.sort(mine["collider"]), np.sort(theirs[:, 14].astype(np.int32)))
    ps.finish()


def _apply_area_events(ps, sc, areas, events, step, gone_bodies, gone_areas):
    for (s, idx, kind, v) in events:
        if s != step:
            continue
        if kind == 6 and idx not in gone_areas:
            areas[idx]["xf"] = v[:12]
            ps.areas_write(idx, areas[idx:idx + 1])
        elif kind == 7 and idx not in gone_areas:
            ps.areas_destroy([idx])
            gone_areas.add(idx)
    _apply_events(ps, [e for e in events if e[2] not in (6, 7)], step, gone_bodies)


def test_areas_overrides_and_monitoring_follow_the_reference():
    """Areas (bradot_area_pair_3d.cpp:36-118, the area loop of integrate_forces bradot_body_3d.cpp:498-585,
    compute_gravity bradot_area_3d.cpp:312-332, call_queries :232-310): scenes.area_zones() free-running beside the reference
    running the same script; BradotArea2Pair3D (:124-206) for the areas that watch areas.  No body ever touches another, so the two runs differ by rounding only: after EVERY step the
    enter / exit events told to the monitor callbacks are the same set, and positions and velocities agree to 1e-4."""
    _need_oracle()
    sc = scenes.area_zones()
    events = scenes.area_zones_script()
    steps = 130
    _, dump = parity.run_oracle(sc, steps, events=events, dump_from=0, dump_every=1)
    ps = parity.new_server(sc)
    areas = sc.areas.copy()
    kin = np.nonzero(sc.bodies["mode"] == scenes.MODE_KINEMATIC)[0]
    gone_b, gone_a = set(), set()
    seen = {0: 0, 1: 0, 2: 0, 3: 0}
    for k in range(1, steps + 1):
        _apply_area_events(ps, sc, areas, events, k, gone_b, gone_a)
        for row in dump.get(("kin_targets", k), ()):
            ps.bodies_write(server.FIELD_KINEMATIC_TARGET, row[1:13].reshape(1, 12), first=int(row[0]))
        ps.step(sc.dt)
        mine = {tuple(int(x) for x in e) for e in ps.area_events()}
        ref = {tuple(int(x) for x in e) for e in dump.get(("area_events", k), np.zeros((0, 3), np.int32))}
        assert mine == ref, (k, sorted(mine - ref), sorted(ref - mine))
        for e in mine:
            seen[e[2]] += 1
        st = ps.bodies_read_state()
        refb = dump[("bodies", k)]
        alive = np.array([i not in gone_b for i in range(len(sc.bodies))])
        assert np.max(np.abs(st[alive, 0:18] - refb[alive, 0:18])) < 1e-4, (k, int(np.argmax(np.abs(st[alive, 0:18] - refb[alive, 0:18]).max(axis=1))))
        assert np.array_equal(st[alive, 18] != 0, refb[alive, 18] != 0), k
        if k == 45:      # the zone was moved onto the sleeping box at step 40: attached to it, and the box sleeps on
            inside, attached = ps.body_areas(13)
            assert list(inside) == [9] and attached.all() and st[13, 18] == 0
    assert seen[0] >= 10 and seen[1] >= 8 and seen[2] >= 7 and seen[3] >= 6, seen     # 2 / 3: an AREA entered / left a watching area
    # what the override did: the ball in the well was pulled off its vertical line, the box under the moved zone rose once woken
    st = ps.bodies_read_state()
    assert abs(st[5, 9] - 10.0) > 0.05 and st[13, 10] > 3.05
    ps.finish()


def test_areas_of_several_shapes_follow_the_reference():
    """An area with several shapes (area_add_shape beyond the first): to the C ABI every shape is an area record, the records of
    one area share a `group`, and a body standing in several shapes of the area has it ONCE in its override walk
    (BradotBody3D::add_area counts references).  scenes.area_groups() free-running beside the reference, where the groups are
    single areas of two shapes: positions and velocities to 1e-4 after every step, and the monitor callbacks hear the same
    (area, body, entered / left) events -- per shape there, per record here."""
    _need_oracle()
    sc = scenes.area_groups()
    steps = 110
    _, dump = parity.run_oracle(sc, steps, dump_from=0, dump_every=1)
    parent = np.arange(len(sc.areas))
    for i in range(1, len(sc.areas)):
        if sc.areas["group"][i] != 0 and sc.areas["group"][i] == sc.areas["group"][i - 1]:
            parent[i] = parent[i - 1]
    ps = parity.new_server(sc)
    heard = 0
    for k in range(1, steps + 1):
        ps.step(sc.dt)
        mine = sorted((int(parent[e[0]]), int(e[1]), int(e[2])) for e in ps.area_events())
        ref = sorted(tuple(int(x) for x in e) for e in dump.get(("area_events", k), np.zeros((0, 3), np.int32)))
        assert mine == ref, (k, mine, ref)
        heard += len(mine)
        st = ps.bodies_read_state()
        refb = dump[("bodies", k)]
        assert np.max(np.abs(st[:, 0:18] - refb[:, 0:18])) < 1e-4, (k, int(np.argmax(np.abs(st[:, 0:18] - refb[:, 0:18]).max(axis=1))))
    assert heard >= 30
    st = ps.bodies_read_state()
    assert np.all(st[0:6, 9] - sc.bodies["xf"][0:6, 9] > 0.05)        # the sideways gravity did act on the first column
    # with the group ignored (two areas instead of one) a body in both boxes is pushed twice: another trajectory
    apart = scenes.area_groups()
    apart.areas["group"] = 0
    ps2 = parity.new_server(apart)
    for k in range(steps):
        ps2.step(sc.dt)
    assert np.max(np.abs(ps2.bodies_read_state()[0:6, 9] - st[0:6, 9])) > 0.02
    ps.finish()
    ps2.finish()


def test_static_triangle_meshes_match_the_reference():
    """Static concave colliders (bradot_shape_3d.cpp:1174-1256,1433-1495, bradot_collision_solver_3d.cpp:297-360,
    the four *_face routines of bradot_collision_solver_3d_sat.cpp): spheres, boxes, capsules and hulls on a bumpy 288-triangle
    terrain and on a two-sided ramp, teacher-forced with the reference's own tree over the faces (dump "concave_bvh"): the pair
    list, which pairs collide and how many contacts they hold are equal, contacts to 1e-5, body state to 1e-4."""
    _need_oracle()
    sc = scenes.terrain_pile()
    w = parity.teacher_forced(sc, 151, [20, 45, 70, 100, 150])
    assert w["colliding"] > 20 and w["contacts"] > 40, w
    assert w["contact"] < CONTACT_TOL and w["state"] < 1e-4, w
    # free-running with a tree built here (other order of faces inside the tree where centres tie): everything comes to rest on the mesh
    ps = parity.new_server(sc)
    for _ in range(240):
        ps.step(sc.dt)
    st = ps.bodies_read_state()
    on_mesh = st[2:, 10] > -1.2
    assert on_mesh.mean() > 0.85 and np.all(np.isfinite(st))
    # a query shape against the mesh: it overlaps the terrain body where it dips into it, not above it
    probe = ps.sphere_shape_create(0.3)
    eye = np.array([1, 0, 0, 0, 1, 0, 0, 0, 1], np.float32)
    counts, bodies = ps.intersect_shapes([probe, probe], np.stack([np.concatenate([eye, [0.0, 3.0, 0.0]]), np.concatenate([eye, [0.0, 0.0, 0.0]])]).astype(np.float32), max_results=8)
    assert 1 not in bodies[0, :counts[0]] and 1 in bodies[1, :counts[1]]
    ps.finish()


def test_height_maps_match_the_reference():
    """HeightMapShape3D (bradot_shape_3d.cpp:2028-2078: the cells under the body's box, two faces each) under spheres, boxes,
    capsules and hulls, teacher-forced: pair list, colliding pairs and contact counts equal, contacts 1e-5, body state 1e-4."""
    _need_oracle()
    sc = scenes.heightmap_pile()
    w = parity.teacher_forced(sc, 151, [20, 45, 70, 100, 150])
    assert w["colliding"] > 20 and w["contacts"] > 40, w
    assert w["contact"] < CONTACT_TOL and w["state"] < 1e-4, w


@pytest.mark.parametrize("name", ["cone_twist", "slider", "sixdof", "chains", "pile"])
def test_solver_bins_and_grid_barrier_path_agree(name):
    """The teacher-forced tests run the grid-barrier solver (a forced order does).  Left to itself a scene of small islands goes
    through the solver bins, where contacts and joints work on bodies held in shared memory (solve_pair<true>, solve_joint<true>,
    solve_joint_ext<true>).  Both paths sweep the same colours, so free-running they must agree bit for bit: the same scene
    stepped by a context with the bins and by one without (BPC_NO_BINS=1, read when the context is made)."""
    sc = {"cone_twist": lambda: scenes.hanging_chains(cone_twist=True), "slider": scenes.slider_rigs, "sixdof": scenes.sixdof_rigs,
          "chains": scenes.hanging_chains, "pile": lambda: scenes.batched_worlds(6, lattice=(4, 3, 4))}[name]()
    steps = 40
    states, on_barrier = [], []
    for no_bins in ("0", "1"):
        os.environ["BPC_NO_BINS"] = no_bins
        try:
            ps = server.PhysicsServerCUDA(0)
        finally:
            os.environ.pop("BPC_NO_BINS", None)
        ps.load_scene(sc)
        for _ in range(steps):
            ps.step(sc.dt)
        states.append(ps.bodies_read_state())
        on_barrier.append(ps.stats()["barrier_constraints"])
        ps.finish()
    assert on_barrier[0] < on_barrier[1], on_barrier        # the bins did take constraints off the grid-barrier path
    assert np.isfinite(states[0]).all()
    assert np.abs(states[0][:, 9:12] - sc.bodies["xf"][:, 9:12]).max() > 0.2
    assert np.array_equal(states[0], states[1]), float(np.abs(states[0] - states[1]).max())


@pytest.mark.parametrize("name", ["cone_twist", "slider", "sixdof"])
def test_new_joint_types_in_islands_with_contacts(name):
    """The rigs of the joint tests above, now colliding: a floor under them that the lower boxes land on and neighbours that touch,
    so that islands hold joints AND contacts (the reference sweeps them in one list, joints by solver priority).  Teacher-forced:
    same pair list, contacts 1e-5, state 1e-4."""
    _need_oracle()
    sc = {"cone_twist": lambda: scenes.hanging_chains(6, 4, cone_twist=True), "slider": lambda: scenes.slider_rigs(8), "sixdof": lambda: scenes.sixdof_rigs(8)}[name]()
    dyn = sc.bodies["mode"] == scenes.MODE_RIGID
    sc.bodies["layer"][dyn] = 1
    sc.bodies["mask"][dyn] = 1
    sc.joints["priority"] = 0      # (an island with contacts and priority-2 joints loses its contacts from the harness's record of the order)
    lowest = float(sc.bodies["xf"][dyn, 10].min())
    sc.add_static_plane((0, 1, 0), lowest - (0.1 if name == "cone_twist" else 0.35))      # (the chains hang: their last link starts in the floor)
    _, dump = parity.run_oracle(sc, 70)
    w = parity.teacher_forced(sc, 70, [1, 10, 20, 35, 50, 68], dump=dump)
    assert w["state"] < STATE_TOL and w["contact"] < CONTACT_TOL and w["contacts"] > 20, w


def test_shape_custom_solver_bias_matches_the_reference():
    """shape_set_custom_solver_bias (BradotBodyPair3D::setup, bradot_body_pair_3d.cpp:344-357): a pair takes the bias of the shape
    that has one, or the mean of both, in place of 0.8.  A settling mixed pile where the spheres, the boxes and the floor
    plane carry their own biases: teacher-forced, contacts 1e-5, state 1e-4; and the biases matter."""
    _need_oracle()
    sc = scenes.mixed_pile(4, half=3.0)
    types = sc.shapes["type"]
    sc.shapes["custom_bias"][types == scenes.SHAPE_SPHERE] = 0.3
    sc.shapes["custom_bias"][types == scenes.SHAPE_BOX] = 0.55
    sc.shapes["custom_bias"][np.nonzero(types == scenes.SHAPE_PLANE)[0][0]] = 0.2
    assert (sc.shapes["custom_bias"] != 0).sum() >= 3 and (sc.shapes["custom_bias"] == 0).sum() >= 2
    _, dump = parity.run_oracle(sc, 90)
    w = parity.teacher_forced(sc, 90, [1, 30, 45, 60, 75, 88], dump=dump)
    assert w["state"] < STATE_TOL and w["contact"] < CONTACT_TOL and w["contacts"] > 100, w
    plain = scenes.mixed_pile(4, half=3.0)
    ps = parity.new_server(plain, dump)
    parity.load_state(ps, dump, 60)
    ps.set_solve_order(parity.solve_order(dump, 61))
    ps.step(sc.dt)
    st = ps.bodies_read_state()
    assert parity.rel_err(st[:, 0:12], dump[("bodies", 61)][:, 0:12], 1.0) > 1e-4
    ps.finish()


def test_slider_joints_match_the_reference():
    """SliderJoint3D (joints/bradot_slider_joint_3d.cpp:74-337): rigs of two boxes on slider joints in turned frames, default and
    own parameters (22 per joint through bpc_joint_params_write), linear and angular limits hit; teacher-forced with the
    reference's sweep order, body state after the step within 1e-4.  The parameters are at work: the same rigs with every
    default swing differently; and a parameter written at run time reads back."""
    _need_oracle()
    sc = scenes.slider_rigs()
    assert (sc.joints["type"] == scenes.JOINT_SLIDER).sum() == 36 and len(sc.joint_ext) >= 24
    _, dump = parity.run_oracle(sc, 80)
    w = parity.teacher_forced(sc, 80, [1, 2, 5, 15, 30, 50, 79], dump=dump)
    assert w["state"] < STATE_TOL, w
    plain = scenes.slider_rigs()
    plain.joint_ext = plain.joint_ext[:0]
    ps = parity.new_server(plain, dump)
    parity.load_state(ps, dump, 30)
    ps.set_solve_order(parity.solve_order(dump, 31))
    ps.step(sc.dt)
    av = ps.bodies_read(server.FIELD_ANGULAR_VELOCITY)
    assert parity.rel_err(av, dump[("bodies", 31)][:, 15:18], 1.0) > 1e-2
    assert np.allclose(ps.joint_params_read(0), [-1, 1, 1, .7, 1, 1, .7, 0, 1, .7, 1, 0, 0, 1, .7, 1, 1, .7, 0, 1, .7, 1])
    ps.joint_params_write(0, 3, [0.25, 0.5])
    assert np.array_equal(ps.joint_params_read(0, 2, 4), np.array([1, 0.25, 0.5, 1], np.float32))
    with pytest.raises(Exception):
        ps.joint_params_write(0, 20, [1, 2, 3])
    ps.finish()


def test_generic_6dof_joints_match_the_reference():
    """Generic6DOFJoint3D (joints/bradot_generic_6dof_joint_3d.cpp:64-449): rigs on 6-DOF joints in turned frames -- locked, limited,
    free and switched-off linear axes, violated angular limits (Euler XYZ of the relative frame), a running motor, clipped limit
    forces, bounce; teacher-forced with the reference's sweep order, body state after the step within 1e-4.  The same rigs with
    every default move differently."""
    _need_oracle()
    sc = scenes.sixdof_rigs()
    assert (sc.joints["type"] == scenes.JOINT_6DOF).sum() == 36
    _, dump = parity.run_oracle(sc, 80)
    w = parity.teacher_forced(sc, 80, [1, 2, 5, 15, 30, 50, 79], dump=dump)
    assert w["state"] < STATE_TOL, w
    plain = scenes.sixdof_rigs()
    plain.joint_ext = plain.joint_ext[:0]
    ps = parity.new_server(plain, dump)
    parity.load_state(ps, dump, 30)
    ps.set_solve_order(parity.solve_order(dump, 31))
    ps.step(sc.dt)
    av = ps.bodies_read(server.FIELD_ANGULAR_VELOCITY)
    assert parity.rel_err(av, dump[("bodies", 31)][:, 15:18], 1.0) > 1e-2
    assert np.array_equal(ps.joint_params_read(0, 28, 28), scenes.g6dof_axis_defaults())
    ps.finish()


def test_cone_twist_joints_match_the_reference():
    """ConeTwistJoint3D (joints/bradot_cone_twist_joint_3d.cpp:67-279: swing ellipse and twist limits with accumulated, clamped
    impulses, the shortest-arc quaternion for the twist reference): chains whose links alternate hinges and cone-twist joints,
    teacher-forced with the reference's sweep order, body state after the step within 1e-4; the limits do engage."""
    _need_oracle()
    sc = scenes.hanging_chains(cone_twist=True)
    assert (sc.joints["type"] == scenes.JOINT_CONE_TWIST).sum() >= 20
    _, dump = parity.run_oracle(sc, 80)
    w = parity.teacher_forced(sc, 80, [1, 5, 15, 30, 50, 79], dump=dump)
    assert w["state"] < STATE_TOL, w
    # the same chains with pins in place of the cone twists swing differently: the limits are at work
    plain = scenes.hanging_chains(cone_twist=True)
    plain.joints["type"][plain.joints["type"] == scenes.JOINT_CONE_TWIST] = scenes.JOINT_PIN
    plain.joints["flags"] &= ~np.uint32(scenes.JOINT_HAS_PARAMS)
    ps = parity.new_server(plain, dump)
    parity.load_state(ps, dump, 30)
    ps.set_solve_order(parity.solve_order(dump, 31))
    ps.step(sc.dt)
    av = ps.bodies_read(server.FIELD_ANGULAR_VELOCITY)
    assert parity.rel_err(av, dump[("bodies", 31)][:, 15:18], 1.0) > 1e-2
    ps.finish()
