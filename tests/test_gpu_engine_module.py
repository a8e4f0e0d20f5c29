"""Drop-in boundary, end to end: the engine core (reference objects) + our module, server obtained BY NAME from
PhysicsServer3DManager, scene built through the public PhysicsServer3D RID API (engine_module/engine_demo.cpp).
The body positions must equal a direct C-ABI run of the same scene bit for bit: same library, same kernels."""
import json
import os
import subprocess

import numpy as np
import pytest

import parity
from bradot_b200 import scenes, server

pytestmark = pytest.mark.gpu
DEMO = os.path.join(parity.ROOT, "engine_module", "_build", "engine_demo")


def test_engine_module_matches_direct_c_abi(tmp_path):
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    sc = scenes.mixed_pile(4, half=3.0)
    path = str(tmp_path / "scene.bscn")
    sc.write(path)
    steps = 80
    r = subprocess.run([DEMO, path, str(steps)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["server"] == "BradotPhysicsCUDA"
    engine = np.array(out["origins"], np.float32)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(steps):
        ps.step(sc.dt)
    direct = ps.bodies_read(server.FIELD_TRANSFORM)[:, 9:12]
    st = ps.stats()
    # PhysicsDirectSpaceState3D::intersect_ray through the RID API == the batched C-ABI query on the same state
    eng_rays = np.array(out["rays"], np.float64)
    n = len(eng_rays)
    ft = np.zeros((n, 6), np.float32)
    ft[:, 0], ft[:, 1], ft[:, 2] = direct[:n, 0], 30.0, direct[:n, 2]
    ft[:, 3], ft[:, 4], ft[:, 5] = direct[:n, 0], -1.0, direct[:n, 2]
    hit, body, pos, _ = ps.intersect_rays(ft)
    # cast_motion (with an exclude list) and rest_info through the RID API == the C-ABI calls on the same state
    casts = np.array(out["casts"], np.float64)
    probe = ps.sphere_shape_create(0.3)
    eye = np.array([1, 0, 0, 0, 1, 0, 0, 0, 1], np.float32)
    for i in range(n):
        ps.set_query_exclude([i])
        fr, info = ps.cast_motions([probe], np.concatenate([eye, [direct[i, 0], 30.0, direct[i, 2]]]).astype(np.float32), [[0, -31.0, 0]])
        assert np.array_equal(casts[i, 0:2].astype(np.float32), fr[0]) and int(casts[i, 2]) == (info["body"][0] if info["hit"][0] else -1), i
        assert 0 < fr[0, 0] < 1 and info["body"][0] != i
        ps.set_query_exclude([])
        rest = ps.rest_infos([probe], np.concatenate([eye, direct[i]]).astype(np.float32), margin=0.05)
        assert int(casts[i, 3]) == (rest["body"][0] if rest["hit"][0] else -1), i
        if rest["hit"][0]:
            assert np.array_equal(casts[i, 4:7].astype(np.float32), rest["point"][0])
    # get_closest_point_to_object_volume through the RID API == bpc_bodies_closest_points on the same state
    closest = np.array(out["closest"], np.float64).astype(np.float32)
    mine_c = ps.bodies_closest_points(np.arange(n), direct[:n] + np.array([0.6, 0.4, -0.3], np.float32))
    assert np.array_equal(closest, mine_c), float(np.abs(closest - mine_c).max())
    # body_test_motion through the RID API == bpc_bodies_test_motion on the same state
    mot = np.array(out["motions"], np.float64)
    prm = np.zeros(n, server.MOTION_PARAMS_DT)
    prm["body"] = np.arange(n)
    prm["from"] = ps.bodies_read_state()[:n, 0:12]
    prm["motion"], prm["margin"], prm["max_collisions"] = (0.05, -0.4, 0.0), 0.02, 4
    mr = ps.bodies_test_motion(prm)
    assert np.array_equal(mot[:, 0] != 0, mr["collided"] != 0) and mr["collided"].any()
    hitm = mr["collided"] != 0
    assert np.array_equal(mot[hitm, 1].astype(np.int32), mr["collisions"][hitm, 0]["collider"])
    assert np.array_equal(mot[hitm, 2].astype(np.int32), mr["collision_count"][hitm])
    assert np.array_equal(mot[:, 3].astype(np.float32), mr["collision_safe_fraction"]) and np.array_equal(mot[:, 4:6].astype(np.float32), mr["travel"][:, 0:2])
    ps.finish()
    assert n == min(16, len(sc.bodies)) and hit.all() and np.array_equal(eng_rays[:, 0] != 0, hit)
    assert np.array_equal(eng_rays[:, 1].astype(np.int64), body.astype(np.int64))
    assert np.array_equal(eng_rays[:, 2:5].astype(np.float32), pos)
    assert engine.shape == direct.shape
    assert np.array_equal(engine, direct), float(np.abs(engine - direct).max())
    assert out["collision_pairs"] == st["collision_pairs"] and out["active_objects"] == st["active_objects"]
    assert st["contact_pairs"] > 0


def _pile_with_disabled_shapes():
    sc = scenes.mixed_pile(4, half=3.0)
    dyn = np.nonzero(sc.bodies["mode"] == scenes.MODE_RIGID)[0]
    sc.bodies["flags"][dyn[::3]] |= scenes.BODY_SHAPE_DISABLED
    return sc


@pytest.mark.parametrize("name", ["pushed", "bullets", "terrain", "cone_twist", "slider", "sixdof", "disabled"])
def test_engine_module_forces_and_ccd_match_direct_c_abi(tmp_path, name):
    """body_apply_central_force / body_apply_torque and body_set_enable_continuous_collision_detection through the RID API
    against the same scene driven through the C ABI directly (force rows written before every step; CCD from the flags)."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    sc = {"pushed": scenes.pushed, "bullets": scenes.bullets, "terrain": scenes.terrain_pile,    # terrain: a ConcavePolygonShape3D
          "cone_twist": lambda: scenes.hanging_chains(cone_twist=True),                          # joint_make_cone_twist / cone_twist_joint_set_param
          "sixdof": scenes.sixdof_rigs,                                                           # joint_make_generic_6dof / generic_6dof_joint_set_param / _set_flag
          "disabled": lambda: _pile_with_disabled_shapes(),                                       # body_set_shape_disabled
          "slider": scenes.slider_rigs}[name]()                                                   # joint_make_slider / slider_joint_set_param
    path = str(tmp_path / "scene.bscn")
    sc.write(path)
    steps = 40
    r = subprocess.run([DEMO, path, str(steps)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    engine = np.array(out["origins"], np.float32)
    push = np.zeros((len(sc.bodies), 6), np.float32)
    for i in np.nonzero(sc.bodies["flags"] & scenes.PUSH_FLAG)[0]:
        push[i] = scenes.pushed_force(int(i))
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(steps):
        if push.any():
            ps.bodies_write(server.FIELD_APPLIED_FORCE, push)
        ps.step(sc.dt)
    direct = ps.bodies_read(server.FIELD_TRANSFORM)[:, 9:12]
    ps.finish()
    if name == "terrain":   # (its hulls are built by the engine's ConvexHullComputer there and by bradot_b200.hull here: same hull, another face order)
        assert np.max(np.abs(engine - direct)) < 5e-3
    else:
        assert np.array_equal(engine, direct), float(np.abs(engine - direct).max())
    if name == "terrain":
        assert np.mean(direct[2:, 1] > -1.2) > 0.7          # most of them lie on the mesh, not on the plane 6 m below it
    if name == "bullets":
        ccd = (sc.bodies["flags"] & 2) != 0
        assert np.all(direct[ccd, 2] < 3.0)        # stopped by the wall / the plates


def test_engine_module_static_body_with_several_shapes(tmp_path):
    """body_add_shape beyond the first on a static body (level geometry): the module keeps the further shapes as records of their
    own on the GPU side.  Against (1) the C ABI stepping the same scene with six static bodies: the same to rounding; (2) the reference with
    ONE static body of six shapes, teacher-forced (its pairs are per shape: the same pair list, contacts 1e-5, state 1e-4)."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")
    sc = scenes.level_geometry()
    steps = 50
    path = str(tmp_path / "scene.bscn")
    sc.write(path)
    r = subprocess.run([DEMO, path, str(steps)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    engine = np.array(out["origins"], np.float32)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(steps):
        ps.step(sc.dt)
    direct = ps.bodies_read(server.FIELD_TRANSFORM)[:, 9:12]
    st = ps.stats()
    ps.finish()
    # (the hull is built by the engine's ConvexHullComputer there and by bradot_b200.hull here: same hull, another face order)
    assert np.max(np.abs(engine - direct)) < 5e-3, float(np.abs(engine - direct).max())
    assert out["collision_pairs"] == st["collision_pairs"] and st["contact_pairs"] > 10
    # the reference steps ONE static body of six shapes; its pairs are listed per shape, which is what the six records are
    w = parity.teacher_forced(sc, 60, [1, 15, 30, 45, 58])
    assert w["state"] < 1e-4 and w["contact"] < 1e-5 and w["contacts"] > 50, w
    dyn = sc.bodies["mode"] == scenes.MODE_RIGID
    assert np.mean(direct[dyn, 1] > -1.0) > 0.6            # most of them came to lie on the level, not on the plane below


def test_engine_module_custom_inertia_and_centre_of_mass(tmp_path):
    """body_set_param(BODY_PARAM_INERTIA / BODY_PARAM_CENTER_OF_MASS / BODY_PARAM_MASS) and body_reset_mass_properties on bodies
    that are in a space, through the RID API (the module hands the values over with bpc_bodies_set_mass_override and describes
    the body anew): the script of test_gpu_features.py::test_custom_inertia_and_centre_of_mass_follow_the_reference, free-running
    against the reference, positions to 2 mm."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")
    sc = scenes.locks_and_damping(24, on_plane=True)
    sc.bodies["flags"] &= ~np.uint32(63 << scenes.BODY_LOCK_SHIFT)
    dyn = np.nonzero(sc.bodies["mode"] == scenes.MODE_RIGID)[0]
    events = []
    for k, b in enumerate(dyn):
        rows = {0: [(8, (0.02 + 0.01 * (k % 3), 0.05, 0.03))], 1: [(8, (0.06, 0.0, 0.0)), (9, (0.05, -0.1, 0.0))], 2: [(9, (0.0, -0.15, 0.05))], 3: []}[k % 4]
        if k % 4 != 3:
            rows.append((11, (float(sc.bodies["mass"][b]),)))
        for kind, v in rows:
            row = np.zeros(12, np.float32)
            row[:len(v)] = v
            events.append((2, int(b), kind, row))
        if k % 4 in (0, 1) and k % 8 < 4:
            events.append((45, int(b), 10, np.zeros(12, np.float32)))
    steps = 90
    path, evp = str(tmp_path / "scene.bscn"), str(tmp_path / "events.bin")
    sc.write(path)
    parity.write_events(evp, events)
    r = subprocess.run([DEMO, path, str(steps), "--events", evp], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    engine = np.array(json.loads(r.stdout.strip().splitlines()[-1])["origins"], np.float32)
    _, dump = parity.run_oracle(sc, steps, events=events, dump_steps=[steps])
    _, plain = parity.run_oracle(sc, steps, dump_steps=[steps])
    ref = dump[("bodies", steps)][:, 9:12]
    assert np.max(np.abs(engine - ref)) < 2e-3, float(np.abs(engine - ref).max())
    assert np.max(np.abs(ref - plain[("bodies", steps)][:, 9:12])) > 0.02


def test_kinematic_body_with_several_shapes(tmp_path):
    """A kinematic platform of three shapes (slab and two rails) carrying boxes and spheres: (1) the C ABI moving three kinematic
    records alike, marked as one object (bpc_bodies_set_same_object: they overlap, and two kinematic records would pair),
    teacher-forced against the reference moving ONE kinematic body of three shapes (its targets: dump "kin_targets"): same
    pair list per shape, state 5e-4 (the bar of test_kinematic_platform_matches_the_reference: acosf); (2) the module
    (body_add_shape on a kinematic body, body_set_state(TRANSFORM) before every step) free-running beside the reference:
    positions to 5 mm after 100 steps."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")
    sc = scenes.kinematic_platform(rails=True)
    steps = 100
    _, dump = parity.run_oracle(sc, steps, dump_from=0, dump_every=1)
    further = [int(f) for f in np.nonzero(sc.bodies["flags"] & scenes.BODY_FURTHER_SHAPE)[0]]
    owner = {}
    for f in further:
        o = f - 1
        while o in owner:
            o = owner[o]
        owner[f] = o
    ps = parity.new_server(sc, dump)
    records = {}
    for f, o in owner.items():
        records.setdefault(o, [o]).append(f)
    for recs in records.values():
        for i in range(len(recs)):
            for j in range(i + 1, len(recs)):
                ps.bodies_set_same_object(recs[i], recs[j])
    errs, collided = [], 0
    for k in range(1, steps - 1):
        parity.load_state(ps, dump, k)
        for row in dump[("kin_targets", k + 1)]:
            for b in records.get(int(row[0]), [int(row[0])]):
                ps.bodies_write(server.FIELD_KINEMATIC_TARGET, row[1:13].reshape(1, 12), first=b)
        ps.set_solve_order(parity.solve_order(dump, k + 1))
        ps.step(sc.dt)
        ref = dump[("bodies", k + 1)]
        st = ps.bodies_read_state()
        errs.append(float(np.max(np.abs(st[:, 0:18] - ref[:, 0:18]))))
        ids, _, _ = ps.pairs_read()
        rids, _, _ = parity.oracle_pairs(dump, k + 1)
        order = np.argsort(parity.pair_keys(rids))
        assert np.array_equal(parity.pair_keys(ids), parity.pair_keys(rids)[order]), k + 1
        collided += int((ids[:, 3] & 1).sum())
    ps.finish()
    errs = np.array(errs)
    assert collided > 300 and np.median(errs) < 1e-6 and errs.max() < 5e-4, (collided, float(np.median(errs)), float(errs.max()))
    path = str(tmp_path / "scene.bscn")
    sc.write(path)
    r = subprocess.run([DEMO, path, str(steps), "--drive-kinematic"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    engine = np.array(json.loads(r.stdout.strip().splitlines()[-1])["origins"], np.float32)
    ref = dump[("bodies", steps)][:, 9:12]
    assert np.max(np.abs(engine - ref)) < 5e-3, float(np.abs(engine - ref).max())
    assert ref[1, 0] > 0.9          # the platform did travel


def test_engine_module_collision_priority(tmp_path):
    """body_set_collision_priority through the RID API (set on bodies that are already in the space): the module's body_test_motion
    answers what bpc_bodies_test_motion answers on the same state with the same priorities -- and not what it answers without."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    sc = scenes.mixed_pile(4, half=3.0)
    nb = len(sc.bodies)
    prio = np.ones(nb, np.float32)
    prio[1::2] = 4.0
    prio[0] = 0.2
    events = [(2, int(b), 12, np.array([prio[b]] + [0] * 11, np.float32)) for b in range(nb) if prio[b] != 1.0]
    steps = 80
    path, evp = str(tmp_path / "scene.bscn"), str(tmp_path / "events.bin")
    sc.write(path)
    parity.write_events(evp, events)
    r = subprocess.run([DEMO, path, str(steps), "--events", evp], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    mot = np.array(out["motions"], np.float64)
    n = len(mot)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(steps):
        ps.step(sc.dt)
    prm = np.zeros(n, server.MOTION_PARAMS_DT)
    prm["body"] = np.arange(n)
    prm["from"] = ps.bodies_read_state()[:n, 0:12]
    prm["motion"], prm["margin"], prm["max_collisions"] = (0.05, -0.4, 0.0), 0.02, 4
    plain = ps.bodies_test_motion(prm)
    ps.bodies_set_collision_priority(np.arange(nb), prio)
    mr = ps.bodies_test_motion(prm)
    ps.finish()
    assert np.array_equal(mot[:, 0] != 0, mr["collided"] != 0)
    assert np.array_equal(mot[:, 3].astype(np.float32), mr["collision_safe_fraction"]) and np.array_equal(mot[:, 4:6].astype(np.float32), mr["travel"][:, 0:2])
    assert not np.array_equal(plain["travel"], mr["travel"])            # the priorities do change the recovery


def test_engine_module_areas_of_several_shapes(tmp_path):
    """area_add_shape beyond the first through the RID API: the module keeps each further shape as an area record of its own that
    shares the area's group (one override per body, events per shape through the area's own callback).  scenes.area_groups()
    against the C ABI stepping the same records: the callbacks heard the same (step, area, body, status) events, bodies bit-equal."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    sc = scenes.area_groups()
    steps = 110
    path = str(tmp_path / "scene.bscn")
    sc.write(path)
    r = subprocess.run([DEMO, path, str(steps)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    heard = sorted(tuple(e) for e in out["area_events"])
    parent = np.arange(len(sc.areas))
    for i in range(1, len(sc.areas)):
        if sc.areas["group"][i] != 0 and sc.areas["group"][i] == sc.areas["group"][i - 1]:
            parent[i] = parent[i - 1]
    ps = parity.new_server(sc)
    mine = []
    for s in range(1, steps + 1):
        ps.step(sc.dt)
        mine += [(s, int(parent[e[0]]), int(e[1]), int(e[2])) for e in ps.area_events()]
    st = ps.bodies_read_state()
    ps.finish()
    assert heard == sorted(mine), (sorted(set(heard) - set(mine)), sorted(set(mine) - set(heard)))
    assert len(heard) >= 30
    engine = np.array(out["origins"], np.float32)
    assert np.array_equal(engine, st[:, 9:12]), float(np.abs(engine - st[:, 9:12]).max())


def test_engine_module_simple_hinges_match_the_reference(tmp_path):
    """joint_make_hinge_simple through the module (the frames built from pivots, axes and body A's transform, as
    joints/bradot_hinge_joint_3d.cpp:69-108 does) against the reference making the same hinges: free-running, positions to 2 mm."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")
    sc = scenes.simple_hinges()
    steps = 50
    path = str(tmp_path / "scene.bscn")
    sc.write(path)
    r = subprocess.run([DEMO, path, str(steps)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    engine = np.array(json.loads(r.stdout.strip().splitlines()[-1])["origins"], np.float32)
    _, dump = parity.run_oracle(sc, steps, dump_steps=[steps])
    ref = dump[("bodies", steps)][:, 9:12]
    assert np.max(np.abs(engine - ref)) < 2e-3, float(np.abs(engine - ref).max())
    assert np.max(np.abs(ref - sc.bodies["xf"][:, 9:12])) > 0.2        # they do swing


def test_engine_module_callbacks_set_state_and_free_match_c_abi_and_reference(tmp_path):
    """A scene driven like a game drives it -- every body has a state-sync callback (what RigidBody3D registers), a script of
    body_set_state / free calls runs between the ticks -- through the RID API of the module, against (1) the same script on
    the C ABI directly: bit-equal; (2) the reference running the same script (oracle_step --events): same sleep state for
    every body, positions to 2 mm; and the transforms the callbacks reported are the ones body_get_state returns."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")
    sc = scenes.scripted_events()
    events = scenes.scripted_events_script()
    steps = 160
    path, evp = str(tmp_path / "scene.bscn"), str(tmp_path / "events.bin")
    sc.write(path)
    parity.write_events(evp, events)
    r = subprocess.run([DEMO, path, str(steps), "--events", evp], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    engine = np.array(out["origins"], np.float32)
    synced = np.array(out["synced"], np.float64)
    sleeping = np.array(out["sleeping"])
    freed = {body for (_, body, kind, _) in events if kind == 3}
    alive = np.array([i not in freed for i in range(len(sc.bodies))])
    dyn = sc.bodies["mode"] >= scenes.MODE_RIGID
    # (1) the C ABI directly
    from test_gpu_features import _apply_events
    ps = parity.new_server(sc)
    gone = set()
    for s in range(1, steps + 1):
        _apply_events(ps, events, s, gone)
        ps.step(sc.dt)
    st = ps.bodies_read_state()
    ps.finish()
    assert np.array_equal(engine[alive], st[alive, 9:12]), float(np.abs(engine[alive] - st[alive, 9:12]).max())
    assert np.array_equal(sleeping[alive & dyn] == 0, st[alive & dyn, 18] != 0)
    # (2) the reference
    _, dump = parity.run_oracle(sc, steps, events=events, dump_steps=[steps])
    ref = dump[("bodies", steps)]
    assert np.max(np.abs(engine[alive] - ref[alive, 9:12])) < 2e-3
    assert np.array_equal(sleeping[alive & dyn] == 0, ref[alive & dyn, 18] != 0)
    # the callbacks: every moving body was told, freed ones are marked, and the last reported transform is the current one
    calls = synced[:, 0]
    assert np.all(calls[~alive] == -1) and np.all(calls[alive & dyn] > 0) and np.all(calls[alive & ~dyn] == 0)
    assert np.array_equal(synced[alive & dyn, 1:4].astype(np.float32), engine[alive & dyn])
    # a body asleep for most of the run was told far fewer times than there were ticks
    assert calls[4] < steps // 2


def test_engine_module_reported_contacts_match_direct_c_abi(tmp_path):
    """body_set_max_contacts_reported + the PhysicsDirectBodyState3D contact getters through the RID API (read inside the
    state-sync callbacks, as RigidBody3D does for its body_entered signals: scene/3d/physics/rigid_body_3d.cpp:455-520)
    against bpc_contacts_read on the same scene: the same records, the collider mapped back to its RID and object id."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    sc = scenes.pyramid(36, base=8)
    flags = sc.bodies["flags"].copy()
    flags[1:] |= np.where(np.arange(len(flags) - 1) % 2 == 0, 6, 0).astype(np.uint32) << scenes.BODY_REPORT_SHIFT
    flags[10] = (flags[10] & 0xFF00FFFF) | (2 << scenes.BODY_REPORT_SHIFT)
    sc.bodies["flags"] = flags
    path = str(tmp_path / "scene.bscn")
    sc.write(path)
    steps = 20
    r = subprocess.run([DEMO, path, str(steps)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(steps):
        ps.step(sc.dt)
    counts, offsets, rec = ps.contacts_read()
    ps.finish()
    seen = 0
    for body, listed in enumerate(out["contacts"]):
        mine = rec[offsets[body]:offsets[body] + counts[body]]
        assert len(listed) == len(mine), body
        if not len(mine):
            continue
        got = np.array(listed, np.float64)
        assert np.all(got[:, 1] == got[:, 0] + 1000)          # collider object id = what the demo attached to that body
        want = np.concatenate([mine["collider"][:, None].astype(np.float64), mine["local_pos"], mine["local_normal"], mine["impulse"],
                               mine["local_velocity"], mine["collider_pos"], mine["collider_velocity"]], axis=1)
        got = np.delete(got, 1, axis=1)
        o1, o2 = np.lexsort(got.T[::-1]), np.lexsort(want.T[::-1])
        assert np.array_equal(got[o1].astype(np.float32), want[o2].astype(np.float32)), body
        seen += len(mine)
    assert seen > 40


def test_engine_module_areas_match_direct_c_abi(tmp_path):
    """area_* through the RID API (areas created like Area3D nodes do, a monitor callback on every monitoring one, the script
    moving and freeing areas and freeing a body inside one) against the same scene and script on the C ABI: the callbacks heard
    the same (step, area, body, status) events, with the right body RID and object id, and the bodies end bit-equal."""
    if not os.path.exists(DEMO):
        pytest.skip("engine_module/_build/engine_demo not built (needs the engine tree: python engine_module/build_demo.py)")
    sc = scenes.area_zones()
    events = scenes.area_zones_script()
    steps = 100
    path, evp = str(tmp_path / "scene.bscn"), str(tmp_path / "events.bin")
    sc.write(path)
    parity.write_events(evp, events)
    r = subprocess.run([DEMO, path, str(steps), "--events", evp], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    heard = sorted(tuple(e) for e in out["area_events"])
    from test_gpu_features import _apply_area_events
    ps = parity.new_server(sc)
    areas = sc.areas.copy()
    gone_b, gone_a = set(), set()
    mine = []
    for s in range(1, steps + 1):
        _apply_area_events(ps, sc, areas, events, s, gone_b, gone_a)
        ps.step(sc.dt)
        mine += [(s, int(e[0]), int(e[1]), int(e[2])) for e in ps.area_events()]
    st = ps.bodies_read_state()
    ps.finish()
    mine = sorted(mine)     # (the body freed before step 55 leaves its zone then: both sides hear it with that step's events)
    assert heard == mine, (sorted(set(heard) - set(mine)), sorted(set(mine) - set(heard)))
    assert len(heard) > 15 and all(e[2] >= 0 for e in heard)
    alive = np.array([i not in gone_b for i in range(len(sc.bodies))])
    engine = np.array(out["origins"], np.float32)
    assert np.array_equal(engine[alive], st[alive, 9:12]), float(np.abs(engine[alive] - st[alive, 9:12]).max())
