"""Edge cases of the C ABI and size-independent properties at BASELINE.json's full sizes.

The reference has no physics tests to mirror; these cover what its server guards with
ERR_FAIL_* (invalid handles return an error, nothing crashes), degenerate inputs (empty context, a single
body, bodies added between steps) and, at the 100 000-body size of the headline configuration, properties
that do not need the oracle: the pair list is sorted / unique / complete against a brute-force overlap
check of the stored boxes, manifolds are well formed, the run is bit-reproducible.
"""
import numpy as np
import pytest

from bradot_b200 import scenes, server

pytestmark = pytest.mark.gpu


def test_empty_context_and_single_body():
    ps = server.PhysicsServerCUDA(0)
    ps.step(1 / 60)                                  # no spaces, no bodies: a no-op, not an error
    assert ps.stats()["collision_pairs"] == 0
    sp = ps.space_create()
    sh = ps.sphere_shape_create(0.5)
    rec = np.zeros(1, scenes.BODY_DT)
    rec["xf"] = scenes.IDENT12
    rec["shape_xf"] = scenes.IDENT12
    rec["mode"], rec["shape"], rec["space"], rec["mass"], rec["flags"] = scenes.MODE_RIGID, sh, sp, 2.0, 1
    rec["gravity_scale"], rec["layer"], rec["mask"] = 1.0, 1, 1
    ps.bodies_create(rec)
    for _ in range(10):
        ps.step(1 / 60)
    v = ps.bodies_read(server.FIELD_LINEAR_VELOCITY)[0]
    assert v[1] < -1.5 and v[0] == 0 and v[2] == 0   # ten steps of gravity with 0.1 damping
    assert ps.stats()["active_objects"] == 1 and ps.stats()["collision_pairs"] == 0
    ps.finish()


def test_invalid_arguments_are_refused_not_fatal():
    ps = server.PhysicsServerCUDA(0)
    sp = ps.space_create()
    sh = ps.box_shape_create((0.5, 0.5, 0.5))
    rec = np.zeros(1, scenes.BODY_DT)
    rec["xf"] = scenes.IDENT12
    rec["shape_xf"] = scenes.IDENT12
    rec["mass"], rec["mode"], rec["shape"], rec["space"] = 1.0, scenes.MODE_RIGID, sh, sp
    for field, bad in (("space", 7), ("shape", 99), ("mode", 9)):
        r = rec.copy()
        r[field] = bad
        with pytest.raises(server.PhysicsError):
            ps.bodies_create(r)
    with pytest.raises(server.PhysicsError):
        ps.shape_create(scenes.SHAPE_CYLINDER, (0.5, 1.0))      # cylinders are a "next" row: refused, never ignored
    with pytest.raises(server.PhysicsError):
        ps.space_set_param(3, server.SPACE_PARAM_SOLVER_ITERATIONS, 8)
    with pytest.raises(server.PhysicsError):
        ps.bodies_read(server.FIELD_TRANSFORM, 0, 5)            # no bodies yet
    with pytest.raises(server.PhysicsError):
        ps.step(0.0)
    ps.bodies_create(rec)                                       # the context is still usable
    ps.step(1 / 60)
    assert ps.body_count() == 1
    ps.finish()


def test_bodies_added_between_steps_keep_existing_state():
    sc = scenes.spheres_container(4, 4, 4, half=1.5)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(40):
        ps.step(sc.dt)
    before = ps.bodies_read(server.FIELD_TRANSFORM)
    extra = sc.bodies[5:6].copy()
    extra["xf"][0, 9:12] = (0.0, 6.0, 0.0)
    first = ps.bodies_create(extra)
    assert first == len(sc.bodies)
    assert np.array_equal(ps.bodies_read(server.FIELD_TRANSFORM, 0, first), before)
    for _ in range(120):
        ps.step(sc.dt)
    y = ps.bodies_read(server.FIELD_TRANSFORM)[first, 10]
    assert 0.2 < y < 3.0, y                                     # it fell onto the pile and stayed in the container
    ps.finish()


def test_sort_edge_sizes():
    ps = server.PhysicsServerCUDA(0)
    rng = np.random.default_rng(5)
    for n in (0, 1, 2, 255, 256, 257, 2047, 2048, 2049, 70001):
        k = rng.integers(0, 2 ** 63, n, dtype=np.uint64)
        assert np.array_equal(ps.sort_u64(k), np.sort(k)), n
    ps.finish()


def _run_c2(steps):
    sc = scenes.spheres_container(50, 40, 50)                    # BASELINE.json configs[1]: 100 000 spheres
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    for _ in range(steps):
        ps.step(sc.dt)
    return sc, ps


def test_full_size_c2_pair_list_and_manifold_properties():
    sc, ps = _run_c2(69)
    boxes = ps.bodies_read(server.FIELD_ITEM_BOX)                # stored boxes the next step will pair on
    ps.step(sc.dt)
    ids, pf, cf = ps.pairs_read()
    st = ps.stats()
    keys = ids[:, 0].astype(np.int64) << 32 | ids[:, 1].astype(np.int64)
    assert len(keys) == st["collision_pairs"] > 100000
    assert np.all(ids[:, 0] < ids[:, 1]) and np.all(np.diff(keys) > 0)          # A < B, sorted, unique
    # completeness and exactness against brute force on a sample of bodies
    mn, mx = boxes[:, 6:9], boxes[:, 9:12]
    static = sc.bodies["mode"] == scenes.MODE_STATIC
    rng = np.random.default_rng(0)
    key_set = set(keys.tolist())
    for a in rng.choice(np.nonzero(~static)[0], 150, replace=False):
        ov = np.all(mn[a] <= mx, axis=1) & np.all(mn <= mx[a], axis=1)
        ov[a] = False
        want = {(min(a, b) << 32) | max(a, b) for b in np.nonzero(ov)[0]}
        got = {k for k in key_set if (k >> 32) == a or (k & 0xffffffff) == a}
        assert got == want, (a, len(got), len(want))
    # manifolds: colliding pairs hold 1..4 contacts with unit normals and finite, non-negative accumulated normal impulses
    col = (ids[:, 3] & 1) > 0
    assert st["contact_pairs"] == int(col.sum()) > 1000
    assert np.all(ids[col, 2] >= 1) and np.all(ids[:, 2] <= 4)
    valid = np.arange(4)[None, :] < ids[:, 2][:, None]
    n = cf[:, :, 6:9][valid]
    # (a contact whose two points coincide gets the zero vector: (p_point_A - p_point_B).normalized(), bradot_body_pair_3d.cpp:60;
    # its depth is 0 and the solver skips it.  In 336k contacts that happens once or not at all.)
    nn = np.linalg.norm(n, axis=1)
    assert np.all((np.abs(nn - 1.0) < 1e-5) | (nn == 0.0)) and int((nn == 0.0).sum()) <= 3
    assert np.all(np.isfinite(cf)) and np.all(cf[:, :, 9][valid] >= 0)
    xf = ps.bodies_read(server.FIELD_TRANSFORM)
    dyn = ~static
    assert np.all(np.isfinite(xf)) and np.all(xf[dyn, 10] > 0.0) and np.all(np.abs(xf[dyn, 9]) < 15.6) and np.all(np.abs(xf[dyn, 11]) < 15.6)
    ps.finish()


def test_full_size_c2_is_bit_reproducible():
    _, a = _run_c2(50)
    xa, va = a.bodies_read(server.FIELD_TRANSFORM), a.bodies_read(server.FIELD_LINEAR_VELOCITY)
    a.finish()
    _, b = _run_c2(50)
    xb, vb = b.bodies_read(server.FIELD_TRANSFORM), b.bodies_read(server.FIELD_LINEAR_VELOCITY)
    b.finish()
    assert np.array_equal(xa, xb) and np.array_equal(va, vb)


def test_mixed_pile_is_bit_reproducible_and_every_awake_body_is_stepped():
    """27 k mixed shapes in one growing island, 120 steps, two contexts stepped side by side: bodies, pair list and
    manifolds stay bit-identical.  This is the case that exposed a race in the island flattening (a body could keep a
    non-root island label for a step and was then skipped as if asleep)."""
    sc = scenes.mixed_pile(30)
    a = server.PhysicsServerCUDA(0)
    b = server.PhysicsServerCUDA(0)
    a.load_scene(sc)
    b.load_scene(sc)
    for s in range(120):
        a.step(sc.dt)
        b.step(sc.dt)
        if s % 20 == 19:
            for f in (server.FIELD_TRANSFORM, server.FIELD_LINEAR_VELOCITY, server.FIELD_ANGULAR_VELOCITY):
                assert np.array_equal(a.bodies_read(f), b.bodies_read(f)), (s, f)
            ia, pa, ca = a.pairs_read()
            ib, pb, cb = b.pairs_read()
            assert ia.shape == ib.shape and np.array_equal(ia, ib) and np.array_equal(pa, pb) and np.array_equal(ca, cb), s
    a.finish()
    b.finish()


def test_pipelined_read_back_matches_the_blocking_read():
    """bpc_bodies_read_begin / _end: the rows queued after step k arrive unchanged although step k + 1 is issued before they are
    awaited; one read at a time; fields other than the transform and the velocities are refused."""
    sc = scenes.mixed_pile(5, half=3.5)
    a, b = server.PhysicsServerCUDA(0), server.PhysicsServerCUDA(0)
    a.load_scene(sc)
    b.load_scene(sc)
    n = len(sc.bodies)
    import torch
    bufs = [torch.zeros((n, 12), dtype=torch.float32).pin_memory().numpy() for _ in range(2)]
    vel = torch.zeros((n, 3), dtype=torch.float32).pin_memory().numpy()
    expect = []
    for k in range(12):
        a.step(sc.dt)
        expect.append(a.bodies_read(server.FIELD_TRANSFORM).copy())
    for k in range(12):
        b.step(sc.dt)
        b.bodies_read_end()
        if k:
            assert np.array_equal(bufs[(k - 1) & 1], expect[k - 1]), k
        b.bodies_read_begin(server.FIELD_TRANSFORM, bufs[k & 1])
    with pytest.raises(Exception):
        b.bodies_read_begin(server.FIELD_TRANSFORM, bufs[0])          # the previous one was not ended
    b.bodies_read_end()
    assert np.array_equal(bufs[1], expect[11])
    b.bodies_read_begin(server.FIELD_LINEAR_VELOCITY, vel)
    b.bodies_read_end()
    assert np.array_equal(vel, a.bodies_read(server.FIELD_LINEAR_VELOCITY))
    with pytest.raises(Exception):
        b.bodies_read_begin(server.FIELD_ACTIVE, vel)
    b.bodies_read_end()                                               # nothing in flight: a no-op
    a.finish()
    b.finish()


def test_edge_cases_of_the_round_2_entry_points():
    """Empty batches, bad handles and refused combinations of the entry points added in round 2 answer with an error message
    (or with nothing), never with a crash or a silent ignore."""
    sc = scenes.spheres_container(2, 2, 2, half=1.5)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    ps.step(sc.dt)
    eye = np.array([1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 1, 0], np.float32)
    probe = ps.sphere_shape_create(0.3)
    # empty batches
    assert len(ps.rest_infos(np.zeros(0, np.int32), np.zeros((0, 12), np.float32))) == 0
    fr, info = ps.cast_motions(np.zeros(0, np.int32), np.zeros((0, 12), np.float32), np.zeros((0, 3), np.float32))
    assert len(fr) == 0 and len(info) == 0
    assert len(ps.bodies_test_motion(np.zeros(0, server.MOTION_PARAMS_DT))) == 0
    assert len(ps.area_events()) == 0
    ps.set_query_exclude([])
    # a zero motion: nothing in the way, fractions (1, 1)
    fr, info = ps.cast_motions([probe], eye, [[0, 0, 0]])
    assert fr[0, 0] == 1 and fr[0, 1] == 1
    # bad handles
    with pytest.raises(server.PhysicsError):
        ps.rest_infos([9999], eye)
    with pytest.raises(server.PhysicsError):
        ps.set_query_exclude([10 ** 6])
    bad = np.zeros(1, server.MOTION_PARAMS_DT)
    bad["body"] = 10 ** 6
    with pytest.raises(server.PhysicsError):
        ps.bodies_test_motion(bad)
    bad["body"], bad["max_collisions"] = 5, 33
    with pytest.raises(server.PhysicsError):
        ps.bodies_test_motion(bad)
    # areas: unknown space, world boundary as area shape, out-of-range override mode, writing a destroyed area
    ar = np.zeros(1, scenes.AREA_DT)
    ar["xf"], ar["shape_xf"], ar["shape"], ar["layer"], ar["mask"] = scenes.IDENT12, scenes.IDENT12, probe, 1, 1
    ar["space"] = 7
    with pytest.raises(server.PhysicsError):
        ps.areas_create(ar)
    ar["space"], ar["shape"] = 0, int(sc.bodies["shape"][0])          # body 0 is a world boundary
    with pytest.raises(server.PhysicsError):
        ps.areas_create(ar)
    ar["shape"], ar["gravity_mode"] = probe, 9
    with pytest.raises(server.PhysicsError):
        ps.areas_create(ar)
    ar["gravity_mode"] = 3
    first = ps.areas_create(ar)
    ps.step(sc.dt)
    ps.areas_destroy([first])
    with pytest.raises(server.PhysicsError):
        ps.areas_write(first, ar)
    ps.areas_destroy([first])                                          # twice: nothing happens
    ps.step(sc.dt)
    # meshes: only static bodies may carry one; an empty mesh and a tree that is not one are refused
    tri = np.array([[[0, 0, 0], [1, 0, 1], [1, 0, 0]]], np.float32)
    mesh = ps.concave_polygon_shape_create(tri)
    body = sc.bodies[-1:].copy()
    body["shape"], body["mode"] = mesh, scenes.MODE_RIGID
    with pytest.raises(server.PhysicsError):
        ps.bodies_create(body)
    body["mode"] = scenes.MODE_STATIC
    ps.bodies_create(body)
    with pytest.raises(server.PhysicsError):
        ps.concave_polygon_shape_create(np.zeros((0, 9), np.float32))
    with pytest.raises(server.PhysicsError):
        ps.concave_polygon_shape_create(tri, bvh=np.array([[0, 0, 0, 1, 0, 1, 5, 7, -1]], np.float32), normals=np.array([[0, 1, 0]], np.float32))
    with pytest.raises(server.PhysicsError):
        ps.heightmap_shape_create(np.zeros(1, np.float32), 1, 1, 0.0, 0.0)
    ps.step(sc.dt)
    assert np.all(np.isfinite(ps.bodies_read(server.FIELD_TRANSFORM)))
    ps.finish()
