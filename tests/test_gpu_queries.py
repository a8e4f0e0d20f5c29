"""Space queries: batched intersect_ray on the GPU-resident LBVH against the reference's
PhysicsDirectSpaceState3D::intersect_ray (oracle/_ref/oracle_step --rays) on the same state."""
import numpy as np
import pytest

import parity
from bradot_b200 import scenes, server

pytestmark = pytest.mark.gpu

POS_TOL = 1e-4   # metres, on positions of order 1-10 (the reference's own fp32 chain: world -> local -> world)
NRM_TOL = 1e-4


def _rays(n, seed, half, top):
    rng = np.random.default_rng(seed)
    frm = rng.uniform(-half, half, (n, 3))
    frm[:, 1] = rng.uniform(0.05, top, n)
    to = frm + rng.normal(size=(n, 3)) * half
    return np.concatenate([frm, to], 1).astype(np.float32)


def _compare(sc, rays, steps, check):
    _, dump = parity.run_oracle(sc, steps, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    seen_hits = 0
    for k in check:
        parity.load_state(ps, dump, k)
        hit, body, pos, nrm = ps.intersect_rays(rays)
        ref = dump[("rays", k)]
        rhit = ref[:, 0] > 0
        # a ray that grazes a shape within rounding may flip; everything else must agree exactly on hit and body
        same = hit == rhit
        assert same.mean() > 0.999, (k, int((~same).sum()))
        both = hit & rhit
        rbody = ref[:, 1].astype(np.int64)
        agree = body[both] == rbody[both]
        assert agree.mean() > 0.999, (k, int((~agree).sum()))
        sel = np.nonzero(both)[0][agree]
        assert np.max(np.abs(pos[sel] - ref[sel, 2:5])) < POS_TOL, k
        assert np.max(np.abs(nrm[sel] - ref[sel, 5:8])) < NRM_TOL, k
        seen_hits += len(sel)
    ps.finish()
    return seen_hits


def test_intersect_rays_mixed_pile_boxes_spheres_capsules_planes():
    parity_needed()
    sc = scenes.mixed_pile(5, half=3.5)
    assert _compare(sc, _rays(4000, 5, 4.0, 9.0), 60, [1, 20, 40, 59]) > 4000


def test_intersect_rays_convex_hulls():
    parity_needed()
    sc = scenes.convex_pile(4)
    assert _compare(sc, _rays(4000, 6, 4.0, 8.0), 50, [1, 25, 49]) > 2000


@pytest.mark.parametrize("name", ["mixed", "hulls"])
def test_closest_point_to_object_volume(name):
    """PhysicsDirectSpaceState3D::get_closest_point_to_object_volume (bradot_space_3d.cpp:573-611; get_closest_point_to of the plane,
    sphere, box, capsule and convex polygon): 3 000 (body, point) pairs per scene at three steps against the reference --
    points inside a shape come back unchanged, the others land on the surface; to 1e-4."""
    parity_needed()
    sc = scenes.mixed_pile(5, half=3.5) if name == "mixed" else scenes.convex_pile(4)
    rays = _rays(3000, 31, 3.6, 5.0)
    nb = len(sc.bodies)
    which = np.arange(len(rays)) % nb
    rays[:400, 0:3] = sc.bodies["xf"][which[:400], 9:12] + np.float32(0.02)      # these start inside their body (at the first step at least)
    _, dump = parity.run_oracle(sc, 40, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    unchanged = moved = 0
    for k in (1, 20, 39):
        parity.load_state(ps, dump, k)
        mine = ps.bodies_closest_points(which, rays[:, 0:3])
        ref = dump[("closest", k)]
        err = np.abs(mine - ref).max(axis=1)
        assert (err < 1e-4).mean() > 0.999 and np.median(err) < 1e-6, (k, float(err.max()), int((err >= 1e-4).sum()))
        same = np.all(ref == rays[:, 0:3], axis=1)
        unchanged += int(same.sum())
        moved += int((~same).sum())
    ps.finish()
    assert unchanged > 50 and moved > 5000


def test_intersect_rays_pick_ray_passes_unpickable_bodies():
    """RayParameters::pick_ray with body_set_ray_pickable(false) on every third body (bradot_space_3d.cpp:134): those bodies are
    passed through by picking rays and hit by ordinary ones, as in the reference."""
    parity_needed()
    sc = scenes.mixed_pile(4, half=3.0)
    dyn = np.nonzero(sc.bodies["mode"] == scenes.MODE_RIGID)[0]
    sc.bodies["flags"][dyn[::3]] |= scenes.BODY_NOT_PICKABLE
    sc.bodies["flags"][0] |= scenes.BODY_NOT_PICKABLE         # and one of the container's planes
    rays = _rays(3000, 9, 3.5, 6.0)
    _, dump = parity.run_oracle(sc, 40, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    differ = 0
    for k in (1, 20, 39):
        parity.load_state(ps, dump, k)
        for name, pick in (("rays", False), ("rays_pick", True)):
            hit, body, pos, nrm = ps.intersect_rays(rays, pick_ray=pick)
            ref = dump[(name, k)]
            rhit = ref[:, 0] > 0
            assert (hit == rhit).mean() > 0.999, (k, name)
            both = hit & rhit
            agree = body[both] == ref[both, 1].astype(np.int64)
            assert agree.mean() > 0.999, (k, name)
            sel = np.nonzero(both)[0][agree]
            assert np.max(np.abs(pos[sel] - ref[sel, 2:5])) < POS_TOL, (k, name)
            if pick:
                assert not np.any((sc.bodies["flags"][body[hit]] & scenes.BODY_NOT_PICKABLE) != 0)
        differ += int((dump[("rays", k)][:, 1] != dump[("rays_pick", k)][:, 1]).sum())
    ps.finish()
    assert differ > 300


def test_rays_and_points_with_areas():
    """collide_with_areas / collide_with_bodies of intersect_ray and intersect_point (_can_collide_with, bradot_space_3d.cpp:45-60):
    rays and points over the area scene answer with bodies and areas, with areas alone (picking rays: only some of the areas are
    pickable), as the reference does; an area's answer is -2 - its index."""
    parity_needed()
    sc = scenes.area_zones()
    pickable = [0, 2, 3, 5, 7]
    sc.areas["flags"][pickable] |= scenes.AREA_PICKABLE
    rng = np.random.default_rng(21)
    n = 3000
    frm = np.stack([rng.uniform(-3, 33, n), rng.uniform(0, 9, n), rng.uniform(-3, 3, n)], -1)
    to = frm + rng.normal(size=(n, 3)) * np.array([6.0, 3.0, 2.0])
    frm[:600] = sc.areas["xf"][rng.integers(0, len(sc.areas), 600), 9:12] + rng.normal(size=(600, 3)) * 0.6     # many start inside an area
    rays = np.concatenate([frm, to], 1).astype(np.float32)
    _, dump = parity.run_oracle(sc, 40, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    area_hits = 0
    for k in (1, 20, 39):
        parity.load_state(ps, dump, k)
        for name, kw in (("rays", {}), ("rays_areas", dict(areas=True)), ("rays_areas_only", dict(areas=True, bodies=False, pick_ray=True))):
            hit, body, pos, nrm = ps.intersect_rays(rays, **kw)
            ref = dump[(name, k)]
            rhit = ref[:, 0] > 0
            assert (hit == rhit).mean() > 0.998, (k, name, int((hit != rhit).sum()))
            both = hit & rhit
            agree = body[both] == ref[both, 1].astype(np.int64)
            assert agree.mean() > 0.998, (k, name, int((~agree).sum()))
            sel = np.nonzero(both)[0][agree]
            assert np.max(np.abs(pos[sel] - ref[sel, 2:5])) < POS_TOL and np.max(np.abs(nrm[sel] - ref[sel, 5:8])) < NRM_TOL, (k, name)
            if name == "rays_areas_only":
                assert np.all(body[hit] <= -2) and np.all(np.isin(-2 - body[hit], pickable))
                area_hits += int(hit.sum())
        counts, listed = ps.intersect_points(rays[:, 0:3], max_results=16, areas=True)
        refp = dump[("points_areas", k)]
        differ = sum(set(listed[i, :counts[i]].tolist()) != set(refp[i, 1:1 + refp[i, 0]].tolist()) for i in range(n))
        assert differ <= 3, (k, differ)
        assert (refp[:, 1:] <= -2).sum() > 400
        # intersect_shape with collide_with_areas: the query shapes of the dump at its transforms
        xf = dump[("shape_query_xf", k)]
        if k == 1:
            shape_of = _shape_queries(ps, rays)
        cs, ls = ps.intersect_shapes(shape_of, xf, max_results=16, areas=True)
        refs = dump[("shape_queries_areas", k)]
        differ = sum(set(ls[i, :cs[i]].tolist()) != set(refs[i, 1:1 + refs[i, 0]].tolist()) for i in range(n))
        assert differ <= n // 500, (k, differ)
        assert (refs[:, 1:] <= -2).sum() > 600
    ps.finish()
    assert area_hits > 500


def test_intersect_rays_height_map():
    """BradotHeightMapShape3D::intersect_segment (bradot_shape_3d.cpp:1950-2008; the cell walk of _intersect_grid_segment,
    :1810-1948): rays over the height-map scene -- steep ones that stay inside one cell, long oblique ones that walk many
    cells, rays from below (back faces: hit only inside one cell), rays along the lanes."""
    parity_needed()
    sc = scenes.heightmap_pile()
    rays = _rays(4000, 43, 6.5, 4.0)
    rng = np.random.default_rng(7)
    rays[:1000, 1] = rng.uniform(1.0, 3.0, 1000)
    rays[:1000, 3:6] = rays[:1000, 0:3] + np.array([0, -5.0, 0]) + rng.normal(size=(1000, 3)) * np.array([0.1, 0, 0.1])      # steep
    rays[1000:1500, 1] = rng.uniform(-3.0, -1.0, 500)
    rays[1000:1500, 3:6] = rays[1000:1500, 0:3] + np.array([0, 5.0, 0]) + rng.normal(size=(500, 3)) * np.array([0.3, 0, 0.3])  # from below
    rays[1500:1700, 0] = np.round(rays[1500:1700, 0])                                                                       # starting on a lane
    rays[1700:1900, 3] = rays[1700:1900, 0]                                                                                 # along z only
    rays = rays.astype(np.float32)
    _, dump = parity.run_oracle(sc, 3, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    parity.load_state(ps, dump, 1)
    hit, body, pos, nrm = ps.intersect_rays(rays)
    ref = dump[("rays", 1)]
    rhit = ref[:, 0] > 0
    same = hit == rhit
    assert same.mean() > 0.998, int((~same).sum())
    both = hit & rhit
    agree = body[both] == ref[both, 1].astype(np.int64)
    assert agree.mean() > 0.998, int((~agree).sum())
    sel = np.nonzero(both)[0][agree]
    close = (np.abs(pos[sel] - ref[sel, 2:5]).max(axis=1) < POS_TOL) & (np.abs(nrm[sel] - ref[sel, 5:8]).max(axis=1) < NRM_TOL)
    assert close.mean() > 0.998, int((~close).sum())
    on_map = body[sel] == 1
    assert on_map.sum() > 1200 and (on_map & (sel < 1000)).sum() > 500
    ps.finish()


def test_intersect_rays_edge_cases():
    sc = scenes.spheres_container(3, 3, 3, half=1.75)
    ps = server.PhysicsServerCUDA(0)
    ps.load_scene(sc)
    hit, body, pos, nrm = ps.intersect_rays(np.zeros((0, 6), np.float32))
    assert len(hit) == 0
    # degenerate segment (from == to), a ray that misses everything, a ray from far above straight down onto the floor
    rays = np.array([[0, 5, 0, 0, 5, 0], [100, 100, 100, 101, 100, 100], [1.6, 50, 1.6, 1.6, -1, 1.6]], np.float32)
    hit, body, pos, nrm = ps.intersect_rays(rays)
    assert not hit[0] and not hit[1] and body[1] == -1
    assert hit[2] and abs(pos[2, 1]) < 1e-5 and np.allclose(nrm[2], [0, 1, 0])
    # collision mask 0 sees nothing; an unknown space is refused
    hit, _, _, _ = ps.intersect_rays(rays, collision_mask=0)
    assert not hit.any()
    with pytest.raises(server.PhysicsError):
        ps.intersect_rays(rays, space=7)
    ps.finish()


def parity_needed():
    if not parity.have_oracle():
        pytest.skip("oracle/_ref/oracle_step not built")


def test_intersect_points_against_the_reference():
    parity_needed()
    sc = scenes.mixed_pile(5, half=3.5)
    rng = np.random.default_rng(9)
    n = 6000
    pts = rng.uniform(-3.4, 3.4, (n, 3))
    pts[:, 1] = rng.uniform(-0.5, 4.0, n)          # some below the floor plane: inside the world boundary
    rays = np.concatenate([pts, pts + 1.0], 1).astype(np.float32)
    _, dump = parity.run_oracle(sc, 60, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    inside = 0
    for k in (1, 30, 59):
        parity.load_state(ps, dump, k)
        counts, bodies = ps.intersect_points(rays[:, 0:3], max_results=8)
        ref = dump[("points", k)]
        differ = 0
        for i in range(n):
            a = set(bodies[i, :counts[i]].tolist())
            b = set(ref[i, 1:1 + ref[i, 0]].tolist())
            differ += a != b
            inside += len(b)
        assert differ <= n // 1000, (k, differ)      # a point within rounding of a surface may flip
    assert inside > 1000
    ps.finish()


def test_intersect_shapes_against_the_reference():
    """intersect_shape: spheres, boxes and capsules placed at 3 000 poses (the harness's, dumped as "shape_query_xf") in a
    mixed pile and in a hull pile; the set of overlapped bodies equals the reference's -- with margin 0 (the plain collision
    routines) and with margin 0.05 (the withMargin ones)."""
    parity_needed()
    for sc, seed in ((scenes.mixed_pile(5, half=3.5), 12), (scenes.convex_pile(4), 13)):
        rays = _rays(3000, seed, 3.4, 4.0)
        _, dump = parity.run_oracle(sc, 60, dump_from=0, dump_every=1, rays=rays)
        ps = parity.new_server(sc, dump)
        handles = np.array([ps.sphere_shape_create(0.35), ps.box_shape_create((0.3, 0.2, 0.25)), ps.capsule_shape_create(0.2, 0.9)], np.int32)
        shape_of = handles[np.arange(len(rays)) % 3]
        overlaps = 0
        for k, margin in ((1, 0.0), (30, 0.0), (59, 0.0), (30, 0.05), (59, 0.05)):
            parity.load_state(ps, dump, k)
            counts, bodies = ps.intersect_shapes(shape_of, dump[("shape_query_xf", k)], max_results=8, margin=margin)
            ref = dump[("shape_queries_margin" if margin else "shape_queries", k)]
            differ = 0
            for i in range(len(rays)):
                if ref[i, 0] >= 8 or counts[i] >= 8:
                    continue                    # the cut at max_results follows each side's own tree order
                a = set(bodies[i, :counts[i]].tolist())
                b = set(ref[i, 1:1 + ref[i, 0]].tolist())
                differ += a != b
                overlaps += len(b)
            assert differ <= len(rays) // 1000, (k, differ)     # a shape touching within rounding may flip
        assert overlaps > 1500
        ps.finish()


def test_collide_shapes_against_the_reference():
    """collide_shape: the contact pairs (point on the query shape, point on the body) for the same query shapes and poses."""
    parity_needed()
    sc = scenes.mixed_pile(5, half=3.5)
    rays = _rays(2000, 21, 3.4, 4.0)
    _, dump = parity.run_oracle(sc, 60, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    handles = np.array([ps.sphere_shape_create(0.35), ps.box_shape_create((0.3, 0.2, 0.25)), ps.capsule_shape_create(0.2, 0.9)], np.int32)
    shape_of = handles[np.arange(len(rays)) % 3]
    pairs_seen = 0
    for k in (1, 30, 59):
        parity.load_state(ps, dump, k)
        counts, pts = ps.collide_shapes(shape_of, dump[("shape_query_xf", k)], max_results=16)
        ref = dump[("shape_collide", k)]
        differ = 0
        for i in range(len(rays)):
            rc = int(ref[i, 0])
            if rc >= 16 or counts[i] >= 16:
                continue                      # which pairs survive a full list depends on the order of the candidates
            if rc != counts[i]:
                differ += 1
                continue
            if rc == 0:
                continue
            a = pts[i, :rc].reshape(rc, 6)
            b = ref[i, 1:1 + rc * 6].reshape(rc, 6)
            # same set of pairs (the candidates come in a different order): match each reference pair to its nearest
            dist = np.abs(a[:, None, :] - b[None, :, :]).max(axis=2)
            if dist.min(axis=0).max() > 1e-4 or dist.min(axis=1).max() > 1e-4:
                differ += 1
            pairs_seen += rc
        assert differ <= len(rays) // 500, (k, differ)
    assert pairs_seen > 1000
    ps.finish()


def _shape_queries(ps, rays):
    handles = np.array([ps.sphere_shape_create(0.35), ps.box_shape_create((0.3, 0.2, 0.25)), ps.capsule_shape_create(0.2, 0.9)], np.int32)
    return handles[np.arange(len(rays)) % 3]


def test_rest_info_against_the_reference():
    """rest_info (bradot_space_3d.cpp:513-569; the withMargin collision routines + the rest callback): same query shapes and
    poses as above, margin 0.04 on every other query and the 0.0001 floor on the rest.  Hit / no hit agrees except where a
    contact is within rounding of the least allowed depth; where both hit the same body, point and normal to 1e-4."""
    parity_needed()
    for sc, seed in ((scenes.mixed_pile(5, half=3.5), 21), (scenes.convex_pile(4), 22)):
        rays = _rays(3000, seed, 3.4, 4.0)
        _, dump = parity.run_oracle(sc, 60, dump_from=0, dump_every=1, rays=rays)
        ps = parity.new_server(sc, dump)
        shape_of = _shape_queries(ps, rays)
        d = rays[:, 3:6] - rays[:, 0:3]
        motions = (d / np.linalg.norm(d, axis=1, keepdims=True) * np.float32(0.01)).astype(np.float32)
        odd = np.arange(len(rays)) % 2 == 1
        hits = 0
        for k in (1, 30, 59):
            parity.load_state(ps, dump, k)
            xf = dump[("shape_query_xf", k)]
            out = np.zeros(len(rays), server.REST_INFO_DT)
            out[odd] = ps.rest_infos(shape_of[odd], xf[odd], motions[odd], margin=0.04)
            out[~odd] = ps.rest_infos(shape_of[~odd], xf[~odd], motions[~odd], margin=0.0)
            ref = dump[("shape_rest", k)]
            rhit = ref[:, 0] > 0
            same = (out["hit"] > 0) == rhit
            assert same.mean() > 0.998, (k, int((~same).sum()))
            both = (out["hit"] > 0) & rhit
            agree = out["body"][both] == ref[both, 1].astype(np.int32)
            assert agree.mean() > 0.99, (k, int((~agree).sum()))       # two bodies equally deep: either may be "the deepest"
            sel = np.nonzero(both)[0][agree]
            close = (np.abs(out["point"][sel] - ref[sel, 2:5]).max(axis=1) < POS_TOL) & (np.abs(out["normal"][sel] - ref[sel, 5:8]).max(axis=1) < NRM_TOL)
            assert close.mean() > 0.995, (k, int((~close).sum()))      # same body, two contacts equally deep (a box lying flat)
            assert np.max(np.abs(out["linear_velocity"][sel][close] - ref[sel, 8:11][close])) < 1e-3
            hits += len(sel)
        assert hits > 1500
        ps.finish()


def test_cast_motion_against_the_reference():
    """cast_motion (bradot_space_3d.cpp:264-381; GJK distance on the swept shape, 8 bisection steps): the (safe, unsafe)
    fractions are dyadic numbers chosen by distance tests, so they either equal the reference's exactly or -- where a test sits
    within GJK's own 1e-4 accuracy of touching -- differ by a bisection step, or by a whole obstacle when that test was the
    "starts inside, ignore it" one; the stopping body and the closest point on it agree."""
    parity_needed()
    for sc, seed in ((scenes.mixed_pile(5, half=3.5), 23), (scenes.convex_pile(4), 24)):
        rays = _rays(3000, seed, 3.4, 4.0)
        _, dump = parity.run_oracle(sc, 60, dump_from=0, dump_every=1, rays=rays)
        ps = parity.new_server(sc, dump)
        shape_of = _shape_queries(ps, rays)
        motions = ((rays[:, 3:6] - rays[:, 0:3]) * np.float32(0.5)).astype(np.float32)
        stopped = 0
        for k in (1, 30, 59):
            parity.load_state(ps, dump, k)
            fr, info = ps.cast_motions(shape_of, dump[("shape_query_xf", k)], motions)
            ref = dump[("shape_cast", k)]
            exact = np.all(fr == ref[:, 0:2], axis=1)
            assert exact.mean() > 0.99, (k, int((~exact).sum()), np.abs(fr - ref[:, 0:2]).max())
            both = exact & (info["hit"] > 0) & (ref[:, 2] > 0)
            assert np.array_equal((info["hit"] > 0)[exact], (ref[:, 2] > 0)[exact])
            agree = info["body"][both] == ref[both, 3].astype(np.int32)
            assert agree.mean() > 0.99, (k, int((~agree).sum()))
            sel = np.nonzero(both)[0][agree]
            close = np.abs(info["point"][sel] - ref[sel, 4:7]).max(axis=1) < 2e-3    # closest points of GJK: its accuracy is 1e-4 relative
            assert close.mean() > 0.98, (k, int((~close).sum()))
            stopped += int(((fr[:, 0] < 1) & exact).sum())
        assert stopped > 500
        ps.finish()


def test_query_exclude_lists_against_the_reference():
    """RayParameters / ShapeParameters ::exclude: every ray and shape query again with the body its first answer named excluded
    (the harness does the same through the reference's exclude sets)."""
    parity_needed()
    sc = scenes.mixed_pile(5, half=3.5)
    rays = _rays(600, 31, 3.4, 4.0)
    _, dump = parity.run_oracle(sc, 40, dump_from=0, dump_every=1, rays=rays)
    ps = parity.new_server(sc, dump)
    shape_of = _shape_queries(ps, rays)
    k = 39
    parity.load_state(ps, dump, k)
    ref_r, ref_s, first_r, first_s = dump[("rays_excl", k)], dump[("shapes_excl", k)], dump[("rays", k)], dump[("shape_queries", k)]
    xf = dump[("shape_query_xf", k)]
    ray_bad = shape_bad = changed = 0
    for i in range(len(rays)):
        ex = [int(first_r[i, 1])] if first_r[i, 0] > 0 and first_r[i, 1] >= 0 else []
        ps.set_query_exclude(ex)
        hit, body, _, _ = ps.intersect_rays(rays[i:i + 1])
        ray_bad += bool(hit[0]) != bool(ref_r[i, 0]) or (hit[0] and body[0] != ref_r[i, 1])
        changed += bool(ex) and (not hit[0] or body[0] != ex[0])
        ex = [int(first_s[i, 1])] if first_s[i, 0] > 0 and first_s[i, 1] >= 0 else []
        ps.set_query_exclude(ex)
        counts, bodies = ps.intersect_shapes(shape_of[i:i + 1], xf[i:i + 1], max_results=8)
        if ref_s[i, 0] < 8 and counts[0] < 8:
            shape_bad += set(bodies[0, :counts[0]].tolist()) != set(ref_s[i, 1:1 + ref_s[i, 0]].tolist())
        assert not ex or ex[0] not in bodies[0, :counts[0]]
    ps.set_query_exclude([])
    assert ray_bad <= 1 and shape_bad <= 1, (ray_bad, shape_bad)
    assert changed > 100
    hit, body, _, _ = ps.intersect_rays(rays)
    assert np.array_equal(hit, first_r[:, 0] > 0) or (hit != (first_r[:, 0] > 0)).sum() <= 1       # the list is cleared again
    ps.finish()


def test_body_test_motion_against_the_reference():
    """body_test_motion (bradot_space_3d.cpp:651-1029: recovery, cast, rest contacts) for every third body of a settled pile,
    each from where it stands -- touching its neighbours, so the recovery step has work -- with margins 0.001 / 0.04, 1 or 4
    collisions asked, recovery_as_collision on a third.  The requests are the harness's (dump "motion_queries").  Collided or
    not agrees; the safe / unsafe fractions are equal except where a distance test sits within GJK's accuracy of touching.  Most
    answers are equal to the last bit (median and 90th percentile of every difference: 0); the recovery is a sequential
    projection over the contacts in tree order, which differs between the two trees, so a body wedged between several
    neighbours can come out a few millimetres apart (travel: 95 % within 5 mm, all within 5 cm) and then name another
    neighbour as its deepest contact."""
    parity_needed()
    for sc, steps in ((scenes.mixed_pile(5, half=3.5), 60), (scenes.convex_pile(4), 60)):
        _, dump = parity.run_oracle(sc, steps, dump_from=0, dump_every=1, motions=True)
        ps = parity.new_server(sc, dump)
        total = hits = 0
        for k in (20, 40, 59):
            parity.load_state(ps, dump, k)
            q, ref = dump[("motion_queries", k)], dump[("motion_results", k)]
            prm = np.zeros(len(q), server.MOTION_PARAMS_DT)
            prm["body"], prm["from"], prm["motion"] = q[:, 0].astype(np.int32), q[:, 1:13], q[:, 13:16]
            prm["margin"], prm["max_collisions"], prm["flags"] = q[:, 16], q[:, 17].astype(np.int32), q[:, 18].astype(np.uint32)
            out = ps.bodies_test_motion(prm)
            same = (out["collided"] > 0) == (ref[:, 0] > 0)
            assert same.mean() > 0.98, (k, int((~same).sum()))
            both = same & (ref[:, 0] > 0)
            exact = (out["collision_safe_fraction"] == ref[:, 9]) & (out["collision_unsafe_fraction"] == ref[:, 10])
            assert exact[same].mean() > 0.9, (k, int((~exact[same]).sum()))
            ok = same & exact
            dtravel = np.abs(out["travel"] - ref[:, 2:5]).max(axis=1)
            assert np.percentile(dtravel[ok], 50) == 0 and np.percentile(dtravel[ok], 95) < 5e-3 and dtravel[ok].max() < 5e-2, (k, dtravel[ok].max())
            sel = np.nonzero(ok & both)[0]
            coll = out["collisions"][sel, 0]
            agree = coll["collider"] == ref[sel, 11].astype(np.int32)
            assert agree.mean() > 0.9, (k, int((~agree).sum()))
            s2 = sel[agree]
            c2 = out["collisions"][s2, 0]
            dpos = np.abs(c2["position"] - ref[s2, 12:15]).max(axis=1)
            dnrm = np.abs(c2["normal"] - ref[s2, 15:18]).max(axis=1)
            assert np.percentile(dpos, 50) == 0 and np.percentile(dpos, 95) < 3e-3 and np.percentile(dnrm, 95) < 3e-3, (k, np.percentile(dpos, 95), np.percentile(dnrm, 95))
            assert np.percentile(np.abs(out["collision_depth"][s2] - ref[s2, 8]), 95) < 2e-3
            many = s2[(ref[s2, 1] > 1)]
            assert len(many) == 0 or np.mean(out["collision_count"][many] == ref[many, 1].astype(np.int32)) >= 0.8
            total += len(q)
            hits += len(sel)
        assert total > 60 and hits > 30, (total, hits)
        ps.finish()


def test_body_test_motion_weighs_obstacles_by_collision_priority():
    """body_set_collision_priority (bradot_space_3d.cpp:748-762): the recovery step of body_test_motion pushes a stuck body away
    from its obstacles in proportion to their priorities.  The settled pile of the test above with priority 4 on every other
    body and 0.2 on the floor: the travel agrees with the reference running the same priorities (the bars of that test), and is
    closer to it than what the same requests give when the priorities are left out."""
    parity_needed()
    sc = scenes.mixed_pile(5, half=3.5)
    nb = len(sc.bodies)
    prio = np.ones(nb, np.float32)
    prio[1::2] = 4.0
    prio[0] = 0.2
    events = [(1, int(b), 12, np.array([prio[b]] + [0] * 11, np.float32)) for b in range(nb) if prio[b] != 1.0]
    steps = 60
    _, dump = parity.run_oracle(sc, steps, dump_from=0, dump_every=1, motions=True, events=events)
    ps = parity.new_server(sc, dump)
    worse = better = 0
    for with_prio in (False, True):
        if with_prio:
            ps.bodies_set_collision_priority(np.arange(nb), prio)
        errs = []
        for k in (20, 40, 59):
            parity.load_state(ps, dump, k)
            q, ref = dump[("motion_queries", k)], dump[("motion_results", k)]
            prm = np.zeros(len(q), server.MOTION_PARAMS_DT)
            prm["body"], prm["from"], prm["motion"] = q[:, 0].astype(np.int32), q[:, 1:13], q[:, 13:16]
            prm["margin"], prm["max_collisions"], prm["flags"] = q[:, 16], q[:, 17].astype(np.int32), q[:, 18].astype(np.uint32)
            out = ps.bodies_test_motion(prm)
            same = (out["collided"] > 0) == (ref[:, 0] > 0)
            exact = (out["collision_safe_fraction"] == ref[:, 9]) & (out["collision_unsafe_fraction"] == ref[:, 10])
            dtravel = np.abs(out["travel"] - ref[:, 2:5]).max(axis=1)
            errs.append(dtravel)
            if with_prio:
                assert same.mean() > 0.98, (k, int((~same).sum()))
                ok = same & exact
                assert np.percentile(dtravel[ok], 50) == 0 and np.percentile(dtravel[ok], 95) < 5e-3, (k, float(np.percentile(dtravel[ok], 95)))
        e = np.concatenate(errs)
        if with_prio:
            better = float(np.mean(e == 0))
        else:
            worse = float(np.mean(e == 0))
    ps.finish()
    assert better > worse + 0.1, (better, worse)       # share of requests whose travel is equal to the last bit


def test_queries_against_a_triangle_mesh():
    """intersect_ray (BradotConcavePolygonShape3D::intersect_segment, bradot_shape_3d.cpp:1350-1423), intersect_shape, collide_shape
    and rest_info (solve_concave + the *_face routines) where the space holds triangle meshes: the terrain scene, the GPU using
    the reference's own tree over the faces.  Same bars as for the other shapes."""
    parity_needed()
    sc = scenes.terrain_pile()
    rays = _rays(3000, 41, 3.8, 4.0)
    rays[:1500, 1] = np.random.default_rng(1).uniform(1.5, 4.0, 1500)       # half of them start above the terrain ...
    rays[:1500, 3:6] = rays[:1500, 0:3] + np.random.default_rng(2).normal(size=(1500, 3)) * 0.5 + np.array([0, -4.0, 0])   # ... and point down at it
    rays = rays.astype(np.float32)
    _, dump = parity.run_oracle(sc, 60, dump_from=0, dump_every=1, rays=rays, motions=True)
    ps = parity.new_server(sc, dump)
    shape_of = _shape_queries(ps, rays)
    mesh_hits = mesh_stops = 0
    for k in (1, 30, 59):
        parity.load_state(ps, dump, k)
        hit, body, pos, nrm = ps.intersect_rays(rays)
        ref = dump[("rays", k)]
        rhit = ref[:, 0] > 0
        same = hit == rhit
        assert same.mean() > 0.999, (k, int((~same).sum()))
        both = hit & rhit
        agree = body[both] == ref[both, 1].astype(np.int64)
        assert agree.mean() > 0.999, (k, int((~agree).sum()))
        sel = np.nonzero(both)[0][agree]
        assert np.max(np.abs(pos[sel] - ref[sel, 2:5])) < POS_TOL and np.max(np.abs(nrm[sel] - ref[sel, 5:8])) < NRM_TOL, k
        mesh_hits += int(np.isin(body[sel], (1, 2)).sum())
        # RayResult::face_index: the face of the mesh that was hit (its index in the shape's face list), -1 for every other shape
        faces, ref_faces = ps.last_ray_faces, dump[("ray_faces", k)]
        on_mesh = np.isin(body[sel], (1, 2))
        assert np.array_equal(faces[sel], ref_faces[sel]) and np.all(faces[sel][on_mesh] >= 0) and np.all(faces[sel][~on_mesh] == -1), k
        xf = dump[("shape_query_xf", k)]
        counts, bodies = ps.intersect_shapes(shape_of, xf, max_results=8)
        refq = dump[("shape_queries", k)]
        differ = sum(set(bodies[i, :counts[i]].tolist()) != set(refq[i, 1:1 + refq[i, 0]].tolist()) for i in range(len(rays)) if refq[i, 0] < 8 and counts[i] < 8)
        assert differ <= len(rays) // 500, (k, differ)
        d = rays[:, 3:6] - rays[:, 0:3]
        motions = (d / np.linalg.norm(d, axis=1, keepdims=True) * np.float32(0.01)).astype(np.float32)
        odd = np.arange(len(rays)) % 2 == 1
        out = np.zeros(len(rays), server.REST_INFO_DT)
        out[odd] = ps.rest_infos(shape_of[odd], xf[odd], motions[odd], margin=0.04)
        out[~odd] = ps.rest_infos(shape_of[~odd], xf[~odd], motions[~odd], margin=0.0)
        refr = dump[("shape_rest", k)]
        sameh = (out["hit"] > 0) == (refr[:, 0] > 0)
        assert sameh.mean() > 0.995, (k, int((~sameh).sum()))
        bothr = (out["hit"] > 0) & (refr[:, 0] > 0)
        agr = out["body"][bothr] == refr[bothr, 1].astype(np.int32)
        assert agr.mean() > 0.98, (k, int((~agr).sum()))
        s2 = np.nonzero(bothr)[0][agr]
        close = (np.abs(out["point"][s2] - refr[s2, 2:5]).max(axis=1) < POS_TOL) & (np.abs(out["normal"][s2] - refr[s2, 5:8]).max(axis=1) < NRM_TOL)
        assert close.mean() > 0.98, (k, int((~close).sum()))
        assert int(np.isin(out["body"][s2], (1, 2)).sum()) > 20
        # cast_motion: the GJK distance to the faces of the mesh (solve_distance's concave branch)
        fr, info = ps.cast_motions(shape_of, xf, ((rays[:, 3:6] - rays[:, 0:3]) * np.float32(0.5)).astype(np.float32))
        refc = dump[("shape_cast", k)]
        exact = np.all(fr == refc[:, 0:2], axis=1)
        assert exact.mean() > 0.99, (k, int((~exact).sum()))
        stop = exact & (info["hit"] > 0) & (refc[:, 2] > 0)
        same_body = info["body"][stop] == refc[stop, 3].astype(np.int32)
        assert same_body.mean() > 0.98
        mesh_stops += int(np.isin(info["body"][stop][same_body], (1, 2)).sum())
        # body_test_motion of the bodies lying on the mesh
        q, refm = dump[("motion_queries", k)], dump[("motion_results", k)]
        prm = np.zeros(len(q), server.MOTION_PARAMS_DT)
        prm["body"], prm["from"], prm["motion"] = q[:, 0].astype(np.int32), q[:, 1:13], q[:, 13:16]
        prm["margin"], prm["max_collisions"], prm["flags"] = q[:, 16], q[:, 17].astype(np.int32), q[:, 18].astype(np.uint32)
        mo = ps.bodies_test_motion(prm)
        assert np.array_equal(mo["collided"] > 0, refm[:, 0] > 0)
        hitm = refm[:, 0] > 0
        assert (not hitm.any() or np.mean(mo["collision_safe_fraction"][hitm] == refm[hitm, 9]) >= 0.8) and np.max(np.abs(mo["travel"] - refm[:, 2:5])) < 2e-2
        motion_hits = motion_hits + int(hitm.sum()) if k > 1 else 0
    assert mesh_hits > 1000 and mesh_stops > 300 and motion_hits > 5
    ps.finish()
