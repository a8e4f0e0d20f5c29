"""Convex hull of a point cloud in the layout the C ABI takes (bpc_shape_create_convex), for host programs that are
not the engine (bench.py, scripts): the engine module gets the same data from Geometry3D::MeshData
(ConvexHullComputer, core/math/convex_hull.cpp, called at modules/bradot_physics_3d/bradot_shape_3d.cpp:1099).

Conventions, as the engine's MeshData has them: plane normals point out of the hull; a face lists its vertices
clockwise when seen from outside; an edge (a, b, face_a, face_b) is walked b->a by face_a and a->b by face_b.
Qhull (scipy.spatial.ConvexHull) finds the facets; coplanar triangles are merged into one polygon.
"""
import numpy as np


def mesh_data(points, merge_tol=1e-6):
    from scipy.spatial import ConvexHull

    pts = np.asarray(points, np.float64).reshape(-1, 3)
    hull = ConvexHull(pts)
    used = np.unique(hull.simplices)
    remap = -np.ones(len(pts), np.int64)
    remap[used] = np.arange(len(used))
    verts = pts[used]
    tris = remap[hull.simplices]
    eq = hull.equations  # n.x + off = 0, n outward
    # orient every triangle clockwise from outside
    for t in range(len(tris)):
        a, b, c = verts[tris[t]]
        if np.dot(np.cross(b - a, c - a), eq[t, :3]) > 0:
            tris[t] = tris[t][[0, 2, 1]]
    # group coplanar triangles
    group = -np.ones(len(tris), np.int64)
    planes = []
    for t in range(len(tris)):
        for g, p in enumerate(planes):
            if np.dot(p[:3], eq[t, :3]) > 1.0 - merge_tol and abs(p[3] - eq[t, 3]) < merge_tol:
                group[t] = g
                break
        else:
            group[t] = len(planes)
            planes.append(eq[t].copy())
    faces, index, out_planes = [], [], []
    owner = {}  # directed edge -> face
    for g, p in enumerate(planes):
        directed = set()
        for t in np.nonzero(group == g)[0]:
            i, j, k = tris[t]
            directed.update([(i, j), (j, k), (k, i)])
        boundary = {a: b for (a, b) in directed if (b, a) not in directed}
        start = min(boundary)
        loop, cur = [start], boundary[start]
        while cur != start:
            loop.append(cur)
            cur = boundary[cur]
        assert len(loop) == len(boundary), "face boundary is not a single loop"
        faces.append((len(index), len(loop)))
        for q in range(len(loop)):
            owner[(loop[q], loop[(q + 1) % len(loop)])] = g
        index.extend(loop)
        out_planes.append((p[0], p[1], p[2], -p[3]))
    edges = []
    for (a, b), f in sorted(owner.items()):
        if a < b:
            edges.append((a, b, owner[(b, a)], f))
    return {"verts": verts.astype(np.float32), "planes": np.asarray(out_planes, np.float32),
            "faces": np.asarray(faces, np.int32), "index": np.asarray(index, np.int32), "edges": np.asarray(edges, np.int32)}
