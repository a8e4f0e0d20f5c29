"""CPU-side tests (no GPU): the C-ABI library loads and exports every symbol the header declares,
scene / dump file formats round-trip, the scene generators are deterministic, and the host-side
multi-rank plumbing of bench.py works over gloo with world_size 2."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from bradot_b200 import scenes, server

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "bradot_physics_cuda.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bpc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(server.LIB_PATH):
        from bradot_b200 import build
        build.build()
    lib = ctypes.CDLL(server.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "symbol %s declared in the header is not exported" % n
    # and the Python stub binds the same set
    assert set(server.EXPORTS) == set(names), set(server.EXPORTS) ^ set(names)


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(server.PhysicsError):
        server.load_library(str(tmp_path / "nope.so"))


def test_no_cuda_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(server.PhysicsError):
        server.PhysicsServerCUDA(0)   # no CPU fallback: construction must raise


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "bradot_b200")):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text, f
            if f.endswith(".py") and f != "build.py":
                # the product's Python may mention the oracle in prose, but must never run or import it
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "subprocess" not in text and "build_ref" not in text and "import parity" not in text, f


def test_scene_file_roundtrip(tmp_path):
    sc = scenes.convex_chains(n_chains=3, chain_len=4, n_hulls=2)
    p = str(tmp_path / "s.bscn")
    sc.write(p)
    back = scenes.Scene.read(p)
    assert back.n_spaces == sc.n_spaces and back.iterations == sc.iterations
    for a, b in ((sc.shapes, back.shapes), (sc.bodies, back.bodies), (sc.joints, back.joints)):
        assert a.tobytes() == b.tobytes()
    assert np.array_equal(sc.verts, back.verts)
    assert scenes.BODY_DT.itemsize == 4 * 4 + 12 * 4 * 2 + 6 * 4 + 6 * 4 + 8   # must match bpc_body_desc / ScnBody


def test_scene_file_carries_the_round_2_sections(tmp_path):
    """Areas (with groups), long joint parameter rows, per-shape solver bias and the body flags added in round 2 survive a
    write / read, the record sizes are those of the C ABI, and the generators place what their docstrings say."""
    for sc in (scenes.slider_rigs(4), scenes.sixdof_rigs(4), scenes.area_groups(), scenes.level_geometry(2), scenes.kinematic_platform(rails=True)):
        p = str(tmp_path / "s.bscn")
        sc.write(p)
        back = scenes.Scene.read(p)
        for name in ("shapes", "bodies", "joints", "areas", "joint_ext"):
            assert getattr(sc, name).tobytes() == getattr(back, name).tobytes(), (sc.name, name)
    assert scenes.JOINT_DT.itemsize == 160 and scenes.AREA_DT.itemsize == 164 and scenes.JOINT_EXT_DT.itemsize == 128 and scenes.SHAPE_DT.itemsize == 32
    six = scenes.sixdof_rigs(6)
    assert (six.joints["type"] == scenes.JOINT_6DOF).all() and set(six.joint_ext["first"]) == {0, 28, 56} and (six.joint_ext["count"] == 28).all()
    assert np.array_equal(scenes.g6dof_axis_defaults()[[2, 3, 4, 10, 11, 15, 18, 22]], np.array([0.7, 0.5, 1.0, -1e30, 1e30, 300.0, 0.1, 1.0], np.float32))
    ag = scenes.area_groups()
    assert list(ag.areas["group"]) == [1, 1, 0, 2, 2]
    lv = scenes.level_geometry(2)
    assert int(((lv.bodies["flags"] & scenes.BODY_FURTHER_SHAPE) != 0).sum()) == 5 and not (scenes.level_geometry(2, merged=False).bodies["flags"] & scenes.BODY_FURTHER_SHAPE).any()
    ch = scenes.hanging_chains(3, 5, cone_twist=True)
    assert (ch.joints["type"] == scenes.JOINT_CONE_TWIST).sum() == 6


def test_scene_generators_are_deterministic_and_sized():
    a, b = scenes.mixed_pile(4), scenes.mixed_pile(4)
    assert a.bodies.tobytes() == b.bodies.tobytes() and a.shapes.tobytes() == b.shapes.tobytes()
    assert len(scenes.pyramid(1000).bodies) == 1001
    c2 = scenes.spheres_container(5, 4, 5)
    assert len(c2.bodies) == 5 + 100 and int(np.sum(c2.bodies["mode"] == scenes.MODE_STATIC)) == 5
    c5 = scenes.batched_worlds(3, lattice=(2, 2, 2))
    assert c5.n_spaces == 3 and len(c5.bodies) == 3 * 9
    assert set(c5.bodies["space"]) == {0, 1, 2}
    c4 = scenes.convex_chains(5, 8, 4)
    assert len(c4.joints) == 5 * 7 and set(c4.joints["type"]) == {0, 1}


def test_bench_space_sharding_plan():
    import bench
    # every space goes to exactly one rank, contiguous blocks, sizes differ by at most one
    for world in (1, 2, 3, 8):
        seen = []
        for r in range(world):
            lo, hi = bench.shard_range(37, r, world)
            seen += list(range(lo, hi))
        assert seen == list(range(37))


_GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, %r)
import torch, torch.distributed as dist
import bench
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
lo, hi = bench.shard_range(10, r, w)
ms = bench.reduce_max_ms(float(100 + 50 * r), device="cpu")
tot = bench.reduce_sum(float(hi - lo), device="cpu")
info = bench.reduce_process_info({"active_objects": 10 + r, "collision_pairs": 100 * (r + 1), "island_count": 3}, device="cpu")
if r == 0:
    print("RESULT", ms, tot, info["active_objects"], info["collision_pairs"], info["island_count"])
dist.destroy_process_group()
"""


def test_gloo_two_ranks_timing_reduction(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29617", str(script)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")][0].split()
    assert float(line[1]) == 150.0    # max over ranks
    assert float(line[2]) == 10.0     # all units accounted for once
    assert [int(x) for x in line[3:6]] == [21, 300, 6]   # get_process_info summed over the ranks


def test_hull_builder_matches_the_reference_hull_topology():
    """bradot_b200.hull (used by bench.py for C4, where no engine builds the hulls) against the meshes the reference's
    ConvexHullComputer produced for the same point clouds (tests/golden/chains12.bdmp): same faces (as clockwise
    vertex loops), same edges with the same face on each side.  Vertices are matched by position: the reference's
    integer hull moves them by ~5e-5."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden
    from bradot_b200 import hull, server
    make, _, _ = make_golden.FIXTURES["chains12"]
    sc = make()
    ref = server.meshes_from_dump(scenes.read_dump(os.path.join(ROOT, "tests", "golden", "chains12.bdmp")))

    def topo(m, vmap):
        fk = {}
        for f, (o, c) in enumerate(m["faces"]):
            loop = [int(vmap[j]) for j in m["index"][o:o + c]]
            k = loop.index(min(loop))
            fk[f] = tuple(loop[k:] + loop[:k])
        edges = set()
        for a, b, fa, fb in m["edges"]:
            a, b = int(vmap[a]), int(vmap[b])
            edges.add((a, b, fk[fa], fk[fb]) if a < b else (b, a, fk[fb], fk[fa]))
        return set(fk.values()), edges

    seen = 0
    for i, s in enumerate(sc.shapes):
        if s["type"] != scenes.SHAPE_CONVEX:
            continue
        m = hull.mesh_data(sc.verts[s["vert_ofs"]:s["vert_ofs"] + s["vert_cnt"]])
        r = ref[i]
        vmap = np.array([np.argmin(np.linalg.norm(r["verts"] - v, axis=1)) for v in m["verts"]])
        assert len(set(vmap)) == len(r["verts"]) == len(m["verts"])
        assert np.max(np.linalg.norm(r["verts"][vmap] - m["verts"], axis=1)) < 1e-3
        assert topo(m, vmap) == topo(r, np.arange(len(r["verts"])))
        for f in range(len(m["planes"])):   # outward normals, d = n.v
            o, c = m["faces"][f]
            assert np.allclose(m["verts"][m["index"][o:o + c]] @ m["planes"][f, :3], m["planes"][f, 3], atol=1e-5)
            assert np.all(m["verts"] @ m["planes"][f, :3] <= m["planes"][f, 3] + 1e-5)
        seen += 1
    assert seen >= 3
