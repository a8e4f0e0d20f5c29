"""Synthetic scene generator and the neutral scene / dump file formats.

One generator feeds both sides: the CUDA backend (through ``bradot_b200.server``)
and the reference harness ``oracle/_ref/oracle_step`` (through the ``.bscn`` file).
The five scenes are the ones BASELINE.json names: C1 pyramid,
C2 spheres in a container, C3 mixed pile, C4 convex chains with pin/hinge joints,
C5 many independent spaces.  Every scene takes a size knob so the parity tests can
run small versions of the same layout.

Shape type codes follow the reference enum PhysicsServer3D::ShapeType
(servers/physics_server_3d.h:247-259); body modes follow BodyMode (:387-392).
"""
import struct

import numpy as np

SHAPE_PLANE, SHAPE_RAY, SHAPE_SPHERE, SHAPE_BOX, SHAPE_CAPSULE, SHAPE_CYLINDER, SHAPE_CONVEX, SHAPE_CONCAVE, SHAPE_HEIGHTMAP = range(9)
MODE_STATIC, MODE_KINEMATIC, MODE_RIGID, MODE_RIGID_LINEAR = range(4)
JOINT_PIN, JOINT_HINGE, JOINT_CONE_TWIST, JOINT_SLIDER, JOINT_6DOF = 0, 1, 2, 3, 4

SHAPE_DT = np.dtype([("type", "<u4"), ("p", "<f4", 4), ("vert_ofs", "<u4"), ("vert_cnt", "<u4"), ("custom_bias", "<f4")])   # custom_bias: shape_set_custom_solver_bias, 0 = none
BODY_DT = np.dtype([
    ("space", "<u4"), ("mode", "<u4"), ("shape", "<u4"), ("flags", "<u4"),
    ("xf", "<f4", 12), ("shape_xf", "<f4", 12), ("linvel", "<f4", 3), ("angvel", "<f4", 3),
    ("mass", "<f4"), ("friction", "<f4"), ("bounce", "<f4"), ("gravity_scale", "<f4"),
    ("lin_damp", "<f4"), ("ang_damp", "<f4"), ("layer", "<u4"), ("mask", "<u4")])
# joint flags: bit0 the two bodies do not collide, bit1 hinge use_limit, bit2 hinge enable_motor, bit3 `params` given
# params: pin = PinJointParam order (bias, damping, impulse clamp); hinge = HingeJointParam order (bias, limit upper, limit
# lower, limit bias, limit softness, limit relaxation, motor target velocity, motor max impulse); priority 0 = default (1)
JOINT_DT = np.dtype([("type", "<u4"), ("body_a", "<u4"), ("body_b", "<u4"), ("flags", "<u4"),
                     ("fa", "<f4", 12), ("fb", "<f4", 12), ("params", "<f4", 8), ("priority", "<u4"), ("pad", "<u4", 3)])
# long parameter table of a joint (bpc_joint_params_write): `count` values from parameter `first` on; a slider joint's 22
# values are in PhysicsServer3D::SliderJointParam order
JOINT_EXT_DT = np.dtype([("joint", "<u4"), ("first", "<u4"), ("count", "<u4"), ("pad", "<u4"), ("values", "<f4", 28)])
# bpc_area_desc as is.  flags: bit0 monitorable, bit1 body monitor callback (events), bit2 gravity is a point; override modes:
# PhysicsServer3D::AreaSpaceOverrideMode (0 disabled, 1 combine, 2 combine-replace, 3 replace, 4 replace-combine)
AREA_DT = np.dtype([("space", "<i4"), ("shape", "<i4"), ("flags", "<u4"), ("priority", "<i4"), ("xf", "<f4", 12), ("shape_xf", "<f4", 12),
                    ("layer", "<u4"), ("mask", "<u4"), ("gravity_mode", "<i4"), ("linear_damp_mode", "<i4"), ("angular_damp_mode", "<i4"),
                    ("gravity", "<f4"), ("gravity_vector", "<f4", 3), ("point_unit_distance", "<f4"), ("linear_damp", "<f4"),
                    ("angular_damp", "<f4"), ("group", "<u4")])   # group: see bpc_area_desc
AREA_MONITORABLE, AREA_MONITOR_CB, AREA_GRAVITY_POINT, AREA_AREA_MONITOR_CB, AREA_PICKABLE = 1, 2, 4, 8, 16
JOINT_NO_COLLIDE, JOINT_USE_LIMIT, JOINT_ENABLE_MOTOR, JOINT_HAS_PARAMS = 1, 2, 4, 8
PIN_DEFAULTS = np.array([0.3, 1.0, 0.0, 0, 0, 0, 0, 0], np.float32)
HINGE_DEFAULTS = np.array([0.3, -np.pi, np.pi, 0.3, 0.9, 1.0, 0.0, 0.0], np.float32)   # upper = -pi, lower = pi: the reference's "no limit"
# body flags: bit0 can_sleep, bit1 continuous cd, bit2 start asleep, bit3 pushed every step, bit4 / bit5 linear / angular damp
# mode REPLACE, bits 8..13 axis locks (PhysicsServer3D::BodyAxis: linear x y z, angular x y z)
# bits 16..23: body_set_max_contacts_reported
BODY_LIN_DAMP_REPLACE, BODY_ANG_DAMP_REPLACE, BODY_LOCK_SHIFT, BODY_REPORT_SHIFT = 16, 32, 8, 16
BODY_NOT_PICKABLE = 128   # bit7: body_set_ray_pickable(false)
BODY_SHAPE_DISABLED = 0x8000   # bit15: body_set_shape_disabled(body, 0, true)
IDENT12 = np.array([1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0], dtype=np.float32)


class Scene:
    """Plain-data scene: structured numpy arrays, no behaviour."""

    def __init__(self, n_spaces=1, dt=1.0 / 60.0, iterations=16, name="scene"):
        self.n_spaces = n_spaces
        self.dt = np.float32(dt)
        self.iterations = iterations
        self.iter_step = 0   # space s sweeps iterations + s * iter_step times (SPACE_PARAM_SOLVER_ITERATIONS is per space)
        self.name = name
        self.shapes = np.zeros(0, SHAPE_DT)
        self.bodies = np.zeros(0, BODY_DT)
        self.joints = np.zeros(0, JOINT_DT)
        self.areas = np.zeros(0, AREA_DT)
        self.joint_ext = np.zeros(0, JOINT_EXT_DT)
        self.verts = np.zeros((0, 3), np.float32)

    # -- builders -------------------------------------------------------------
    def add_shape(self, type_, p=(0, 0, 0, 0), verts=None):
        rec = np.zeros(1, SHAPE_DT)
        rec["type"] = type_
        pp = np.zeros(4, np.float32)
        pp[:len(p)] = p
        rec["p"] = pp
        if verts is not None:
            verts = np.asarray(verts, np.float32).reshape(-1, 3)
            rec["vert_ofs"] = len(self.verts)
            rec["vert_cnt"] = len(verts)
            self.verts = np.concatenate([self.verts, verts])
        self.shapes = np.concatenate([self.shapes, rec])
        return len(self.shapes) - 1

    def add_bodies(self, n, **fields):
        rec = np.zeros(n, BODY_DT)
        rec["xf"] = IDENT12
        rec["shape_xf"] = IDENT12
        rec["mode"] = MODE_RIGID
        rec["flags"] = 1
        rec["mass"] = 1.0
        rec["friction"] = 1.0
        rec["gravity_scale"] = 1.0
        rec["layer"] = 1
        rec["mask"] = 1
        for k, v in fields.items():
            rec[k] = v
        first = len(self.bodies)
        self.bodies = np.concatenate([self.bodies, rec])
        return first

    def add_joints(self, n, **fields):
        rec = np.zeros(n, JOINT_DT)
        rec["fa"] = IDENT12
        rec["fb"] = IDENT12
        for k, v in fields.items():
            rec[k] = v
        first = len(self.joints)
        self.joints = np.concatenate([self.joints, rec])
        return first

    def set_joint_params(self, joint, first, values):
        """Values of a joint's long parameter table (slider), from parameter `first` on."""
        rec = np.zeros(1, JOINT_EXT_DT)
        rec["joint"], rec["first"], rec["count"] = joint, first, len(values)
        rec["values"][0, :len(values)] = values
        self.joint_ext = np.concatenate([self.joint_ext, rec])

    def add_areas(self, n, **fields):
        rec = np.zeros(n, AREA_DT)
        rec["xf"] = IDENT12
        rec["shape_xf"] = IDENT12
        rec["layer"] = 1
        rec["mask"] = 1
        rec["gravity"] = 9.80665          # BradotArea3D defaults, bradot_area_3d.h:50-57
        rec["gravity_vector"] = (0, -1, 0)
        rec["linear_damp"] = 0.1
        rec["angular_damp"] = 0.1
        for k, v in fields.items():
            rec[k] = v
        first = len(self.areas)
        self.areas = np.concatenate([self.areas, rec])
        return first

    def add_static_plane(self, normal, d, space=0):
        s = self.add_shape(SHAPE_PLANE, (normal[0], normal[1], normal[2], d))
        return self.add_bodies(1, mode=MODE_STATIC, shape=s, space=space, mass=1.0)

    # -- file I/O ---------------------------------------------------------------
    def write(self, path):
        hdr = struct.pack("<4sIIIIIIfI7I", b"BSCN", 2, self.n_spaces, len(self.shapes), len(self.bodies),
                          len(self.joints), len(self.verts), float(self.dt), self.iterations, self.iter_step, len(self.areas), len(self.joint_ext), *([0] * 4))
        with open(path, "wb") as f:
            f.write(hdr)
            f.write(self.shapes.tobytes())
            f.write(self.bodies.tobytes())
            f.write(self.joints.tobytes())
            f.write(np.ascontiguousarray(self.verts, np.float32).tobytes())
            f.write(self.areas.tobytes())
            f.write(self.joint_ext.tobytes())

    @staticmethod
    def read(path):
        with open(path, "rb") as f:
            raw = f.read()
        magic, ver, nsp, nsh, nb, nj, nv, dt, it, its, na, nje = struct.unpack_from("<4sIIIIIIfIIII", raw, 0)
        assert magic == b"BSCN" and ver == 2
        off = struct.calcsize("<4sIIIIIIfI7I")
        sc = Scene(nsp, dt, it)
        sc.iter_step = its
        sc.shapes = np.frombuffer(raw, SHAPE_DT, nsh, off).copy()
        off += nsh * SHAPE_DT.itemsize
        sc.bodies = np.frombuffer(raw, BODY_DT, nb, off).copy()
        off += nb * BODY_DT.itemsize
        sc.joints = np.frombuffer(raw, JOINT_DT, nj, off).copy()
        off += nj * JOINT_DT.itemsize
        sc.verts = np.frombuffer(raw, "<f4", nv * 3, off).reshape(-1, 3).copy()
        off += nv * 12
        sc.areas = np.frombuffer(raw, AREA_DT, na, off).copy()
        off += na * AREA_DT.itemsize
        sc.joint_ext = np.frombuffer(raw, JOINT_EXT_DT, nje, off).copy()
        return sc


def read_dump(path):
    """Read a ``.bdmp`` written by oracle_step: {(name, step): ndarray}."""
    out = {}
    with open(path, "rb") as f:
        raw = f.read()
    assert raw[:4] == b"BDMP"
    off = 4
    while off < len(raw):
        name = raw[off:off + 32].split(b"\0")[0].decode()
        off += 32
        step, = struct.unpack_from("<i", raw, off)
        off += 4
        dt = raw[off:off + 4].split(b"\0")[0].decode()
        off += 4
        nd, = struct.unpack_from("<I", raw, off)
        off += 4
        shape = struct.unpack_from("<%dQ" % nd, raw, off)
        off += 8 * nd
        nbytes, = struct.unpack_from("<Q", raw, off)
        off += 8
        arr = np.frombuffer(raw, "<" + dt, nbytes // 4, off).reshape(shape).copy()
        off += nbytes
        out[(name, step)] = arr
    return out


def _quat_to_rows(q):
    """unit quaternions (n,4 xyzw) -> row-major 3x3 (n,9), float32."""
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    m = np.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
        2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
        2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1)
    return m.astype(np.float32)


def _random_rotations(rng, n):
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return _quat_to_rows(q)


def _container(sc, half, space=0):
    """Floor + 4 walls as static world-boundary planes (normal points inward)."""
    sc.add_static_plane((0, 1, 0), 0.0, space)
    sc.add_static_plane((1, 0, 0), -half, space)
    sc.add_static_plane((-1, 0, 0), -half, space)
    sc.add_static_plane((0, 0, 1), -half, space)
    sc.add_static_plane((0, 0, -1), -half, space)


# ------------------------------------------------------------------------ C1
def pyramid(n_boxes=1000, base=45):
    """C1: 2-D wall pyramid of unit boxes on a static plane; no RNG."""
    sc = Scene(name="C1_pyramid_%d" % n_boxes)
    sc.add_static_plane((0, 1, 0), 0.0)
    box = sc.add_shape(SHAPE_BOX, (0.5, 0.5, 0.5))
    pos = []
    row, y = base, 0.5 + 0.001
    while len(pos) < n_boxes and row > 0:
        x0 = -(row - 1) * 1.05 * 0.5
        for i in range(row):
            if len(pos) < n_boxes:
                pos.append((x0 + i * 1.05, y, 0.0))
        row -= 1
        y += 1.0 + 0.001
    pos = np.array(pos, np.float32)
    xf = np.tile(IDENT12, (len(pos), 1))
    xf[:, 9:12] = pos
    sc.add_bodies(len(pos), shape=box, xf=xf, mass=1.0, friction=1.0)
    return sc


# ------------------------------------------------------------------------ C2
def spheres_container(nx=50, ny=40, nz=50, seed=42, half=15.0):
    """C2: spheres r=0.25 on a lattice (spacing 0.55, jitter +-0.05) falling into 5 static planes."""
    sc = Scene(name="C2_spheres_%d" % (nx * ny * nz))
    _container(sc, half)
    sph = sc.add_shape(SHAPE_SPHERE, (0.25,))
    rng = np.random.default_rng(seed)
    gx, gy, gz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    pos = np.stack([(gx - (nx - 1) * 0.5) * 0.55, 1.0 + gy * 0.55, (gz - (nz - 1) * 0.5) * 0.55], -1).reshape(-1, 3)
    pos = pos + rng.uniform(-0.05, 0.05, pos.shape)
    xf = np.tile(IDENT12, (len(pos), 1))
    xf[:, 9:12] = pos.astype(np.float32)
    sc.add_bodies(len(pos), shape=sph, xf=xf, mass=1.0, friction=1.0)
    return sc


# ------------------------------------------------------------------------ C3
def _mixed_bodies(sc, rng, pos, space=0):
    n = len(pos)
    kind = np.arange(n) % 3
    he = rng.uniform(0.2, 0.4, (n, 3)).astype(np.float32)
    rs = rng.uniform(0.2, 0.4, n).astype(np.float32)
    cr = rng.uniform(0.15, 0.3, n).astype(np.float32)
    ch = rng.uniform(0.8, 1.4, n).astype(np.float32)
    shape_ids = np.zeros(n, np.uint32)
    mass = np.zeros(n, np.float32)
    first_shape = len(sc.shapes)
    rec = np.zeros(n, SHAPE_DT)
    for i in range(n):
        pass
    rec["type"] = np.where(kind == 0, SHAPE_BOX, np.where(kind == 1, SHAPE_SPHERE, SHAPE_CAPSULE))
    p = np.zeros((n, 4), np.float32)
    p[kind == 0, :3] = he[kind == 0]
    p[kind == 1, 0] = rs[kind == 1]
    p[kind == 2, 0] = cr[kind == 2]
    p[kind == 2, 1] = ch[kind == 2]
    rec["p"] = p
    sc.shapes = np.concatenate([sc.shapes, rec])
    shape_ids[:] = first_shape + np.arange(n)
    vol_box = 8 * he[:, 0] * he[:, 1] * he[:, 2]
    vol_sph = 4.0 / 3.0 * np.pi * rs ** 3
    vol_cap = np.pi * cr ** 2 * (ch - 2 * cr) + 4.0 / 3.0 * np.pi * cr ** 3
    mass = np.where(kind == 0, vol_box, np.where(kind == 1, vol_sph, vol_cap)).astype(np.float32)
    xf = np.zeros((n, 12), np.float32)
    xf[:, :9] = _random_rotations(rng, n)
    xf[:, 9:12] = pos.astype(np.float32)
    return sc.add_bodies(n, shape=shape_ids, xf=xf, mass=mass, friction=1.0, space=space)


def mixed_pile(n_side=100, seed=7, half=None, spacing=1.1):
    """C3: box / sphere / capsule round-robin on an n^3 lattice dropped into a container."""
    n = n_side ** 3
    sc = Scene(name="C3_mixed_%d" % n)
    if half is None:
        half = max(60.0 * n_side / 100.0, n_side * spacing * 0.5 + 2.0)
    _container(sc, half)
    rng = np.random.default_rng(seed)
    g = np.stack(np.meshgrid(np.arange(n_side), np.arange(n_side), np.arange(n_side), indexing="ij"), -1).reshape(-1, 3)
    pos = np.stack([(g[:, 0] - (n_side - 1) * 0.5) * spacing, 1.0 + g[:, 1] * spacing,
                    (g[:, 2] - (n_side - 1) * 0.5) * spacing], -1)
    _mixed_bodies(sc, rng, pos)
    return sc


# ------------------------------------------------------------------------ C4
def convex_chains(n_chains=25000, chain_len=8, n_hulls=64, seed=11):
    """C4: convex-hull debris in chains linked alternately by pin and hinge joints."""
    sc = Scene(name="C4_convex_%d" % (n_chains * chain_len))
    sc.add_static_plane((0, 1, 0), 0.0)
    rng = np.random.default_rng(seed)
    hull_ids = []
    for _ in range(n_hulls):
        semi = rng.uniform(0.3, 0.6, 3)
        d = rng.normal(size=(14, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        hull_ids.append(sc.add_shape(SHAPE_CONVEX, verts=(d * semi).astype(np.float32)))
    hull_ids = np.array(hull_ids, np.uint32)
    side = int(np.ceil(np.sqrt(n_chains)))
    n = n_chains * chain_len
    c = np.arange(n_chains)
    cx = (c % side - (side - 1) * 0.5) * 1.6
    cz = (c // side - (side - 1) * 0.5) * 1.6
    k = np.arange(chain_len)
    pos = np.zeros((n_chains, chain_len, 3))
    pos[:, :, 0] = cx[:, None]
    pos[:, :, 2] = cz[:, None]
    pos[:, :, 1] = 1.0 + k[None, :] * 1.3
    pos = pos.reshape(-1, 3)
    xf = np.tile(IDENT12, (n, 1))
    xf[:, 9:12] = pos.astype(np.float32)
    shape = hull_ids[rng.integers(0, n_hulls, n)]
    first = sc.add_bodies(n, shape=shape, xf=xf, mass=1.0, friction=1.0)
    ja = (first + (c[:, None] * chain_len + k[None, :-1])).reshape(-1)
    jb = ja + 1
    nj = len(ja)
    jt = np.tile(np.arange(chain_len - 1) % 2, n_chains).astype(np.uint32)
    fa = np.tile(IDENT12, (nj, 1))
    fb = np.tile(IDENT12, (nj, 1))
    fa[:, 10] = 0.65
    fb[:, 10] = -0.65
    sc.add_joints(nj, type=jt, body_a=ja.astype(np.uint32), body_b=jb.astype(np.uint32), flags=1, fa=fa, fb=fb)
    return sc


# ------------------------------------------------------------------------ C5
def batched_worlds(n_spaces=4096, seed0=1000, lattice=(8, 4, 8)):
    """C5: independent spaces, each a plane + a small mixed pile (seed0 + space index per space).
    Records are collected per space and concatenated once, so 4096 spaces build in seconds."""
    sc = Scene(n_spaces=n_spaces, name="C5_worlds_%dx%d" % (n_spaces, lattice[0] * lattice[1] * lattice[2]))
    g = np.stack(np.meshgrid(np.arange(lattice[0]), np.arange(lattice[1]), np.arange(lattice[2]), indexing="ij"),
                 -1).reshape(-1, 3)
    pos = np.stack([(g[:, 0] - (lattice[0] - 1) * 0.5) * 1.1, 1.0 + g[:, 1] * 1.1,
                    (g[:, 2] - (lattice[2] - 1) * 0.5) * 1.1], -1)
    shape_parts, body_parts = [], []
    n_shapes = 0
    for s in range(n_spaces):
        tmp = Scene()
        tmp.add_static_plane((0, 1, 0), 0.0)
        _mixed_bodies(tmp, np.random.default_rng(seed0 + s), pos)
        tmp.bodies["space"] = s
        tmp.bodies["shape"] += n_shapes
        n_shapes += len(tmp.shapes)
        shape_parts.append(tmp.shapes)
        body_parts.append(tmp.bodies)
    sc.shapes = np.concatenate(shape_parts)
    sc.bodies = np.concatenate(body_parts)
    return sc


def convex_pile(n_side=4, n_hulls=8, seed=11, with_primitives=True):
    """Convex-hull debris (C4's hulls, no joints) dropped on a plane, optionally mixed with
    boxes / spheres / capsules so every hull-vs-primitive narrowphase pair occurs."""
    sc = Scene(name="convex_pile_%d" % n_side ** 3)
    _container(sc, n_side * 0.6 + 1.5)
    rng = np.random.default_rng(seed)
    hull_ids = []
    for _ in range(n_hulls):
        semi = rng.uniform(0.3, 0.6, 3)
        d = rng.normal(size=(14, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        hull_ids.append(sc.add_shape(SHAPE_CONVEX, verts=(d * semi).astype(np.float32)))
    prim = [sc.add_shape(SHAPE_BOX, (0.3, 0.25, 0.35)), sc.add_shape(SHAPE_SPHERE, (0.35,)), sc.add_shape(SHAPE_CAPSULE, (0.2, 1.0))]
    n = n_side ** 3
    g = np.stack(np.meshgrid(np.arange(n_side), np.arange(n_side), np.arange(n_side), indexing="ij"), -1).reshape(-1, 3)
    pos = np.stack([(g[:, 0] - (n_side - 1) * 0.5) * 1.2, 1.0 + g[:, 1] * 1.2, (g[:, 2] - (n_side - 1) * 0.5) * 1.2], -1)
    shape = np.array(hull_ids, np.uint32)[rng.integers(0, n_hulls, n)]
    if with_primitives:
        sel = np.arange(n) % 4 == 3
        shape[sel] = np.array(prim, np.uint32)[rng.integers(0, 3, int(sel.sum()))]
    xf = np.zeros((n, 12), np.float32)
    xf[:, :9] = _random_rotations(rng, n)
    xf[:, 9:12] = pos.astype(np.float32)
    sc.add_bodies(n, shape=shape, xf=xf, mass=1.0, friction=1.0)
    return sc


def free_fall(n=64, seed=3):
    """Contact-free scene: tumbling mixed bodies far apart (integration parity)."""
    sc = Scene(name="free_fall_%d" % n)
    rng = np.random.default_rng(seed)
    pos = np.stack([np.arange(n) * 5.0, np.full(n, 100.0), np.zeros(n)], -1)
    first = _mixed_bodies(sc, rng, pos)
    sc.bodies["linvel"][first:] = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    sc.bodies["angvel"][first:] = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    return sc


def sleepers(n_boxes=6):
    """Sleep / wake test: separate boxes dropped from small heights (each its own island, asleep one after the other),
    a two-box stack, and a sphere released from high up that lands on a sleeping box and wakes its island."""
    sc = Scene(name="sleepers")
    sc.add_static_plane((0, 1, 0), 0.0)
    box = sc.add_shape(SHAPE_BOX, (0.4, 0.4, 0.4, 0))
    ball = sc.add_shape(SHAPE_SPHERE, (0.3, 0, 0, 0))
    xf = np.tile(IDENT12, (n_boxes, 1))
    xf[:, 9] = (np.arange(n_boxes) - (n_boxes - 1) * 0.5) * 2.0
    xf[:, 10] = 0.41 + 0.05 * np.arange(n_boxes)
    sc.add_bodies(n_boxes, shape=box, xf=xf, mass=1.0)
    top = IDENT12.copy()
    top[9], top[10] = xf[0, 9], 1.25          # second box resting on box 0
    sc.add_bodies(1, shape=box, xf=top, mass=1.0)
    hi = IDENT12.copy()
    hi[9], hi[10] = xf[2, 9] + 0.05, 14.0     # lands on box 2 after ~100 steps, when it is asleep
    sc.add_bodies(1, shape=ball, xf=hi, mass=2.0)
    return sc


def kinematic_platform(rails=False):
    """A kinematic platform (moved by body_set_state(TRANSFORM) every step: 0.6 m/s along x, 0.3 rad/s about y, scripted
    through its scene velocities) carrying boxes and spheres over a ground plane; a second kinematic box stands still.
    rails: the platform has two further shapes, rails along two of its edges (records flagged BODY_FURTHER_SHAPE: one kinematic
    body of three shapes to the harness and the engine demo, three kinematic records moved alike and marked as one object
    to the C ABI)."""
    sc = Scene(name="kinematic_platform")
    sc.add_static_plane((0, 1, 0), 0.0)
    slab = sc.add_shape(SHAPE_BOX, (2.0, 0.2, 2.0, 0))
    box = sc.add_shape(SHAPE_BOX, (0.25, 0.25, 0.25, 0))
    ball = sc.add_shape(SHAPE_SPHERE, (0.25, 0, 0, 0))
    x = IDENT12.copy()
    x[10] = 1.0
    sc.add_bodies(1, mode=MODE_KINEMATIC, shape=slab, xf=x, linvel=(0.6, 0.0, 0.0), angvel=(0.0, 0.3, 0.0))
    if rails:
        rail = sc.add_shape(SHAPE_BOX, (0.1, 0.3, 2.0, 0))
        for side in (-1.0, 1.0):
            sx = IDENT12.copy()
            sx[9], sx[10] = 1.9 * side, 0.5
            sc.add_bodies(1, mode=MODE_KINEMATIC, shape=rail, xf=x, shape_xf=sx, linvel=(0.6, 0.0, 0.0), angvel=(0.0, 0.3, 0.0), flags=1 | BODY_FURTHER_SHAPE)
    x2 = IDENT12.copy()
    x2[9], x2[10] = 6.0, 0.5
    sc.add_bodies(1, mode=MODE_KINEMATIC, shape=slab, xf=x2)   # never moved: stays inactive
    g = np.stack(np.meshgrid(np.arange(3), np.arange(3), indexing="ij"), -1).reshape(-1, 2)
    xf = np.tile(IDENT12, (9, 1))
    xf[:, 9] = (g[:, 0] - 1) * 1.1
    xf[:, 11] = (g[:, 1] - 1) * 1.1
    xf[:, 10] = 1.5
    shape = np.where(np.arange(9) % 2 == 0, box, ball).astype(np.uint32)
    sc.add_bodies(9, shape=shape, xf=xf, mass=1.0)
    return sc


def bullets(n=12):
    """Continuous collision detection: small fast spheres and boxes (60 m/s = 1 m per step, five times their size) fired at a
    thin static wall (every second one has the CCD flag, the others tunnel) and, one CCD bullet each, at three thin
    free-floating plates (a plate has no other contact, so the order of its pre-solve among others cannot matter)."""
    sc = Scene(name="bullets")
    sc.add_static_plane((0, 1, 0), -5.0)
    wall = sc.add_shape(SHAPE_BOX, (3.0, 3.0, 0.02, 0))
    plate = sc.add_shape(SHAPE_BOX, (1.0, 1.0, 0.03, 0))
    ball = sc.add_shape(SHAPE_SPHERE, (0.1, 0, 0, 0))
    cube = sc.add_shape(SHAPE_BOX, (0.08, 0.08, 0.08, 0))
    sc.add_bodies(1, mode=MODE_STATIC, shape=wall)
    px = np.array([6.0, 9.0, 12.0])
    xp = np.tile(IDENT12, (3, 1))
    xp[:, 9] = px
    sc.add_bodies(3, shape=plate, xf=xp, mass=5.0, gravity_scale=0.0)
    rng = np.random.default_rng(4)
    xf = np.tile(IDENT12, (n, 1))
    xf[:, 9] = rng.uniform(-2.0, 2.0, n)
    xf[:, 10] = rng.uniform(-1.5, 1.5, n)
    xf[:, 11] = -rng.uniform(2.3, 7.7, n)
    shape = np.where(np.arange(n) % 3 == 0, cube, ball).astype(np.uint32)
    flags = np.where(np.arange(n) % 2 == 0, 3, 1).astype(np.uint32)   # bit1 = continuous cd
    lin = np.zeros((n, 3), np.float32)
    lin[:, 2] = 60.0
    sc.add_bodies(n, shape=shape, xf=xf, flags=flags, linvel=lin, mass=0.2, gravity_scale=0.0)
    xb = np.tile(IDENT12, (3, 1))
    xb[:, 9] = px + np.array([0.3, -0.2, 0.45])
    xb[:, 10] = np.array([0.2, -0.4, 0.1])
    xb[:, 11] = -np.array([3.3, 5.6, 4.45])
    sc.add_bodies(3, shape=np.array([ball, cube, ball], np.uint32), xf=xb, flags=3, linvel=(0.0, 0.0, 60.0), mass=0.2, gravity_scale=0.0)
    return sc


def mesh_bullets(n=12):
    """Continuous collision detection against static meshes: the bullets of bullets() fired at a wall made of two triangles
    (a ConcavePolygonShape3D whose faces look at the bullets), at a second, two-sided one behind a gap in the first, and, from
    above, at a height map.  Every second bullet has the CCD flag."""
    sc = Scene(name="mesh_bullets")
    sc.add_static_plane((0, 1, 0), -12.0)
    quad = lambda x0, x1: np.array([[(x0, -3, 0), (x1, -3, 0), (x0, 3, 0)], [(x1, -3, 0), (x1, 3, 0), (x0, 3, 0)]], np.float32)   # normals -z
    wall = sc.add_shape(SHAPE_CONCAVE, (0.0,), verts=quad(-3.0, 0.0).reshape(-1, 3))
    wall2 = sc.add_shape(SHAPE_CONCAVE, (1.0,), verts=quad(0.0, 3.0)[:, ::-1].copy().reshape(-1, 3))                         # normals +z, two-sided
    sc.add_bodies(1, mode=MODE_STATIC, shape=wall)
    x2 = IDENT12.copy()
    x2[11] = 0.5
    sc.add_bodies(1, mode=MODE_STATIC, shape=wall2, xf=x2)
    w = d = 9
    gx, gz = np.meshgrid(np.arange(w) - 4.0, np.arange(d) - 4.0, indexing="xy")
    h = (0.3 * np.sin(gx * 0.8) * np.cos(gz * 0.6)).astype(np.float32)
    verts = np.zeros((w * d, 3), np.float32)
    verts[:, 0] = h.reshape(-1)
    hm = sc.add_shape(SHAPE_HEIGHTMAP, (w, d, float(h.min()) - 0.1, float(h.max()) + 0.1), verts=verts)
    xh = IDENT12.copy()
    xh[9], xh[10] = 20.0, -6.0
    sc.add_bodies(1, mode=MODE_STATIC, shape=hm, xf=xh)
    ball = sc.add_shape(SHAPE_SPHERE, (0.1, 0, 0, 0))
    cube = sc.add_shape(SHAPE_BOX, (0.08, 0.08, 0.08, 0))
    rng = np.random.default_rng(6)
    xf = np.tile(IDENT12, (n, 1))
    xf[:, 9] = rng.uniform(-2.6, 2.6, n)
    xf[:, 10] = rng.uniform(-2.5, 2.5, n)
    xf[:, 11] = -rng.uniform(2.3, 7.7, n)
    shape = np.where(np.arange(n) % 3 == 0, cube, ball).astype(np.uint32)
    flags = np.where(np.arange(n) % 2 == 0, 3, 1).astype(np.uint32)
    lin = np.zeros((n, 3), np.float32)
    lin[:, 2] = 60.0
    sc.add_bodies(n, shape=shape, xf=xf, flags=flags, linvel=lin, mass=0.2, gravity_scale=0.0)
    m = 6
    xd = np.tile(IDENT12, (m, 1))
    xd[:, 9] = 20.0 + rng.uniform(-3.0, 3.0, m)
    xd[:, 10] = rng.uniform(-3.5, 0.5, m)
    xd[:, 11] = rng.uniform(-3.0, 3.0, m)
    lin = np.zeros((m, 3), np.float32)
    lin[:, 1] = -60.0
    lin[:, 0] = rng.uniform(-4.0, 4.0, m)
    sc.add_bodies(m, shape=np.where(np.arange(m) % 2 == 0, ball, cube).astype(np.uint32), xf=xd, flags=np.where(np.arange(m) % 3 != 2, 3, 1).astype(np.uint32),
                  linvel=lin, mass=0.2, gravity_scale=0.0)
    return sc


PUSH_FLAG = 8   # scene flag bit3: the harness applies a force and a torque to the body before every step


def pushed_force(i):
    """The push oracle/ref_harness.cpp (push_bodies) gives body i: (force xyz, torque xyz)."""
    return np.array([0.5 * ((i % 3) - 1), 2.0, 0.3, 0.1, 0.0, 0.05 * (i % 2)], np.float32)


def pushed(n=24):
    """body_apply_central_force / body_apply_torque every step on boxes and spheres resting on or floating above a plane."""
    sc = Scene(name="pushed")
    sc.add_static_plane((0, 1, 0), 0.0)
    box = sc.add_shape(SHAPE_BOX, (0.3, 0.2, 0.25, 0))
    ball = sc.add_shape(SHAPE_SPHERE, (0.25, 0, 0, 0))
    xf = np.tile(IDENT12, (n, 1))
    xf[:, 9] = (np.arange(n) % 6 - 2.5) * 1.5
    xf[:, 11] = (np.arange(n) // 6 - 1.5) * 1.5
    xf[:, 10] = np.where(np.arange(n) % 4 == 0, 2.0, 0.26)
    shape = np.where(np.arange(n) % 2 == 0, box, ball).astype(np.uint32)
    flags = np.where(np.arange(n) % 5 == 4, 1, 1 | PUSH_FLAG).astype(np.uint32)
    sc.add_bodies(n, shape=shape, xf=xf, flags=flags, mass=1.5)
    return sc


def hanging_chains(n_chains=12, chain_len=5, seed=5, cone_twist=False):
    """Joint parameters, hinge limits and motors, solver priorities: chains of boxes hanging from static anchors in empty
    space (the links are on no collision layer, so every island consists of joints only).  Chain c: links joined
    alternately by hinge and pin, the top link pinned to the anchor.  Hinges with an even index use a limit of
    [-0.35, 0.5] rad, every third one also runs a motor; pins with an index divisible by 3 carry their own bias / damping /
    impulse clamp.  All joints of chain c have solver priority 2 when c % 3 == 2 (one priority per island: the reference
    compacts its island lists in place between priority rounds, which a mixed island would not survive as a dump).
    cone_twist: the pins between links become cone-twist joints (ConeTwistJointParam order: swing span, twist span, bias,
    softness, relaxation) with spans small enough for both limits to engage; every third one keeps the untouched defaults."""
    sc = Scene(name="hanging_chains_%d" % n_chains)
    anchor = sc.add_shape(SHAPE_BOX, (0.1, 0.1, 0.1))
    link = sc.add_shape(SHAPE_BOX, (0.1, 0.3, 0.15))
    rng = np.random.default_rng(seed)
    hinge_i = pin_i = 0
    for c in range(n_chains):
        x0 = c * 10.0
        xa = IDENT12.copy()
        xa[9], xa[10] = x0, 10.0
        a = sc.add_bodies(1, mode=MODE_STATIC, shape=anchor, xf=xa, layer=0, mask=0)
        xf = np.tile(IDENT12, (chain_len, 1))
        xf[:, 9] = x0
        xf[:, 10] = 10.0 - 0.45 - 0.7 * np.arange(chain_len)
        lin = rng.uniform(-1.5, 1.5, (chain_len, 3)).astype(np.float32)
        ang = rng.uniform(-2.0, 2.0, (chain_len, 3)).astype(np.float32)
        first = sc.add_bodies(chain_len, shape=link, xf=xf, mass=0.5 + 0.25 * (c % 3), layer=0, mask=0, linvel=lin, angvel=ang, flags=0)
        prio = 2 if c % 3 == 2 else 0
        # anchor -- link 0: pin
        fa, fb = IDENT12.copy(), IDENT12.copy()
        fa[10], fb[10] = -0.1, 0.35
        sc.add_joints(1, type=JOINT_PIN, body_a=a, body_b=first, flags=JOINT_NO_COLLIDE, fa=fa, fb=fb, priority=prio)
        for k in range(chain_len - 1):
            fa, fb = IDENT12.copy(), IDENT12.copy()
            fa[10], fb[10] = -0.35, 0.35
            if k % 2 == 0:
                flags, prm = JOINT_NO_COLLIDE, HINGE_DEFAULTS.copy()
                if hinge_i % 2 == 0:
                    flags |= JOINT_USE_LIMIT | JOINT_HAS_PARAMS
                    prm[1], prm[2] = 0.5, -0.35            # upper, lower
                    prm[3], prm[5] = 0.4, 0.9              # limit bias, relaxation
                if hinge_i % 3 == 0:
                    flags |= JOINT_ENABLE_MOTOR | JOINT_HAS_PARAMS
                    prm[6], prm[7] = 1.5, 0.05             # target velocity, max impulse
                if hinge_i % 5 == 4:
                    flags |= JOINT_HAS_PARAMS
                    prm[0] = 0.45                          # bias
                sc.add_joints(1, type=JOINT_HINGE, body_a=first + k, body_b=first + k + 1, flags=flags, fa=fa, fb=fb, params=prm, priority=prio)
                hinge_i += 1
            elif cone_twist:
                flags, prm = JOINT_NO_COLLIDE, np.zeros(8, np.float32)
                if pin_i % 3 != 2:
                    flags |= JOINT_HAS_PARAMS
                    prm[0:5] = (0.3 + 0.1 * (pin_i % 2), 0.25, 0.3 + 0.1 * (pin_i % 3), 0.8, 0.9 + 0.1 * (pin_i % 2))
                    prm[5] = prm[0]              # the second swing span: set_param(SWING_SPAN) writes both
                ca, cb = fa.copy(), fb.copy()
                ang = 0.2 * (pin_i % 4)          # the frames are turned about z and y, so that swing and twist axes are not body axes
                ca[0], ca[1], ca[3], ca[4] = np.cos(ang), -np.sin(ang), np.sin(ang), np.cos(ang)
                cb[0], cb[2], cb[6], cb[8] = np.cos(ang), np.sin(ang), -np.sin(ang), np.cos(ang)
                sc.add_joints(1, type=JOINT_CONE_TWIST, body_a=first + k, body_b=first + k + 1, flags=flags, fa=ca, fb=cb, params=prm, priority=prio)
                pin_i += 1
            else:
                flags, prm = JOINT_NO_COLLIDE, PIN_DEFAULTS.copy()
                if pin_i % 3 == 0:
                    flags |= JOINT_HAS_PARAMS
                    prm[0], prm[1], prm[2] = 0.5, 0.8, 0.05
                sc.add_joints(1, type=JOINT_PIN, body_a=first + k, body_b=first + k + 1, flags=flags, fa=fa, fb=fb, params=prm, priority=prio)
                pin_i += 1
    return sc


BODY_FURTHER_SHAPE = 0x4000   # scene flag bit14 (harness / engine demo): this record is a further shape of the static body before it


def level_geometry(n_side=4, seed=19, merged=True):
    """A static body with several shapes -- a floor slab, two tilted ramps, a wall, a hull and a sphere bump, the body itself
    turned and shifted -- under a shower of mixed bodies; a world boundary far below catches what falls off.  In the scene file
    the further shapes are records of their own flagged BODY_FURTHER_SHAPE (same transform as the body): the harness and the
    engine demo add them to the body before (body_add_shape), the C ABI sees six static bodies -- which is the same thing to
    the step (one broadphase entry and one pair object per shape in the reference).  merged=False drops the flag."""
    sc = Scene(name="level_geometry")
    rng = np.random.default_rng(seed)
    xf = IDENT12.copy()
    th = 0.15
    xf[0], xf[2], xf[6], xf[8] = np.cos(th), np.sin(th), -np.sin(th), np.cos(th)
    xf[9], xf[10], xf[11] = 0.5, -0.2, -0.3

    def tilt(angle, axis, pos):
        f = IDENT12.copy()
        c, s = np.cos(angle), np.sin(angle)
        if axis == 0:
            f[4], f[5], f[7], f[8] = c, -s, s, c
        else:
            f[0], f[1], f[3], f[4] = c, -s, s, c
        f[9:12] = pos
        return f
    parts = [(sc.add_shape(SHAPE_BOX, (4.0, 0.2, 4.0)), tilt(0.0, 0, (0, 0, 0))),
             (sc.add_shape(SHAPE_BOX, (1.5, 0.1, 1.0)), tilt(0.45, 1, (-2.0, 0.8, 1.0))),
             (sc.add_shape(SHAPE_BOX, (1.0, 0.1, 1.5)), tilt(-0.4, 0, (1.5, 0.7, -1.5))),
             (sc.add_shape(SHAPE_BOX, (0.15, 1.0, 4.0)), tilt(0.0, 0, (3.9, 1.2, 0))),
             (sc.add_shape(SHAPE_CONVEX, verts=rng.normal(size=(16, 3)) * np.array([0.7, 0.35, 0.7])), tilt(0.2, 1, (0.5, 0.5, 1.5))),
             (sc.add_shape(SHAPE_SPHERE, (0.8,)), tilt(0.0, 0, (-1.0, 0.0, -1.5)))]
    for k, (sh, sxf) in enumerate(parts):
        sc.add_bodies(1, mode=MODE_STATIC, shape=sh, xf=xf, shape_xf=sxf, friction=0.8, flags=(BODY_FURTHER_SHAPE if (k and merged) else 0))
    sc.add_static_plane((0, 1, 0), -8.0)
    g = np.stack(np.meshgrid(np.arange(n_side), np.arange(3), np.arange(n_side), indexing="ij"), -1).reshape(-1, 3)
    pos = np.stack([(g[:, 0] - (n_side - 1) * 0.5) * 1.3 + 0.4, 2.2 + g[:, 1] * 1.2, (g[:, 2] - (n_side - 1) * 0.5) * 1.3], -1)
    _mixed_bodies(sc, rng, pos)
    return sc


JOINT_HINGE_SIMPLE = 16   # scene flag (harness / engine demo): make the hinge with joint_make_hinge_simple from the frames' origins and z axes


def simple_hinges(n=12, seed=17):
    """joint_make_hinge_simple: boxes hanging from static anchors by hinges given as pivots and axes (the z columns of the
    scene's frames; the other columns are ignored).  The bodies start turned, the two axes differ, every second hinge has a
    limit -- all of which enter the frames the reference builds from body A's transform at that moment."""
    sc = Scene(name="simple_hinges_%d" % n)
    anchor = sc.add_shape(SHAPE_BOX, (0.1, 0.1, 0.1))
    link = sc.add_shape(SHAPE_BOX, (0.3, 0.12, 0.2))
    rng = np.random.default_rng(seed)
    for r in range(n):
        xa = IDENT12.copy()
        th = 0.3 * (r % 4)
        xa[0], xa[2], xa[6], xa[8] = np.cos(th), np.sin(th), -np.sin(th), np.cos(th)      # anchors turned about y
        xa[9], xa[10] = r * 6.0, 10.0
        a = sc.add_bodies(1, mode=MODE_STATIC, shape=anchor, xf=xa, layer=0, mask=0)
        xb = IDENT12.copy()
        xb[9], xb[10] = r * 6.0 + 0.5, 9.7
        b = sc.add_bodies(1, shape=link, xf=xb, mass=0.8, layer=0, mask=0, linvel=rng.uniform(-1, 1, 3).astype(np.float32),
                          angvel=rng.uniform(-2, 2, 3).astype(np.float32), flags=0)
        ax_a = np.array([0.2 * (r % 3), 0.1, 1.0]) if r % 5 else np.array([1.0, 0.0, 0.0])     # r % 5 == 0: along the anchor's x (the special case)
        ax_b = np.array([0.1, 0.15 * (r % 2), 1.0])
        ax_a, ax_b = ax_a / np.linalg.norm(ax_a), ax_b / np.linalg.norm(ax_b)
        fa, fb = IDENT12.copy(), IDENT12.copy()
        fa[[2, 5, 8]], fb[[2, 5, 8]] = ax_a, ax_b
        fa[9:12], fb[9:12] = (0.2, -0.1, 0.0), (-0.3, 0.2, 0.0)
        flags, prm = JOINT_NO_COLLIDE | JOINT_HINGE_SIMPLE, HINGE_DEFAULTS.copy()
        if r % 2:
            flags |= JOINT_USE_LIMIT | JOINT_HAS_PARAMS
            prm[1], prm[2] = 0.4, -0.3
        sc.add_joints(1, type=JOINT_HINGE, body_a=a, body_b=b, flags=flags, fa=fa, fb=fb, params=prm)
    return sc


def slider_rigs(n_rigs=18, seed=11):
    """Slider joints: rig r is a static anchor, a box hanging from it by a slider joint and a second box on the first by
    another one; no collisions (joint-only islands, as in hanging_chains).  The joint frames are turned so that the slider
    axis is no body axis; rig r % 3 == 0 keeps every default (no linear limit: lower > upper; angular limit [0, 0], always
    engaged), the others get linear limits narrow enough to be hit, an angular range, and their own softness / restitution /
    damping values for every branch; rigs with r % 4 == 3 run at solver priority 2."""
    sc = Scene(name="slider_rigs_%d" % n_rigs)
    anchor = sc.add_shape(SHAPE_BOX, (0.1, 0.1, 0.1))
    link = sc.add_shape(SHAPE_BOX, (0.3, 0.12, 0.2))
    rng = np.random.default_rng(seed)
    for r in range(n_rigs):
        x0 = r * 8.0
        xa = IDENT12.copy()
        xa[9], xa[10] = x0, 10.0
        a = sc.add_bodies(1, mode=MODE_STATIC, shape=anchor, xf=xa, layer=0, mask=0)
        xf = np.tile(IDENT12, (2, 1))
        xf[:, 9] = x0 + np.array([0.5, 1.3])
        xf[:, 10] = 10.0 - np.array([0.3, 0.7])
        lin = rng.uniform(-1.5, 1.5, (2, 3)).astype(np.float32)
        ang = rng.uniform(-1.5, 1.5, (2, 3)).astype(np.float32)
        first = sc.add_bodies(2, shape=link, xf=xf, mass=0.6 + 0.3 * (r % 3), layer=0, mask=0, linvel=lin, angvel=ang, flags=0)
        prio = 2 if r % 4 == 3 else 0
        for k, (ba, bb) in enumerate(((a, first), (first, first + 1))):
            th, ph = 0.25 * ((r + k) % 4), 0.3 * ((r + 2 * k) % 3)
            c, s, c2, s2 = np.cos(th), np.sin(th), np.cos(ph), np.sin(ph)
            rz = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
            rx = np.array([[1, 0, 0], [0, c2, -s2], [0, s2, c2]])
            fa, fb = IDENT12.copy(), IDENT12.copy()
            fa[0:9] = (rz @ rx).reshape(-1)
            fb[0:9] = (rz @ rx @ np.array([[1, 0, 0], [0, np.cos(0.15 * k), -np.sin(0.15 * k)], [0, np.sin(0.15 * k), np.cos(0.15 * k)]])).reshape(-1)
            fa[9:12] = (0.4, -0.2, 0.0)
            fb[9:12] = (-0.35, 0.1, 0.05)
            j = sc.add_joints(1, type=JOINT_SLIDER, body_a=ba, body_b=bb, flags=JOINT_NO_COLLIDE, fa=fa, fb=fb, priority=prio)
            if r % 3 != 0:
                v = np.array([0.08 + 0.05 * k, -0.12, 0.9, 0.6, 0.8, 0.95, 0.65, 0.1 * (r % 2), 0.9, 0.75, 0.85,
                              0.3, -0.2 - 0.1 * k, 0.85, 0.6, 0.9, 0.9, 0.5, 0.2 * (r % 2), 0.95, 0.65, 0.8], np.float32)
                if r % 3 == 2:
                    sc.set_joint_params(j, 0, v)
                else:           # only some of them: the linear limits and the angular block
                    sc.set_joint_params(j, 0, v[0:2])
                    sc.set_joint_params(j, 11, v[11:22])
    return sc


def g6dof_axis_defaults():
    """One axis block of a 6-DOF joint's long parameter table (include/bradot_physics_cuda.h): 22 G6DOFJointAxisParam values,
    6 flags; the members of the reference's joint as constructed."""
    q = np.zeros(28, np.float32)
    q[2], q[3], q[4] = 0.7, 0.5, 1.0
    q[10], q[11], q[12], q[13], q[15], q[16], q[18] = -1e30, 1e30, 0.5, 1.0, 300.0, 0.5, 0.1
    q[22] = 1.0
    return q


def sixdof_rigs(n_rigs=18, seed=13):
    """Generic 6-DOF joints: rig r is a static anchor, a box on it and a second box on the first, both by 6-DOF joints in turned
    frames; no collisions.  r % 3 == 0: every default (three locked linear axes, no angular limits).  r % 3 == 1: x limited
    to a range that is hit, y locked, z free; angular limits on x and y narrow enough to be violated, z free with its motor
    running.  r % 3 == 2: own softness / restitution / damping / ERP values, a limit force small enough to clip, bounce, the
    y linear limit switched off.  Rigs with r % 4 == 3 run at solver priority 2."""
    sc = Scene(name="sixdof_rigs_%d" % n_rigs)
    anchor = sc.add_shape(SHAPE_BOX, (0.1, 0.1, 0.1))
    link = sc.add_shape(SHAPE_BOX, (0.3, 0.12, 0.2))
    rng = np.random.default_rng(seed)
    for r in range(n_rigs):
        x0 = r * 8.0
        xa = IDENT12.copy()
        xa[9], xa[10] = x0, 10.0
        a = sc.add_bodies(1, mode=MODE_STATIC, shape=anchor, xf=xa, layer=0, mask=0)
        xf = np.tile(IDENT12, (2, 1))
        xf[:, 9] = x0 + np.array([0.5, 1.3])
        xf[:, 10] = 10.0 - np.array([0.3, 0.7])
        lin = rng.uniform(-1.5, 1.5, (2, 3)).astype(np.float32)
        ang = rng.uniform(-2.5, 2.5, (2, 3)).astype(np.float32)
        first = sc.add_bodies(2, shape=link, xf=xf, mass=0.6 + 0.3 * (r % 3), layer=0, mask=0, linvel=lin, angvel=ang, flags=0)
        prio = 2 if r % 4 == 3 else 0
        for k, (ba, bb) in enumerate(((a, first), (first, first + 1))):
            th, ph = 0.25 * ((r + k) % 4), 0.3 * ((r + 2 * k) % 3)
            c, s, c2, s2 = np.cos(th), np.sin(th), np.cos(ph), np.sin(ph)
            rz = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
            rx = np.array([[1, 0, 0], [0, c2, -s2], [0, s2, c2]])
            ry = np.array([[np.cos(0.2 * k), 0, np.sin(0.2 * k)], [0, 1, 0], [-np.sin(0.2 * k), 0, np.cos(0.2 * k)]])
            fa, fb = IDENT12.copy(), IDENT12.copy()
            fa[0:9] = (rz @ rx).reshape(-1)
            fb[0:9] = (rz @ rx @ ry).reshape(-1)
            fa[9:12] = (0.4, -0.2, 0.0)
            fb[9:12] = (-0.35, 0.1, 0.05)
            j = sc.add_joints(1, type=JOINT_6DOF, body_a=ba, body_b=bb, flags=JOINT_NO_COLLIDE, fa=fa, fb=fb, priority=prio)
            if r % 3 == 0:
                continue
            for ax in range(3):
                q = g6dof_axis_defaults()
                if r % 3 == 1:
                    if ax == 0:
                        q[0], q[1] = -0.05, 0.08
                    elif ax == 2:
                        q[0], q[1] = 1.0, -1.0
                    q[23] = 1.0
                    if ax == 0:
                        q[10], q[11] = -0.1, 0.15
                    elif ax == 1:
                        q[10], q[11] = -0.05, 0.1
                    else:
                        q[10], q[11] = 1.0, -1.0
                        q[26], q[17] = 1.0, 1.5
                else:
                    q[0], q[1] = (-0.03, 0.04) if ax != 1 else (0.0, 0.0)
                    q[2], q[3], q[4] = 0.8, 0.6 - 0.1 * ax, 0.9
                    q[22] = 0.0 if ax == 1 else 1.0
                    q[23] = 1.0
                    q[10], q[11] = -0.08 - 0.02 * ax, 0.06
                    q[12], q[13], q[14], q[15], q[16] = 0.6, 0.9, 0.2, 2.0 + ax, 0.4
                sc.set_joint_params(j, ax * 28, q)
    return sc


def locks_and_damping(n=48, seed=9, on_plane=False):
    """body_set_axis_lock and the body damp modes: tumbling mixed bodies, each with its own lock bits (every other body),
    damping values and COMBINE / REPLACE modes; far apart in the air, or (on_plane) dropped on a plane from a small height."""
    sc = Scene(name="locks_and_damping_%d" % n)
    rng = np.random.default_rng(seed)
    if on_plane:
        sc.add_static_plane((0, 1, 0), 0.0)
        pos = np.stack([(np.arange(n) % 8) * 2.5, 0.8 + 0.1 * (np.arange(n) % 3), (np.arange(n) // 8) * 2.5], -1)
    else:
        pos = np.stack([np.arange(n) * 5.0, np.full(n, 100.0), np.zeros(n)], -1)
    first = _mixed_bodies(sc, rng, pos)
    b = sc.bodies
    idx = np.arange(n)
    b["linvel"][first:] = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    b["angvel"][first:] = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    b["lin_damp"][first:] = rng.uniform(0.0, 0.8, n).astype(np.float32)
    b["ang_damp"][first:] = rng.uniform(0.0, 1.5, n).astype(np.float32)
    locks = np.where(idx % 2 == 0, rng.integers(1, 64, n), 0).astype(np.uint32)
    flags = np.ones(n, np.uint32) | (locks << BODY_LOCK_SHIFT)
    flags |= np.where(idx % 3 == 0, BODY_LIN_DAMP_REPLACE, 0).astype(np.uint32)
    flags |= np.where(idx % 4 == 1, BODY_ANG_DAMP_REPLACE, 0).astype(np.uint32)
    b["flags"][first:] = flags
    return sc


def scripted_events():
    """body_set_state and free() at run time.  Boxes resting (and falling asleep) on the ground and on a static platform; the
    script (`events` below, applied before the step of that number by both sides): a sleeping box is teleported upwards, the
    platform is moved away from under its boxes (they are woken as its neighbours), a sleeping box is given a velocity, a
    moving body is put to sleep, a box is told it may not sleep, bodies are freed."""
    sc = Scene(name="scripted_events")
    sc.add_static_plane((0, 1, 0), 0.0)                                   # body 0
    slab = sc.add_shape(SHAPE_BOX, (2.0, 0.1, 0.6))
    box = sc.add_shape(SHAPE_BOX, (0.25, 0.25, 0.25))
    ball = sc.add_shape(SHAPE_SPHERE, (0.3,))
    xs = IDENT12.copy()
    xs[9], xs[10], xs[11] = 0.0, 1.0, 4.0
    sc.add_bodies(1, mode=MODE_STATIC, shape=slab, xf=xs)                   # body 1: platform
    n = 6
    xf = np.tile(IDENT12, (n, 1))
    xf[:, 9] = (np.arange(n) - 2.5) * 1.5
    xf[:, 10] = 0.27
    sc.add_bodies(n, shape=box, xf=xf, mass=1.0)                            # bodies 2..7 on the ground
    xp = np.tile(IDENT12, (3, 1))
    xp[:, 9] = np.array([-1.2, 0.0, 1.2])
    xp[:, 10] = 1.1 + 0.26
    xp[:, 11] = 4.0
    sc.add_bodies(3, shape=box, xf=xp, mass=1.0)                            # bodies 8..10 on the platform
    xb = IDENT12.copy()
    xb[9], xb[10], xb[11] = 6.0, 0.31, 0.0
    sc.add_bodies(1, shape=ball, xf=xb, mass=2.0, linvel=(-0.5, 0.0, 0.0))  # body 11: a rolling ball
    return sc


def scripted_events_script():
    """(step, body, kind, 12 floats) records for scripted_events(); kinds as in oracle/ref_harness.cpp."""
    ev = []

    def add(step, body, kind, v):
        row = np.zeros(12, np.float32)
        row[:len(v)] = v
        ev.append((step, body, kind, row))
    up = IDENT12.copy()
    up[0], up[2], up[6], up[8] = np.cos(0.3), np.sin(0.3), -np.sin(0.3), np.cos(0.3)   # turned about y, not quite orthonormal in float
    up[9], up[10], up[11] = -2.25, 1.5, 0.0
    add(70, 3, 0, up * np.float32(1.0005))          # teleport a sleeping box (scaled basis: it is orthonormalised on the way in)
    moved = IDENT12.copy()
    moved[9], moved[10], moved[11] = 0.0, 0.4, 4.0
    add(80, 1, 0, moved)                             # the platform drops by 0.6: its boxes are woken and fall after it
    add(90, 5, 1, (0.0, 3.0, 0.5))                   # a sleeping box is thrown upwards
    add(90, 5, 2, (0.0, 2.0, 0.0))
    add(95, 11, 4, (1.0,))                           # the rolling ball is put to sleep: its velocities are cleared
    add(100, 6, 3, ())                               # free a sleeping box
    add(100, 9, 3, ())                               # free a box that is falling / resting on the platform
    add(110, 2, 5, (0.0,))                           # a sleeping box may no longer sleep: it wakes
    add(120, 11, 4, (0.0,))                          # wake the ball again
    return ev


def area_zones():
    """Areas: bodies in free flight -- nothing ever touches, so a free run follows the reference to rounding --
    (all masks are 0) falling through zones that override gravity and damping by priority, a point-gravity well, monitoring areas that report who
    enters and leaves, a kinematic body carried through one, static bodies in a monitorable and in a plain area, a body on a
    layer the areas do not see, a sleeping body an area is moved onto.  Script: area_zones_script()."""
    sc = Scene(name="area_zones")
    ball = sc.add_shape(SHAPE_SPHERE, (0.2,))
    box = sc.add_shape(SHAPE_BOX, (0.2, 0.15, 0.25))
    cap = sc.add_shape(SHAPE_CAPSULE, (0.15, 0.7))
    zone = sc.add_shape(SHAPE_BOX, (1.0, 0.75, 1.0))        # area shapes
    slab = sc.add_shape(SHAPE_BOX, (1.0, 0.3, 1.0))
    well = sc.add_shape(SHAPE_SPHERE, (2.0,))
    tube = sc.add_shape(SHAPE_CAPSULE, (0.8, 3.0))

    def at(x, y, z, rot=0.0):
        f = IDENT12.copy()
        f[0], f[2], f[6], f[8] = np.cos(rot), np.sin(rot), -np.sin(rot), np.cos(rot)
        f[9], f[10], f[11] = x, y, z
        return f
    # column A (x = 0): three shapes fall through an anti-gravity zone with a damping zone inside it
    for k, (sh, y) in enumerate(((ball, 6.0), (box, 7.5), (cap, 9.0))):
        sc.add_bodies(1, shape=sh, xf=at(0.9 * k - 0.9, y, 0.1 * k, 0.2 * k), angvel=(0.5 * k, 0.2, 0.1), flags=0)
    sc.add_areas(1, shape=zone, xf=at(0.0, 3.0, 0.0), gravity_mode=3, gravity=4.0, gravity_vector=(0, 1, 0), flags=AREA_MONITOR_CB, priority=0)
    sc.add_areas(1, shape=slab, xf=at(0.0, 3.2, 0.0, 0.3), linear_damp_mode=1, linear_damp=2.0, angular_damp_mode=3, angular_damp=3.0, priority=1)
    # column B (x = 5): combine-replace sideways gravity of high priority over a replace zone of low priority
    for k, (sh, y) in enumerate(((box, 6.0), (ball, 7.0))):
        sc.add_bodies(1, shape=sh, xf=at(5.0, y, 0.3 * k), flags=0)
    sc.add_areas(1, shape=zone, xf=at(5.0, 3.0, 0.0), gravity_mode=3, gravity=1.0, gravity_vector=(0, -1, 0), priority=0, flags=AREA_MONITOR_CB)
    sc.add_areas(1, shape=zone, xf=at(5.0, 3.5, 0.0), gravity_mode=2, gravity=3.0, gravity_vector=(1, 0, 0), priority=5)
    # column C (x = 10): a point-gravity well (replace-combine: the default gravity is added on top), unit distance 1
    for k in range(3):
        sc.add_bodies(1, shape=ball, xf=at(10.0 + 0.9 * k, 7.0 + k, 0.5 * k), linvel=(0.3, 0.0, 0.0), flags=0)
    sc.add_areas(1, shape=well, xf=at(10.0, 3.0, 0.0), gravity_mode=4, gravity=6.0, gravity_vector=(0.0, 0.5, 0.0), point_unit_distance=1.0,
                 flags=AREA_GRAVITY_POINT | AREA_MONITOR_CB)
    # column D (x = 15): a monitorable, monitoring area with no override: a kinematic box is carried through it, a static box
    # stands in it; next to it a plain monitoring area with a static box in it (they never pair)
    sc.add_bodies(1, mode=MODE_KINEMATIC, shape=box, xf=at(12.0, 3.0, 0.0), linvel=(1.5, 0.0, 0.0), angvel=(0.0, 0.4, 0.0))
    sc.add_bodies(1, mode=MODE_STATIC, shape=box, xf=at(15.3, 3.2, 0.2))
    sc.add_bodies(1, mode=MODE_STATIC, shape=box, xf=at(20.0, 3.0, 0.0))
    sc.add_areas(1, shape=tube, xf=at(15.0, 3.0, 0.0), flags=AREA_MONITORABLE | AREA_MONITOR_CB)
    sc.add_areas(1, shape=zone, xf=at(20.0, 3.0, 0.0), flags=AREA_MONITOR_CB)
    # column E (x = 25): a ball on layer 2 that the zone's mask does not see, one it does; a zone that only sees layer 4
    sc.add_bodies(1, shape=ball, xf=at(25.0, 5.0, 0.0), layer=2, mask=2, flags=0)
    sc.add_bodies(1, shape=ball, xf=at(25.5, 5.5, 0.0), flags=0)
    sc.add_areas(1, shape=zone, xf=at(25.0, 3.0, 0.0), gravity_mode=3, gravity=0.0, flags=AREA_MONITOR_CB)
    sc.add_areas(1, shape=zone, xf=at(25.0, 3.0, 0.0), mask=4, gravity_mode=3, gravity=50.0, flags=AREA_MONITOR_CB, priority=9)
    # column F (x = 30): a body the script puts to sleep in mid-air at step 2 (the reference wakes every body on the first
    # step after its shapes were added, so "asleep from the start" would not hold there); the script moves a zone onto it
    sc.add_bodies(1, shape=box, xf=at(30.0, 3.0, 0.0), flags=1)
    sc.add_areas(1, shape=zone, xf=at(30.0, 9.0, 0.0), gravity_mode=3, gravity=2.0, gravity_vector=(0, 1, 0), flags=AREA_MONITOR_CB)
    # column G (x = 35): a ball that is freed while inside a monitoring zone, and a zone that is freed with a ball inside
    sc.add_bodies(1, shape=ball, xf=at(35.0, 4.5, 0.0), flags=0)
    sc.add_areas(1, shape=zone, xf=at(35.0, 3.0, 0.0), linear_damp_mode=3, linear_damp=8.0, flags=AREA_MONITOR_CB)
    sc.add_bodies(1, shape=ball, xf=at(40.0, 4.5, 0.0), flags=0)
    sc.add_areas(1, shape=zone, xf=at(40.0, 3.0, 0.0), gravity_mode=3, gravity=0.5, gravity_vector=(0, 1, 0), flags=AREA_MONITOR_CB)
    # column H (x = 50): areas watching areas.  12 watches and can be watched, 13 can only be watched, 14 only watches, 15 does
    # both but on a layer nobody's mask meets; 16 overlaps 12 from the start.  The script moves 13, 14 and 15 onto 12.
    both = AREA_MONITORABLE | AREA_AREA_MONITOR_CB
    sc.add_areas(1, shape=zone, xf=at(50.0, 3.0, 0.0), flags=both)
    sc.add_areas(1, shape=slab, xf=at(50.0, 9.0, 0.0), flags=AREA_MONITORABLE)
    sc.add_areas(1, shape=well, xf=at(56.0, 3.0, 0.0), flags=AREA_AREA_MONITOR_CB)
    sc.add_areas(1, shape=zone, xf=at(50.0, -6.0, 0.0), flags=both, layer=8, mask=8)
    sc.add_areas(1, shape=tube, xf=at(50.8, 3.0, 0.3, 0.4), flags=both)
    sc.bodies["mask"] = 0          # no body collides with any other: free flight only (an area looks at the body's LAYER)
    return sc


def area_groups():
    """Areas of several shapes: in the scene file (and to the C ABI) every further shape is an area record of its own carrying the
    same non-zero `group` as the record before it; the harness adds it to the area before (area_add_shape).  Bodies in free
    flight (no masks, nothing touches): a column falls through an area of two overlapping boxes that adds a sideways gravity
    (a body in both boxes must feel it once), another through an area of two overlapping spheres that adds linear damping and watches areas (its own shapes must not be told of each other),
    with a plain replace-gravity zone of higher priority cutting through the first."""
    sc = Scene(name="area_groups")
    ball = sc.add_shape(SHAPE_SPHERE, (0.2,))
    box = sc.add_shape(SHAPE_BOX, (0.2, 0.15, 0.25))
    slab = sc.add_shape(SHAPE_BOX, (1.5, 0.6, 1.5))
    orb = sc.add_shape(SHAPE_SPHERE, (0.9,))
    zone = sc.add_shape(SHAPE_BOX, (0.5, 0.3, 2.0))

    def at(x, y, z):
        f = IDENT12.copy()
        f[9], f[10], f[11] = x, y, z
        return f
    for k in range(6):
        sc.add_bodies(1, shape=(ball if k % 2 else box), xf=at(-1.0 + 0.4 * k, 5.0 + 0.5 * k, 0.2 * k), angvel=(0.3, 0.1 * k, 0.0), flags=0, mask=0)
    for k in range(4):
        sc.add_bodies(1, shape=ball, xf=at(8.0 + (2.2 if k % 2 else 0.0), 5.0 + 0.7 * k, 0.0), flags=0, mask=0)
    # area 0 + 1: one area of two overlapping boxes (group 1)
    common = dict(gravity_mode=1, gravity=3.0, gravity_vector=(1, 0, 0), flags=AREA_MONITOR_CB, priority=0, group=1)
    sc.add_areas(1, shape=slab, xf=at(0.0, 2.5, 0.0), shape_xf=at(0.0, 0.4, 0.0), **common)
    sc.add_areas(1, shape=slab, xf=at(0.0, 2.5, 0.0), shape_xf=at(0.3, -0.4, 0.0), **common)
    # area 2: a plain zone of higher priority through the middle of it: replaces gravity, the walk ends there for gravity
    sc.add_areas(1, shape=zone, xf=at(0.0, 2.5, 0.0), gravity_mode=3, gravity=2.0, gravity_vector=(0, -1, 0), priority=3, flags=AREA_MONITOR_CB)
    # area 3 + 4: one area of two separate spheres (group 2): combine linear damping
    common = dict(linear_damp_mode=1, linear_damp=1.5, flags=AREA_MONITOR_CB | AREA_MONITORABLE | AREA_AREA_MONITOR_CB, priority=1, group=2)   # (its two shapes overlap nothing but could be told of each other: they must not)
    sc.add_areas(1, shape=orb, xf=at(9.0, 2.5, 0.0), shape_xf=at(-0.8, 0.0, 0.0), **common)
    sc.add_areas(1, shape=orb, xf=at(9.0, 2.5, 0.0), shape_xf=at(0.8, 0.3, 0.0), **common)
    return sc


def area_zones_script():
    """(step, index, kind, 12 floats): kind 6 = area_set_transform(area index), 7 = free(area), 3 = free(body), 4 = sleeping."""
    ev = []

    def add(step, idx, kind, v=()):
        row = np.zeros(12, np.float32)
        row[:len(v)] = v
        ev.append((step, idx, kind, row))
    moved = IDENT12.copy()
    moved[9], moved[10], moved[11] = 30.0, 3.3, 0.2
    add(2, 13, 4, (1.0,))         # the box of column F is put to sleep where it is
    add(40, 9, 6, moved)          # the zone of column F comes down onto the sleeping box: it is told although the box sleeps
    add(60, 13, 4, (0.0,))        # ... and only when the box is woken does the override move it
    add(55, 14, 3)                # free the ball of column G inside its zone: the zone hears it leave
    add(50, 11, 7)                # free the last zone with its ball inside: the ball falls on under default gravity
    away = IDENT12.copy()
    away[9], away[10], away[11] = 5.0, 30.0, 0.0
    add(70, 3, 6, away)           # the sideways zone of column B is taken away while bodies may still be in it

    def place(x, y, z):
        f = IDENT12.copy()
        f[9], f[10], f[11] = x, y, z
        return f
    add(20, 13, 6, place(50.0, 3.5, 0.0))     # the watched-only slab comes down into 12 (and 16): both hear it, it hears nothing
    add(30, 14, 6, place(52.5, 3.0, 0.0))     # the watch-only well reaches 12, 13 and 16: it hears all three, none hears it
    add(45, 15, 6, place(50.0, 3.0, 0.0))     # layer 8 meets nobody's mask: silence both ways
    add(60, 13, 6, place(50.0, 9.0, 0.0))     # the slab leaves again
    add(90, 16, 7)                            # the tube is freed inside 12: 12 and the well hear it leave
    add(100, 12, 6, place(50.0, 20.0, 0.0))   # 12 itself moves off: it leaves the well
    return ev


def terrain_faces(n=12, size=8.0, height=0.6, seed=3):
    """A bumpy square of 2 n^2 triangles (vertices jittered so that no two triangle centres tie on an axis), facing up."""
    rng = np.random.default_rng(seed)
    xs = np.linspace(-size / 2, size / 2, n + 1)
    gx, gz = np.meshgrid(xs, xs, indexing="ij")
    gx = gx + rng.uniform(-0.05, 0.05, gx.shape)
    gz = gz + rng.uniform(-0.05, 0.05, gz.shape)
    gy = height * (np.sin(gx * 0.9) * np.cos(gz * 0.7) + 0.3 * rng.uniform(-1, 1, gx.shape))
    p = np.stack([gx, gy, gz], -1).astype(np.float32)
    faces = []
    for i in range(n):
        for j in range(n):
            a, b, c, d = p[i, j], p[i + 1, j], p[i + 1, j + 1], p[i, j + 1]
            faces.append((a, b, c))        # winding: the reference's face normal (v0 - v2) x (v0 - v1) points up, the solid side is below
            faces.append((a, c, d))
    return np.array(faces, np.float32).reshape(-1, 3, 3)


def heightmap_pile(n=5, seed=12):
    """terrain_pile() with a HeightMapShape3D (13 x 13 heights, one unit apart) in place of the triangle mesh."""
    sc = terrain_pile(n, seed)
    rng = np.random.default_rng(seed + 1)
    w = d = 13
    gx, gz = np.meshgrid(np.arange(w) - 6.0, np.arange(d) - 6.0, indexing="xy")
    h = (0.5 * np.sin(gx * 0.9) * np.cos(gz * 0.7) + 0.15 * rng.uniform(-1, 1, gx.shape)).astype(np.float32)
    verts = np.zeros((w * d, 3), np.float32)
    verts[:, 0] = h.reshape(-1)               # row by row: index = z * width + x
    hm = sc.add_shape(SHAPE_HEIGHTMAP, (w, d, float(h.min()) - 0.1, float(h.max()) + 0.1), verts=verts)
    sc.bodies["shape"][1] = hm                # body 1 was the triangle-mesh terrain
    sc.name = "heightmap_pile"
    return sc


def terrain_pile(n=5, seed=11, backface=False):
    """Static concave colliders: spheres, boxes, capsules and hulls dropped onto a bumpy triangle mesh with a
    second, tilted mesh (a ramp made of two big triangles) above one corner; world boundaries far below catch what falls off."""
    sc = Scene(name="terrain_pile")
    sc.add_static_plane((0, 1, 0), -6.0)
    tf = terrain_faces()
    terrain = sc.add_shape(SHAPE_CONCAVE, (1.0 if backface else 0.0,), verts=tf.reshape(-1, 3))
    ramp_faces = np.array([[(-1.5, 0, -1.5), (1.5, 0, 1.5), (1.5, 0, -1.5)], [(-1.5, 0, -1.5), (-1.5, 0, 1.5), (1.5, 0, 1.5)]], np.float32)
    ramp = sc.add_shape(SHAPE_CONCAVE, (1.0,), verts=ramp_faces.reshape(-1, 3))
    sc.add_bodies(1, mode=MODE_STATIC, shape=terrain)
    x = IDENT12.copy()
    c, s = np.cos(0.35), np.sin(0.35)
    x[0], x[1], x[3], x[4] = c, -s, s, c          # tilted about z
    x[9], x[10], x[11] = 2.0, 2.2, 2.0
    sc.add_bodies(1, mode=MODE_STATIC, shape=ramp, xf=x)
    rng = np.random.default_rng(seed)
    hulls = [sc.add_shape(SHAPE_CONVEX, verts=rng.normal(size=(14, 3)) * 0.22) for _ in range(3)]
    g = np.stack(np.meshgrid(np.arange(n), np.arange(n), indexing="ij"), -1).reshape(-1, 2)
    for k, (i, j) in enumerate(g):
        kind = k % 4
        if kind == 0:
            sh = sc.add_shape(SHAPE_SPHERE, (rng.uniform(0.18, 0.3),))
        elif kind == 1:
            sh = sc.add_shape(SHAPE_BOX, tuple(rng.uniform(0.15, 0.3, 3)))
        elif kind == 2:
            sh = sc.add_shape(SHAPE_CAPSULE, (rng.uniform(0.12, 0.2), rng.uniform(0.6, 0.9)))
        else:
            sh = hulls[k % 3]
        f = IDENT12.copy()
        ang = rng.uniform(0, 6.28)
        f[0], f[2], f[6], f[8] = np.cos(ang), np.sin(ang), -np.sin(ang), np.cos(ang)
        f[9], f[10], f[11] = (i - (n - 1) / 2) * 1.3 + rng.uniform(-0.1, 0.1), 2.0 + 0.9 * (k % 3) + (2.5 if (i >= n - 2 and j >= n - 2) else 0.0), (j - (n - 1) / 2) * 1.3 + rng.uniform(-0.1, 0.1)
        sc.add_bodies(1, shape=sh, xf=f, angvel=tuple(rng.uniform(-1, 1, 3)), mass=1.0)
    return sc
