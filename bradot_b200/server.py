"""Host-side mirror of the PhysicsServer3D calls the rigid-body step needs, over the C ABI.

This is the Python face of ``libbradot_physics_cuda.so`` (include/bradot_physics_cuda.h) used by
the tests and ``bench.py``.  Method names follow the reference's server
(servers/physics_server_3d.h:263-814: space_create, space_set_param, *_shape_create,
body_* state access, joint_make_pin / joint_make_hinge, step, get_process_info); bodies are created
in bulk from the scene records because one RID call per body is host-bound at 1 M bodies
(hard part 8).

There is no CPU fallback: if the CUDA library is missing or no GPU is usable the constructor raises.
"""
import ctypes
import os

import numpy as np

from . import scenes

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BPC_LIBRARY", os.path.join(_HERE, "libbradot_physics_cuda.so"))  # override: A/B builds of the same ABI

FIELD_TRANSFORM, FIELD_LINEAR_VELOCITY, FIELD_ANGULAR_VELOCITY, FIELD_ACTIVE = 0, 1, 2, 3
FIELD_MASS_PROPS, FIELD_ITEM_BOX, FIELD_APPLIED_FORCE, FIELD_STILL_TIME = 4, 5, 6, 7
FIELD_KINEMATIC_TARGET, FIELD_KINEMATIC_VELOCITY = 8, 9   # write only, kinematic bodies
FIELD_CONSTANT_FORCE, FIELD_MASS_WORLD = 10, 11
_FIELD_WIDTH = {0: 12, 1: 3, 2: 3, 3: 1, 4: 16, 5: 14, 6: 6, 7: 1, 8: 12, 9: 6, 10: 6, 11: 13}

SPACE_PARAM_CONTACT_RECYCLE_RADIUS, SPACE_PARAM_CONTACT_MAX_SEPARATION = 0, 1
SPACE_PARAM_CONTACT_MAX_ALLOWED_PENETRATION, SPACE_PARAM_CONTACT_DEFAULT_BIAS = 2, 3
SPACE_PARAM_BODY_LINEAR_VELOCITY_SLEEP_THRESHOLD, SPACE_PARAM_BODY_ANGULAR_VELOCITY_SLEEP_THRESHOLD = 4, 5
SPACE_PARAM_BODY_TIME_TO_SLEEP, SPACE_PARAM_SOLVER_ITERATIONS = 6, 7
SPACE_GRAVITY_X, SPACE_GRAVITY_Y, SPACE_GRAVITY_Z, SPACE_LINEAR_DAMP, SPACE_ANGULAR_DAMP = 100, 101, 102, 103, 104

INFO_ACTIVE_OBJECTS, INFO_COLLISION_PAIRS, INFO_ISLAND_COUNT = 0, 1, 2  # PhysicsServer3D::ProcessInfo
BODY_STATE_TRANSFORM, BODY_STATE_LINEAR_VELOCITY, BODY_STATE_ANGULAR_VELOCITY, BODY_STATE_SLEEPING, BODY_STATE_CAN_SLEEP = range(5)
_STATE_WIDTH = {0: 12, 1: 3, 2: 3, 3: 1, 4: 1}

EXPORTS = [
    "bpc_create", "bpc_destroy", "bpc_last_error", "bpc_set_stream", "bpc_space_create", "bpc_space_set_param",
    "bpc_shape_create", "bpc_shape_create_convex", "bpc_shape_create_concave", "bpc_shape_create_heightmap", "bpc_bodies_create", "bpc_bodies_write", "bpc_bodies_read",
    "bpc_body_count", "bpc_joints_create", "bpc_step", "bpc_synchronize", "bpc_get_stats", "bpc_pair_count",
    "bpc_pairs_read", "bpc_pairs_write", "bpc_set_solve_order", "bpc_sort_u64", "bpc_set_profiling", "bpc_get_stage_times",
    "bpc_space_set_active", "bpc_space_destroy", "bpc_bodies_create_slots", "bpc_bodies_destroy", "bpc_bodies_update",
    "bpc_bodies_set_state", "bpc_bodies_read_slots", "bpc_bodies_write_slots", "bpc_bodies_read_state", "bpc_reserve",
    "bpc_shape_set_custom_bias", "bpc_bodies_set_collision_priority", "bpc_bodies_closest_points", "bpc_bodies_set_same_object", "bpc_bodies_set_mass_override", "bpc_bodies_read_begin", "bpc_bodies_read_end", "bpc_joints_write", "bpc_joints_destroy", "bpc_joint_count", "bpc_joint_params_write", "bpc_joint_params_read", "bpc_test_segment_math", "bpc_bodies_set_exception",
    "bpc_bodies_set_max_contacts_reported", "bpc_body_read_contacts", "bpc_contacts_capacity", "bpc_contacts_read",
    "bpc_sort_u64_noalloc", "bpc_noalloc_last_error", "bpc_space_intersect_rays", "bpc_space_intersect_points", "bpc_space_intersect_shapes", "bpc_space_collide_shapes",
    "bpc_space_set_query_exclude", "bpc_space_rest_infos", "bpc_space_cast_motions",
    "bpc_bodies_test_motion", "bpc_areas_create", "bpc_areas_write", "bpc_areas_destroy", "bpc_area_count", "bpc_area_events_read", "bpc_body_read_areas",
]
STAGES = ("integrate_forces", "broadphase", "islands", "narrowphase", "schedule", "solve", "integrate_velocities")


class Stats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in ("active_objects", "collision_pairs", "island_count", "contact_pairs",
                                              "colors", "overflow", "steps", "kernel_launches", "active_contacts",
                                              "full_pairs", "barrier_constraints")]


class PhysicsError(RuntimeError):
    pass


CONTACT_DT = np.dtype([("local_pos", "<f4", 3), ("depth", "<f4"), ("local_normal", "<f4", 3), ("collider", "<i4"), ("local_velocity", "<f4", 3),
                       ("collider_pos", "<f4", 3), ("collider_velocity", "<f4", 3), ("impulse", "<f4", 3)])
MOTION_PARAMS_DT = np.dtype([("body", "<i4"), ("from", "<f4", 12), ("motion", "<f4", 3), ("margin", "<f4"), ("max_collisions", "<i4"), ("flags", "<u4")])
MOTION_COLLISION_DT = np.dtype([("position", "<f4", 3), ("normal", "<f4", 3), ("collider_velocity", "<f4", 3), ("collider_angular_velocity", "<f4", 3),
                                ("depth", "<f4"), ("collider", "<i4")])
MOTION_RESULT_DT = np.dtype([("collided", "<i4"), ("collision_count", "<i4"), ("travel", "<f4", 3), ("remainder", "<f4", 3), ("collision_depth", "<f4"),
                             ("collision_safe_fraction", "<f4"), ("collision_unsafe_fraction", "<f4"), ("collisions", MOTION_COLLISION_DT, 32)])
REST_INFO_DT = np.dtype([("hit", "<i4"), ("body", "<i4"), ("point", "<f4", 3), ("normal", "<f4", 3), ("linear_velocity", "<f4", 3)])
RAY_RESULT_DT = np.dtype([("hit", "<i4"), ("body", "<i4"), ("position", "<f4", 3), ("normal", "<f4", 3), ("face_index", "<i4")])


def load_library(path=LIB_PATH):
    """dlopen the CUDA backend; raises if it has not been built (no fallback exists)."""
    if not os.path.exists(path):
        raise PhysicsError("CUDA backend %s is missing: run `python -m bradot_b200.build` (there is no CPU fallback)" % path)
    lib = ctypes.CDLL(path)
    vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    lib.bpc_create.restype = vp
    lib.bpc_create.argtypes = [ci]
    lib.bpc_destroy.restype = None
    lib.bpc_destroy.argtypes = [vp]
    lib.bpc_last_error.restype = ctypes.c_char_p
    lib.bpc_last_error.argtypes = [vp]
    lib.bpc_set_stream.argtypes = [vp, vp]
    lib.bpc_space_create.argtypes = [vp]
    lib.bpc_space_set_param.argtypes = [vp, ci, ci, cf]
    lib.bpc_shape_create.argtypes = [vp, ci, vp]
    lib.bpc_shape_create_convex.argtypes = [vp, vp, ci, vp, vp, ci, vp, ci, vp, ci]
    lib.bpc_shape_create_concave.argtypes = [vp, vp, ci, ci, vp, ci, vp]
    lib.bpc_shape_create_heightmap.argtypes = [vp, vp, ci, ci, cf, cf]
    lib.bpc_bodies_create.argtypes = [vp, vp, ci]
    lib.bpc_bodies_write.argtypes = [vp, ci, ci, ci, vp]
    lib.bpc_bodies_read.argtypes = [vp, ci, ci, ci, vp]
    lib.bpc_body_count.argtypes = [vp]
    lib.bpc_joints_create.argtypes = [vp, vp, ci]
    lib.bpc_joints_write.argtypes = [vp, ci, ci, vp]
    lib.bpc_joints_destroy.argtypes = [vp, ci, ci]
    lib.bpc_shape_set_custom_bias.argtypes = [vp, ci, ctypes.c_float]
    lib.bpc_bodies_read_begin.argtypes = [vp, ci, ci, ci, vp]
    lib.bpc_bodies_set_collision_priority.argtypes = [vp, vp, ci, vp]
    lib.bpc_bodies_closest_points.argtypes = [vp, ci, vp, vp, vp]
    lib.bpc_bodies_set_same_object.argtypes = [vp, ci, ci, ci]
    lib.bpc_bodies_set_mass_override.argtypes = [vp, vp, ci, vp, vp]
    lib.bpc_bodies_read_end.argtypes = [vp]
    lib.bpc_joint_params_write.argtypes = [vp, ci, ci, ci, vp]
    lib.bpc_joint_params_read.argtypes = [vp, ci, ci, ci, vp]
    lib.bpc_joint_count.argtypes = [vp]
    lib.bpc_space_set_active.argtypes = [vp, ci, ci]
    lib.bpc_space_destroy.argtypes = [vp, ci]
    lib.bpc_bodies_create_slots.argtypes = [vp, vp, ci, vp]
    lib.bpc_bodies_destroy.argtypes = [vp, vp, ci]
    lib.bpc_bodies_update.argtypes = [vp, vp, ci, vp, ci]
    lib.bpc_bodies_set_state.argtypes = [vp, ci, vp, ci, vp]
    lib.bpc_bodies_read_slots.argtypes = [vp, ci, vp, ci, vp]
    lib.bpc_bodies_write_slots.argtypes = [vp, ci, vp, ci, vp]
    lib.bpc_bodies_read_state.argtypes = [vp, ci, ci, vp]
    lib.bpc_reserve.argtypes = [vp, ci, ci]
    lib.bpc_bodies_set_exception.argtypes = [vp, ci, ci, ci]
    lib.bpc_bodies_set_max_contacts_reported.argtypes = [vp, vp, ci, vp]
    lib.bpc_body_read_contacts.argtypes = [vp, ci, ci, vp]
    lib.bpc_contacts_capacity.argtypes = [vp]
    lib.bpc_contacts_read.argtypes = [vp, vp, vp, vp]
    lib.bpc_test_segment_math.argtypes = [vp, ci, ci, vp, vp]
    lib.bpc_step.argtypes = [vp, cf]
    lib.bpc_space_intersect_rays.argtypes = [vp, ci, ci, vp, ctypes.c_uint, ci, vp]
    lib.bpc_space_intersect_points.argtypes = [vp, ci, ci, vp, ctypes.c_uint, ci, ci, vp, vp]
    lib.bpc_space_intersect_shapes.argtypes = [vp, ci, ci, vp, vp, ctypes.c_uint, ci, cf, ci, vp, vp]
    lib.bpc_space_collide_shapes.argtypes = [vp, ci, ci, vp, vp, ctypes.c_uint, cf, ci, vp, vp]
    lib.bpc_space_set_query_exclude.argtypes = [vp, vp, ci]
    lib.bpc_bodies_test_motion.argtypes = [vp, ci, vp, vp]
    lib.bpc_areas_create.argtypes = [vp, vp, ci]
    lib.bpc_areas_write.argtypes = [vp, ci, ci, vp]
    lib.bpc_areas_destroy.argtypes = [vp, vp, ci]
    lib.bpc_area_count.argtypes = [vp]
    lib.bpc_area_events_read.argtypes = [vp, vp, ci]
    lib.bpc_body_read_areas.argtypes = [vp, ci, vp, ci]
    lib.bpc_space_rest_infos.argtypes = [vp, ci, ci, vp, vp, vp, cf, ctypes.c_uint, vp]
    lib.bpc_space_cast_motions.argtypes = [vp, ci, ci, vp, vp, vp, cf, ctypes.c_uint, vp, vp]
    lib.bpc_synchronize.argtypes = [vp]
    lib.bpc_get_stats.argtypes = [vp, ctypes.POINTER(Stats)]
    lib.bpc_pair_count.argtypes = [vp]
    lib.bpc_pairs_read.argtypes = [vp, vp, vp, vp]
    lib.bpc_pairs_write.argtypes = [vp, ci, vp, vp, vp]
    lib.bpc_set_solve_order.argtypes = [vp, ci, vp]
    lib.bpc_sort_u64.argtypes = [vp, vp, vp, ci]
    lib.bpc_sort_u64_noalloc.argtypes = [vp, vp, ci, vp]
    lib.bpc_sort_u64_noalloc.restype = ci
    lib.bpc_noalloc_last_error.argtypes = []
    lib.bpc_noalloc_last_error.restype = ctypes.c_char_p
    lib.bpc_set_profiling.argtypes = [vp, ci]
    lib.bpc_get_stage_times.argtypes = [vp, vp, vp]
    for name in EXPORTS:
        if name not in ("bpc_create", "bpc_destroy", "bpc_last_error", "bpc_noalloc_last_error"):
            getattr(lib, name).restype = ci
    return lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class PhysicsServerCUDA:
    """One context on one GPU ("BradotPhysicsCUDA")."""

    def __init__(self, device=0, stream=None):
        self.lib = load_library()
        self.ctx = self.lib.bpc_create(device)
        if not self.ctx:
            raise PhysicsError("bpc_create(%d) failed: no usable CUDA device (there is no CPU fallback)" % device)
        if stream:
            self._check(self.lib.bpc_set_stream(self.ctx, ctypes.c_void_p(stream)))
        self.mesh_source = None

    # -- plumbing -----------------------------------------------------------------------------
    def _check(self, rc):
        if rc < 0:
            raise PhysicsError(self.lib.bpc_last_error(self.ctx).decode())
        return rc

    def finish(self):
        if getattr(self, "ctx", None):
            self.lib.bpc_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.finish()
        except Exception:
            pass

    # -- spaces -------------------------------------------------------------------------------
    def space_create(self):
        return self._check(self.lib.bpc_space_create(self.ctx))

    def space_set_param(self, space, param, value):
        self._check(self.lib.bpc_space_set_param(self.ctx, space, param, float(value)))

    # -- shapes -------------------------------------------------------------------------------
    def shape_create(self, type_, params):
        p = np.zeros(4, np.float32)
        p[:len(params)] = params
        return self._check(self.lib.bpc_shape_create(self.ctx, int(type_), _ptr(p)))

    def world_boundary_shape_create(self, normal, d):
        return self.shape_create(scenes.SHAPE_PLANE, (normal[0], normal[1], normal[2], d))

    def sphere_shape_create(self, radius):
        return self.shape_create(scenes.SHAPE_SPHERE, (radius,))

    def box_shape_create(self, half_extents):
        return self.shape_create(scenes.SHAPE_BOX, tuple(half_extents))

    def capsule_shape_create(self, radius, height):
        return self.shape_create(scenes.SHAPE_CAPSULE, (radius, height))

    def convex_polygon_shape_create(self, mesh):
        """mesh: dict with verts (n,3) f4, planes (f,4) f4, faces (f,2) i4, index (k,) i4, edges (e,4) i4."""
        v = np.ascontiguousarray(mesh["verts"], np.float32)
        pl = np.ascontiguousarray(mesh["planes"], np.float32)
        fa = np.ascontiguousarray(mesh["faces"], np.int32)
        ix = np.ascontiguousarray(mesh["index"], np.int32)
        ed = np.ascontiguousarray(mesh["edges"], np.int32)
        return self._check(self.lib.bpc_shape_create_convex(self.ctx, _ptr(v), len(v), _ptr(pl), _ptr(fa), len(fa), _ptr(ix), len(ix),
                                                            _ptr(ed), len(ed)))

    def concave_polygon_shape_create(self, faces, backface_collision=False, bvh=None, normals=None):
        """faces (n, 3, 3) or (n, 9) float32; bvh (nodes, 9) + normals (n, 3): a ready tree (the reference's, from a dump)."""
        f = np.ascontiguousarray(faces, np.float32).reshape(-1, 9)
        b = None if bvh is None else np.ascontiguousarray(bvh, np.float32).reshape(-1, 9)
        nr = None if normals is None else np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        return self._check(self.lib.bpc_shape_create_concave(self.ctx, _ptr(f), len(f), 1 if backface_collision else 0, None if b is None else _ptr(b),
                                                             0 if b is None else len(b), None if nr is None else _ptr(nr)))

    def heightmap_shape_create(self, heights, width, depth, min_height, max_height):
        h = np.ascontiguousarray(heights, np.float32).reshape(-1)
        assert len(h) == width * depth
        return self._check(self.lib.bpc_shape_create_heightmap(self.ctx, _ptr(h), int(width), int(depth), ctypes.c_float(min_height), ctypes.c_float(max_height)))

    # -- bodies -------------------------------------------------------------------------------
    def bodies_create(self, records):
        rec = np.ascontiguousarray(records, scenes.BODY_DT)
        return self._check(self.lib.bpc_bodies_create(self.ctx, _ptr(rec), len(rec)))

    def body_count(self):
        return self._check(self.lib.bpc_body_count(self.ctx))

    def bodies_read(self, field, first=0, n=None, out=None):
        """Read `n` bodies' `field`; `out` may be a caller-owned float32 array (e.g. pinned host memory)."""
        n = self.body_count() - first if n is None else n
        if out is None:
            out = np.zeros((n, _FIELD_WIDTH[field]), np.float32)
        assert out.dtype == np.float32 and out.size == n * _FIELD_WIDTH[field] and out.flags["C_CONTIGUOUS"]
        self._check(self.lib.bpc_bodies_read(self.ctx, field, first, n, _ptr(out)))
        return out

    def bodies_write(self, field, values, first=0):
        v = np.ascontiguousarray(values, np.float32).reshape(-1, _FIELD_WIDTH[field])
        self._check(self.lib.bpc_bodies_write(self.ctx, field, first, len(v), _ptr(v)))

    def bodies_create_slots(self, records):
        """Create bodies, reusing the slots of destroyed ones first; returns their indices."""
        rec = np.ascontiguousarray(records, scenes.BODY_DT)
        slots = np.zeros(len(rec), np.int32)
        self._check(self.lib.bpc_bodies_create_slots(self.ctx, _ptr(rec), len(rec), _ptr(slots)))
        return slots

    def bodies_destroy(self, slots):
        """free(RID) of bodies (servers/physics_server_3d.h: free)."""
        sl = np.ascontiguousarray(slots, np.int32).reshape(-1)
        self._check(self.lib.bpc_bodies_destroy(self.ctx, _ptr(sl), len(sl)))

    def bodies_update(self, slots, records, keep_state=True):
        """body_set_mode / body_set_param / body_set_collision_layer ... on bodies that are already in a space."""
        sl = np.ascontiguousarray(slots, np.int32).reshape(-1)
        rec = np.ascontiguousarray(records, scenes.BODY_DT)
        assert len(sl) == len(rec)
        self._check(self.lib.bpc_bodies_update(self.ctx, _ptr(sl), len(sl), _ptr(rec), 1 if keep_state else 0))

    def body_set_state(self, slots, state, values):
        """body_set_state with the reference's side effects (wake-ups, orthonormalised transform, broadphase box)."""
        sl = np.ascontiguousarray(slots, np.int32).reshape(-1)
        v = np.ascontiguousarray(values, np.float32).reshape(len(sl), _STATE_WIDTH[state])
        self._check(self.lib.bpc_bodies_set_state(self.ctx, int(state), _ptr(sl), len(sl), _ptr(v)))

    def bodies_read_slots(self, field, slots):
        sl = np.ascontiguousarray(slots, np.int32).reshape(-1)
        out = np.zeros((len(sl), _FIELD_WIDTH[field]), np.float32)
        self._check(self.lib.bpc_bodies_read_slots(self.ctx, field, _ptr(sl), len(sl), _ptr(out)))
        return out

    def bodies_write_slots(self, field, slots, values):
        sl = np.ascontiguousarray(slots, np.int32).reshape(-1)
        v = np.ascontiguousarray(values, np.float32).reshape(len(sl), _FIELD_WIDTH[field])
        self._check(self.lib.bpc_bodies_write_slots(self.ctx, field, _ptr(sl), len(sl), _ptr(v)))

    def bodies_read_state(self, first=0, n=None, out=None):
        """(n, 19): transform, linear velocity, angular velocity, active flag of every body in one copy."""
        n = self.body_count() - first if n is None else n
        if out is None:
            out = np.zeros((n, 19), np.float32)
        self._check(self.lib.bpc_bodies_read_state(self.ctx, first, n, _ptr(out)))
        return out

    def test_segment_math(self, op, rows):
        r = np.ascontiguousarray(rows, np.float32).reshape(-1, 12 if op == 0 else 9)
        out = np.zeros((len(r), 6 if op == 0 else 3), np.float32)
        self._check(self.lib.bpc_test_segment_math(self.ctx, int(op), len(r), _ptr(r), _ptr(out)))
        return out

    def body_set_collision_exception(self, a, b, excepted=True):
        self._check(self.lib.bpc_bodies_set_exception(self.ctx, int(a), int(b), 1 if excepted else 0))

    def bodies_set_max_contacts_reported(self, slots, counts):
        sl = np.ascontiguousarray(slots, np.int32).reshape(-1)
        c = np.ascontiguousarray(counts, np.int32).reshape(-1)
        self._check(self.lib.bpc_bodies_set_max_contacts_reported(self.ctx, _ptr(sl), len(sl), _ptr(c)))

    def body_contacts(self, slot, max_records=64):
        """Contacts reported to one body in the last step (structured array, CONTACT_DT)."""
        out = np.zeros(max_records, CONTACT_DT)
        n = self._check(self.lib.bpc_body_read_contacts(self.ctx, int(slot), max_records, _ptr(out)))
        return out[:min(n, max_records)]

    def contacts_read(self):
        """(counts, offsets, records) of every body's reported contacts in one copy."""
        n = self.body_count()
        cap = self._check(self.lib.bpc_contacts_capacity(self.ctx))
        counts, offsets = np.zeros(n, np.int32), np.zeros(n, np.int32)
        rec = np.zeros(max(cap, 1), CONTACT_DT)
        self._check(self.lib.bpc_contacts_read(self.ctx, _ptr(counts), _ptr(offsets), _ptr(rec)))
        return counts, offsets, rec[:cap]

    def reserve(self, bodies, pairs):
        self._check(self.lib.bpc_reserve(self.ctx, int(bodies), int(pairs)))

    def space_set_active(self, space, active):
        self._check(self.lib.bpc_space_set_active(self.ctx, int(space), 1 if active else 0))

    def space_destroy(self, space):
        self._check(self.lib.bpc_space_destroy(self.ctx, int(space)))

    def joints_create(self, records):
        rec = np.ascontiguousarray(records, scenes.JOINT_DT)
        return self._check(self.lib.bpc_joints_create(self.ctx, _ptr(rec), len(rec)))

    def joints_write(self, first, records):
        rec = np.ascontiguousarray(records, scenes.JOINT_DT)
        self._check(self.lib.bpc_joints_write(self.ctx, int(first), len(rec), _ptr(rec)))

    def bodies_closest_points(self, bodies, points):
        """Batched get_closest_point_to_object_volume: for each (body, world point) the nearest point of the body's shape."""
        b = np.ascontiguousarray(bodies, np.int32).reshape(-1)
        p = np.ascontiguousarray(points, np.float32).reshape(len(b), 3)
        out = np.zeros((len(b), 3), np.float32)
        self._check(self.lib.bpc_bodies_closest_points(self.ctx, len(b), _ptr(b), _ptr(p), _ptr(out)))
        return out

    def bodies_set_collision_priority(self, slots, priorities):
        """body_set_collision_priority: the obstacles' weights when body_test_motion frees a stuck body."""
        s = np.ascontiguousarray(slots, np.int32).reshape(-1)
        v = np.ascontiguousarray(priorities, np.float32).reshape(len(s))
        self._check(self.lib.bpc_bodies_set_collision_priority(self.ctx, _ptr(s), len(s), _ptr(v)))

    def bodies_set_same_object(self, a, b, same=True):
        """The two records are shapes of one object: they never enter the pair list."""
        self._check(self.lib.bpc_bodies_set_same_object(self.ctx, int(a), int(b), 1 if same else 0))

    def bodies_set_mass_override(self, slots, inertia=None, center_of_mass=None):
        """body_set_param(INERTIA / CENTER_OF_MASS): rows of 3 per listed body, None = leave as it is; both None = back to the
        values computed from the shape.  Takes effect when the bodies are described next (bodies_update)."""
        s = np.ascontiguousarray(slots, np.int32).reshape(-1)
        i = None if inertia is None else np.ascontiguousarray(inertia, np.float32).reshape(len(s), 3)
        c = None if center_of_mass is None else np.ascontiguousarray(center_of_mass, np.float32).reshape(len(s), 3)
        self._check(self.lib.bpc_bodies_set_mass_override(self.ctx, _ptr(s), len(s), None if i is None else _ptr(i), None if c is None else _ptr(c)))

    def bodies_read_begin(self, field, out, first=0):
        """Queue a pipelined read of TRANSFORM / a velocity of all bodies from `first` on into `out` (a pinned float32 array);
        bodies_read_end() waits for it.  The copy overlaps whatever is issued in between, the next step included."""
        n = out.shape[0]
        self._check(self.lib.bpc_bodies_read_begin(self.ctx, int(field), int(first), int(n), _ptr(out)))

    def bodies_read_end(self):
        self._check(self.lib.bpc_bodies_read_end(self.ctx))

    def shape_set_custom_bias(self, shape, bias):
        """shape_set_custom_solver_bias: the contact bias of pairs with this shape (0 = the default 0.8)."""
        self._check(self.lib.bpc_shape_set_custom_bias(self.ctx, int(shape), float(bias)))

    def joint_params_write(self, joint, first, values):
        """slider_joint_set_param: values of the joint's long parameter table from parameter `first` on."""
        v = np.ascontiguousarray(values, np.float32).reshape(-1)
        self._check(self.lib.bpc_joint_params_write(self.ctx, int(joint), int(first), len(v), _ptr(v)))

    def joint_params_read(self, joint, first=0, n=22):   # (a 6-DOF joint has 84: three blocks of 28)
        v = np.zeros(n, np.float32)
        self._check(self.lib.bpc_joint_params_read(self.ctx, int(joint), int(first), int(n), _ptr(v)))
        return v

    def joints_destroy(self, first, n=1):
        self._check(self.lib.bpc_joints_destroy(self.ctx, int(first), int(n)))

    # -- stepping -----------------------------------------------------------------------------
    def step(self, dt):
        self._check(self.lib.bpc_step(self.ctx, float(dt)))

    def sync(self):
        self._check(self.lib.bpc_synchronize(self.ctx))

    def stats(self):
        s = Stats()
        self._check(self.lib.bpc_get_stats(self.ctx, ctypes.byref(s)))
        return {n: getattr(s, n) for n, _ in Stats._fields_}

    def set_profiling(self, on):
        self._check(self.lib.bpc_set_profiling(self.ctx, 1 if on else 0))

    def stage_times(self):
        """({stage: total ms}, steps) accumulated since set_profiling(True)."""
        ms = np.zeros(len(STAGES), np.float64)
        n = ctypes.c_int(0)
        self._check(self.lib.bpc_get_stage_times(self.ctx, _ptr(ms), ctypes.byref(n)))
        return dict(zip(STAGES, ms.tolist())), n.value

    def get_process_info(self, info):
        s = self.stats()
        return [s["active_objects"], s["collision_pairs"], s["island_count"]][info]

    # -- pair / manifold snapshot ---------------------------------------------------------------
    def pairs_read(self):
        n = self._check(self.lib.bpc_pair_count(self.ctx))
        ids = np.zeros((n, 4), np.int32)
        pf = np.zeros((n, 3), np.float32)
        cf = np.zeros((n, 4, 16), np.float32)
        if n:
            self._check(self.lib.bpc_pairs_read(self.ctx, _ptr(ids), _ptr(pf), _ptr(cf)))
        return ids, pf, cf

    def pairs_write(self, ids, pf, cf):
        ids = np.ascontiguousarray(ids, np.int32).reshape(-1, 4)
        pf = np.ascontiguousarray(pf, np.float32).reshape(-1, 3)
        cf = np.ascontiguousarray(cf, np.float32).reshape(-1, 4, 16)
        self._check(self.lib.bpc_pairs_write(self.ctx, len(ids), _ptr(ids), _ptr(pf), _ptr(cf)))

    def intersect_rays(self, from_to, space=0, collision_mask=0xFFFFFFFF, hit_from_inside=False, hit_back_faces=True, pick_ray=False, areas=False, bodies=True):
        """Batched PhysicsDirectSpaceState3D.intersect_ray: returns (hit bool[n], body int32[n], position (n,3), normal (n,3))."""
        r = np.ascontiguousarray(from_to, np.float32).reshape(-1, 6)
        out = np.zeros(len(r), RAY_RESULT_DT)
        self._check(self.lib.bpc_space_intersect_rays(self.ctx, int(space), len(r), _ptr(r), ctypes.c_uint(collision_mask),
                                                      (1 if hit_from_inside else 0) | (2 if hit_back_faces else 0) | (4 if pick_ray else 0) | (8 if areas else 0) | (0 if bodies else 16), _ptr(out)))
        self.last_ray_faces = out["face_index"].copy()      # RayResult::face_index of the rays just cast (-1: no mesh face)
        return out["hit"] != 0, out["body"].copy(), out["position"].copy(), out["normal"].copy()

    def intersect_points(self, points, space=0, collision_mask=0xFFFFFFFF, max_results=32, areas=False, bodies=True):
        """Batched PhysicsDirectSpaceState3D.intersect_point: (counts int32[n], bodies int32[n, max_results], -1 padded);
        with `areas`, areas are listed too, as -2 - area index."""
        with_bodies = bodies
        p = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
        counts = np.zeros(len(p), np.int32)
        bodies = np.full((len(p), max_results), -1, np.int32)
        self._check(self.lib.bpc_space_intersect_points(self.ctx, int(space), len(p), _ptr(p), ctypes.c_uint(collision_mask),
                                                        (8 if areas else 0) | (0 if with_bodies else 16), int(max_results), _ptr(counts), _ptr(bodies)))
        return counts, bodies

    def intersect_shapes(self, shapes, xforms, space=0, collision_mask=0xFFFFFFFF, max_results=32, margin=0.0, areas=False, bodies=True):
        """Batched PhysicsDirectSpaceState3D.intersect_shape: shape handles (n) placed at xforms (n, 12); with `areas`, areas are
        listed too, as -2 - area index."""
        with_bodies = bodies
        sh = np.ascontiguousarray(shapes, np.int32).reshape(-1)
        xf = np.ascontiguousarray(xforms, np.float32).reshape(-1, 12)
        assert len(sh) == len(xf)
        counts = np.zeros(len(sh), np.int32)
        bodies = np.full((len(sh), max_results), -1, np.int32)
        self._check(self.lib.bpc_space_intersect_shapes(self.ctx, int(space), len(sh), _ptr(sh), _ptr(xf), ctypes.c_uint(collision_mask),
                                                        (8 if areas else 0) | (0 if with_bodies else 16), ctypes.c_float(margin), int(max_results), _ptr(counts), _ptr(bodies)))
        return counts, bodies

    def collide_shapes(self, shapes, xforms, space=0, collision_mask=0xFFFFFFFF, max_results=16, margin=0.0):
        """Batched PhysicsDirectSpaceState3D.collide_shape: (counts int32[n], points float32[n, max_results, 2, 3])."""
        sh = np.ascontiguousarray(shapes, np.int32).reshape(-1)
        xf = np.ascontiguousarray(xforms, np.float32).reshape(-1, 12)
        counts = np.zeros(len(sh), np.int32)
        pts = np.zeros((len(sh), max_results, 2, 3), np.float32)
        self._check(self.lib.bpc_space_collide_shapes(self.ctx, int(space), len(sh), _ptr(sh), _ptr(xf), ctypes.c_uint(collision_mask),
                                                      ctypes.c_float(margin), int(max_results), _ptr(counts), _ptr(pts)))
        return counts, pts

    # -- areas ------------------------------------------------------------------------------------
    def areas_create(self, records):
        rec = np.ascontiguousarray(records, scenes.AREA_DT)
        return self._check(self.lib.bpc_areas_create(self.ctx, _ptr(rec), len(rec)))

    def areas_write(self, first, records):
        """Any area_set_* on areas [first, first + n): the records again; the areas count as moved for the next step."""
        rec = np.ascontiguousarray(records, scenes.AREA_DT)
        self._check(self.lib.bpc_areas_write(self.ctx, int(first), len(rec), _ptr(rec)))

    def areas_destroy(self, areas):
        a = np.ascontiguousarray(areas, np.int32).reshape(-1)
        self._check(self.lib.bpc_areas_destroy(self.ctx, _ptr(a), len(a)))

    def area_events(self, max_events=65536):
        """Enter / exit events since the last call: int32 (n, 3) = area, body, status (0 added, 1 removed)."""
        out = np.zeros((max_events, 3), np.int32)
        n = self._check(self.lib.bpc_area_events_read(self.ctx, _ptr(out), max_events))
        return out[:min(n, max_events)]

    def body_areas(self, slot):
        """(areas the body stands in, which of them are attached as space overrides)."""
        out = np.zeros(8, np.int32)
        n = self._check(self.lib.bpc_body_read_areas(self.ctx, int(slot), _ptr(out), 8))
        v = out[:min(n, 8)]
        return v & 0x3FFFFFFF, (v & 0x40000000) != 0

    def bodies_test_motion(self, params):
        """Batched PhysicsServer3D.body_test_motion: params MOTION_PARAMS_DT[n] -> MOTION_RESULT_DT[n]."""
        p = np.ascontiguousarray(params, MOTION_PARAMS_DT)
        out = np.zeros(len(p), MOTION_RESULT_DT)
        self._check(self.lib.bpc_bodies_test_motion(self.ctx, len(p), _ptr(p), _ptr(out)))
        return out

    def set_query_exclude(self, bodies):
        """Bodies every following space query skips (the `exclude` list of the query parameters); empty clears."""
        b = np.ascontiguousarray(bodies, np.int32).reshape(-1)
        self._check(self.lib.bpc_space_set_query_exclude(self.ctx, _ptr(b) if len(b) else None, len(b)))

    def rest_infos(self, shapes, xforms, motions=None, margin=0.0, space=0, collision_mask=0xFFFFFFFF):
        """Batched PhysicsDirectSpaceState3D.rest_info: structured array REST_INFO_DT, one record per query."""
        sh = np.ascontiguousarray(shapes, np.int32).reshape(-1)
        xf = np.ascontiguousarray(xforms, np.float32).reshape(-1, 12)
        mo = None if motions is None else np.ascontiguousarray(motions, np.float32).reshape(-1, 3)
        out = np.zeros(len(sh), REST_INFO_DT)
        self._check(self.lib.bpc_space_rest_infos(self.ctx, int(space), len(sh), _ptr(sh), _ptr(xf), None if mo is None else _ptr(mo),
                                                  ctypes.c_float(margin), ctypes.c_uint(collision_mask), _ptr(out)))
        return out

    def cast_motions(self, shapes, xforms, motions, margin=0.0, space=0, collision_mask=0xFFFFFFFF):
        """Batched PhysicsDirectSpaceState3D.cast_motion: (fractions float32[n, 2] = closest safe / unsafe, REST_INFO_DT[n])."""
        sh = np.ascontiguousarray(shapes, np.int32).reshape(-1)
        xf = np.ascontiguousarray(xforms, np.float32).reshape(-1, 12)
        mo = np.ascontiguousarray(motions, np.float32).reshape(-1, 3)
        fr = np.zeros((len(sh), 2), np.float32)
        info = np.zeros(len(sh), REST_INFO_DT)
        self._check(self.lib.bpc_space_cast_motions(self.ctx, int(space), len(sh), _ptr(sh), _ptr(xf), _ptr(mo), ctypes.c_float(margin),
                                                    ctypes.c_uint(collision_mask), _ptr(fr), _ptr(info)))
        return fr, info

    def set_solve_order(self, order):
        o = np.ascontiguousarray(order, np.int32).reshape(-1, 3)
        self._check(self.lib.bpc_set_solve_order(self.ctx, len(o), _ptr(o)))

    def sort_u64(self, keys):
        k = np.ascontiguousarray(keys, np.uint64)
        out = np.zeros_like(k)
        self._check(self.lib.bpc_sort_u64(self.ctx, _ptr(k), _ptr(out), len(k)))
        return out

    # -- scene loading --------------------------------------------------------------------------
    def load_scene(self, sc, meshes=None):
        """Build every space / shape / body / joint of a scenes.Scene.  ``meshes`` maps a convex
        shape index to its hull (Geometry3D::MeshData layout), e.g. the reference's own from an oracle dump; a convex
        shape without an entry gets its hull from bradot_b200.hull (same conventions, Qhull underneath)."""
        for k in range(sc.n_spaces):
            sp = self.space_create()
            self.space_set_param(sp, SPACE_PARAM_SOLVER_ITERATIONS, sc.iterations + k * getattr(sc, "iter_step", 0))
        for i, s in enumerate(sc.shapes):
            if s["type"] == scenes.SHAPE_HEIGHTMAP:     # p = (width, depth, min height, max height); the heights are the x of its vertices
                hs = sc.verts[int(s["vert_ofs"]):int(s["vert_ofs"]) + int(s["vert_cnt"]), 0]
                self.heightmap_shape_create(hs, int(s["p"][0]), int(s["p"][1]), float(s["p"][2]), float(s["p"][3]))
            elif s["type"] == scenes.SHAPE_CONCAVE:
                tree = (meshes or {}).get(("concave", i))
                faces = sc.verts[int(s["vert_ofs"]):int(s["vert_ofs"]) + int(s["vert_cnt"])]
                self.concave_polygon_shape_create(faces, s["p"][0] != 0, None if tree is None else tree[0], None if tree is None else tree[1])
            elif s["type"] == scenes.SHAPE_CONVEX:
                if meshes is None or i not in meshes:
                    from . import hull
                    key = (int(s["vert_ofs"]), int(s["vert_cnt"]))
                    mesh = hull.mesh_data(sc.verts[key[0]:key[0] + key[1]])
                else:
                    mesh = meshes[i]
                self.convex_polygon_shape_create(mesh)
            else:
                self.shape_create(int(s["type"]), s["p"])
        for i, s in enumerate(sc.shapes):
            if s["custom_bias"] != 0:
                self.shape_set_custom_bias(i, float(s["custom_bias"]))
        bodies = sc.bodies
        kin = bodies["mode"] == scenes.MODE_KINEMATIC
        if kin.any():   # in a scene file a kinematic body's velocities are its motion script (oracle/ref_harness.cpp), not body state
            bodies = bodies.copy()
            bodies["linvel"][kin] = 0
            bodies["angvel"][kin] = 0
        first = self.bodies_create(bodies) if len(bodies) else 0
        want = (bodies["flags"] >> scenes.BODY_REPORT_SHIFT) & 0xFF     # body_set_max_contacts_reported
        if want.any():
            idx = np.nonzero(want)[0]
            self.bodies_set_max_contacts_reported(first + idx, want[idx])
        if len(sc.joints):
            self.joints_create(sc.joints)
            for e in sc.joint_ext:
                self.joint_params_write(int(e["joint"]), int(e["first"]), e["values"][:int(e["count"])])
        if len(getattr(sc, "areas", ())):
            self.areas_create(sc.areas)
        return first


def meshes_from_dump(dump):
    """Collect the convex hulls an oracle dump holds (oracle/ref_harness.cpp dump_meshes)."""
    out = {}
    for (name, idx), arr in dump.items():
        if name == "concave_bvh":
            out[("concave", idx)] = (arr, dump[("concave_normals", idx)])
        if name == "mesh_verts":
            out[idx] = {"verts": arr, "planes": dump[("mesh_planes", idx)], "faces": dump[("mesh_faces", idx)],
                        "index": dump[("mesh_index", idx)], "edges": dump[("mesh_edges", idx)]}
    return out
