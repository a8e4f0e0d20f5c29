/* bradot_physics_cuda.h -- C ABI of the B200 rigid-body step ("BradotPhysicsCUDA").
 *
 * This is the boundary the engine-side PhysicsServer3D backend binds: plain C
 * structs, caller-owned host buffers, int return codes (0 = ok, <0 = error, text from
 * bpc_last_error).  No exceptions and no C++/torch types cross it.  One context per GPU; calls on
 * a context must be externally serialised (the engine has a single physics thread).
 *
 * Citations are relative to the reference tree; P3D = modules/bradot_physics_3d.
 * Numeric values of shape types, body modes and body parameters are those of the reference enums
 * (servers/physics_server_3d.h:247-259, 387-392, 439-451).
 */
#ifndef BRADOT_PHYSICS_CUDA_H
#define BRADOT_PHYSICS_CUDA_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bpc_context bpc_context;

enum { BPC_SHAPE_WORLD_BOUNDARY = 0, BPC_SHAPE_SPHERE = 2, BPC_SHAPE_BOX = 3, BPC_SHAPE_CAPSULE = 4, BPC_SHAPE_CONVEX_POLYGON = 6 };
enum { BPC_BODY_MODE_STATIC = 0, BPC_BODY_MODE_KINEMATIC = 1, BPC_BODY_MODE_RIGID = 2, BPC_BODY_MODE_RIGID_LINEAR = 3 };
enum { BPC_JOINT_PIN = 0, BPC_JOINT_HINGE = 1, BPC_JOINT_CONE_TWIST = 2, BPC_JOINT_SLIDER = 3, BPC_JOINT_6DOF = 4 };
#define BPC_SLIDER_PARAMS 22 /* PhysicsServer3D::SLIDER_JOINT_MAX */
#define BPC_G6DOF_AXIS_STRIDE 28 /* per axis: PhysicsServer3D::G6DOF_JOINT_MAX (22) values, then G6DOF_JOINT_FLAG_MAX (6) flags */
#define BPC_G6DOF_PARAMS 84

/* Bulk body description; one record per body.  Same layout as the record of the neutral scene
 * file (bradot_b200/scenes.py BODY_DT).  Replaces the per-body call sequence body_create /
 * body_set_mode / body_add_shape / body_set_param / body_set_state / body_set_space
 * (P3D/bradot_physics_server_3d.cpp:438-760). */
typedef struct bpc_body_desc {
    /* flags: bit0 can_sleep, bit1 continuous collision detection, bit2 start asleep, bit4 / bit5 linear / angular damp mode
     * REPLACE (BODY_PARAM_*_DAMP_MODE, P3D/bradot_body_3d.cpp:586-606), bit6 omit force integration, bit7 NOT ray pickable
     * (body_set_ray_pickable(false); only rays cast with BPC_RAY_PICK care), bit15 the shape is disabled (body_set_shape_disabled:
     * nothing collides with the body, no area or query sees it, and its mass data are those of a body without shapes), bits 8..13 axis locks
     * in PhysicsServer3D::BodyAxis order (body_set_axis_lock, P3D/bradot_body_3d.cpp:685-699).
     * bit2 is kept as given: the body sleeps until something wakes it.  (The reference would wake it on its first step:
     * every body whose shapes were just added has a pending shape update, and BradotBody3D::_shapes_changed ends in
     * wakeup(), P3D/bradot_body_3d.cpp:343-347 -- there a body can only be put to sleep after that step.) */
    uint32_t space, mode, shape, flags;
    float xf[12];                       /* basis rows (9) then origin (3) */
    float shape_xf[12];                 /* local transform of the single shape */
    float linvel[3], angvel[3];
    float mass, friction, bounce, gravity_scale, lin_damp, ang_damp;
    uint32_t layer, mask;
} bpc_body_desc;

typedef struct bpc_joint_desc {
    /* flags: bit0 the two bodies do not collide with each other (joint_disable_collisions_between_bodies), bit1 hinge
     * HINGE_JOINT_FLAG_USE_LIMIT, bit2 hinge HINGE_JOINT_FLAG_ENABLE_MOTOR, bit3 `params` holds values (otherwise the
     * reference's defaults: pin 0.3 / 1 / 0, hinge P3D/joints/bradot_hinge_joint_3d.h:73-88, cone twist
     * P3D/joints/bradot_cone_twist_joint_3d.h:80-86) */
    uint32_t type, body_a, body_b, flags;
    float frame_a[12], frame_b[12]; /* pin: origins only (local anchors); hinge, cone twist: local frames */
    /* pin: PinJointParam order (bias, damping, impulse clamp; P3D/joints/bradot_pin_joint_3d.cpp:143-155); hinge:
     * HingeJointParam order (bias, limit upper, limit lower, limit bias, limit softness, limit relaxation, motor target
     * velocity, motor max impulse; P3D/joints/bradot_hinge_joint_3d.cpp:362-392); cone twist: ConeTwistJointParam order
     * (swing span, twist span, bias, softness, relaxation; P3D/joints/bradot_cone_twist_joint_3d.cpp:274-296), then in
     * params[5] the SECOND swing span: cone_twist_joint_set_param(SWING_SPAN) writes both spans, so pass the same value
     * twice -- except for a joint whose swing span was never set, which has tau / 8 and 0 */
    float params[8];
    uint32_t priority; /* joint_set_solver_priority (P3D/bradot_physics_server_3d.cpp:1376): 0 or 1 = default */
    uint32_t reserved[3];
} bpc_joint_desc;

/* PhysicsServer3D::SpaceParameter (servers/physics_server_3d.h:283-292) plus the default-area
 * values a space carries (P3D/bradot_area_3d.h:50-55). */
enum {
    BPC_SPACE_CONTACT_RECYCLE_RADIUS = 0, BPC_SPACE_CONTACT_MAX_SEPARATION = 1, BPC_SPACE_CONTACT_MAX_ALLOWED_PENETRATION = 2,
    BPC_SPACE_CONTACT_DEFAULT_BIAS = 3, BPC_SPACE_BODY_LINEAR_VELOCITY_SLEEP_THRESHOLD = 4,
    BPC_SPACE_BODY_ANGULAR_VELOCITY_SLEEP_THRESHOLD = 5, BPC_SPACE_BODY_TIME_TO_SLEEP = 6, BPC_SPACE_SOLVER_ITERATIONS = 7,
    BPC_SPACE_GRAVITY_X = 100, BPC_SPACE_GRAVITY_Y = 101, BPC_SPACE_GRAVITY_Z = 102, BPC_SPACE_LINEAR_DAMP = 103, BPC_SPACE_ANGULAR_DAMP = 104
};

/* Per-body state arrays for bulk read/write; `n` consecutive bodies starting at `first`. */
enum {
    BPC_FIELD_TRANSFORM = 0,        /* 12 floats: basis rows + origin          (body_get/set_state BODY_STATE_TRANSFORM) */
    BPC_FIELD_LINEAR_VELOCITY = 1,  /* 3 floats                                (BODY_STATE_LINEAR_VELOCITY) */
    BPC_FIELD_ANGULAR_VELOCITY = 2, /* 3 floats                                (BODY_STATE_ANGULAR_VELOCITY) */
    BPC_FIELD_ACTIVE = 3,           /* 1 float (0/1)                           (BODY_STATE_SLEEPING, inverted) */
    BPC_FIELD_MASS_PROPS = 4,       /* 16 floats: inv_mass, inv_inertia(3), com_local(3), principal axes rows(9) */
    BPC_FIELD_ITEM_BOX = 5,         /* 14 floats: aabb_cache pos(3) size(3), stored min(3) max(3), num_pairs, id */
    BPC_FIELD_APPLIED_FORCE = 6,    /* 6 floats: force, torque accumulated for the next step (body_apply_force/torque); a non-zero row wakes its body */
    BPC_FIELD_STILL_TIME = 7,       /* 1 float */
    /* kinematic bodies (write only; ignored for bodies of other modes in the range):
     * TARGET = the transform the body is to have after the next step: what body_set_state(BODY_STATE_TRANSFORM) means for a
     * kinematic body (new_transform, modules/bradot_physics_3d/bradot_body_3d.cpp:354-364); it wakes the body.  The step
     * derives the body's velocities from the move (bradot_body_3d.cpp:611-627).  BPC_FIELD_TRANSFORM on a kinematic body
     * places it (current transform and target).
     * VELOCITY = constant linear + angular velocity (6 floats) added to the derived ones: what BODY_STATE_LINEAR_VELOCITY /
     * ANGULAR_VELOCITY mean for a kinematic body (constant_linear_velocity, bradot_body_3d.cpp:383-393). */
    BPC_FIELD_KINEMATIC_TARGET = 8,
    BPC_FIELD_KINEMATIC_VELOCITY = 9,
    BPC_FIELD_CONSTANT_FORCE = 10,  /* 6 floats: constant force, constant torque (body_set_constant_force / _torque, P3D/bradot_body_3d.h:257-281); a non-zero row wakes its body */
    BPC_FIELD_MASS_WORLD = 11       /* read only, 13 floats: world inverse inertia tensor rows (9), world-oriented centre of mass (3), inverse mass */
};

typedef struct bpc_stats {
    int32_t active_objects;  /* INFO_ACTIVE_OBJECTS   (P3D/bradot_physics_server_3d.cpp:1712-1726) */
    int32_t collision_pairs; /* INFO_COLLISION_PAIRS */
    int32_t island_count;    /* INFO_ISLAND_COUNT */
    int32_t contact_pairs;   /* pairs whose narrowphase reported a collision this step */
    int32_t colors;          /* sweep groups used by the solver this step */
    int32_t overflow;        /* constraints that did not fit the colour budget (solved serially) */
    int32_t steps;
    int32_t kernel_launches; /* kernels launched by the last bpc_step */
    int32_t active_contacts; /* contacts with positive depth handed to the solver this step */
    int32_t full_pairs;      /* pairs that went through the full narrowphase pass (contact generation) this step */
    int32_t barrier_constraints; /* constraints swept on the solver's grid-barrier path (the rest ran in solver bins) */
} bpc_stats;

/* ---- lifecycle: BradotPhysicsServer3D::init / finish (P3D/bradot_physics_server_3d.cpp:1628-1630, 1704-1706) */
bpc_context *bpc_create(int device);
void bpc_destroy(bpc_context *ctx);
const char *bpc_last_error(bpc_context *ctx);
/* CUDA stream all work is enqueued on (cudaStream_t as void*); 0 = default stream. */
int bpc_set_stream(bpc_context *ctx, void *cuda_stream);

/* ---- spaces: space_create / space_set_param / space_set_active / free(space) (P3D/bradot_physics_server_3d.cpp:144-191,
 * 1603-1611).  Each space has its own solver iteration count (1..255).  An inactive space is not stepped: its bodies keep
 * their state and their cached pairs and are not counted by bpc_get_stats.  Destroying a space destroys its bodies and their joints; its index is
 * not reused. */
int bpc_space_create(bpc_context *ctx);
int bpc_space_set_param(bpc_context *ctx, int space, int param, float value);
int bpc_space_set_active(bpc_context *ctx, int space, int active);
int bpc_space_destroy(bpc_context *ctx, int space);

/* ---- shapes: shape_create + shape_set_data (P3D/bradot_physics_server_3d.cpp:50-142; P3D/bradot_shape_3d.cpp _setup) */
int bpc_shape_create(bpc_context *ctx, int type, const float params[4]);
/* shape_set_custom_solver_bias (P3D/bradot_physics_server_3d.cpp:110-116): a shape's own contact bias; a pair uses it in place
 * of 0.8 -- the one that is set, or the mean of both (BradotBodyPair3D::setup, P3D/bradot_body_pair_3d.cpp:344-357).  0 = none. */
int bpc_shape_set_custom_bias(bpc_context *ctx, int shape, float bias);
/* Convex hull in the layout of Geometry3D::MeshData (core/math/geometry_3d.h): the engine-side
 * module runs ConvexHullComputer (host, set-up time only, P3D/bradot_shape_3d.cpp:1099) and hands
 * the result over.  faces: per face (first index, index count) into `indices`; edges: per edge
 * (vertex_a, vertex_b, face_a, face_b); planes: per face (nx, ny, nz, d). */
int bpc_shape_create_convex(bpc_context *ctx, const float *verts, int n_verts, const float *planes, const int32_t *faces, int n_faces,
                            const int32_t *indices, int n_indices, const int32_t *edges, int n_edges);
/* ConcavePolygonShape3D (static triangle meshes; P3D/bradot_shape_3d.cpp:1289-1690): faces = n_faces x 9 floats
 * (three vertices each).  Only STATIC bodies may carry one.  The tree over the faces is built here as the reference builds its
 * own (median split along the longest axis, P3D/bradot_shape_3d.cpp:1541-1586) unless one is handed over: bvh_nodes = n_nodes x 9
 * floats (box position, box size, left, right, face index or -1) in pre-order, normals = n_faces x 3 -- what the tests do with
 * the reference's own tree, so that faces are met in its order.  Against a mesh: the step (all four body shapes), intersect_shape,
 * collide_shape, rest_info, body_test_motion's recovery and rest steps, intersect_ray (nearest face; a point is never inside a
 * mesh), cast_motion and the cast of body_test_motion (GJK distance face by face). */
int bpc_shape_create_concave(bpc_context *ctx, const float *faces, int n_faces, int backface_collision, const float *bvh_nodes, int n_nodes,
                             const float *normals);
/* HeightMapShape3D (P3D/bradot_shape_3d.cpp:2028-2078, 2159-2178): width x depth heights row by row (z major), one unit apart,
 * centred on the shape's origin, solid on both sides; static bodies only.  Collides in the step and answers intersect_shape,
 * collide_shape, rest_info, cast_motion and body_test_motion like a triangle mesh; intersect_ray and the casts of CCD walk
 * the cells under the segment in crossing order (BradotHeightMapShape3D::intersect_segment, :1950-2008). */
int bpc_shape_create_heightmap(bpc_context *ctx, const float *heights, int width, int depth, float min_height, float max_height);

/* ---- bodies */
int bpc_bodies_create(bpc_context *ctx, const bpc_body_desc *descs, int n); /* returns first body index */
int bpc_bodies_write(bpc_context *ctx, int field, int first, int n, const float *host);
int bpc_bodies_read(bpc_context *ctx, int field, int first, int n, float *host);
int bpc_body_count(bpc_context *ctx);
/* Like bpc_bodies_create, but slots of destroyed bodies are handed out first (most recently freed first, as the reference's
 * PooledList does for broadphase ids, core/templates/pooled_list.h:100-121); slots_out receives one body index per record. */
int bpc_bodies_create_slots(bpc_context *ctx, const bpc_body_desc *descs, int n, int32_t *slots_out);
/* free(RID) of a body (P3D/bradot_physics_server_3d.cpp:1572-1590): the body leaves the broadphase, its pairs are gone with
 * the next step, joints attached to it are destroyed with it.  As in the reference, nothing is woken (the body is unpaired
 * before its shapes go, so wakeup_neighbours finds no neighbours): a sleeping body resting on a freed one sleeps on.
 * Slots stay valid indices (reads return the tombstone) until reused. */
int bpc_bodies_destroy(bpc_context *ctx, const int32_t *slots, int n);
/* Re-describe existing bodies: everything body_set_mode / body_set_param / body_set_collision_layer|mask / body_set_shape /
 * body_set_axis_lock / body_set_enable_continuous_collision_detection change after a body is in a space
 * (P3D/bradot_physics_server_3d.cpp:479-760).  keep_state != 0: the records' transform and velocities are ignored, the
 * body keeps the ones it has (and its sleep state, unless the new mode forbids it). */
int bpc_bodies_update(bpc_context *ctx, const int32_t *slots, int n, const bpc_body_desc *descs, int keep_state);

/* body_set_state with the reference's side effects (BradotBody3D::set_state, P3D/bradot_body_3d.cpp:352-417), for a list
 * of bodies.  The BPC_FIELD_* writes above restore raw state (snapshots, tests) and have none of these side effects.
 *   BPC_STATE_TRANSFORM        12 floats.  Rigid: orthonormalised, mass data and broadphase box follow (the box through the
 *                              usual hysteresis, as a move), the body wakes.  Static: placed, bodies it touches are woken.
 *                              Kinematic: becomes the target of the next step (= BPC_FIELD_KINEMATIC_TARGET).
 *   BPC_STATE_LINEAR_VELOCITY  3 floats; wakes the body (kinematic: its constant velocity)
 *   BPC_STATE_ANGULAR_VELOCITY 3 floats; wakes the body
 *   BPC_STATE_SLEEPING         1 float (0/1); rigid bodies only; going to sleep zeroes the velocities
 *   BPC_STATE_CAN_SLEEP        1 float (0/1); clearing it wakes a sleeping rigid body
 * Numeric values are PhysicsServer3D::BodyState (servers/physics_server_3d.h:472-478). */
enum { BPC_STATE_TRANSFORM = 0, BPC_STATE_LINEAR_VELOCITY = 1, BPC_STATE_ANGULAR_VELOCITY = 2, BPC_STATE_SLEEPING = 3, BPC_STATE_CAN_SLEEP = 4 };
int bpc_bodies_set_state(bpc_context *ctx, int state, const int32_t *slots, int n, const float *values);
/* Gather / scatter variants of the bulk I/O for a list of bodies (state sync of the active bodies only, forces of the pushed
 * bodies only): one staging copy and one kernel whatever n is. */
/* body_set_collision_priority (P3D/bradot_physics_server_3d.cpp:611-616): how much of the push an obstacle gets when
 * body_test_motion frees a stuck body (the only place the reference consults it, P3D/bradot_space_3d.cpp:748-762).  Default 1. */
int bpc_bodies_set_collision_priority(bpc_context *ctx, const int32_t *slots, int n, const float *priorities);
/* Two body records that are shapes of ONE object (the engine module's records for the further shapes of a static or kinematic
 * body): they never enter the pair list -- in the reference the shapes of one body are not paired with each other
 * (the pair test of its broadphase, `p_a == p_b`), whereas two separate kinematic bodies are.  same = 0 takes the mark back. */
int bpc_bodies_set_same_object(bpc_context *ctx, int body_a, int body_b, int same);
/* body_set_param(BODY_PARAM_INERTIA / BODY_PARAM_CENTER_OF_MASS) and body_reset_mass_properties (P3D/bradot_body_3d.cpp:209-229,
 * update_mass_properties :56-150): a body's own principal inertia and / or local centre of mass in place of the values
 * computed from its shape.  inertia: n x 3 or NULL (left as it is); all three components > 0 = the body's inertia about
 * its own axes, otherwise the positive ones replace the diagonal of the computed tensor before it is diagonalised, as in
 * the reference.  center_of_mass: n x 3 or NULL (left as it is).  Both NULL = body_reset_mass_properties: the centre of mass
 * is computed again, and so is the inertia -- except that, as in the reference, the positive components given before go
 * on replacing the diagonal of the computed tensor.  (The reference recomputes the inertia about a new centre of mass only
 * at its next mass update -- set mass, shape change --; here it is recomputed when the body is described.)  The override
 * is kept with the slot and takes effect when the body is described next (bpc_bodies_create_slots / bpc_bodies_update). */
int bpc_bodies_set_mass_override(bpc_context *ctx, const int32_t *slots, int n, const float *inertia, const float *center_of_mass);
/* The same read, pipelined: _begin queues the read of rows [first, first + n) as they stand after the steps issued so far and
 * returns; the copy to `out` (pinned host memory, or it will not overlap) runs beside whatever is issued next -- the next
 * bpc_step included; _end waits until `out` holds the rows.  One read at a time.  TRANSFORM and the two velocities only.
 * (What flush_queries would use to hand last tick's transforms to the scene while this tick's step runs.) */
int bpc_bodies_read_begin(bpc_context *ctx, int field, int first, int n, float *out);
int bpc_bodies_read_end(bpc_context *ctx);
int bpc_bodies_read_slots(bpc_context *ctx, int field, const int32_t *slots, int n, float *host);
int bpc_bodies_write_slots(bpc_context *ctx, int field, const int32_t *slots, int n, const float *host);
/* Transform (12), linear velocity (3), angular velocity (3), active (1) of every body in one call: 19 floats per body. */
int bpc_bodies_read_state(bpc_context *ctx, int first, int n, float *host);

/* body_add_collision_exception / body_remove_collision_exception (P3D/bradot_physics_server_3d.cpp:690-720): the pair
 * (a, b) is skipped by the narrowphase as long as the exception stands (joints that disable collisions between their bodies
 * add theirs on their own). */
int bpc_bodies_set_exception(bpc_context *ctx, int body_a, int body_b, int excepted);

/* Contact reporting: body_set_max_contacts_reported + the PhysicsDirectBodyState3D contact getters
 * (P3D/bradot_body_3d.h:353-395, P3D/bradot_body_pair_3d.cpp:417-428).  A body that asks for up to N contacts gets, each
 * step, the N deepest contacts it takes part in (the reference keeps the N deepest too; their order in the list is not
 * defined here).  Positions and velocities are world-space as the reference reports them; `impulse` is the contact's
 * accumulated impulse at pre-solve time; `collider` is the other body's index.  A pair none of whose bodies would otherwise
 * collide (static against kinematic) is still tested when one of them reports contacts, as in the reference. */
typedef struct bpc_contact {
    float local_pos[3], depth;
    float local_normal[3];
    int32_t collider;
    float local_velocity[3], collider_pos[3], collider_velocity[3], impulse[3];
} bpc_contact;
int bpc_bodies_set_max_contacts_reported(bpc_context *ctx, const int32_t *slots, int n, const int32_t *max_contacts);
/* Contacts of one body from the last step: at most max_records are written, the number available is returned. */
int bpc_body_read_contacts(bpc_context *ctx, int slot, int max_records, bpc_contact *out);
/* All reported contacts of the last step in one copy: counts[n_bodies] (contacts of each body), offsets[n_bodies] (where its
 * records start in `records`), records[bpc_contacts_capacity()]. */
int bpc_contacts_capacity(bpc_context *ctx);
int bpc_contacts_read(bpc_context *ctx, int32_t *counts, int32_t *offsets, bpc_contact *records);

/* ---- areas: PhysicsServer3D::area_* (servers/physics_server_3d.h:296-358; P3D/bradot_area_3d.{h,cpp},
 * P3D/bradot_area_pair_3d.cpp:36-118, the area loop of integrate_forces P3D/bradot_body_3d.cpp:498-585).  An area is a shape
 * in a space that bodies overlap: one with a space override changes the gravity / damping of the bodies standing in it
 * (by priority, with the reference's combine / replace rules), one with a monitor callback is told who enters and leaves.
 * flags: bit0 monitorable, bit1 a body monitor callback is set (events are raised), bit2 the gravity vector is a point
 * (in the area's space) bodies are pulled towards, bit3 an area monitor callback is set (it is told which monitorable
 * areas it overlaps: BradotArea2Pair3D, P3D/bradot_area_pair_3d.cpp:124-206; looked at in steps where one of the two moved),
 * bit4 ray pickable (area_set_ray_pickable(true); without it rays cast with BPC_RAY_PICK pass the area by, which is how the
 * reference's areas start out, P3D/bradot_area_3d.cpp:343).
 * Override modes are
 * PhysicsServer3D::AreaSpaceOverrideMode: 0 disabled, 1 combine, 2 combine-replace, 3 replace, 4 replace-combine.
 * Not built: wind (soft bodies only). */
typedef struct bpc_area_desc {
    int32_t space, shape;
    uint32_t flags;
    int32_t priority;
    float xf[12], shape_xf[12];
    uint32_t layer, mask;
    int32_t gravity_mode, linear_damp_mode, angular_damp_mode;
    float gravity, gravity_vector[3], point_unit_distance, linear_damp, angular_damp;
    uint32_t group; /* 0, or: the areas of a space with the same group are the shapes of ONE area -- a body standing in several of
                     * them has that area once in its override walk (BradotBody3D::add_area counts references, P3D/bradot_body_3d.h) */
} bpc_area_desc;
typedef struct bpc_area_event {
    int32_t area, body, status; /* status 0 / 1: body `body` entered / left the area; 2 / 3: AREA `body` entered / left it */
} bpc_area_event;
int bpc_areas_create(bpc_context *ctx, const bpc_area_desc *descs, int n);              /* -> index of the first */
int bpc_areas_write(bpc_context *ctx, int first, int n, const bpc_area_desc *descs);    /* any area_set_*: also "the area moved" */
int bpc_areas_destroy(bpc_context *ctx, const int32_t *areas, int n);
int bpc_area_count(bpc_context *ctx);
/* Enter / exit events raised since the last call (BradotArea3D::call_queries hands them to the monitor callback); returns how
 * many there were, writes at most max_events (the rest stay for the next call). */
int bpc_area_events_read(bpc_context *ctx, bpc_area_event *out, int max_events);
/* The areas a body stands in right now (bit 30 set: attached as a space override).  At most 8 are tracked per body. */
int bpc_body_read_areas(bpc_context *ctx, int slot, int32_t *out, int max_areas);

/* Size every per-body and per-pair buffer for `bodies` bodies and `pairs` broadphase pairs up front, so that bpc_step does
 * not allocate while the counts stay below them (a pair burst beyond `pairs` still grows the buffers, once). */
int bpc_reserve(bpc_context *ctx, int bodies, int pairs);

/* ---- joints: joint_make_pin / joint_make_hinge / joint_disable_collisions_between_bodies
 * (P3D/bradot_physics_server_3d.cpp:1224-1246, 1388-1406) */
int bpc_joints_create(bpc_context *ctx, const bpc_joint_desc *descs, int n);
/* pin_joint_set_param / hinge_joint_set_param / hinge_joint_set_flag / joint_set_solver_priority /
 * joint_disable_collisions_between_bodies at run time (P3D/bradot_physics_server_3d.cpp:1248-1406): rewrite n joint records
 * starting at `first`; type and bodies must stay what they are. */
int bpc_joints_write(bpc_context *ctx, int first, int n, const bpc_joint_desc *descs);
/* joint_clear / free(joint) (P3D/bradot_physics_server_3d.cpp:1229-1246, 1592-1601): the joint stops acting, the collision
 * exception it made is dropped; indices of other joints do not move. */
int bpc_joints_destroy(bpc_context *ctx, int first, int n);
/* The long parameter table of a joint whose parameters do not fit `bpc_joint_desc.params`: a slider joint's 22 values in
 * PhysicsServer3D::SliderJointParam order (slider_joint_set_param / slider_joint_get_param,
 * P3D/joints/bradot_slider_joint_3d.cpp:349-479; defaults P3D/joints/bradot_slider_joint_3d.h:80-110, which a new slider
 * joint starts with -- its `params` and the `params holds values` flag are ignored).  Write / read `n` values starting at
 * parameter `first`.  Slider frames: the joint slides along the x axis of frame A.
 * A generic 6-DOF joint (P3D/joints/bradot_generic_6dof_joint_3d.cpp) has 84 values: for each axis x, y, z a block of
 * BPC_G6DOF_AXIS_STRIDE, the G6DOFJointAxisParam values first, then the G6DOFJointAxisFlag flags as 0 / 1.  The table holds the
 * MEMBERS of the reference's joint (defaults bradot_generic_6dof_joint_3d.h:67-112: linear limits [0, 0] enabled, angular
 * limits disabled): the entries the reference ignores (linear motors, springs) are ignored here, and
 * ANGULAR_MOTOR_FORCE_LIMIT is the motor's force limit -- which the reference's own set_param never reaches (its case
 * writes the LIMIT force, :497-499; the engine module mirrors that). */
int bpc_joint_params_write(bpc_context *ctx, int joint, int first, int n, const float *values);
int bpc_joint_params_read(bpc_context *ctx, int joint, int first, int n, float *values);
int bpc_joint_count(bpc_context *ctx);

/* ---- the step: BradotPhysicsServer3D::step -> BradotStep3D::step for every space at once
 * (P3D/bradot_physics_server_3d.cpp:1632-1648, P3D/bradot_step_3d.cpp:186-410) */
int bpc_step(bpc_context *ctx, float dt);
int bpc_synchronize(bpc_context *ctx);
int bpc_get_stats(bpc_context *ctx, bpc_stats *out);
/* Per-stage device time, the counterpart of the reference's five ELAPSED_TIME_* phases
 * (P3D/bradot_space_3d.h:66-74, P3D/bradot_step_3d.cpp:202-404): CUDA events on the step's stream.
 * Stages: 0 integrate forces, 1 broadphase, 2 islands, 3 narrowphase (constraint setup),
 * 4 sweep scheduling, 5 pre-solve + solve, 6 integrate velocities + sleep.  ms_out[7] are totals
 * since profiling was switched on, over *steps_out steps. */
int bpc_set_profiling(bpc_context *ctx, int on);
int bpc_get_stage_times(bpc_context *ctx, double *ms_out, int *steps_out);

/* ---- state snapshot of the pair/manifold cache (debug dumps, differential tests, checkpoint/restore).
 * Pair record: body_a, body_b, contact_count, flags(bit0 collided, bit1 collide_A, bit2 collide_B).
 * Pair floats: sep_axis(3).  Contact floats (4 per pair, 16 each):
 *   local_A(3) local_B(3) normal(3) acc_normal acc_tangent(3) acc_bias acc_bias_com used          */
int bpc_pair_count(bpc_context *ctx);
int bpc_pairs_read(bpc_context *ctx, int32_t *pair_ids, float *pair_f, float *contacts);
int bpc_pairs_write(bpc_context *ctx, int n_pairs, const int32_t *pair_ids, const float *pair_f, const float *contacts);
/* Teacher-forced sweep order for differential tests: constraints are solved in exactly this
 * sequence (sequential Gauss-Seidel semantics, P3D/bradot_step_3d.cpp:134-160) instead of the
 * colouring.  order: n records (kind, a, b): kind 0 = body pair (a, b), kind 1 = joint index a.
 * n = 0 returns to the default scheduling. */
int bpc_set_solve_order(bpc_context *ctx, int n, const int32_t *order);

/* ---- space queries
 * PhysicsDirectSpaceState3D::intersect_ray (servers/physics_server_3d.h:164-186; implementation
 * modules/bradot_physics_3d/bradot_space_3d.cpp:109-207), batched: one call answers n rays against the current state of
 * one space -- closest hit along each segment, bodies whose collision layer meets collision_mask, a shape the ray starts
 * inside is ignored unless BPC_RAY_HIT_FROM_INSIDE.  from_to: n x (from xyz, to xyz), host; out: n records, host.
 * `body` is the body slot (what the engine module maps back to a RID / ObjectID), -1 for a miss.
 * Exclude lists: bpc_space_set_query_exclude below. */
typedef struct bpc_ray_result {
    int32_t hit;
    int32_t body;
    float position[3];
    float normal[3];
    int32_t face_index; /* RayResult::face_index: the face of a triangle mesh that was hit, else -1 */
} bpc_ray_result;
#define BPC_RAY_HIT_FROM_INSIDE 1
#define BPC_QUERY_AREAS 8      /* collide_with_areas (rays and points): areas answer as well; an area's `body` is -2 - its index */
#define BPC_QUERY_NO_BODIES 16 /* collide_with_bodies = false */
#define BPC_RAY_PICK 4 /* RayParameters::pick_ray: bodies that are not ray pickable are passed through (P3D/bradot_space_3d.cpp:134) */
#define BPC_RAY_HIT_BACK_FACES 2 /* RayParameters::hit_back_faces (default on in the engine): a two-sided triangle mesh is also hit from behind */
int bpc_space_intersect_rays(bpc_context *ctx, int space, int n, const float *from_to, unsigned collision_mask, int flags,
                             bpc_ray_result *out);

/* PhysicsDirectSpaceState3D::intersect_point (servers/physics_server_3d.h:188-203; implementation
 * modules/bradot_physics_3d/bradot_space_3d.cpp:62-107), batched: for each of n points (xyz, host) the bodies whose shape
 * holds it.  out_counts: n; out_bodies: n x max_results body slots, -1 padded.  When more than max_results bodies hold a
 * point, which ones are listed follows this backend's tree order (the reference's follows its own BVH).
 * flags: BPC_QUERY_AREAS / BPC_QUERY_NO_BODIES (PointParameters::collide_with_areas / collide_with_bodies); areas are listed
 * after the bodies, as -2 - area index. */
int bpc_space_intersect_points(bpc_context *ctx, int space, int n, const float *points, unsigned collision_mask, int flags,
                               int max_results, int32_t *out_counts, int32_t *out_bodies);

/* PhysicsDirectSpaceState3D::get_closest_point_to_object_volume (servers/physics_server_3d.h:207; implementation
 * modules/bradot_physics_3d/bradot_space_3d.cpp:573-611 with the shapes' get_closest_point_to), batched: for each (body record,
 * world point) the point of the record's shape nearest to it -- the point itself when it is inside.  points / out: n x 3.
 * (An object of several shapes is several records: the caller keeps the nearest answer, as the reference's loop does.) */
int bpc_bodies_closest_points(bpc_context *ctx, int n, const int32_t *bodies, const float *points, float *out);

/* PhysicsDirectSpaceState3D::intersect_shape (servers/physics_server_3d.h:183-205; implementation
 * modules/bradot_physics_3d/bradot_space_3d.cpp:208-262), batched: query i is shape handle shapes[i] (from bpc_shape_create*)
 * placed at xforms[i] (12 floats, basis rows + origin), grown by `margin` (0 = the plain routines, else the reference's withMargin
 * variants); the answer lists the bodies whose shape it overlaps.
 * Output layout and flags (BPC_QUERY_AREAS / BPC_QUERY_NO_BODIES) as bpc_space_intersect_points. */
int bpc_space_intersect_shapes(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, unsigned collision_mask,
                               int flags, float margin, int max_results, int32_t *out_counts, int32_t *out_bodies);

/* PhysicsDirectSpaceState3D::collide_shape (servers/physics_server_3d.h; implementation
 * modules/bradot_physics_3d/bradot_space_3d.cpp:383-431), batched: for each query shape (grown by `margin`) the contact pairs with
 * the bodies it overlaps.  out_points: n x max_results x 6 (point on the query shape, point on the body), max_results <= 32;
 * once full, a deeper pair replaces the least deep one (the reference's collecting callback). */
int bpc_space_collide_shapes(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, unsigned collision_mask,
                             float margin, int max_results, int32_t *out_counts, float *out_points);

/* The bodies every space query of this context skips from now on (RayParameters / PointParameters / ShapeParameters ::exclude,
 * servers/physics_server_3d.h:139-201; P3D/bradot_space_3d.cpp:86,151,233,294,408,540).  n = 0 clears the list. */
int bpc_space_set_query_exclude(bpc_context *ctx, const int32_t *bodies, int n);

/* PhysicsDirectSpaceState3D::rest_info (P3D/bradot_space_3d.cpp:513-569), batched: the deepest contact of query shape i, grown
 * by max(margin, 0.0001), with the bodies it touches -- through the margin variants of the collision routines
 * (P3D/bradot_collision_solver_3d_sat.cpp, withMargin) and the reference's rest callback (:453-511).  motions (n x 3, may be
 * NULL) only sets the least depth a contact needs to count: min(|motion|, 0.05 margin).  `point` is on the body, `normal`
 * points from the query shape to it, `linear_velocity` is the body's at that point. */
typedef struct bpc_rest_info {
    int32_t hit;
    int32_t body;
    float point[3];
    float normal[3];
    float linear_velocity[3];
} bpc_rest_info;
int bpc_space_rest_infos(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, const float *motions, float margin,
                         unsigned collision_mask, bpc_rest_info *out);

/* PhysicsDirectSpaceState3D::cast_motion (P3D/bradot_space_3d.cpp:264-381), batched: query shape i at xforms[i] travels along
 * motions[i] (n x 3, world); out_fractions[i] = (closest safe, closest unsafe) fraction of the motion, (1, 1) when nothing is
 * in the way; bodies the shape starts inside of are ignored.  The distance test is the reference's GJK (P3D/gjk_epa.cpp:216-542,
 * 888-927) on the swept shape, eight bisection steps per candidate.  out_info (may be NULL) is the ShapeRestInfo the reference
 * fills alongside: the body that stops the motion first, the closest point on it, the normal towards the query shape. */
int bpc_space_cast_motions(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, const float *motions, float margin,
                           unsigned collision_mask, float *out_fractions, bpc_rest_info *out_info);

/* PhysicsServer3D::body_test_motion (servers/physics_server_3d.h:525-571; P3D/bradot_space_3d.cpp:651-1029), batched: what
 * CharacterBody3D's move_and_slide / move_and_collide ask.  Body `body`, placed at `from` instead of where the step left it,
 * is (1) pushed out of what it overlaps (up to 4 rounds, 40 % of the excess depth each), (2) cast along `motion` (GJK distance
 * on the swept shape, 8 bisection steps per obstacle) and (3) asked for up to max_collisions contacts where it stops (the
 * margin variants of the collision routines, the reference's rest callback).  Obstacles: bodies of its space whose layer
 * meets its mask, minus its collision exceptions and the query exclude list (bpc_space_set_query_exclude).
 * flags: bit1 recovery_as_collision.  The order in which overlapping obstacles are visited is this backend's tree order; the
 * recovery of step 1 is a sequential projection, so with several simultaneous contacts its result can differ from the
 * reference's in the last digits of a millimetre. */
#define BPC_MAX_MOTION_COLLISIONS 32
typedef struct bpc_motion_params {
    int32_t body;
    float from[12];
    float motion[3];
    float margin;
    int32_t max_collisions;
    uint32_t flags;
} bpc_motion_params;
typedef struct bpc_motion_collision {
    float position[3], normal[3], collider_velocity[3], collider_angular_velocity[3];
    float depth;
    int32_t collider;
} bpc_motion_collision;
typedef struct bpc_motion_result {
    int32_t collided, collision_count;
    float travel[3], remainder[3];
    float collision_depth, collision_safe_fraction, collision_unsafe_fraction;
    bpc_motion_collision collisions[BPC_MAX_MOTION_COLLISIONS];
} bpc_motion_result;
int bpc_bodies_test_motion(bpc_context *ctx, int n, const bpc_motion_params *params, bpc_motion_result *out);

/* ---- utilities exposed for tests and tools */
/* Sort n 64-bit keys on the device (the radix sort the broadphase uses); host in, host out. */
int bpc_sort_u64(bpc_context *ctx, const uint64_t *keys_in, uint64_t *keys_out, int n);

/* The device forms of Geometry3D::get_closest_points_between_segments (op 0: 12 floats p0 p1 q0 q1 -> 6 floats ps qt) and
 * Geometry3D::get_closest_point_to_segment (op 1: 9 floats point s0 s1 -> 3 floats), which the capsule routines use
 * (core/math/geometry_3d.h:86-129, 331-348); the tests run the reference's own vectors through them
 * (tests/core/math/test_geometry_3d.h:39-48, 133). */
int bpc_test_segment_math(bpc_context *ctx, int op, int n, const float *in, float *out);

/* Allocation-free variant of the key sort.  It has no counterpart in the reference; it exists for
 * hosts that hand over device buffers of their own and may not cause a cudaMalloc / cudaFree while
 * they run (a plugin confined to a fixed device arena, a frame or physics loop with a "no driver
 * allocation after start-up" rule, anything that preallocates and then must stay free of the
 * device-wide stalls an allocation can bring).  It needs no context and allocates nothing, on the
 * device or through any library: its scratch is a fixed array inside the loaded module, about
 * 0.56 MB that every GPU using the library holds until the process ends.
 *   d_keys_in / d_keys_out : DEVICE pointers to n keys each (caller-owned, may not alias)
 *   n                      : at most BPC_SORT_NOALLOC_MAX_KEYS; larger sizes are rejected with an error
 *   cuda_stream            : cudaStream_t the kernels are enqueued on; the result is ready once that
 *                            stream has reached this point (the call itself does not wait)
 * The first call on each device also looks up where the scratch lives on that device; that lookup
 * is not a stream operation, so make the first call before starting any stream capture.
 * All calls share the one scratch array: calls that may overlap on the GPU (other streams, other
 * threads) race on it and must be serialised by the caller.  The array keeps what the previous call
 * left in it; no part of it has to be cleared, because each pass writes every cell it later reads.
 * Returns 0, or -1 with the reason available from bpc_noalloc_last_error() (per host thread). */
#define BPC_SORT_NOALLOC_MAX_KEYS 65536
int bpc_sort_u64_noalloc(const uint64_t *d_keys_in, uint64_t *d_keys_out, int n, void *cuda_stream);
const char *bpc_noalloc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
