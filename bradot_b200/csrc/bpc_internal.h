// Internal layout of the CUDA rigid-body backend: struct-of-arrays state resident in HBM.
// Public entry points are in include/bradot_physics_cuda.h; nothing here crosses the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "bpc_math.cuh"

// ---- shape / body enums: numeric values of the reference enums -------------------------------
// servers/physics_server_3d.h:247-259 (ShapeType) and :387-392 (BodyMode)
enum { SH_PLANE = 0, SH_RAY = 1, SH_SPHERE = 2, SH_BOX = 3, SH_CAPSULE = 4, SH_CYLINDER = 5, SH_CONVEX = 6, SH_CONCAVE = 7, SH_HEIGHTMAP = 8 };
enum { MODE_STATIC = 0, MODE_KINEMATIC = 1, MODE_RIGID = 2, MODE_RIGID_LINEAR = 3 };

// body flag bits (meta.x = mode | flags << 8)
#define BF_CAN_SLEEP 1u
#define BF_CCD 2u
#define BF_ACTIVE 4u
#define BF_OMIT_FORCE 8u
#define BF_FROZEN 16u      // the body's space is inactive (space_set_active(false)) or the body was destroyed: no stage touches it
#define BF_LIN_REPLACE 32u // BODY_DAMP_MODE_REPLACE for the linear / angular damping (bradot_body_3d.cpp:586-606)
#define BF_ANG_REPLACE 64u
#define BF_NOT_PICKABLE 128u // body_set_ray_pickable(false)
#define BF_LOCK_SHIFT 8 // 6 axis-lock bits, same order as PhysicsServer3D::BodyAxis
#define BF_SHAPE_IDENT (1u << 14)  // the shape's local transform is the identity (no product needed)
#define BF_ISLAND_AWAKE (1u << 15) // refreshed every step: the body's island holds an awake body
// meta.x bits 24..31: shape type of the body's shape (copy of ShapeRec.meta.x, saves a dependent gather)

// 64-byte records: one gather = two 32-byte sectors, all of it used.
struct __align__(16) BodyXf {
    float4 r0, r1, r2; // basis rows (w unused)
    float4 o;          // origin
};
struct __align__(16) BodyVel {
    float4 lin, ang;   // linear / angular velocity
    float4 blin, bang; // bias ("split impulse") velocities
};
struct __align__(16) BodyMassW {
    float4 i0, i1, i2; // world inverse inertia tensor rows; w = world-oriented centre of mass x,y,z
    float4 m;          // inv_mass, friction, bounce, (mode|flags as bits)
};
struct __align__(16) BodyProps {
    float4 com_mass;   // com_local.xyz, mass
    float4 iinv_gs;    // principal inverse inertia xyz, gravity_scale
    float4 a0, a1, a2; // principal axes (local) rows; w = linear_damp, angular_damp, still_time
    float4 prev_lin, prev_ang; // velocities saved before force integration (restitution)
    float4 pad;
};
struct __align__(16) BodyForces {
    float4 force, torque, cforce, ctorque; // applied (cleared each step) and constant
};
struct __align__(16) ItemBox {
    float4 cache_pos, cache_size; // exact shape AABB of the last update (aabb_cache)
    float4 smin, smax;            // stored (expanded, hysteretic) broadphase box
};
struct __align__(16) ShapeRec {
    float4 p;         // plane: n.xyz,d | sphere: r | box: half extents | capsule: radius,height
    float4 aabb_pos;  // local AABB
    float4 aabb_size; // .w = custom bias
    uint4 meta;       // type, mesh index (convex), 0, 0
};
struct MeshRec { // convex hull, offsets into the shared pools
    uint32_t vert_ofs, vert_cnt, face_ofs, face_cnt, edge_ofs, edge_cnt, index_ofs, scan_all;
};

// Persistent manifold: 4 contacts per pair.
struct __align__(16) ContactGeo {
    float4 la; // local_A.xyz, w: bit0 used
    float4 lb; // local_B.xyz
    float4 n;  // normal.xyz
};
struct __align__(16) ContactAcc {
    float4 t_n; // acc_tangent.xyz, acc_normal
    float4 b;   // acc_bias, acc_bias_com, 0, 0
};
struct __align__(16) PairHdr {
    // cached separating axis xyz ; w = bits: contact_count | collided<<8 | collide_A<<9 | collide_B<<10 | check_ccd<<11 |
    // (sweep colour + 1)<<16 (0 = none; kept across steps while the pair stays a constraint)
    float4 sep;
};
#define PAIR_COLOUR_SHIFT 16
#define PAIR_COLOUR_MASK 0x7f0000u
// Solver-facing contact written by pre_solve in sweep order.
struct __align__(16) SolveContact {
    float4 n_mn;   // normal, mass_normal
    float4 ra_b;   // rA, bias
    float4 rb_bo;  // rB, bounce
    float4 t_n;    // acc_tangent, acc_normal
    float4 b;      // acc_bias, acc_bias_com, friction, active(1/0)
};

// A contact reported to a body that asked for contacts (body_set_max_contacts_reported): what BradotBody3D::add_contact keeps
// (bradot_body_3d.h:353-395).  Same layout as bpc_contact.
struct __align__(16) ReportRec {
    float local_pos[3], depth;
    float local_normal[3];
    int32_t collider; // body index of the other body
    float local_velocity[3], collider_pos[3], collider_velocity[3], impulse[3];
};

template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t cap = 0;
    void reserve(size_t n, bool keep = false, cudaStream_t st = 0);
    void free_();
};

struct SpaceParams {
    float gravity[3] = { 0.f, -9.80665f, 0.f }; // BradotArea3D defaults, bradot_area_3d.h:50-55
    float lin_damp = 0.1f, ang_damp = 0.1f;
    float sleep_lin = 0.1f, sleep_ang = 0.13962634f, time_to_sleep = 0.5f; // servers/physics_server_3d.cpp:1135-1137
    float recycle_radius = 0.01f, max_separation = 0.05f, max_penetration = 0.01f; // :1139-1141
    int iterations = 16;
    int active = 1; // space_set_active
    int alive = 1;  // 0 after bpc_space_destroy
};

// An area (BradotArea3D, bradot_area_3d.h:46-70): a shape that bodies overlap, not collide with.  `xf` is the area's own
// transform (point gravity is given in its space), `sxf` the world transform of its shape.
#define AREA_SLOTS 8u
#define AREA_ATTACHED 0x80000000u
#define AF_MONITORABLE 1u
#define AF_MONITOR_CB 2u
#define AF_GRAVITY_POINT 4u
#define AF_AREA_MONITOR_CB 8u
#define AF_PICKABLE 16u // area_set_ray_pickable(true); an area is not pickable until told so (bradot_area_3d.cpp:343)
#define AF_ALIVE 0x100u
#define AF_MOVED 0x200u
#define AREA_EVENT_CAP (1u << 20)
struct AreaRec {
    float4 x0, x1, x2, xo;   // area transform: basis rows, origin
    float4 s0, s1, s2, so;   // area transform * shape transform
    float4 shape_p;          // shape parameters (hull: mesh index in x)
    uint32_t shape_type, space, layer, mask;
    uint32_t flags;
    int32_t priority;
    uint32_t gravity_mode, lin_damp_mode;
    uint32_t ang_damp_mode;
    float gravity, point_unit_distance, lin_damp;
    uint32_t group, pad_[3]; // non-zero: the areas of a space that share it are the shapes of ONE area (counted once by a body)
    float4 gravity_vector;   // w = angular damp
};

// Solver bins: islands whose root index shares a block of BIN_BLOCK body indices; one small CTA sweeps a bin whose dynamic bodies
// span fewer than BIN_MAX_BODIES indices (their velocity records sit in shared memory) and that has at most BIN_MAX_CONS
// constraints.
#ifndef BIN_BLOCK
#define BIN_BLOCK 256u
#define BIN_MAX_BODIES 704u
#define BIN_MAX_CONS 2048u
#endif

// The exclude list of the space queries (ShapeParameters::exclude and friends, bradot_space_3d.cpp:86,151,233,294,408,540):
// sorted body indices, looked up per candidate.
struct QExclude {
    const uint32_t *ids;
    uint32_t n;
};
#ifdef __CUDACC__
__device__ __forceinline__ bool q_excluded(const QExclude &e, uint32_t body) {
    uint32_t lo = 0, hi = e.n;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (e.ids[mid] < body) lo = mid + 1;
        else hi = mid;
    }
    return lo < e.n && e.ids[lo] == body;
}
#endif

// stages timed by bpc_set_profiling: forces, broadphase, islands, narrowphase, schedule, solve, integrate+sleep
#define BPC_N_STAGES 7

// Same layout as bpc_joint_desc.  flags: bit0 bodies do not collide, bit1 hinge use_limit, bit2 hinge enable_motor,
// bit3 prm[] given (else defaults), bit31 destroyed.  prm: pin = bias, damping, impulse clamp; hinge = HingeJointParam order.
struct JointRec {
    uint32_t type, a, b, flags;
    float fa[12], fb[12];
    float prm[8];
    uint32_t priority, pad[3];
};
#define JF_USE_LIMIT 2u
#define JF_ENABLE_MOTOR 4u
#define JF_HAS_PARAMS 8u
#define JF_DEAD 0x80000000u
#define JEXT_STRIDE 84 // long parameter table of a joint (slider: 22 values, 6-DOF: 3 x 28)

// Per-step joint data written by k_joint_setup (the Jacobian diagonals of BradotJacobianEntry3D and
// everything about the pivots / axes that stays constant while the solver sweeps).
struct __align__(16) JointSetup {
    float4 n0, n1, n2; // linear row directions xyz, w = 1 / Jacobian diagonal
    float4 pa;         // pivot A in world minus origin A ; w = flags (bit0 dynamic_A, bit1 dynamic_B, bit2 awake)
    float4 pb;         // pivot B in world minus origin B
    float4 dpos;       // pivotAInW - pivotBInW
    float4 axis_a;     // hinge axis of A in world ; w = kHinge
    float4 axis_b;     // hinge axis of B in world ; w = limit sign (0 = limit not violated)
    float4 lim;        // hinge: limit correction, accumulated limit impulse (cleared by the setup), motor target velocity, motor max impulse
    float4 prm;        // pin: bias, damping, impulse clamp ; hinge: bias, relaxation, limit bias
};

// Per-step data of a generic 6-DOF joint, beside its JointSetup (pa / pb = anchor minus origin A / B): k_joint_setup decides
// which of the six axes act this step and with what values (bradot_generic_6dof_joint_3d.cpp:316-370 and the tests the
// two solve routines make on constant data, :86-96, :186-206); the solver keeps the accumulated impulses in it.
struct __align__(16) G6Setup {
    float4 lin_a[3];  // axis xyz, 1 / Jacobian diagonal
    float4 lin_b[3];  // depth (limit taken off), softness, restitution, damping
    float4 lin_acc;   // accumulated impulses xyz ; w = bits: 2 per axis (0 skip, 1 lower clamp 0, 2 upper clamp 0, 3 no clamp)
    float4 ang_a[3];  // axis xyz, 1 / Jacobian diagonal
    float4 ang_b[3];  // target velocity, max impulse (force * dt), softness, damping
    float4 ang_c;     // 1 + bounce xyz ; w = bits: 1 per axis (acts)
    float4 ang_acc;   // accumulated impulses xyz
    float4 pad;
};

struct Stats {
    int active_objects, collision_pairs, island_count, contact_pairs, colors, contacts, overflow, steps;
};

// Kinematic bodies only: where the host wants the body after the next step, and its constant velocities.
struct KinState {
    BodyXf target;
    float4 clin, cang;
};

struct bpc_context {
    int device = 0;
    cudaStream_t stream = 0;
    // pipelined state read-back (bpc_bodies_read_begin / _end): packed on the main stream, copied on a side stream
    cudaStream_t copy_stream = nullptr, copy_stream_in = nullptr; // device to host / host to device: one must not queue behind the other
    cudaEvent_t ev_packed = nullptr, ev_read_done = nullptr;
    bool read_inflight = false;
    DevBuf<float> async_staging;
    // bulk force rows go up the same way: copied on the side stream into a buffer of their own, unpacked on the main one
    cudaEvent_t ev_h2d = nullptr, ev_unpacked = nullptr;
    bool unpack_recorded = false;
    DevBuf<float> force_staging;
    std::string err;

    // host mirrors of small tables
    std::vector<SpaceParams> spaces;
    std::vector<ShapeRec> h_shapes;
    std::vector<MeshRec> h_meshes;
    std::vector<float4> h_mverts, h_mplanes;
    std::vector<uint4> h_medges;
    std::vector<uint32_t> h_mindex; // per face: [count, idx...] packed; face_ofs indexes h_mfaces
    std::vector<uint2> h_mfaces;    // (index_ofs, count)
    std::vector<JointRec> h_joints;
    struct MassOverride { float inertia[3]; float com[3]; bool has_inertia, has_com, partial; }; // partial: after a reset (see bpc_bodies_set_mass_override)
    std::unordered_map<uint32_t, MassOverride> h_mass_override; // body_set_param(INERTIA / CENTER_OF_MASS), by slot
    std::vector<uint64_t> h_same; // sorted (a << 32 | b), a < b: records that are shapes of one object and never pair
    DevBuf<uint64_t> d_same;
    bool same_dirty = false;
    std::vector<float> h_coll_priority; // body_set_collision_priority per body; empty until one is set (body_test_motion only)
    DevBuf<float> d_coll_priority;
    std::vector<float> h_shape_bias; // shape_set_custom_solver_bias per shape; empty until one is set
    DevBuf<float> d_shape_bias;
    std::vector<float> h_jext; // JEXT_STRIDE values per joint, present once a joint with a long parameter table exists
    DevBuf<float> d_jext;
    bool shapes_dirty = true, joints_dirty = true, spaces_dirty = true;
    uint32_t *h_counters = nullptr; // pinned mirror of the device counters of the last step
    uint32_t launches = 0;          // kernels launched by the current / last step
    // optional per-stage device timing (CUDA events on the step's stream)
    bool profiling = false, ev_pending = false;
    cudaEvent_t ev[BPC_N_STAGES + 1] = {};
    double stage_ms[BPC_N_STAGES] = {};
    uint32_t stage_steps = 0;
    DevBuf<JointRec> d_joints;
    DevBuf<JointSetup> jsetup;
    DevBuf<G6Setup> jsetup6; // one per joint, present once a 6-DOF joint exists
    bool have_g6dof = false;
    bool have_ext_joints = false; // a cone twist, slider or 6-DOF joint exists: the solve kernels that know them run
    DevBuf<uint32_t> a_range_cur; // same for the current pair list (vertex-centric island pass)
    DevBuf<uint32_t> a_range; // [2 * n_bodies]: begin / end of each body's run (as A) in the previous pair list
    DevBuf<uint32_t> io_slots; // body indices of a gather / scatter call
    DevBuf<float> staging;    // packed per-body floats on their way to / from the host (bulk state I/O)
    DevBuf<uint32_t> q_exclude; // bodies the space queries skip (bpc_space_set_query_exclude), sorted
    uint32_t n_q_exclude = 0;
    DevBuf<uint32_t> np_work; // broadphase pairs that need the full narrowphase pass this step
    DevBuf<uint32_t> np_cwork; // hull-against-hull pairs handed to the warp-per-pair kernel
    DevBuf<uint64_t> forced_keys;

    uint32_t n_bodies = 0;
    // body SoA
    DevBuf<BodyXf> xf, shape_xf;
    DevBuf<BodyVel> vel;
    DevBuf<BodyMassW> massw;
    DevBuf<BodyProps> props;
    DevBuf<BodyForces> forces;
    DevBuf<uint4> meta; // mode|flags<<8, layer, mask, space
    DevBuf<uint32_t> shape_id;
    DevBuf<float4> shape_p;    // per-body copy of the shape parameters (convex: x = mesh index bits)
    DevBuf<ItemBox> items;
    DevBuf<uint32_t> npairs;   // pairs per item after the last broadphase update (drives the margin)
    DevBuf<KinState> kin;        // indexed by body; allocated when the first kinematic body is created
    DevBuf<uint32_t> kin_ids;    // the kinematic bodies
    std::vector<uint32_t> h_kin_ids;
    DevBuf<uint32_t> ccd_ids;    // bodies with continuous collision detection
    std::vector<uint32_t> h_ccd_ids;
    DevBuf<uint32_t> tree_items; // body ids that live in the LBVH (everything except unbounded statics)
    uint32_t n_tree = 0;
    DevBuf<uint32_t> island; // connected-component label per body
    DevBuf<uint32_t> island_flag;
    // shapes
    DevBuf<ShapeRec> shapes;
    DevBuf<MeshRec> meshes;
    DevBuf<float4> mverts, mplanes;
    DevBuf<uint4> medges;
    DevBuf<uint2> mfaces;
    DevBuf<uint32_t> mindex;
    DevBuf<SpaceParams> d_spaces;

    // broadphase
    DevBuf<uint64_t> keys, keys_alt;
    DevBuf<uint32_t> vals, vals_alt;
    DevBuf<uint32_t> hist;
    DevBuf<float4> leaf_min, leaf_max; // sorted leaves: min.w = item id (bits), max.w = space (bits)
    DevBuf<float4> nodes;              // 4 float4 per internal node
    DevBuf<uint32_t> node_parent, node_flag;
    DevBuf<uint32_t> node_maxleaf;     // per internal node and side: highest leaf index below it (all ones if a static leaf is below)
    DevBuf<uint32_t> big_items;        // static items with unbounded boxes (planes), brute-forced; grouped by space
    DevBuf<uint32_t> big_start;        // [n_spaces + 1] offsets of each space's run in big_items
    std::vector<uint32_t> h_space_of;  // host mirror: space of each body
    DevBuf<float4> bounds;             // scene bounds reduction result
    DevBuf<uint32_t> counters;         // device counters (see CNT_*)
    uint32_t n_big = 0;
    bool statics_dirty = true;

    // pairs (double buffered: cur = this step, old = previous step)
    DevBuf<uint64_t> pair_keys[2], pair_keys_raw;
    DevBuf<PairHdr> pair_hdr[2];
    DevBuf<ContactGeo> cgeo[2];
    DevBuf<ContactAcc> cacc[2];
    uint32_t n_pairs[2] = { 0, 0 };
    int cur = 0;
    size_t pair_cap = 0;

    // body pairs excepted from colliding (joint_disable_collisions_between_bodies / body_add_collision_exception)
    DevBuf<uint64_t> except_keys;
    std::vector<uint64_t> h_except;
    std::vector<uint64_t> h_except_user; // body_add_collision_exception (sorted); h_except = these + the live joints' 
    uint32_t n_except = 0;
    bool except_dirty = false;
    uint32_t shape_type_mask = 0; // shape types present among the bodies (bit = ShapeType)
    std::vector<uint8_t> h_big;  // 1 = unbounded static item (world boundary), kept out of the LBVH
    std::vector<uint8_t> h_mode; // host mirror of the body modes (scheduling of teacher-forced sweeps)
    std::vector<uint8_t> h_ccd;  // host mirror: continuous collision detection asked for
    std::vector<uint8_t> h_dead; // 1 = slot of a destroyed body (reused by bpc_bodies_create_slots)
    std::vector<uint32_t> free_slots;
    bool lists_dirty = false;    // tree / unbounded-static lists have to be rebuilt before the next step or query
    uint32_t max_priority = 1;   // highest joint solver priority

    // contact reporting: per body the number of contacts it wants (0 = none), the start of its records, how many it got
    // this step (may exceed the cap: the rest went to the overflow list and was merged by k_report_overflow)
    DevBuf<uint32_t> report_cap, report_ofs, report_count;
    DevBuf<ReportRec> reports, report_overflow;
    DevBuf<uint32_t> report_overflow_body;
    std::vector<uint32_t> h_report_cap;
    uint32_t report_total = 0, n_reporting = 0;
    // areas (k_areas.cu): the table, the areas of each space, what each body stands in
    std::vector<AreaRec> h_areas;
    DevBuf<AreaRec> d_areas;
    DevBuf<uint32_t> area_list, area_start; // area indices grouped by space; offsets per space (n_spaces + 1)
    DevBuf<uint32_t> body_area_cnt, body_area; // per body: how many areas it overlaps, which (AREA_SLOTS each, bit 31 = attached)
    DevBuf<uint4> area_events;                // (area, body, status, 0) of the last step
    std::vector<uint4> h_area_events_pending; // events raised outside a step (a body or an area went away)
    std::vector<uint32_t> area_reset_slots;   // body slots reused since the last step: their area lists start empty
    uint32_t body_area_cap = 0;               // bodies the per-body arrays are sized for
    bool areas_dirty = false, areas_moved = false;
    // area against area (BradotArea2Pair3D): candidates of the coming / last step, their overlap bytes, who overlaps whom
    struct AreaPairCand { uint32_t a, b, flags_a, flags_b, layer_a, mask_a, layer_b, mask_b; };
    std::vector<AreaPairCand> aa_cand;
    DevBuf<uint2> aa_pairs;
    DevBuf<uint32_t> aa_result;
    bool aa_launched = false;
    std::vector<uint64_t> aa_colliding; // sorted (a << 32 | b): a's mask meets b's layer and the shapes overlap
    uint32_t n_areas_alive = 0;
    bool report_dirty = false;

    // constraints / solver
    DevBuf<uint32_t> con_list, con_color, con_sorted, con_sorted_alt;
    DevBuf<uint4> con_info; // per scheduled constraint: body a, body b, flags|count, pair index
    DevBuf<uint64_t> body_best;
    DevBuf<uint64_t> body_colmask;
    DevBuf<uint32_t> color_start; // 66 entries + count
    // solver bins (k_graph.cu): bin of every constraint; per-bin count / lowest body / highest body / fits-one-CTA;
    // per (bin, colour) histogram, offsets and scatter cursors
    DevBuf<uint32_t> con_bin, bin_tab, bin_base;
    DevBuf<uint16_t> con_sorted_color; // colour * 4 + work class of each entry of con_sorted
    DevBuf<uint8_t> con_cls;           // work class (cached contacts - 1) of each entry of con_list
    DevBuf<uint8_t> joint_color;       // sweep colour + 1 each joint had last step (0 = none)
    uint32_t colors_at_fresh = 0;      // colours the last from-scratch colouring needed
    bool last_color_fresh = false;
    bool recolor = true;               // ignore the kept colours at the next step (bodies re-described, joints changed, too many colours)
    uint32_t n_bins = 0;
    bool use_bins = true;
    DevBuf<SolveContact> scontacts;
    DevBuf<uint32_t> forced_order; // teacher-forced sweep order (pair indices), optional
    DevBuf<uint32_t> forced_level_start;
    uint32_t n_forced = 0, n_forced_levels = 0;
    std::vector<uint64_t> h_forced_keys;
    std::vector<uint32_t> h_forced_joints;

    int sm_count = 148;
    Stats stats = {};
    uint32_t step_count = 0;
    float last_dt = 0.f;
};

enum { CNT_PAIRS = 0, CNT_CONS = 1, CNT_COLORS = 2, CNT_UNCOLORED = 3, CNT_ACTIVE = 4, CNT_ISLANDS = 5, CNT_CONTACTS = 6,
       CNT_OVERFLOW = 7, CNT_GROUPS = 8 /* colours on the grid-barrier path */, CNT_BARRIER_CONS = 9, CNT_ACTIVE_CONTACTS = 10, CNT_WORK = 11, CNT_CWORK = 12, CNT_REPORT_OVERFLOW = 13, CNT_AREA_EVENTS = 14, CNT_AREA_OVERFLOW = 15, CNT_ROUND = 20 /* 3 slots */, CNT_N = 32 };

#define REPORT_OVERFLOW_CAP 65536u // reported contacts beyond their bodies' caps held per step for the deepest-N merge

#define BPC_CHECK(x)                                                                                  \
    do {                                                                                              \
        cudaError_t e_ = (x);                                                                         \
        if (e_ != cudaSuccess) {                                                                      \
            ctx->err = std::string(#x) + ": " + cudaGetErrorString(e_);                               \
            return -1;                                                                                \
        }                                                                                             \
    } while (0)

// Grows to at least n elements.  On failure (out of memory, or a failed copy of the old contents) the buffer is left EMPTY
// (p == nullptr, cap == 0), so that the `!buf.p` test every caller makes also catches a failed growth -- never the old,
// smaller allocation that the following kernels would overrun.
template <typename T>
void DevBuf<T>::reserve(size_t n, bool keep, cudaStream_t st) {
    if (n <= cap) return;
    size_t ncap = n + n / 4 + 256;
    T *np = nullptr;
    bool ok = cudaMalloc(&np, ncap * sizeof(T)) == cudaSuccess;
    if (ok) ok = cudaMemsetAsync(np, 0, ncap * sizeof(T), st) == cudaSuccess;
    if (ok && keep && p && cap) ok = cudaMemcpyAsync(np, p, cap * sizeof(T), cudaMemcpyDeviceToDevice, st) == cudaSuccess;
    if (p) {
        cudaStreamSynchronize(st);
        cudaFree(p);
    }
    if (!ok) {
        cudaGetLastError(); // cudaErrorMemoryAllocation is not sticky: clear it so that it is not blamed on a later call
        if (np) cudaFree(np);
        np = nullptr;
        ncap = 0;
    }
    p = np;
    cap = ncap;
}
template <typename T>
void DevBuf<T>::free_() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}

// bpc_bodies_set_state: one thread per listed body (kernel in k_integrate.cu, next to the item / mass-data helpers it uses)
struct SetStateArgs {
    uint32_t n;
    int state;
    const uint32_t *slots;
    const float *values; // n rows
    BodyXf *xf;
    BodyVel *vel;
    BodyMassW *massw;
    BodyProps *props;
    uint4 *meta;
    KinState *kin;
    const BodyXf *shape_xf;
    const uint32_t *shape_id;
    const ShapeRec *shapes;
    ItemBox *items;
    const uint32_t *npairs;
    uint32_t *moved_static; // bit per body: a static body was moved (its neighbours are woken afterwards)
};
int bpc_launch_set_state(bpc_context *ctx, const SetStateArgs &a);

// ---- kernel launchers (one per .cu file) ------------------------------------------------------
int bpc_launch_integrate_forces(bpc_context *ctx, float dt);
int bpc_launch_integrate_velocities(bpc_context *ctx, float dt);
// passes > 0: creation (pass 0 stores the exact box, later ones re-store it as moves); -1: world mass data only;
// -2: world mass data + the broadphase item moved as _update_shapes does (hysteresis with the item's pair count)
int bpc_launch_refresh_bodies(bpc_context *ctx, uint32_t first, uint32_t n, int passes);
int bpc_rebuild_lists(bpc_context *ctx);
int bpc_resolve_step(bpc_context *ctx);
int bpc_launch_broadphase(bpc_context *ctx);
int bpc_build_tree(bpc_context *ctx); // LBVH over the current stored boxes (also used by the space queries)
int bpc_flush_tables(bpc_context *ctx); // bpc_api.cu: shapes, joints, exceptions, spaces to the device when they changed
int bpc_launch_islands(bpc_context *ctx);
int bpc_launch_narrowphase(bpc_context *ctx);
int bpc_flush_area_tables(bpc_context *ctx);                                      // k_areas.cu
int bpc_launch_areas(bpc_context *ctx);
int bpc_areas_body_gone(bpc_context *ctx, const std::vector<uint32_t> &slots);
int bpc_launch_solver(bpc_context *ctx, float dt);
int bpc_launch_joint_setup(bpc_context *ctx, float dt);
int bpc_launch_sleep(bpc_context *ctx, float dt);
// radix sort of 64-bit keys (+ optional 32-bit values) on the 8-bit digits that intersect `bitmask`;
// result is left in *keys/*vals (the pointers are swapped with their alternates as needed)
int bpc_radix_sort(bpc_context *ctx, uint64_t **keys, uint64_t **keys_alt, uint32_t **vals, uint32_t **vals_alt, uint32_t n, uint64_t bitmask);
