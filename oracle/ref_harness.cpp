// oracle_step -- drives the UNMODIFIED reference 3D physics (BradotPhysicsServer3D)
// through its public RID API on a neutral scene file and dumps per-step state.
//
// TEST INFRASTRUCTURE ONLY: this is the checker and the timed CPU baseline, never
// part of the product path (bradot_b200/).  It is our own code; it includes the
// reference's headers from /root/reference and links the reference's objects
// (oracle/build_ref.py).  Compiled with -fno-access-control so private state of
// BradotBodyPair3D / BVH can be READ for the dumps (no reference source is edited).
//
// Reference entry points exercised (relative to /root/reference):
//   modules/bradot_physics_3d/bradot_physics_server_3d.cpp:1632  BradotPhysicsServer3D::step
//   modules/bradot_physics_3d/bradot_step_3d.cpp:186              BradotStep3D::step
//
// usage: oracle_step <scene.bscn> [--steps N] [--dump out.bdmp] [--dump-from A] [--dump-every K | --dump-steps a,b,c]
//                    [--slim] [--rays rays.bin]
//   rays.bin: uint32 n, then n x (from xyz, to xyz) float32; every dumped step also holds "rays" (n,8): hit, body
//   index, position xyz, normal xyz of PhysicsDirectSpaceState3D::intersect_ray in space 0 (mask 0xffffffff), and
//   "points" (n,9): count and up to 8 body indices of intersect_point at each ray's origin
//                    [--threads T] [--warmup W] [--state state.bin] [--events events.bin]
//   state.bin: uint32 n_bodies, then n x 18 float32 (transform rows + origin, linear velocity, angular velocity): applied
//   to every non-static body through body_set_state after the scene is built (a settled state handed over by the other
//   side, so that the CPU run does not pay for settling it again).
//   events.bin: uint32 n, then n x (int32 step, int32 body, int32 kind, float32[12]): applied before the step of that
//   number.  kind 0 body_set_state(TRANSFORM), 1 LINEAR_VELOCITY, 2 ANGULAR_VELOCITY, 3 free(body), 4 SLEEPING(v[0] != 0),
//   5 CAN_SLEEP(v[0] != 0).
// Scenes with several spaces are stepped space by space with the reference's own stepper (what BradotPhysicsServer3D::step
// does, bradot_physics_server_3d.cpp:1632-1648) whenever a dump is written, so that each space's sweep order can be recorded.
// Timing JSON goes to stdout (one line).

#include "core/config/engine.h"
#include "core/config/project_settings.h"
#include "core/io/file_access_pack.h"
#include "core/object/worker_thread_pool.h"
#include "core/os/main_loop.h"
#include "core/register_core_types.h"
#include "core/string/translation_server.h"
#include "drivers/unix/os_unix.h"
#include "modules/bradot_physics_3d/bradot_body_pair_3d.h"
#include "modules/bradot_physics_3d/bradot_broad_phase_3d_bvh.h"
#include "modules/bradot_physics_3d/bradot_physics_server_3d.h"
#include "modules/bradot_physics_3d/bradot_step_3d.h"
#include "modules/bradot_physics_3d/joints/bradot_hinge_joint_3d.h"
#include "modules/bradot_physics_3d/joints/bradot_pin_joint_3d.h"

#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

class OSHeadless : public OS_Unix {
public:
	MainLoop *ml = nullptr;
	void initialize() override { initialize_core(); }
	void initialize_joypads() override {}
	void set_main_loop(MainLoop *p) override { ml = p; }
	void delete_main_loop() override {}
	void finalize() override {}
	bool _check_internal_feature_support(const String &) override { return false; }
	MainLoop *get_main_loop() const override { return ml; }
};

// ---------------------------------------------------------------- scene file

#pragma pack(push, 1)
struct ScnHeader {
	char magic[4]; // "BSCN"
	uint32_t version; // 2
	uint32_t n_spaces, n_shapes, n_bodies, n_joints, n_verts;
	float dt;
	uint32_t iterations;
	uint32_t iter_step; // space s sweeps iterations + s * iter_step times
	uint32_t n_areas; // area records after the vertices
	uint32_t n_joint_ext; // long joint parameter rows after the areas
	uint32_t reserved[4];
};
struct ScnJointExt { // `count` values of joint `joint` from parameter `first` on (slider: SliderJointParam order)
	uint32_t joint, first, count, pad;
	float values[28];
};
struct ScnShape {
	uint32_t type; // PhysicsServer3D::ShapeType
	float p[4]; // plane: n.xyz,d | sphere: r | box: half extents | capsule: radius,height
	uint32_t vert_ofs, vert_cnt; // convex
	float custom_bias; // shape_set_custom_solver_bias, 0 = none
};
struct ScnBody {
	// flags bit0 can_sleep, bit1 ccd, bit2 start asleep, bit3 pushed every step (see push_bodies),
	// bit4 / bit5 linear / angular damp mode REPLACE, bits 8..13 axis locks (PhysicsServer3D::BodyAxis),
	// bits 16..23 body_set_max_contacts_reported
	uint32_t space, mode, shape, flags;
	float xf[12]; // basis rows (9) + origin
	float shape_xf[12];
	float linvel[3], angvel[3];
	float mass, friction, bounce, gravity_scale, lin_damp, ang_damp;
	uint32_t layer, mask;
};
struct ScnJoint {
	uint32_t type; // 0 pin, 1 hinge
	uint32_t body_a, body_b, flags; // bit0 disable collisions between bodies
	float fa[12], fb[12]; // pin: origin of fa/fb only; hinge: local frames
	// flags bit1: hinge use_limit, bit2: hinge enable_motor, bit3: params are given (else the reference defaults stay)
	// pin: params[0..2] = PinJointParam bias, damping, impulse clamp; hinge: params[0..7] = HingeJointParam order
	float params[8];
	uint32_t priority; // joint_set_solver_priority (0 = leave the default 1)
	uint32_t pad[3];
};
struct Event {
	int32_t step, body, kind; // kinds 6 / 7 address an AREA: 6 area_set_transform, 7 free
	float v[12];
};
struct ScnArea { // = bpc_area_desc
	int32_t space, shape;
	uint32_t flags; // bit0 monitorable, bit1 body monitor callback, bit2 gravity is a point
	int32_t priority;
	float xf[12], shape_xf[12];
	uint32_t layer, mask;
	int32_t gravity_mode, linear_damp_mode, angular_damp_mode;
	float gravity, gravity_vector[3], point_unit_distance, linear_damp, angular_damp;
	uint32_t group; // non-zero and equal to the record before: a further shape of that area, not an area of its own
};
#pragma pack(pop)

// what the areas' monitor callbacks were told since the last dump: (area, body, status)
static std::vector<int32_t> g_area_events;
static std::map<uint64_t, int> g_body_of_rid;
static void on_area_monitor(int p_status, const RID &p_body, ObjectID p_instance, int p_body_shape, int p_area_shape, int p_area) {
	auto it = g_body_of_rid.find(p_body.get_id());
	g_area_events.push_back(p_area);
	g_area_events.push_back(it == g_body_of_rid.end() ? -1 : it->second);
	g_area_events.push_back(p_status);
}
// ... and the area monitor callbacks: (area, other area, 2 + status)
static std::map<uint64_t, int> g_area_of_rid;
static void on_area_area_monitor(int p_status, const RID &p_other, ObjectID p_instance, int p_other_shape, int p_area_shape, int p_area) {
	auto it = g_area_of_rid.find(p_other.get_id());
	g_area_events.push_back(p_area);
	g_area_events.push_back(it == g_area_of_rid.end() ? -1 : it->second);
	g_area_events.push_back(2 + p_status);
}

static Transform3D xf_from(const float *f) {
	Transform3D t;
	t.basis = Basis(f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8]);
	t.origin = Vector3(f[9], f[10], f[11]);
	return t;
}

// ---------------------------------------------------------------- dump file
struct Dump {
	FILE *f = nullptr;
	void arr(const char *name, int step, const char *dtype, const std::vector<uint64_t> &shape, const void *data, size_t bytes) {
		char nm[32];
		memset(nm, 0, sizeof(nm));
		snprintf(nm, sizeof(nm), "%s", name);
		fwrite(nm, 1, 32, f);
		int32_t st = step;
		fwrite(&st, 4, 1, f);
		char dt[4] = { 0, 0, 0, 0 };
		snprintf(dt, 4, "%s", dtype);
		fwrite(dt, 1, 4, f);
		uint32_t nd = shape.size();
		fwrite(&nd, 4, 1, f);
		for (uint64_t s : shape) {
			fwrite(&s, 8, 1, f);
		}
		uint64_t b = bytes;
		fwrite(&b, 8, 1, f);
		if (bytes) {
			fwrite(data, 1, bytes, f);
		}
	}
};

static void put3(std::vector<float> &v, const Vector3 &p) {
	v.push_back(p.x);
	v.push_back(p.y);
	v.push_back(p.z);
}

struct World {
	BradotPhysicsServer3D *ps = nullptr;
	std::vector<RID> spaces, shapes, bodies, joints, areas;
	std::vector<char> area_freed;
	bool any_unpickable = false; // some body has body_set_ray_pickable(false): the ray dumps also hold "rays_pick"
	std::vector<BradotBody3D *> body_ptr;
	std::map<const BradotBody3D *, int> body_index;
	std::vector<uint32_t> body_space;
	// kinematic bodies: every step the harness sets transform(t) = (rotation about y by angvel.y t) * basis0, origin0 + linvel t
	std::vector<int> kin_body;
	std::vector<Transform3D> kin_xf0;
	std::vector<Vector3> kin_lin;
	std::vector<float> kin_wy;
	std::vector<uint32_t> body_flags;
	std::vector<int> body_shape; // which shape of its body a scene record stands for (0; 1.. for the further shapes of a static body)
	std::vector<char> body_freed;
	// sweep order of the last step, recorded space by space: (island counted over all spaces, constraint)
	std::vector<std::pair<int, const BradotConstraint3D *>> last_order;
	int32_t last_stats[4] = { 0, 0, 0, 0 };
	bool order_valid = false;
	bool slim = false; // --slim: contact records only for the pairs that hold contacts ("contacts_idx" names them)
};

// One tick.  With `record`: what BradotPhysicsServer3D::step does (bradot_physics_server_3d.cpp:1632-1648), space by space
// through the reference's stepper, noting after each space the constraints its islands were solved in.
static void step_world(World &w, real_t dt, bool record) {
	BradotPhysicsServer3D *ps = w.ps;
	if (!record) {
		ps->step(dt);
		w.order_valid = false;
		return;
	}
	ps->_update_shapes();
	ps->island_count = 0;
	ps->active_objects = 0;
	ps->collision_pairs = 0;
	w.last_order.clear();
	int island_base = 0;
	for (RID sp : w.spaces) {
		BradotSpace3D *space = ps->space_owner.get_or_null(sp);
		if (!ps->active_spaces.has(space)) {
			continue;
		}
		ps->stepper->step(space, dt);
		ps->island_count += space->get_island_count();
		ps->active_objects += space->get_active_objects();
		ps->collision_pairs += space->get_collision_pairs();
		uint32_t ic = space->get_island_count();
		for (uint32_t i = 0; i < ic && i < ps->stepper->constraint_islands.size(); i++) {
			for (BradotConstraint3D *c : ps->stepper->constraint_islands[i]) {
				w.last_order.push_back({ island_base + (int)i, c });
			}
		}
		island_base += (int)ic;
	}
	w.last_stats[0] = ps->island_count;
	w.last_stats[1] = ps->active_objects;
	w.last_stats[2] = ps->collision_pairs;
	w.order_valid = true;
}

#define PAIR_STRIDE 12
#define CONTACT_STRIDE 32

static void dump_step(Dump &d, World &w, int step) {
	int nb = w.bodies.size();
	std::vector<float> bs;
	bs.reserve(nb * 20);
	std::vector<float> items;
	items.reserve(nb * 14);
	for (int i = 0; i < nb; i++) {
		if (w.body_freed[i]) { // a freed body keeps its row (all zeros) so that indices stay put
			bs.insert(bs.end(), 20, 0.f);
			items.insert(items.end(), 14, 0.f);
			continue;
		}
		BradotBody3D *b = w.body_ptr[i];
		const Transform3D &t = b->get_transform();
		for (int r = 0; r < 3; r++) {
			put3(bs, t.basis[r]);
		}
		put3(bs, t.origin);
		put3(bs, b->get_linear_velocity());
		put3(bs, b->get_angular_velocity());
		bs.push_back(b->is_active() ? 1.0f : 0.0f);
		bs.push_back(b->still_time);
		// broadphase item of the shape this record stands for
		const BradotCollisionObject3D::Shape &s = b->shapes[w.body_shape[i]];
		put3(items, s.aabb_cache.position);
		put3(items, s.aabb_cache.size);
		BradotSpace3D *space = b->get_space();
		BradotBroadPhase3DBVH *bp = static_cast<BradotBroadPhase3DBVH *>(space->get_broadphase());
		if (s.bpid == 0) { // a disabled shape is in no broadphase
			for (int q = 0; q < 8; q++) {
				items.push_back(0.f);
			}
			continue;
		}
		uint32_t id = s.bpid - 1;
		BVHHandle h;
		h.set_id(id);
		auto abb = bp->bvh.tree._nodes[0].aabb; // type only; overwritten below
		bp->bvh.tree.item_get_ABB(h, abb);
		put3(items, abb.min);
		put3(items, -abb.neg_max);
		items.push_back((float)bp->bvh.tree._pairs[id].num_pairs);
		items.push_back((float)id);
	}
	d.arr("bodies", step, "f4", { (uint64_t)nb, 20 }, bs.data(), bs.size() * 4);
	// contacts reported to the bodies that asked for them (BradotBody3D::contacts, filled by add_contact in pre_solve):
	// body, local_pos(3), local_normal(3), depth, local_velocity(3), collider_pos(3), collider body, collider_velocity(3), impulse(3)
	{
		std::map<RID, int> rid_index;
		for (int i = 0; i < nb; i++) {
			rid_index[w.bodies[i]] = i;
		}
		std::vector<float> rep;
		for (int i = 0; i < nb; i++) {
			if (w.body_freed[i]) {
				continue;
			}
			BradotBody3D *b = w.body_ptr[i];
			for (int k = 0; k < b->contact_count; k++) {
				const BradotBody3D::Contact &c = b->contacts[k];
				rep.push_back((float)i);
				put3(rep, c.local_pos);
				put3(rep, c.local_normal);
				rep.push_back(c.depth);
				put3(rep, c.local_velocity_at_pos);
				put3(rep, c.collider_pos);
				auto it = rid_index.find(c.collider);
				rep.push_back(it == rid_index.end() ? -1.f : (float)it->second);
				put3(rep, c.collider_velocity_at_pos);
				put3(rep, c.impulse);
			}
		}
		if (!rep.empty()) {
			d.arr("reports", step, "f4", { (uint64_t)rep.size() / 21, 21 }, rep.data(), rep.size() * 4);
		}
	}
	d.arr("items", step, "f4", { (uint64_t)nb, 14 }, items.data(), items.size() * 4);

	// pairs: walk constraint maps, emit each BradotBodyPair3D once (from its A side)
	std::vector<int32_t> pi;
	std::vector<float> pf;
	std::vector<float> cf;
	std::vector<int32_t> cidx;
	std::map<const BradotConstraint3D *, int> pair_index;
	for (int i = 0; i < nb; i++) {
		if (w.body_freed[i]) {
			continue;
		}
		BradotBody3D *b = w.body_ptr[i];
		if (w.body_shape[i] != 0) {
			continue; // a further shape of the body before it: its pairs were listed there
		}
		for (const KeyValue<BradotConstraint3D *, int> &E : b->get_constraint_map()) {
			BradotBodyPair3D *p = dynamic_cast<BradotBodyPair3D *>(E.key);
			if (!p || p->A != b) {
				continue;
			}
			pair_index[p] = pi.size() / 4;
			pi.push_back(w.body_index[p->A] + p->shape_A); // (the further shapes of a static body follow it in the scene)
			pi.push_back(w.body_index[p->B] + p->shape_B);
			pi.push_back(p->contact_count);
			pi.push_back((p->collided ? 1 : 0) | (p->collide_A ? 2 : 0) | (p->collide_B ? 4 : 0) | (p->check_ccd ? 8 : 0));
			put3(pf, p->sep_axis);
			put3(pf, p->offset_B);
			if (w.slim) {
				if (p->contact_count == 0) {
					continue;
				}
				cidx.push_back((int32_t)(pi.size() / 4 - 1));
			}
			for (int c = 0; c < 4; c++) {
				const BradotBodyPair3D::Contact &ct = p->contacts[c];
				size_t base = cf.size();
				put3(cf, ct.local_A);
				put3(cf, ct.local_B);
				put3(cf, ct.normal);
				cf.push_back(ct.depth);
				cf.push_back(ct.acc_normal_impulse);
				put3(cf, ct.acc_tangent_impulse);
				cf.push_back(ct.acc_bias_impulse);
				cf.push_back(ct.acc_bias_impulse_center_of_mass);
				cf.push_back(ct.mass_normal);
				cf.push_back(ct.bias);
				cf.push_back(ct.bounce);
				cf.push_back(ct.active ? 1.0f : 0.0f);
				cf.push_back(ct.used ? 1.0f : 0.0f);
				put3(cf, ct.rA);
				put3(cf, ct.rB);
				cf.push_back(c < p->contact_count ? 1.0f : 0.0f);
				while (cf.size() - base < CONTACT_STRIDE) {
					cf.push_back(0.0f);
				}
			}
		}
	}
	uint64_t np = pi.size() / 4;
	d.arr("pair_ids", step, "i4", { np, 4 }, pi.data(), pi.size() * 4);
	d.arr("pair_f", step, "f4", { np, 6 }, pf.data(), pf.size() * 4);
	if (w.slim) {
		d.arr("contacts_idx", step, "i4", { (uint64_t)cidx.size() }, cidx.data(), cidx.size() * 4);
		d.arr("contacts", step, "f4", { (uint64_t)cidx.size(), 4, CONTACT_STRIDE }, cf.data(), cf.size() * 4);
	} else {
		d.arr("contacts", step, "f4", { np, 4, CONTACT_STRIDE }, cf.data(), cf.size() * 4);
	}

	// solve order actually used this step: after _pre_solve_island/_solve_island the island vectors
	// hold the constraints that were solved, in sweep order (bradot_step_3d.cpp:121-160); step_world noted them.
	std::vector<int32_t> order; // (island, kind, index) kind 0 = pair (index into pair_ids), 1 = joint (index into scene joints)
	if (w.order_valid) {
		std::map<const BradotConstraint3D *, int> joint_index;
		for (size_t j = 0; j < w.joints.size(); j++) {
			joint_index[w.ps->joint_owner.get_or_null(w.joints[j])] = j;
		}
		for (const std::pair<int, const BradotConstraint3D *> &oc : w.last_order) {
			auto it = pair_index.find(oc.second);
			if (it != pair_index.end()) {
				order.push_back(oc.first);
				order.push_back(0);
				order.push_back(it->second);
			} else {
				auto jt = joint_index.find(oc.second);
				if (jt != joint_index.end()) {
					order.push_back(oc.first);
					order.push_back(1);
					order.push_back(jt->second);
				}
			}
		}
	} else {
		w.last_stats[0] = w.last_stats[1] = w.last_stats[2] = 0;
		for (RID sp : w.spaces) {
			BradotSpace3D *space = w.ps->space_owner.get_or_null(sp);
			w.last_stats[0] += space->get_island_count();
			w.last_stats[1] += space->get_active_objects();
			w.last_stats[2] += space->get_collision_pairs();
		}
	}
	d.arr("stats", step, "i4", { 4 }, w.last_stats, sizeof(w.last_stats));
	d.arr("solve_order", step, "i4", { (uint64_t)order.size() / 3, 3 }, order.data(), order.size() * 4);
}

// Bodies with flag bit3 are pushed the way a game pushes a vehicle: body_apply_central_force + body_apply_torque before
// every step (accumulators cleared by the step, bradot_body_3d.cpp:661-662), with values that depend on the body index
// only: F = (0.5 (i % 3 - 1), 2, 0.3) N, T = (0.1, 0, 0.05 (i % 2)) N m.
static void push_bodies(World &w, const std::vector<uint32_t> &flags) {
	for (size_t i = 0; i < w.bodies.size(); i++) {
		if (!(flags[i] & 8) || w.body_freed[i]) {
			continue;
		}
		w.ps->body_apply_central_force(w.bodies[i], Vector3(0.5f * (float)((int)(i % 3) - 1), 2.0f, 0.3f));
		w.ps->body_apply_torque(w.bodies[i], Vector3(0.1f, 0.0f, 0.05f * (float)(i % 2)));
	}
}

// Move the kinematic bodies the way a game does: body_set_state(TRANSFORM) before the step (bradot_body_3d.cpp:354-364).
// The targets are dumped ("kin_targets", step): (n_kinematic, 13) = body index, then the 12 floats of the transform.
static void drive_kinematic(Dump &d, World &w, int step, float dt, bool dump) {
	if (w.kin_body.empty()) {
		return;
	}
	std::vector<float> out;
	for (size_t i = 0; i < w.kin_body.size(); i++) {
		if (w.body_freed[w.kin_body[i]]) {
			continue;
		}
		real_t t = (real_t)step * dt;
		Transform3D x;
		x.basis = Basis(Vector3(0, 1, 0), w.kin_wy[i] * t) * w.kin_xf0[i].basis;
		x.origin = w.kin_xf0[i].origin + w.kin_lin[i] * t;
		w.ps->body_set_state(w.bodies[w.kin_body[i]], PhysicsServer3D::BODY_STATE_TRANSFORM, x);
		out.push_back((float)w.kin_body[i]);
		for (int r = 0; r < 3; r++) {
			put3(out, x.basis.rows[r]);
		}
		put3(out, x.origin);
	}
	if (dump && d.f) {
		d.arr("kin_targets", step, "f4", { (uint64_t)w.kin_body.size(), 13 }, out.data(), out.size() * 4);
	}
}

// PhysicsDirectSpaceState3D::intersect_ray (bradot_space_3d.cpp:109-191) for every ray of the file, space 0
static void dump_rays(Dump &d, World &w, int step, const std::vector<float> &rays) {
	if (rays.empty()) {
		return;
	}
	PhysicsDirectSpaceState3D *ds = w.ps->space_get_direct_state(w.spaces[0]);
	std::map<RID, int> rid_index;
	for (size_t i = 0; i < w.bodies.size(); i++) {
		rid_index[w.bodies[i]] = (int)i;
	}
	size_t n = rays.size() / 6;
	std::vector<float> out(n * 8, 0.f);
	std::vector<int32_t> ray_faces(n, -1); // RayResult::face_index of the plain rays
	std::map<RID, int> area_index;
	for (size_t i = 0; i < w.areas.size(); i++) {
		area_index[w.areas[i]] = (int)i;
	}
	// variants: "rays"; "rays_pick" (RayParameters::pick_ray, when some object is not pickable); with areas in the scene
	// "rays_areas" (collide_with_areas; an area answers as -2 - its index) and "rays_areas_only" (no bodies, pick_ray)
	for (int variant = 0; variant < 4; variant++) {
	const int pick = variant == 1 || variant == 3;
	if ((variant == 1 && !w.any_unpickable) || (variant >= 2 && w.areas.empty())) {
		continue;
	}
	std::fill(out.begin(), out.end(), 0.f);
	for (size_t i = 0; i < n; i++) {
		PhysicsDirectSpaceState3D::RayParameters rp;
		rp.from = Vector3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
		rp.to = Vector3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
		rp.collision_mask = 0xffffffffu;
		rp.pick_ray = pick != 0;
		rp.collide_with_areas = variant >= 2;
		rp.collide_with_bodies = variant != 3;
		PhysicsDirectSpaceState3D::RayResult rr;
		bool hit = ds && ds->intersect_ray(rp, rr);
		out[i * 8 + 0] = hit ? 1.f : 0.f;
		out[i * 8 + 1] = -1.f;
		if (hit) {
			auto it = rid_index.find(rr.rid);
			out[i * 8 + 1] = it == rid_index.end() ? -1.f : (float)it->second;
			auto ita = area_index.find(rr.rid);
			if (ita != area_index.end()) {
				out[i * 8 + 1] = (float)(-2 - ita->second);
			}
			out[i * 8 + 2] = rr.position.x;
			out[i * 8 + 3] = rr.position.y;
			out[i * 8 + 4] = rr.position.z;
			out[i * 8 + 5] = rr.normal.x;
			out[i * 8 + 6] = rr.normal.y;
			out[i * 8 + 7] = rr.normal.z;
			if (variant == 0) {
				ray_faces[i] = rr.face_index;
			}
		}
	}
	static const char *names[4] = { "rays", "rays_pick", "rays_areas", "rays_areas_only" };
	d.arr(names[variant], step, "f4", { (uint64_t)n, 8 }, out.data(), out.size() * 4);
	if (variant == 0) {
		d.arr("ray_faces", step, "i4", { (uint64_t)n }, ray_faces.data(), ray_faces.size() * 4);
	}
	}
	// intersect_point (bradot_space_3d.cpp:62-107) at every ray's origin: count, then up to 8 body indices
	std::vector<int32_t> pts(n * 9, -1);
	for (size_t i = 0; i < n; i++) {
		PhysicsDirectSpaceState3D::PointParameters pp;
		pp.position = Vector3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
		pp.collision_mask = 0xffffffffu;
		PhysicsDirectSpaceState3D::ShapeResult sr[8];
		int cc = ds ? ds->intersect_point(pp, sr, 8) : 0;
		pts[i * 9] = cc;
		for (int k = 0; k < cc; k++) {
			auto it = rid_index.find(sr[k].rid);
			pts[i * 9 + 1 + k] = it == rid_index.end() ? -1 : it->second;
		}
	}
	d.arr("points", step, "i4", { (uint64_t)n, 9 }, pts.data(), pts.size() * 4);
	{ // get_closest_point_to_object_volume (bradot_space_3d.cpp:573-611): ray i's origin against body i % bodies
		std::vector<float> cl(n * 3, 0.f);
		for (size_t i = 0; i < n && ds; i++) {
			size_t bi = i % w.bodies.size();
			if (w.body_freed[bi]) {
				continue;
			}
			Vector3 c = ds->get_closest_point_to_object_volume(w.bodies[bi], Vector3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]));
			cl[i * 3 + 0] = c.x;
			cl[i * 3 + 1] = c.y;
			cl[i * 3 + 2] = c.z;
		}
		d.arr("closest", step, "f4", { (uint64_t)n, 3 }, cl.data(), cl.size() * 4);
	}
	if (!w.areas.empty()) { // "points_areas": the same with collide_with_areas, up to 16 results, areas as -2 - index
		std::vector<int32_t> pa(n * 17, -1);
		for (size_t i = 0; i < n; i++) {
			PhysicsDirectSpaceState3D::PointParameters pp;
			pp.position = Vector3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
			pp.collision_mask = 0xffffffffu;
			pp.collide_with_areas = true;
			PhysicsDirectSpaceState3D::ShapeResult sr[16];
			int cc = ds ? ds->intersect_point(pp, sr, 16) : 0;
			pa[i * 17] = cc;
			for (int k = 0; k < cc; k++) {
				auto it = rid_index.find(sr[k].rid);
				auto ita = area_index.find(sr[k].rid);
				pa[i * 17 + 1 + k] = it != rid_index.end() ? it->second : (ita != area_index.end() ? -2 - ita->second : -1);
			}
		}
		d.arr("points_areas", step, "i4", { (uint64_t)n, 17 }, pa.data(), pa.size() * 4);
	}
	// intersect_shape (bradot_space_3d.cpp:208-262): query i uses shape i % 3 of {sphere 0.35, box (0.3, 0.2, 0.25), capsule r 0.2
	// h 0.9} at the ray's origin, turned about the ray's direction by 0.37 i.  The transforms are dumped so that the other
	// side asks with the same floats.
	static RID qshape[3];
	if (!qshape[0].is_valid()) {
		qshape[0] = w.ps->sphere_shape_create();
		w.ps->shape_set_data(qshape[0], 0.35f);
		qshape[1] = w.ps->box_shape_create();
		w.ps->shape_set_data(qshape[1], Vector3(0.3f, 0.2f, 0.25f));
		qshape[2] = w.ps->capsule_shape_create();
		Dictionary cd;
		cd["radius"] = 0.2f;
		cd["height"] = 0.9f;
		w.ps->shape_set_data(qshape[2], cd);
	}
	std::vector<int32_t> sq(n * 9, -1);
	std::vector<float> sxf;
	for (size_t i = 0; i < n; i++) {
		Vector3 from(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]), to(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
		Vector3 axis = (to - from).normalized();
		if (axis == Vector3()) {
			axis = Vector3(0, 1, 0);
		}
		PhysicsDirectSpaceState3D::ShapeParameters sp;
		sp.shape_rid = qshape[i % 3];
		sp.transform = Transform3D(Basis(axis, 0.37f * (float)i), from);
		sp.collision_mask = 0xffffffffu;
		PhysicsDirectSpaceState3D::ShapeResult sr[8];
		int cc = ds ? ds->intersect_shape(sp, sr, 8) : 0;
		sq[i * 9] = cc;
		for (int k = 0; k < cc; k++) {
			auto it = rid_index.find(sr[k].rid);
			sq[i * 9 + 1 + k] = it == rid_index.end() ? -1 : it->second;
		}
		for (int r = 0; r < 3; r++) {
			put3(sxf, sp.transform.basis.rows[r]);
		}
		put3(sxf, sp.transform.origin);
	}
	// collide_shape (bradot_space_3d.cpp:383-431) with the same shapes and poses, up to 16 contact pairs each
	std::vector<float> cs(n * (1 + 16 * 6), 0.f);
	for (size_t i = 0; i < n; i++) {
		PhysicsDirectSpaceState3D::ShapeParameters sp;
		sp.shape_rid = qshape[i % 3];
		const float *f = &sxf[i * 12];
		sp.transform = xf_from(f); // rows, as dumped
		sp.collision_mask = 0xffffffffu;
		Vector3 res[32];
		int cnt = 0;
		if (ds) {
			ds->collide_shape(sp, res, 16, cnt);
		}
		float *o = &cs[i * (1 + 16 * 6)];
		o[0] = (float)cnt;
		for (int k = 0; k < cnt * 2; k++) {
			o[1 + k * 3 + 0] = res[k].x;
			o[1 + k * 3 + 1] = res[k].y;
			o[1 + k * 3 + 2] = res[k].z;
		}
	}
	// rest_info (bradot_space_3d.cpp:513-569) with the same shapes and poses, margin 0.04 for every other query (the rest use the
	// default 0 -> the 0.0001 floor), motion = 0.01 along the ray; cast_motion (:264-381) along half the ray.  Per query:
	// rest: hit, body, point(3), normal(3), velocity(3); cast: safe, unsafe, hit, body, point(3), normal(3), velocity(3)
	std::vector<float> rest(n * 11, 0.f), cast(n * 13, 0.f);
	for (size_t i = 0; i < n; i++) {
		Vector3 from(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]), to(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
		PhysicsDirectSpaceState3D::ShapeParameters sp;
		sp.shape_rid = qshape[i % 3];
		sp.transform = xf_from(&sxf[i * 12]);
		sp.collision_mask = 0xffffffffu;
		sp.margin = (i % 2) ? 0.04f : 0.0f;
		sp.motion = (to - from).normalized() * 0.01f;
		PhysicsDirectSpaceState3D::ShapeRestInfo ri;
		float *o = &rest[i * 11];
		o[1] = -1.f;
		if (ds && ds->rest_info(sp, &ri)) {
			auto it = rid_index.find(ri.rid);
			o[0] = 1.f;
			o[1] = it == rid_index.end() ? -1.f : (float)it->second;
			o[2] = ri.point.x; o[3] = ri.point.y; o[4] = ri.point.z;
			o[5] = ri.normal.x; o[6] = ri.normal.y; o[7] = ri.normal.z;
			o[8] = ri.linear_velocity.x; o[9] = ri.linear_velocity.y; o[10] = ri.linear_velocity.z;
		}
		sp.margin = 0.0f;
		sp.motion = (to - from) * 0.5f;
		real_t safe = 1, unsafe = 1;
		PhysicsDirectSpaceState3D::ShapeRestInfo ci;
		ci.rid = RID();
		float *c = &cast[i * 13];
		c[3] = -1.f;
		if (ds) {
			ds->cast_motion(sp, safe, unsafe, &ci);
		}
		c[0] = safe;
		c[1] = unsafe;
		if (ci.rid.is_valid()) {
			auto it = rid_index.find(ci.rid);
			c[2] = 1.f;
			c[3] = it == rid_index.end() ? -1.f : (float)it->second;
			c[4] = ci.point.x; c[5] = ci.point.y; c[6] = ci.point.z;
			c[7] = ci.normal.x; c[8] = ci.normal.y; c[9] = ci.normal.z;
			c[10] = ci.linear_velocity.x; c[11] = ci.linear_velocity.y; c[12] = ci.linear_velocity.z;
		}
	}
	d.arr("shape_rest", step, "f4", { (uint64_t)n, 11 }, rest.data(), rest.size() * 4);
	d.arr("shape_cast", step, "f4", { (uint64_t)n, 13 }, cast.data(), cast.size() * 4);
	// exclude lists (RayParameters / ShapeParameters ::exclude): the rays and the shape intersections again with every body that
	// the first pass hit excluded -- rays: hit, body; shapes: count + up to 8 bodies
	std::vector<int32_t> ex_r(n * 2, -1), ex_s(n * 9, -1);
	for (size_t i = 0; i < n; i++) {
		PhysicsDirectSpaceState3D::RayParameters rp;
		rp.from = Vector3(rays[i * 6 + 0], rays[i * 6 + 1], rays[i * 6 + 2]);
		rp.to = Vector3(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]);
		rp.collision_mask = 0xffffffffu;
		if (out[i * 8 + 0] != 0.f && out[i * 8 + 1] >= 0.f) {
			rp.exclude.insert(w.bodies[(int)out[i * 8 + 1]]);
		}
		PhysicsDirectSpaceState3D::RayResult rr;
		bool hit = ds && ds->intersect_ray(rp, rr);
		ex_r[i * 2] = hit ? 1 : 0;
		if (hit) {
			auto it = rid_index.find(rr.rid);
			ex_r[i * 2 + 1] = it == rid_index.end() ? -1 : it->second;
		}
		PhysicsDirectSpaceState3D::ShapeParameters sp;
		sp.shape_rid = qshape[i % 3];
		sp.transform = xf_from(&sxf[i * 12]);
		sp.collision_mask = 0xffffffffu;
		if (sq[i * 9] > 0 && sq[i * 9 + 1] >= 0) {
			sp.exclude.insert(w.bodies[sq[i * 9 + 1]]);
		}
		PhysicsDirectSpaceState3D::ShapeResult sr[8];
		int cc = ds ? ds->intersect_shape(sp, sr, 8) : 0;
		ex_s[i * 9] = cc;
		for (int k = 0; k < cc; k++) {
			auto it = rid_index.find(sr[k].rid);
			ex_s[i * 9 + 1 + k] = it == rid_index.end() ? -1 : it->second;
		}
	}
	// intersect_shape again with ShapeParameters::margin = 0.05 (the withMargin routines as a yes / no test)
	std::vector<int32_t> sqm(n * 9, -1);
	for (size_t i = 0; i < n; i++) {
		PhysicsDirectSpaceState3D::ShapeParameters sp;
		sp.shape_rid = qshape[i % 3];
		sp.transform = xf_from(&sxf[i * 12]);
		sp.collision_mask = 0xffffffffu;
		sp.margin = 0.05f;
		PhysicsDirectSpaceState3D::ShapeResult sr[8];
		int cc = ds ? ds->intersect_shape(sp, sr, 8) : 0;
		sqm[i * 9] = cc;
		for (int k = 0; k < cc; k++) {
			auto it = rid_index.find(sr[k].rid);
			sqm[i * 9 + 1 + k] = it == rid_index.end() ? -1 : it->second;
		}
	}
	d.arr("shape_queries_margin", step, "i4", { (uint64_t)n, 9 }, sqm.data(), sqm.size() * 4);
	if (!w.areas.empty()) { // "shape_queries_areas": intersect_shape with collide_with_areas, up to 16 results, areas as -2 - index
		std::map<RID, int> aidx;
		for (size_t i = 0; i < w.areas.size(); i++) {
			if (aidx.find(w.areas[i]) == aidx.end()) {
				aidx[w.areas[i]] = (int)i;
			}
		}
		std::vector<int32_t> sqa(n * 17, -1);
		for (size_t i = 0; i < n; i++) {
			PhysicsDirectSpaceState3D::ShapeParameters sp;
			sp.shape_rid = qshape[i % 3];
			sp.transform = xf_from(&sxf[i * 12]);
			sp.collision_mask = 0xffffffffu;
			sp.collide_with_areas = true;
			PhysicsDirectSpaceState3D::ShapeResult sr[16];
			int cc = ds ? ds->intersect_shape(sp, sr, 16) : 0;
			sqa[i * 17] = cc;
			for (int k = 0; k < cc; k++) {
				auto it = rid_index.find(sr[k].rid);
				auto ita = aidx.find(sr[k].rid);
				sqa[i * 17 + 1 + k] = it != rid_index.end() ? it->second : (ita != aidx.end() ? -2 - ita->second : -1);
			}
		}
		d.arr("shape_queries_areas", step, "i4", { (uint64_t)n, 17 }, sqa.data(), sqa.size() * 4);
	}
	d.arr("rays_excl", step, "i4", { (uint64_t)n, 2 }, ex_r.data(), ex_r.size() * 4);
	d.arr("shapes_excl", step, "i4", { (uint64_t)n, 9 }, ex_s.data(), ex_s.size() * 4);
	d.arr("shape_collide", step, "f4", { (uint64_t)n, 1 + 16 * 6 }, cs.data(), cs.size() * 4);
	d.arr("shape_queries", step, "i4", { (uint64_t)n, 9 }, sq.data(), sq.size() * 4);
	d.arr("shape_query_xf", step, "f4", { (uint64_t)n, 12 }, sxf.data(), sxf.size() * 4);
}

// body_test_motion (bradot_space_3d.cpp:651-1029) for every third moving body of space 0, from where it stands (so it touches
// its neighbours: the recovery step has work), along a direction and with parameters that depend on its index.  The requests
// are dumped ("motion_queries": body, from(12), motion(3), margin, max_collisions, flags) with the answers ("motion_results":
// collided, count, travel(3), remainder(3), depth, safe, unsafe, then 4 x (collider, position(3), normal(3), depth, velocity(3))).
static void dump_motions(Dump &d, World &w, int step) {
	std::map<RID, int> rid_index;
	for (size_t i = 0; i < w.bodies.size(); i++) {
		rid_index[w.bodies[i]] = (int)i;
	}
	std::vector<float> q, res;
	uint64_t n = 0;
	for (size_t i = 0; i < w.bodies.size() && n < 600; i += 3) {
		if (w.body_freed[i] || w.body_space[i] != 0 || w.body_ptr[i]->get_mode() < PhysicsServer3D::BODY_MODE_KINEMATIC) {
			continue;
		}
		PhysicsServer3D::MotionParameters mp;
		mp.from = w.body_ptr[i]->get_transform();
		float ang = 0.7f * (float)i;
		float len = 0.05f + 0.15f * (float)(i % 5);
		mp.motion = Vector3(Math::cos(ang) * len, ((i % 4) == 0 ? 0.5f : -0.5f) * len, Math::sin(ang) * len);
		mp.margin = (i % 2) ? 0.04f : 0.001f;
		mp.max_collisions = (i % 6) < 3 ? 1 : 4;
		mp.recovery_as_collision = (i % 9) < 3;
		PhysicsServer3D::MotionResult mr;
		bool collided = w.ps->body_test_motion(w.bodies[i], mp, &mr);
		q.push_back((float)i);
		for (int r = 0; r < 3; r++) {
			put3(q, mp.from.basis.rows[r]);
		}
		put3(q, mp.from.origin);
		put3(q, mp.motion);
		q.push_back(mp.margin);
		q.push_back((float)mp.max_collisions);
		q.push_back(mp.recovery_as_collision ? 2.f : 0.f);
		res.push_back(collided ? 1.f : 0.f);
		res.push_back((float)mr.collision_count);
		put3(res, mr.travel);
		put3(res, mr.remainder);
		res.push_back(mr.collision_depth);
		res.push_back(mr.collision_safe_fraction);
		res.push_back(mr.collision_unsafe_fraction);
		for (int c = 0; c < 4; c++) {
			const PhysicsServer3D::MotionCollision &mc = mr.collisions[c];
			auto it = rid_index.find(mc.collider);
			res.push_back(c < mr.collision_count && it != rid_index.end() ? (float)it->second : -1.f);
			put3(res, mc.position);
			put3(res, mc.normal);
			res.push_back(mc.depth);
			put3(res, mc.collider_velocity);
		}
		n++;
	}
	if (n) {
		d.arr("motion_queries", step, "f4", { n, 19 }, q.data(), q.size() * 4);
		d.arr("motion_results", step, "f4", { n, 11 + 4 * 11 }, res.data(), res.size() * 4);
	}
}

static void dump_meshes(Dump &d, World &w) {
	// convex hull meshes as the reference built them (ConvexHullComputer, bradot_shape_3d.cpp:1099)
	for (size_t i = 0; i < w.shapes.size(); i++) {
		BradotShape3D *sh = w.ps->shape_owner.get_or_null(w.shapes[i]);
		if (sh->get_type() == PhysicsServer3D::SHAPE_CONCAVE_POLYGON) {
			// the tree over the faces and the face normals as the reference built them (bradot_shape_3d.cpp:1614-1672):
			// "concave_bvh" (n_nodes, 9) = box position, box size, left, right, face; "concave_normals" (n_faces, 3)
			BradotConcavePolygonShape3D *cc = static_cast<BradotConcavePolygonShape3D *>(sh);
			std::vector<float> nodes, normals;
			for (int k = 0; k < cc->bvh.size(); k++) {
				put3(nodes, cc->bvh[k].aabb.position);
				put3(nodes, cc->bvh[k].aabb.size);
				nodes.push_back((float)cc->bvh[k].left);
				nodes.push_back((float)cc->bvh[k].right);
				nodes.push_back((float)cc->bvh[k].face_index);
			}
			for (int k = 0; k < cc->faces.size(); k++) {
				put3(normals, cc->faces[k].normal);
			}
			d.arr("concave_bvh", (int)i, "f4", { (uint64_t)cc->bvh.size(), 9 }, nodes.data(), nodes.size() * 4);
			d.arr("concave_normals", (int)i, "f4", { (uint64_t)cc->faces.size(), 3 }, normals.data(), normals.size() * 4);
			continue;
		}
		if (sh->get_type() != PhysicsServer3D::SHAPE_CONVEX_POLYGON) {
			continue;
		}
		BradotConvexPolygonShape3D *cv = static_cast<BradotConvexPolygonShape3D *>(sh);
		const Geometry3D::MeshData &m = cv->get_mesh();
		std::vector<float> v;
		for (uint32_t k = 0; k < m.vertices.size(); k++) {
			put3(v, m.vertices[k]);
		}
		std::vector<int32_t> e;
		for (uint32_t k = 0; k < m.edges.size(); k++) {
			e.push_back(m.edges[k].vertex_a);
			e.push_back(m.edges[k].vertex_b);
			e.push_back(m.edges[k].face_a);
			e.push_back(m.edges[k].face_b);
		}
		std::vector<float> fp;
		std::vector<int32_t> fi; // per face: start, count ; then index blob
		std::vector<int32_t> idx;
		for (uint32_t k = 0; k < m.faces.size(); k++) {
			put3(fp, m.faces[k].plane.normal);
			fp.push_back(m.faces[k].plane.d);
			fi.push_back(idx.size());
			fi.push_back(m.faces[k].indices.size());
			for (uint32_t q = 0; q < m.faces[k].indices.size(); q++) {
				idx.push_back(m.faces[k].indices[q]);
			}
		}
		std::vector<int32_t> ext;
		for (uint32_t k = 0; k < cv->extreme_vertices.size(); k++) {
			ext.push_back(cv->extreme_vertices[k]);
		}
		d.arr("mesh_verts", i, "f4", { (uint64_t)m.vertices.size(), 3 }, v.data(), v.size() * 4);
		d.arr("mesh_edges", i, "i4", { (uint64_t)m.edges.size(), 4 }, e.data(), e.size() * 4);
		d.arr("mesh_planes", i, "f4", { (uint64_t)m.faces.size(), 4 }, fp.data(), fp.size() * 4);
		d.arr("mesh_faces", i, "i4", { (uint64_t)m.faces.size(), 2 }, fi.data(), fi.size() * 4);
		d.arr("mesh_index", i, "i4", { (uint64_t)idx.size() }, idx.data(), idx.size() * 4);
		d.arr("mesh_extreme", i, "i4", { (uint64_t)ext.size() }, ext.data(), ext.size() * 4);
	}
	// mass properties as the reference derived them (bradot_body_3d.cpp:56-166)
	int nb = w.bodies.size();
	std::vector<float> mp;
	for (int i = 0; i < nb; i++) {
		BradotBody3D *b = w.body_ptr[i];
		mp.push_back(b->_inv_mass);
		put3(mp, b->_inv_inertia);
		put3(mp, b->center_of_mass_local);
		for (int r = 0; r < 3; r++) {
			put3(mp, b->principal_inertia_axes_local[r]);
		}
	}
	d.arr("mass_props", -1, "f4", { (uint64_t)nb, 16 }, mp.data(), mp.size() * 4);
}

int main(int argc, char **argv) {
	if (argc < 2) {
		fprintf(stderr, "usage: oracle_step scene.bscn [--steps N] [--dump F] [--dump-from A] [--dump-every K] [--threads T] [--warmup W]\n");
		return 2;
	}
	const char *scene_path = argv[1];
	int steps = 1, dump_from = 0, dump_every = 1, threads = -1, warmup = 0;
	const char *dump_path = nullptr, *rays_path = nullptr, *state_path = nullptr, *events_path = nullptr;
	bool slim = false;
	bool motions = false; // --motions: body_test_motion requests on every dumped step (dump_motions)
	std::set<int> dump_steps;
	for (int i = 2; i < argc; i++) {
		std::string a = argv[i];
		if (a == "--steps") {
			steps = atoi(argv[++i]);
		} else if (a == "--dump") {
			dump_path = argv[++i];
		} else if (a == "--dump-from") {
			dump_from = atoi(argv[++i]);
		} else if (a == "--dump-every") {
			dump_every = atoi(argv[++i]);
		} else if (a == "--threads") {
			threads = atoi(argv[++i]);
		} else if (a == "--warmup") {
			warmup = atoi(argv[++i]);
		} else if (a == "--rays") {
			rays_path = argv[++i];
		} else if (a == "--state") {
			state_path = argv[++i];
		} else if (a == "--events") {
			events_path = argv[++i];
		} else if (a == "--motions") {
			motions = true;
		} else if (a == "--slim") {
			slim = true;
		} else if (a == "--dump-steps") { // comma-separated step numbers: dump exactly these (overrides --dump-from / --dump-every)
			for (char *tok = strtok(argv[++i], ","); tok; tok = strtok(nullptr, ",")) {
				dump_steps.insert(atoi(tok));
			}
		}
	}

	std::vector<float> rays;
	if (rays_path) {
		FILE *rf = fopen(rays_path, "rb");
		uint32_t nr = 0;
		if (!rf || fread(&nr, 4, 1, rf) != 1) {
			fprintf(stderr, "cannot read %s\n", rays_path);
			return 2;
		}
		rays.resize((size_t)nr * 6);
		if (fread(rays.data(), 4, rays.size(), rf) != rays.size()) {
			fprintf(stderr, "short ray file\n");
			return 2;
		}
		fclose(rf);
	}
	FILE *sf = fopen(scene_path, "rb");
	if (!sf) {
		fprintf(stderr, "cannot open %s\n", scene_path);
		return 2;
	}
	ScnHeader hd;
	if (fread(&hd, sizeof(hd), 1, sf) != 1 || memcmp(hd.magic, "BSCN", 4) != 0 || hd.version != 2) {
		fprintf(stderr, "bad scene file\n");
		return 2;
	}
	std::vector<ScnShape> sshapes(hd.n_shapes);
	std::vector<ScnBody> sbodies(hd.n_bodies);
	std::vector<ScnJoint> sjoints(hd.n_joints);
	std::vector<float> sverts(hd.n_verts * 3);
	if (hd.n_shapes && fread(sshapes.data(), sizeof(ScnShape), hd.n_shapes, sf) != hd.n_shapes) return 2;
	if (hd.n_bodies && fread(sbodies.data(), sizeof(ScnBody), hd.n_bodies, sf) != hd.n_bodies) return 2;
	if (hd.n_joints && fread(sjoints.data(), sizeof(ScnJoint), hd.n_joints, sf) != hd.n_joints) return 2;
	if (hd.n_verts && fread(sverts.data(), 12, hd.n_verts, sf) != hd.n_verts) return 2;
	std::vector<ScnArea> sareas(hd.n_areas);
	if (hd.n_areas && fread(sareas.data(), sizeof(ScnArea), hd.n_areas, sf) != hd.n_areas) return 2;
	std::vector<ScnJointExt> sjext(hd.n_joint_ext);
	if (hd.n_joint_ext && fread(sjext.data(), sizeof(ScnJointExt), hd.n_joint_ext, sf) != hd.n_joint_ext) return 2;
	fclose(sf);

	// ---- minimal engine bring-up (order follows Main::test_setup, main/main.cpp:705-747)
	OSHeadless os;
	Thread::make_main_thread();
	set_current_thread_safe_for_nodes(true);
	os.initialize();
	Engine *engine = memnew(Engine);
	register_core_types();
	PackedData *packed_data = memnew(PackedData);
	ProjectSettings *globals = memnew(ProjectSettings);
	register_core_settings();
	WorkerThreadPool::get_singleton()->init(threads, 0.3);
	(void)engine;
	(void)packed_data;
	(void)globals;

	World w;
	w.slim = slim;
	w.ps = memnew(BradotPhysicsServer3D(false));
	w.ps->init();
	BradotPhysicsServer3D *ps = w.ps;

	for (uint32_t i = 0; i < hd.n_spaces; i++) {
		RID sp = ps->space_create();
		ps->space_set_active(sp, true);
		ps->space_set_param(sp, PhysicsServer3D::SPACE_PARAM_SOLVER_ITERATIONS, hd.iterations + i * hd.iter_step);
		w.spaces.push_back(sp);
	}
	for (uint32_t i = 0; i < hd.n_shapes; i++) {
		const ScnShape &s = sshapes[i];
		RID sh;
		switch (s.type) {
			case PhysicsServer3D::SHAPE_WORLD_BOUNDARY:
				sh = ps->world_boundary_shape_create();
				ps->shape_set_data(sh, Plane(Vector3(s.p[0], s.p[1], s.p[2]), s.p[3]));
				break;
			case PhysicsServer3D::SHAPE_SPHERE:
				sh = ps->sphere_shape_create();
				ps->shape_set_data(sh, s.p[0]);
				break;
			case PhysicsServer3D::SHAPE_BOX:
				sh = ps->box_shape_create();
				ps->shape_set_data(sh, Vector3(s.p[0], s.p[1], s.p[2]));
				break;
			case PhysicsServer3D::SHAPE_CAPSULE: {
				sh = ps->capsule_shape_create();
				Dictionary dct;
				dct["radius"] = s.p[0];
				dct["height"] = s.p[1];
				ps->shape_set_data(sh, dct);
			} break;
			case PhysicsServer3D::SHAPE_CONVEX_POLYGON: {
				sh = ps->convex_polygon_shape_create();
				Vector<Vector3> pts;
				for (uint32_t k = 0; k < s.vert_cnt; k++) {
					const float *v = &sverts[(s.vert_ofs + k) * 3];
					pts.push_back(Vector3(v[0], v[1], v[2]));
				}
				ps->shape_set_data(sh, pts);
			} break;
			case PhysicsServer3D::SHAPE_CONCAVE_POLYGON: { // vertices: three per face; p[0] != 0: backface_collision
				sh = ps->concave_polygon_shape_create();
				Vector<Vector3> pts;
				for (uint32_t k = 0; k < s.vert_cnt; k++) {
					const float *v = &sverts[(s.vert_ofs + k) * 3];
					pts.push_back(Vector3(v[0], v[1], v[2]));
				}
				Dictionary dct;
				dct["faces"] = pts;
				dct["backface_collision"] = s.p[0] != 0.f;
				ps->shape_set_data(sh, dct);
			} break;
			case PhysicsServer3D::SHAPE_HEIGHTMAP: { // p = (width, depth, min height, max height); heights = x of its vertices
				sh = ps->heightmap_shape_create();
				PackedFloat32Array hs;
				for (uint32_t k = 0; k < s.vert_cnt; k++) {
					hs.push_back(sverts[(s.vert_ofs + k) * 3]);
				}
				Dictionary dct;
				dct["width"] = (int)s.p[0];
				dct["depth"] = (int)s.p[1];
				dct["heights"] = hs;
				dct["min_height"] = s.p[2];
				dct["max_height"] = s.p[3];
				ps->shape_set_data(sh, dct);
			} break;
			default:
				fprintf(stderr, "unsupported shape type %u\n", s.type);
				return 2;
		}
		if (s.custom_bias != 0.0f) {
			ps->shape_set_custom_solver_bias(sh, s.custom_bias);
		}
		w.shapes.push_back(sh);
	}
	for (uint32_t i = 0; i < hd.n_bodies; i++) {
		const ScnBody &b = sbodies[i];
		if ((b.flags & 0x4000) && i > 0) { // bit14: a further shape of the (static) body before it, not a body of its own
			RID prev = w.bodies[i - 1]; // (its shape was added when the body was made, before it entered the space: see below)
			w.bodies.push_back(prev);
			w.body_ptr.push_back(w.body_ptr[i - 1]);
			w.body_shape.push_back(w.body_shape[i - 1] + 1);
			w.body_space.push_back(b.space);
			w.body_flags.push_back(b.flags);
			w.body_freed.push_back(0);
			continue;
		}
		RID body = ps->body_create();
		ps->body_set_mode(body, (PhysicsServer3D::BodyMode)b.mode);
		ps->body_add_shape(body, w.shapes[b.shape], xf_from(b.shape_xf), false);
		for (uint32_t j = i + 1; j < hd.n_bodies && (sbodies[j].flags & 0x4000); j++) { // its further shapes, in order
			ps->body_add_shape(body, w.shapes[sbodies[j].shape], xf_from(sbodies[j].shape_xf), false);
		}
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_MASS, b.mass);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_FRICTION, b.friction);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_BOUNCE, b.bounce);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_GRAVITY_SCALE, b.gravity_scale);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_LINEAR_DAMP, b.lin_damp);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_ANGULAR_DAMP, b.ang_damp);
		ps->body_set_collision_layer(body, b.layer);
		ps->body_set_collision_mask(body, b.mask);
		ps->body_set_state(body, PhysicsServer3D::BODY_STATE_TRANSFORM, xf_from(b.xf));
		if (b.mode != PhysicsServer3D::BODY_MODE_KINEMATIC) { // a kinematic body's scene velocities are its motion script (below)
			ps->body_set_state(body, PhysicsServer3D::BODY_STATE_LINEAR_VELOCITY, Vector3(b.linvel[0], b.linvel[1], b.linvel[2]));
			ps->body_set_state(body, PhysicsServer3D::BODY_STATE_ANGULAR_VELOCITY, Vector3(b.angvel[0], b.angvel[1], b.angvel[2]));
		}
		ps->body_set_state(body, PhysicsServer3D::BODY_STATE_CAN_SLEEP, (b.flags & 1) != 0);
		if (b.flags & 2) {
			ps->body_set_enable_continuous_collision_detection(body, true);
		}
		if (b.flags & 0x8000) { // bit15: the shape is disabled
			ps->body_set_shape_disabled(body, 0, true);
		}
		if (b.flags & 128) { // bit7: not ray pickable
			ps->body_set_ray_pickable(body, false);
			w.any_unpickable = true;
		}
		if (b.flags & 16) {
			ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_LINEAR_DAMP_MODE, PhysicsServer3D::BODY_DAMP_MODE_REPLACE);
		}
		if (b.flags & 32) {
			ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_ANGULAR_DAMP_MODE, PhysicsServer3D::BODY_DAMP_MODE_REPLACE);
		}
		for (int ax = 0; ax < 6; ax++) {
			if ((b.flags >> 8) & (1u << ax)) {
				ps->body_set_axis_lock(body, (PhysicsServer3D::BodyAxis)(1 << ax), true);
			}
		}
		if ((b.flags >> 16) & 0xff) {
			ps->body_set_max_contacts_reported(body, (b.flags >> 16) & 0xff);
		}
		ps->body_set_space(body, w.spaces[b.space]);
		if (b.flags & 4) {
			ps->body_set_state(body, PhysicsServer3D::BODY_STATE_SLEEPING, true);
		}
		w.bodies.push_back(body);
		BradotBody3D *bp = ps->body_owner.get_or_null(body);
		w.body_ptr.push_back(bp);
		w.body_index[bp] = i;
		w.body_shape.push_back(0);
		w.body_space.push_back(b.space);
		w.body_flags.push_back(b.flags);
		w.body_freed.push_back(0);
		if (b.mode == PhysicsServer3D::BODY_MODE_KINEMATIC) {
			w.kin_body.push_back(i);
			w.kin_xf0.push_back(xf_from(b.xf));
			w.kin_lin.push_back(Vector3(b.linvel[0], b.linvel[1], b.linvel[2]));
			w.kin_wy.push_back(b.angvel[1]);
		}
	}
	for (uint32_t i = 0; i < hd.n_joints; i++) {
		const ScnJoint &j = sjoints[i];
		RID jr = ps->joint_create();
		if (j.type == 0) {
			ps->joint_make_pin(jr, w.bodies[j.body_a], Vector3(j.fa[9], j.fa[10], j.fa[11]), w.bodies[j.body_b], Vector3(j.fb[9], j.fb[10], j.fb[11]));
		} else if (j.type == 2) {
			ps->joint_make_cone_twist(jr, w.bodies[j.body_a], xf_from(j.fa), w.bodies[j.body_b], xf_from(j.fb));
		} else if (j.type == 3) {
			ps->joint_make_slider(jr, w.bodies[j.body_a], xf_from(j.fa), w.bodies[j.body_b], xf_from(j.fb));
		} else if (j.type == 4) {
			ps->joint_make_generic_6dof(jr, w.bodies[j.body_a], xf_from(j.fa), w.bodies[j.body_b], xf_from(j.fb));
		} else if (j.flags & 16) { // joint_make_hinge_simple: pivots = the frame origins, axes = the z columns of the frames
			ps->joint_make_hinge_simple(jr, w.bodies[j.body_a], Vector3(j.fa[9], j.fa[10], j.fa[11]), Vector3(j.fa[2], j.fa[5], j.fa[8]), w.bodies[j.body_b],
					Vector3(j.fb[9], j.fb[10], j.fb[11]), Vector3(j.fb[2], j.fb[5], j.fb[8]));
		} else {
			ps->joint_make_hinge(jr, w.bodies[j.body_a], xf_from(j.fa), w.bodies[j.body_b], xf_from(j.fb));
		}
		if (j.flags & 1) {
			ps->joint_disable_collisions_between_bodies(jr, true);
		}
		if (j.type == 3) {
			for (const ScnJointExt &e : sjext) {
				for (uint32_t k = 0; e.joint == i && k < e.count; k++) {
					ps->slider_joint_set_param(jr, (PhysicsServer3D::SliderJointParam)(e.first + k), e.values[k]);
				}
			}
		} else if (j.type == 4) {
			// blocks of 28 per axis: the G6DOFJointAxisParam values, then the flags (include/bradot_physics_cuda.h)
			for (const ScnJointExt &e : sjext) {
				for (uint32_t k = 0; e.joint == i && k < e.count; k++) {
					uint32_t idx = e.first + k, axis = idx / 28, prm = idx % 28;
					if (prm < (uint32_t)PhysicsServer3D::G6DOF_JOINT_MAX) {
						if (prm != (uint32_t)PhysicsServer3D::G6DOF_JOINT_ANGULAR_MOTOR_FORCE_LIMIT) { // (that member cannot be reached through set_param)
							ps->generic_6dof_joint_set_param(jr, (Vector3::Axis)axis, (PhysicsServer3D::G6DOFJointAxisParam)prm, e.values[k]);
						}
					} else {
						ps->generic_6dof_joint_set_flag(jr, (Vector3::Axis)axis, (PhysicsServer3D::G6DOFJointAxisFlag)(prm - 22), e.values[k] != 0.0f);
					}
				}
			}
		} else if (j.flags & 8) {
			if (j.type == 0) {
				for (int k = 0; k < 3; k++) {
					ps->pin_joint_set_param(jr, (PhysicsServer3D::PinJointParam)k, j.params[k]);
				}
			} else if (j.type == 2) {
				for (int k = 0; k < (int)PhysicsServer3D::CONE_TWIST_MAX; k++) {
					ps->cone_twist_joint_set_param(jr, (PhysicsServer3D::ConeTwistJointParam)k, j.params[k]);
				}
			} else {
				for (int k = 0; k < (int)PhysicsServer3D::HINGE_JOINT_MAX; k++) {
					ps->hinge_joint_set_param(jr, (PhysicsServer3D::HingeJointParam)k, j.params[k]);
				}
			}
		}
		if (j.type == 1) {
			ps->hinge_joint_set_flag(jr, PhysicsServer3D::HINGE_JOINT_FLAG_USE_LIMIT, (j.flags & 2) != 0);
			ps->hinge_joint_set_flag(jr, PhysicsServer3D::HINGE_JOINT_FLAG_ENABLE_MOTOR, (j.flags & 4) != 0);
		}
		if (j.priority) {
			ps->joint_set_solver_priority(jr, (int)j.priority);
		}
		w.joints.push_back(jr);
	}

	// areas (servers/physics_server_3d.h:296-358)
	for (size_t i = 0; i < w.bodies.size(); i++) {
		g_body_of_rid[w.bodies[i].get_id()] = (int)i;
	}
	for (uint32_t i = 0; i < hd.n_areas; i++) {
		const ScnArea &a = sareas[i];
		if (i > 0 && a.group != 0 && sareas[i - 1].group == a.group) { // a further shape of the area before it (added when that one was made)
			w.areas.push_back(w.areas[i - 1]);
			w.area_freed.push_back(0);
			continue;
		}
		RID ar = ps->area_create();
		ps->area_add_shape(ar, w.shapes[a.shape], xf_from(a.shape_xf));
		for (uint32_t j = i + 1; a.group != 0 && j < hd.n_areas && sareas[j].group == a.group; j++) {
			ps->area_add_shape(ar, w.shapes[sareas[j].shape], xf_from(sareas[j].shape_xf));
		}
		ps->area_set_transform(ar, xf_from(a.xf));
		ps->area_set_collision_layer(ar, a.layer);
		ps->area_set_collision_mask(ar, a.mask);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_GRAVITY_OVERRIDE_MODE, a.gravity_mode);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_LINEAR_DAMP_OVERRIDE_MODE, a.linear_damp_mode);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_ANGULAR_DAMP_OVERRIDE_MODE, a.angular_damp_mode);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_GRAVITY, a.gravity);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_GRAVITY_VECTOR, Vector3(a.gravity_vector[0], a.gravity_vector[1], a.gravity_vector[2]));
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_GRAVITY_IS_POINT, (a.flags & 4) != 0);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_GRAVITY_POINT_UNIT_DISTANCE, a.point_unit_distance);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_LINEAR_DAMP, a.linear_damp);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_ANGULAR_DAMP, a.angular_damp);
		ps->area_set_param(ar, PhysicsServer3D::AREA_PARAM_PRIORITY, a.priority);
		ps->area_set_monitorable(ar, (a.flags & 1) != 0);
		if (a.flags & 16) { // bit4: ray pickable (areas start out not pickable)
			ps->area_set_ray_pickable(ar, true);
		}
		if (a.flags & 2) {
			ps->area_set_monitor_callback(ar, callable_mp_static(&on_area_monitor).bind((int)i));
		}
		if (a.flags & 8) {
			ps->area_set_area_monitor_callback(ar, callable_mp_static(&on_area_area_monitor).bind((int)i));
		}
		g_area_of_rid[ar.get_id()] = (int)i;
		ps->area_set_space(ar, w.spaces[a.space]);
		w.areas.push_back(ar);
		w.area_freed.push_back(0);
	}

	// a state handed over by the other side: transforms and velocities of the moving bodies
	if (state_path) {
		FILE *stf = fopen(state_path, "rb");
		uint32_t ns = 0;
		if (!stf || fread(&ns, 4, 1, stf) != 1 || ns != hd.n_bodies) {
			fprintf(stderr, "cannot read %s (or its body count differs from the scene's)\n", state_path);
			return 2;
		}
		std::vector<float> stv((size_t)ns * 18);
		if (fread(stv.data(), 4, stv.size(), stf) != stv.size()) {
			fprintf(stderr, "short state file\n");
			return 2;
		}
		fclose(stf);
		for (uint32_t i = 0; i < ns; i++) {
			if (sbodies[i].mode < PhysicsServer3D::BODY_MODE_RIGID) {
				continue;
			}
			const float *f = &stv[(size_t)i * 18];
			ps->body_set_state(w.bodies[i], PhysicsServer3D::BODY_STATE_TRANSFORM, xf_from(f));
			ps->body_set_state(w.bodies[i], PhysicsServer3D::BODY_STATE_LINEAR_VELOCITY, Vector3(f[12], f[13], f[14]));
			ps->body_set_state(w.bodies[i], PhysicsServer3D::BODY_STATE_ANGULAR_VELOCITY, Vector3(f[15], f[16], f[17]));
		}
	}
	std::vector<Event> events;
	if (events_path) {
		FILE *ef = fopen(events_path, "rb");
		uint32_t ne = 0;
		if (!ef || fread(&ne, 4, 1, ef) != 1) {
			fprintf(stderr, "cannot read %s\n", events_path);
			return 2;
		}
		events.resize(ne);
		if (ne && fread(events.data(), sizeof(Event), ne, ef) != ne) {
			fprintf(stderr, "short events file\n");
			return 2;
		}
		fclose(ef);
	}

	Dump d;
	if (dump_path) {
		d.f = fopen(dump_path, "wb");
		fwrite("BDMP", 1, 4, d.f);
	}

	// Step 0 dump = state after construction + mass properties (space->setup() runs inside the first step,
	// so force it here to expose inv_mass / inertia; it is idempotent: bradot_space_3d.cpp:1178-1184).
	for (RID sp : w.spaces) {
		ps->space_owner.get_or_null(sp)->setup();
	}
	if (d.f) {
		dump_meshes(d, w);
		if (dump_steps.empty() ? dump_from <= 0 : dump_steps.count(0) > 0) {
			dump_step(d, w, 0);
		}
	}

	for (int s = 0; s < warmup; s++) {
		ps->step(hd.dt);
	}
	const bool record = d.f != nullptr;
	double phase_us[5] = { 0, 0, 0, 0, 0 };
	auto t0 = std::chrono::steady_clock::now();
	for (int s = 1; s <= steps; s++) {
		for (const Event &e : events) {
			if (e.step == s && (e.kind == 6 || e.kind == 7)) {
				if (e.body >= 0 && e.body < (int)w.areas.size() && !w.area_freed[e.body]) {
					if (e.kind == 6) {
						ps->area_set_transform(w.areas[e.body], xf_from(e.v));
					} else {
						ps->free(w.areas[e.body]);
						w.area_freed[e.body] = 1;
					}
				}
				continue;
			}
			if (e.step != s || e.body < 0 || e.body >= (int)w.bodies.size() || w.body_freed[e.body]) {
				continue;
			}
			RID body = w.bodies[e.body];
			switch (e.kind) {
				case 0:
					ps->body_set_state(body, PhysicsServer3D::BODY_STATE_TRANSFORM, xf_from(e.v));
					break;
				case 1:
					ps->body_set_state(body, PhysicsServer3D::BODY_STATE_LINEAR_VELOCITY, Vector3(e.v[0], e.v[1], e.v[2]));
					break;
				case 2:
					ps->body_set_state(body, PhysicsServer3D::BODY_STATE_ANGULAR_VELOCITY, Vector3(e.v[0], e.v[1], e.v[2]));
					break;
				case 3:
					ps->free(body);
					w.body_freed[e.body] = 1;
					break;
				case 4:
					ps->body_set_state(body, PhysicsServer3D::BODY_STATE_SLEEPING, e.v[0] != 0.f);
					break;
				case 5:
					ps->body_set_state(body, PhysicsServer3D::BODY_STATE_CAN_SLEEP, e.v[0] != 0.f);
					break;
				case 8:
					ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_INERTIA, Vector3(e.v[0], e.v[1], e.v[2]));
					break;
				case 9:
					ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_CENTER_OF_MASS, Vector3(e.v[0], e.v[1], e.v[2]));
					break;
				case 10:
					ps->body_reset_mass_properties(body);
					break;
				case 11:
					ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_MASS, e.v[0]);
					break;
				case 12:
					ps->body_set_collision_priority(body, e.v[0]);
					break;
			}
		}
		drive_kinematic(d, w, s, hd.dt, true);
		push_bodies(w, w.body_flags);
		const bool dump_this = dump_steps.empty() ? (s >= dump_from && ((s - dump_from) % dump_every) == 0) : dump_steps.count(s) > 0;
		step_world(w, hd.dt, record && dump_this);
		for (RID sp : w.spaces) {
			BradotSpace3D *space = ps->space_owner.get_or_null(sp);
			for (int k = 0; k < 5; k++) {
				phase_us[k] += space->get_elapsed_time((BradotSpace3D::ElapsedTime)k);
			}
		}
		if (!w.areas.empty()) {
			ps->flush_queries(); // the areas' monitor callbacks (BradotSpace3D::call_queries, bradot_space_3d.cpp:1164-1176)
			if (d.f) { // every step that had events, whether it is a dump step or not
				if (!g_area_events.empty()) {
					d.arr("area_events", s, "i4", { (uint64_t)g_area_events.size() / 3, 3 }, g_area_events.data(), g_area_events.size() * 4);
				}
			}
			g_area_events.clear();
		}
		if (d.f && dump_this) {
			dump_step(d, w, s);
			dump_rays(d, w, s, rays);
			if (motions) {
				dump_motions(d, w, s);
			}
		}
	}
	auto t1 = std::chrono::steady_clock::now();
	double wall = std::chrono::duration<double>(t1 - t0).count();
	if (d.f) {
		fclose(d.f);
	}
	uint64_t active = 0;
	for (size_t i = 0; i < w.body_ptr.size(); i++) {
		if (!w.body_freed[i] && w.body_ptr[i]->get_mode() >= PhysicsServer3D::BODY_MODE_KINEMATIC) {
			active++;
		}
	}
	printf("{\"steps\": %d, \"warmup\": %d, \"wall_s\": %.6f, \"bodies\": %u, \"dynamic_bodies\": %llu, \"threads\": %d, "
		   "\"collision_pairs\": %d, \"island_count\": %d, \"active_objects\": %d, "
		   "\"phase_us\": {\"integrate_forces\": %.0f, \"generate_islands\": %.0f, \"setup_constraints\": %.0f, "
		   "\"solve_constraints\": %.0f, \"integrate_velocities\": %.0f}}\n",
			steps, warmup, wall, hd.n_bodies, (unsigned long long)active, WorkerThreadPool::get_singleton()->get_thread_count(),
			ps->get_process_info(PhysicsServer3D::INFO_COLLISION_PAIRS), ps->get_process_info(PhysicsServer3D::INFO_ISLAND_COUNT),
			ps->get_process_info(PhysicsServer3D::INFO_ACTIVE_OBJECTS),
			phase_us[0], phase_us[1], phase_us[2], phase_us[3], phase_us[4]);
	fflush(stdout);
	// Skip the engine teardown: process exit reclaims everything and the reference's
	// unregister path expects the full Main lifecycle.
	_exit(0);
}
