// engine_demo -- proves the drop-in path end to end: brings up the engine core (the reference's own objects,
// linked from oracle/_ref/libbradot_ref.a), registers "BradotPhysicsCUDA" with PhysicsServer3DManager, asks the
// manager for that server BY NAME, builds a scene file through the public PhysicsServer3D RID API and steps it.
// Every body gets a state-sync callback (what RigidBody3D registers, scene/3d/physics/rigid_body_3d.cpp:822); a script of
// body_set_state / free events (same records as oracle/ref_harness.cpp --events) can be applied through the RID API.
// Body transforms after the last step -- as body_get_state returns them and as the callbacks last reported them -- go to
// stdout as JSON for comparison with a direct C-ABI run and with the reference's run of the same script.
// Host code only (no CUDA here): it reaches the GPU solely through libbradot_physics_cuda.so.
#include "core/config/engine.h"
#include "core/config/project_settings.h"
#include "core/io/file_access_pack.h"
#include "core/object/worker_thread_pool.h"
#include "core/register_core_types.h"
#include "drivers/unix/os_unix.h"
#include "servers/physics_server_3d.h"

#include "bradot_physics_cuda/register_types.h"

#include <unistd.h>
#include <cstdio>
#include <cstring>
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

#pragma pack(push, 1)
struct ScnHeader { char magic[4]; uint32_t version, n_spaces, n_shapes, n_bodies, n_joints, n_verts; float dt; uint32_t iterations, iter_step, n_areas, n_joint_ext, reserved[4]; };
struct ScnJointExt { uint32_t joint, first, count, pad; float values[28]; };
struct ScnShape { uint32_t type; float p[4]; uint32_t vert_ofs, vert_cnt; float custom_bias; };
struct ScnBody { uint32_t space, mode, shape, flags; float xf[12], shape_xf[12], linvel[3], angvel[3], mass, friction, bounce, gravity_scale, lin_damp, ang_damp; uint32_t layer, mask; };
struct ScnJoint { uint32_t type, body_a, body_b, flags; float fa[12], fb[12], params[8]; uint32_t priority, pad[3]; };
struct Event { int32_t step, body, kind; float v[12]; }; // kinds 6 / 7 address an area: area_set_transform / free
struct ScnArea { int32_t space, shape; uint32_t flags; int32_t priority; float xf[12], shape_xf[12]; uint32_t layer, mask; int32_t gravity_mode, linear_damp_mode, angular_damp_mode;
	float gravity, gravity_vector[3], point_unit_distance, linear_damp, angular_damp; uint32_t group; };
#pragma pack(pop)

// what the areas' monitor callbacks were told: (step whose results were flushed, area, body, status)
static std::vector<int> g_area_events;
static std::vector<RID> *g_body_rids = nullptr;
static int g_flushed_step = 0;
static void on_area_monitor(int p_status, const RID &p_body, ObjectID p_instance, int p_body_shape, int p_area_shape, int p_area) {
	int idx = -1;
	for (size_t k = 0; g_body_rids && k < g_body_rids->size(); k++) {
		if ((*g_body_rids)[k] == p_body) {
			idx = (int)k;
		}
	}
	if (idx >= 0 && (uint64_t)p_instance != (uint64_t)(1000 + idx)) {
		idx = -2; // the object id handed over is not the one attached to that body
	}
	g_area_events.push_back(g_flushed_step);
	g_area_events.push_back(p_area);
	g_area_events.push_back(idx);
	g_area_events.push_back(p_status);
}

static std::vector<RID> *g_area_rids = nullptr;
static void on_area_area_monitor(int p_status, const RID &p_other, ObjectID p_instance, int p_other_shape, int p_area_shape, int p_area) {
	int idx = -1;
	for (size_t k = 0; g_area_rids && k < g_area_rids->size(); k++) {
		if ((*g_area_rids)[k] == p_other) {
			idx = (int)k;
		}
	}
	g_area_events.push_back(g_flushed_step);
	g_area_events.push_back(p_area);
	g_area_events.push_back(idx);
	g_area_events.push_back(2 + p_status);
}

// what the state-sync callbacks reported: last transform per body, number of calls
static std::vector<Transform3D> g_synced;
static std::vector<int> g_sync_calls;
// ... and the contacts its direct state listed at that call (bodies with max_contacts_reported > 0)
struct SeenContact { RID collider; ObjectID collider_id; Vector3 pos, normal, impulse, vel, collider_pos, collider_vel; };
static std::vector<std::vector<SeenContact>> g_contacts;
static void on_state_sync(Object *p_state, int p_index) {
	PhysicsDirectBodyState3D *st = Object::cast_to<PhysicsDirectBodyState3D>(p_state);
	if (st && p_index >= 0 && p_index < (int)g_synced.size()) {
		g_synced[p_index] = st->get_transform();
		g_sync_calls[p_index]++;
		g_contacts[p_index].clear();
		for (int i = 0; i < st->get_contact_count(); i++) {
			g_contacts[p_index].push_back({ st->get_contact_collider(i), st->get_contact_collider_id(i), st->get_contact_local_position(i), st->get_contact_local_normal(i),
					st->get_contact_impulse(i), st->get_contact_local_velocity_at_position(i), st->get_contact_collider_position(i), st->get_contact_collider_velocity_at_position(i) });
		}
	}
}

static Transform3D xf_from(const float *f) {
	return Transform3D(Basis(f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8]), Vector3(f[9], f[10], f[11]));
}

int main(int argc, char **argv) {
	if (argc < 3) {
		fprintf(stderr, "usage: engine_demo scene.bscn steps [--events events.bin] [--drive-kinematic]\n");
		return 2;
	}
	int steps = atoi(argv[2]);
	std::vector<Event> events;
	// --drive-kinematic: kinematic bodies are moved by body_set_state(TRANSFORM) before every step, their scene velocities being
	// the script (rotation about y by angvel.y t, origin + linvel t), exactly as oracle/ref_harness.cpp drives them
	bool drive_kinematic = false;
	for (int i = 3; i < argc; i++) {
		drive_kinematic = drive_kinematic || strcmp(argv[i], "--drive-kinematic") == 0;
	}
	for (int i = 3; i + 1 < argc; i++) {
		if (strcmp(argv[i], "--events") == 0) {
			FILE *ef = fopen(argv[i + 1], "rb");
			uint32_t ne = 0;
			if (!ef || fread(&ne, 4, 1, ef) != 1) {
				fprintf(stderr, "cannot read events\n");
				return 2;
			}
			events.resize(ne);
			if (ne && fread(events.data(), sizeof(Event), ne, ef) != ne) {
				return 2;
			}
			fclose(ef);
		}
	}
	FILE *sf = fopen(argv[1], "rb");
	ScnHeader hd;
	if (!sf || fread(&hd, sizeof(hd), 1, sf) != 1 || memcmp(hd.magic, "BSCN", 4) != 0 || hd.version != 2) {
		fprintf(stderr, "bad scene file\n");
		return 2;
	}
	std::vector<ScnShape> shapes(hd.n_shapes);
	std::vector<ScnBody> bodies(hd.n_bodies);
	std::vector<ScnJoint> joints(hd.n_joints);
	std::vector<float> verts(hd.n_verts * 3);
	if (hd.n_shapes && fread(shapes.data(), sizeof(ScnShape), hd.n_shapes, sf) != hd.n_shapes) return 2;
	if (hd.n_bodies && fread(bodies.data(), sizeof(ScnBody), hd.n_bodies, sf) != hd.n_bodies) return 2;
	if (hd.n_joints && fread(joints.data(), sizeof(ScnJoint), hd.n_joints, sf) != hd.n_joints) return 2;
	if (hd.n_verts && fread(verts.data(), 12, hd.n_verts, sf) != hd.n_verts) return 2;
	std::vector<ScnArea> sareas(hd.n_areas);
	if (hd.n_areas && fread(sareas.data(), sizeof(ScnArea), hd.n_areas, sf) != hd.n_areas) return 2;
	std::vector<ScnJointExt> sjext(hd.n_joint_ext);
	if (hd.n_joint_ext && fread(sjext.data(), sizeof(ScnJointExt), hd.n_joint_ext, sf) != hd.n_joint_ext) return 2;
	fclose(sf);

	OSHeadless os;
	Thread::make_main_thread();
	set_current_thread_safe_for_nodes(true);
	os.initialize();
	memnew(Engine);
	register_core_types();
	memnew(PackedData);
	memnew(ProjectSettings);
	register_core_settings();
	WorkerThreadPool::get_singleton()->init(-1, 0.3);

	// what Main::setup / initialize_physics do (main/main.cpp:319-337): registry, module init, server by name
	PhysicsServer3DManager *manager = memnew(PhysicsServer3DManager);
	initialize_bradot_physics_cuda_module(MODULE_INITIALIZATION_LEVEL_SERVERS);
	PhysicsServer3D *ps = manager->new_server("BradotPhysicsCUDA");
	if (!ps) {
		fprintf(stderr, "PhysicsServer3DManager::new_server(\"BradotPhysicsCUDA\") returned null\n");
		return 1;
	}
	ps->init();

	std::vector<RID> space_r, shape_r, body_r;
	for (uint32_t i = 0; i < hd.n_spaces; i++) {
		RID sp = ps->space_create();
		ps->space_set_active(sp, true);
		ps->space_set_param(sp, PhysicsServer3D::SPACE_PARAM_SOLVER_ITERATIONS, hd.iterations + i * hd.iter_step);
		space_r.push_back(sp);
	}
	for (const ScnShape &s : shapes) {
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
				Dictionary d;
				d["radius"] = s.p[0];
				d["height"] = s.p[1];
				ps->shape_set_data(sh, d);
			} break;
			case PhysicsServer3D::SHAPE_CONVEX_POLYGON: {
				sh = ps->convex_polygon_shape_create();
				Vector<Vector3> pts;
				for (uint32_t k = 0; k < s.vert_cnt; k++) {
					const float *v = &verts[(s.vert_ofs + k) * 3];
					pts.push_back(Vector3(v[0], v[1], v[2]));
				}
				ps->shape_set_data(sh, pts);
			} break;
			case PhysicsServer3D::SHAPE_CONCAVE_POLYGON: {
				sh = ps->concave_polygon_shape_create();
				Vector<Vector3> pts;
				for (uint32_t k = 0; k < s.vert_cnt; k++) {
					const float *v = &verts[(s.vert_ofs + k) * 3];
					pts.push_back(Vector3(v[0], v[1], v[2]));
				}
				Dictionary d;
				d["faces"] = pts;
				d["backface_collision"] = s.p[0] != 0.f;
				ps->shape_set_data(sh, d);
			} break;
			default:
				fprintf(stderr, "unsupported shape\n");
				return 2;
		}
		if (s.custom_bias != 0.0f) {
			ps->shape_set_custom_solver_bias(sh, s.custom_bias);
		}
		shape_r.push_back(sh);
	}
	std::vector<int> kin_body;
	std::vector<Transform3D> kin_xf0;
	std::vector<Vector3> kin_lin;
	std::vector<float> kin_wy;
	for (const ScnBody &b : bodies) {
		if ((b.flags & 0x4000) && !body_r.empty()) { // bit14: a further shape of the (static) body before it
			ps->body_add_shape(body_r.back(), shape_r[b.shape], xf_from(b.shape_xf), false);
			body_r.push_back(body_r.back());
			continue;
		}
		RID body = ps->body_create();
		ps->body_set_mode(body, (PhysicsServer3D::BodyMode)b.mode);
		ps->body_add_shape(body, shape_r[b.shape], xf_from(b.shape_xf), false);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_MASS, b.mass);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_FRICTION, b.friction);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_BOUNCE, b.bounce);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_GRAVITY_SCALE, b.gravity_scale);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_LINEAR_DAMP, b.lin_damp);
		ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_ANGULAR_DAMP, b.ang_damp);
		ps->body_set_collision_layer(body, b.layer);
		ps->body_set_collision_mask(body, b.mask);
		ps->body_set_state(body, PhysicsServer3D::BODY_STATE_TRANSFORM, xf_from(b.xf));
		if (drive_kinematic && b.mode == (uint32_t)PhysicsServer3D::BODY_MODE_KINEMATIC) {
			kin_body.push_back((int)body_r.size());
			kin_xf0.push_back(xf_from(b.xf));
			kin_lin.push_back(Vector3(b.linvel[0], b.linvel[1], b.linvel[2]));
			kin_wy.push_back(b.angvel[1]);
		} else {
			ps->body_set_state(body, PhysicsServer3D::BODY_STATE_LINEAR_VELOCITY, Vector3(b.linvel[0], b.linvel[1], b.linvel[2]));
			ps->body_set_state(body, PhysicsServer3D::BODY_STATE_ANGULAR_VELOCITY, Vector3(b.angvel[0], b.angvel[1], b.angvel[2]));
		}
		ps->body_set_state(body, PhysicsServer3D::BODY_STATE_CAN_SLEEP, (b.flags & 1) != 0);
		if (b.flags & 2) {
			ps->body_set_enable_continuous_collision_detection(body, true);
		}
		if (b.flags & 0x8000) {
			ps->body_set_shape_disabled(body, 0, true);
		}
		if (b.flags & 128) {
			ps->body_set_ray_pickable(body, false);
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
		if ((b.flags >> 16) & 0xFF) {
			ps->body_set_max_contacts_reported(body, (b.flags >> 16) & 0xFF);
		}
		ps->body_attach_object_instance_id(body, ObjectID((uint64_t)(1000 + body_r.size())));
		ps->body_set_space(body, space_r[b.space]);
		ps->body_set_state_sync_callback(body, callable_mp_static(&on_state_sync).bind((int)body_r.size()));
		body_r.push_back(body);
	}
	g_synced.resize(body_r.size());
	g_contacts.resize(body_r.size());
	g_sync_calls.assign(body_r.size(), 0);
	std::vector<char> freed(body_r.size(), 0);
	// scene flag bit3: pushed before every step, the same push as oracle/ref_harness.cpp push_bodies()
	auto push = [&]() {
		for (size_t i = 0; i < body_r.size(); i++) {
			if ((bodies[i].flags & 8) && !freed[i]) {
				ps->body_apply_central_force(body_r[i], Vector3(0.5f * (float)((int)(i % 3) - 1), 2.0f, 0.3f));
				ps->body_apply_torque(body_r[i], Vector3(0.1f, 0.0f, 0.05f * (float)(i % 2)));
			}
		}
	};
	std::vector<RID> joint_r;
	for (const ScnJoint &j : joints) {
		RID jr = ps->joint_create();
		if (j.type == 0) {
			ps->joint_make_pin(jr, body_r[j.body_a], Vector3(j.fa[9], j.fa[10], j.fa[11]), body_r[j.body_b], Vector3(j.fb[9], j.fb[10], j.fb[11]));
			if (j.flags & 8) {
				for (int k = 0; k < 3; k++) {
					ps->pin_joint_set_param(jr, (PhysicsServer3D::PinJointParam)k, j.params[k]);
				}
			}
		} else if (j.type == 3) {
			ps->joint_make_slider(jr, body_r[j.body_a], xf_from(j.fa), body_r[j.body_b], xf_from(j.fb));
			for (const ScnJointExt &e : sjext) {
				for (uint32_t k = 0; e.joint == (uint32_t)joint_r.size() && k < e.count; k++) {
					ps->slider_joint_set_param(jr, (PhysicsServer3D::SliderJointParam)(e.first + k), e.values[k]);
				}
			}
		} else if (j.type == 4) {
			ps->joint_make_generic_6dof(jr, body_r[j.body_a], xf_from(j.fa), body_r[j.body_b], xf_from(j.fb));
			for (const ScnJointExt &e : sjext) {
				for (uint32_t k = 0; e.joint == (uint32_t)joint_r.size() && k < e.count; k++) {
					uint32_t idx = e.first + k, axis = idx / 28, prm = idx % 28;
					if (prm < (uint32_t)PhysicsServer3D::G6DOF_JOINT_MAX) {
						if (prm != (uint32_t)PhysicsServer3D::G6DOF_JOINT_ANGULAR_MOTOR_FORCE_LIMIT) {
							ps->generic_6dof_joint_set_param(jr, (Vector3::Axis)axis, (PhysicsServer3D::G6DOFJointAxisParam)prm, e.values[k]);
						}
					} else {
						ps->generic_6dof_joint_set_flag(jr, (Vector3::Axis)axis, (PhysicsServer3D::G6DOFJointAxisFlag)(prm - 22), e.values[k] != 0.0f);
					}
				}
			}
		} else if (j.type == 2) {
			ps->joint_make_cone_twist(jr, body_r[j.body_a], xf_from(j.fa), body_r[j.body_b], xf_from(j.fb));
			if (j.flags & 8) {
				for (int k = 0; k < (int)PhysicsServer3D::CONE_TWIST_MAX; k++) {
					ps->cone_twist_joint_set_param(jr, (PhysicsServer3D::ConeTwistJointParam)k, j.params[k]);
				}
			}
		} else {
			if (j.flags & 16) { // joint_make_hinge_simple: pivots = the frame origins, axes = the z columns of the frames
				ps->joint_make_hinge_simple(jr, body_r[j.body_a], Vector3(j.fa[9], j.fa[10], j.fa[11]), Vector3(j.fa[2], j.fa[5], j.fa[8]), body_r[j.body_b],
						Vector3(j.fb[9], j.fb[10], j.fb[11]), Vector3(j.fb[2], j.fb[5], j.fb[8]));
			} else {
				ps->joint_make_hinge(jr, body_r[j.body_a], xf_from(j.fa), body_r[j.body_b], xf_from(j.fb));
			}
			if (j.flags & 8) {
				for (int k = 0; k < (int)PhysicsServer3D::HINGE_JOINT_MAX; k++) {
					ps->hinge_joint_set_param(jr, (PhysicsServer3D::HingeJointParam)k, j.params[k]);
				}
			}
			ps->hinge_joint_set_flag(jr, PhysicsServer3D::HINGE_JOINT_FLAG_USE_LIMIT, (j.flags & 2) != 0);
			ps->hinge_joint_set_flag(jr, PhysicsServer3D::HINGE_JOINT_FLAG_ENABLE_MOTOR, (j.flags & 4) != 0);
		}
		if (j.flags & 1) {
			ps->joint_disable_collisions_between_bodies(jr, true);
		}
		if (j.priority) {
			ps->joint_set_solver_priority(jr, (int)j.priority);
		}
		joint_r.push_back(jr);
	}
	// areas through the RID API, every monitoring one with a callback
	std::vector<RID> area_r;
	std::vector<char> area_freed;
	g_body_rids = &body_r;
	g_area_rids = &area_r;
	for (uint32_t i = 0; i < hd.n_areas; i++) {
		const ScnArea &a = sareas[i];
		if (i > 0 && a.group != 0 && sareas[i - 1].group == a.group) { // a further shape of the area before it
			ps->area_add_shape(area_r.back(), shape_r[a.shape], xf_from(a.shape_xf));
			area_r.push_back(area_r.back());
			continue;
		}
		RID ar = ps->area_create();
		ps->area_add_shape(ar, shape_r[a.shape], xf_from(a.shape_xf));
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
		if (a.flags & 16) {
			ps->area_set_ray_pickable(ar, true);
		}
		if (a.flags & 2) {
			ps->area_set_monitor_callback(ar, callable_mp_static(&on_area_monitor).bind((int)i));
		}
		if (a.flags & 8) {
			ps->area_set_area_monitor_callback(ar, callable_mp_static(&on_area_area_monitor).bind((int)i));
		}
		ps->area_set_space(ar, space_r[a.space]);
		area_r.push_back(ar);
		area_freed.push_back(0);
	}
	// the engine's tick (main/main.cpp:4368-4397): sync -> flush_queries (callbacks) -> [user code] -> end_sync -> step
	for (int s = 1; s <= steps; s++) {
		ps->sync();
		g_flushed_step = s - 1;
		ps->flush_queries();
		for (const Event &e : events) { // "user code": the script, through the RID API
			if (e.step == s && (e.kind == 6 || e.kind == 7)) {
				if (e.body >= 0 && e.body < (int)area_r.size() && !area_freed[e.body]) {
					if (e.kind == 6) {
						ps->area_set_transform(area_r[e.body], xf_from(e.v));
					} else {
						ps->free(area_r[e.body]);
						area_freed[e.body] = 1;
					}
				}
				continue;
			}
			if (e.step != s || e.body < 0 || e.body >= (int)body_r.size() || freed[e.body]) {
				continue;
			}
			RID body = body_r[e.body];
			switch (e.kind) {
				case 0: ps->body_set_state(body, PhysicsServer3D::BODY_STATE_TRANSFORM, xf_from(e.v)); break;
				case 1: ps->body_set_state(body, PhysicsServer3D::BODY_STATE_LINEAR_VELOCITY, Vector3(e.v[0], e.v[1], e.v[2])); break;
				case 2: ps->body_set_state(body, PhysicsServer3D::BODY_STATE_ANGULAR_VELOCITY, Vector3(e.v[0], e.v[1], e.v[2])); break;
				case 3:
					ps->free(body);
					freed[e.body] = 1;
					break;
				case 4: ps->body_set_state(body, PhysicsServer3D::BODY_STATE_SLEEPING, e.v[0] != 0.f); break;
				case 5: ps->body_set_state(body, PhysicsServer3D::BODY_STATE_CAN_SLEEP, e.v[0] != 0.f); break;
				case 8: ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_INERTIA, Vector3(e.v[0], e.v[1], e.v[2])); break;
				case 9: ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_CENTER_OF_MASS, Vector3(e.v[0], e.v[1], e.v[2])); break;
				case 10: ps->body_reset_mass_properties(body); break;
				case 11: ps->body_set_param(body, PhysicsServer3D::BODY_PARAM_MASS, e.v[0]); break;
				case 12: ps->body_set_collision_priority(body, e.v[0]); break;
			}
		}
		for (size_t k = 0; k < kin_body.size(); k++) {
			if (freed[kin_body[k]]) {
				continue;
			}
			real_t t = (real_t)s * hd.dt;
			Transform3D x;
			x.basis = Basis(Vector3(0, 1, 0), kin_wy[k] * t) * kin_xf0[k].basis;
			x.origin = kin_xf0[k].origin + kin_lin[k] * t;
			ps->body_set_state(body_r[kin_body[k]], PhysicsServer3D::BODY_STATE_TRANSFORM, x);
		}
		ps->end_sync();
		push();
		ps->step(hd.dt);
	}
	ps->sync();
	g_flushed_step = steps;
	ps->flush_queries(); // the callbacks of the last step
	ps->end_sync();
	ps->sync();
	printf("{\"server\": \"BradotPhysicsCUDA\", \"steps\": %d, \"active_objects\": %d, \"collision_pairs\": %d, \"island_count\": %d, \"origins\": [",
			steps, ps->get_process_info(PhysicsServer3D::INFO_ACTIVE_OBJECTS), ps->get_process_info(PhysicsServer3D::INFO_COLLISION_PAIRS),
			ps->get_process_info(PhysicsServer3D::INFO_ISLAND_COUNT));
	for (size_t i = 0; i < body_r.size(); i++) {
		Transform3D t = freed[i] ? Transform3D() : Transform3D(ps->body_get_state(body_r[i], PhysicsServer3D::BODY_STATE_TRANSFORM));
		printf("%s[%.9g, %.9g, %.9g]", i ? ", " : "", t.origin.x, t.origin.y, t.origin.z);
	}
	printf("], \"synced\": [");
	for (size_t i = 0; i < body_r.size(); i++) { // last transform each body's state-sync callback reported, number of calls
		const Transform3D &t = g_synced[i];
		printf("%s[%d, %.9g, %.9g, %.9g]", i ? ", " : "", freed[i] ? -1 : g_sync_calls[i], t.origin.x, t.origin.y, t.origin.z);
	}
	printf("], \"sleeping\": [");
	for (size_t i = 0; i < body_r.size(); i++) {
		printf("%s%d", i ? ", " : "", freed[i] ? -1 : (int)(bool)ps->body_get_state(body_r[i], PhysicsServer3D::BODY_STATE_SLEEPING));
	}
	printf("], \"contacts\": [");
	for (size_t i = 0; i < body_r.size(); i++) { // per body: [collider index, collider object id, 18 floats] of every contact last listed
		printf("%s[", i ? ", " : "");
		for (size_t k = 0; k < g_contacts[i].size(); k++) {
			const SeenContact &c = g_contacts[i][k];
			int idx = -1;
			for (size_t q = 0; q < body_r.size(); q++) {
				if (body_r[q] == c.collider) {
					idx = (int)q;
					break;
				}
			}
			const Vector3 *v[6] = { &c.pos, &c.normal, &c.impulse, &c.vel, &c.collider_pos, &c.collider_vel };
			printf("%s[%d, %lld", k ? ", " : "", idx, (long long)(uint64_t)c.collider_id);
			for (int q = 0; q < 6; q++) {
				printf(", %.9g, %.9g, %.9g", v[q]->x, v[q]->y, v[q]->z);
			}
			printf("]");
		}
		printf("]");
	}
	printf("], \"rays\": [");
	// PhysicsDirectSpaceState3D::intersect_ray through the RID API: straight down onto the first bodies
	PhysicsDirectSpaceState3D *ds = ps->space_get_direct_state(space_r[0]);
	int n_rays = body_r.size() < 16 ? (int)body_r.size() : 16;
	for (int i = 0; ds && i < n_rays; i++) {
		Transform3D t = freed[i] ? Transform3D() : Transform3D(ps->body_get_state(body_r[i], PhysicsServer3D::BODY_STATE_TRANSFORM));
		PhysicsDirectSpaceState3D::RayParameters rp;
		rp.from = Vector3(t.origin.x, 30.0, t.origin.z);
		rp.to = Vector3(t.origin.x, -1.0, t.origin.z);
		PhysicsDirectSpaceState3D::RayResult rr;
		bool hit = ds->intersect_ray(rp, rr);
		int idx = -1;
		for (size_t k = 0; hit && k < body_r.size(); k++) {
			if (body_r[k] == rr.rid) {
				idx = (int)k;
				break;
			}
		}
		printf("%s[%d, %d, %.9g, %.9g, %.9g]", i ? ", " : "", hit ? 1 : 0, idx, hit ? rr.position.x : 0.0, hit ? rr.position.y : 0.0, hit ? rr.position.z : 0.0);
	}
	// get_closest_point_to_object_volume through the RID API: a point beside each of the first bodies
	printf("], \"closest\": [");
	for (int i = 0; ds && i < n_rays; i++) {
		Transform3D t = freed[i] ? Transform3D() : Transform3D(ps->body_get_state(body_r[i], PhysicsServer3D::BODY_STATE_TRANSFORM));
		Vector3 c = freed[i] ? Vector3() : ds->get_closest_point_to_object_volume(body_r[i], t.origin + Vector3(0.6f, 0.4f, -0.3f));
		printf("%s[%.9g, %.9g, %.9g]", i ? ", " : "", c.x, c.y, c.z);
	}
	// cast_motion and rest_info through the RID API: a sphere dropped from above each of the first bodies with that body excluded
	// (so it lands on whatever lies under it), and the same sphere resting at the body's own place
	printf("], \"area_events\": [");
	for (size_t i = 0; i + 3 < g_area_events.size(); i += 4) {
		printf("%s[%d, %d, %d, %d]", i ? ", " : "", g_area_events[i], g_area_events[i + 1], g_area_events[i + 2], g_area_events[i + 3]);
	}
	printf("], \"casts\": [");
	RID probe = ps->sphere_shape_create();
	ps->shape_set_data(probe, 0.3f);
	for (int i = 0; ds && i < n_rays; i++) {
		Transform3D t = freed[i] ? Transform3D() : Transform3D(ps->body_get_state(body_r[i], PhysicsServer3D::BODY_STATE_TRANSFORM));
		PhysicsDirectSpaceState3D::ShapeParameters sp;
		sp.shape_rid = probe;
		sp.transform = Transform3D(Basis(), Vector3(t.origin.x, 30.0, t.origin.z));
		sp.motion = Vector3(0, -31.0, 0);
		if (!freed[i]) {
			sp.exclude.insert(body_r[i]);
		}
		real_t safe = 1, unsafe = 1;
		PhysicsDirectSpaceState3D::ShapeRestInfo ci;
		ds->cast_motion(sp, safe, unsafe, &ci);
		int idx = -1, ridx = -1;
		for (size_t k = 0; k < body_r.size(); k++) {
			if (ci.rid.is_valid() && body_r[k] == ci.rid) {
				idx = (int)k;
			}
		}
		PhysicsDirectSpaceState3D::ShapeParameters rp;
		rp.shape_rid = probe;
		rp.transform = Transform3D(Basis(), t.origin);
		rp.margin = 0.05;
		PhysicsDirectSpaceState3D::ShapeRestInfo ri;
		bool rest = ds->rest_info(rp, &ri);
		for (size_t k = 0; rest && k < body_r.size(); k++) {
			if (body_r[k] == ri.rid) {
				ridx = (int)k;
			}
		}
		printf("%s[%.9g, %.9g, %d, %d, %.9g, %.9g, %.9g]", i ? ", " : "", (double)safe, (double)unsafe, idx, ridx, rest ? ri.point.x : 0.0, rest ? ri.point.y : 0.0, rest ? ri.point.z : 0.0);
	}
	// body_test_motion through the RID API: each of the first bodies pushed 0.4 m down from where it stands
	printf("], \"motions\": [");
	for (int i = 0; i < n_rays; i++) {
		if (freed[i]) {
			printf("%s[0, -1, 0, 0, 0, 0]", i ? ", " : "");
			continue;
		}
		PhysicsServer3D::MotionParameters mp(Transform3D(ps->body_get_state(body_r[i], PhysicsServer3D::BODY_STATE_TRANSFORM)), Vector3(0.05, -0.4, 0.0), 0.02);
		mp.max_collisions = 4;
		PhysicsServer3D::MotionResult mr;
		bool hit = ps->body_test_motion(body_r[i], mp, &mr);
		int idx = -1;
		for (size_t k = 0; hit && k < body_r.size(); k++) {
			if (body_r[k] == mr.collisions[0].collider) {
				idx = (int)k;
			}
		}
		printf("%s[%d, %d, %d, %.9g, %.9g, %.9g]", i ? ", " : "", hit ? 1 : 0, idx, mr.collision_count, (double)mr.collision_safe_fraction, (double)mr.travel.x, (double)mr.travel.y);
	}
	printf("]}\n");
	fflush(stdout);
	ps->finish();
	_exit(0);
}
