// Engine-side backend: a PhysicsServer3D that forwards the rigid-body path to libbradot_physics_cuda.so
// through the C ABI (include/bradot_physics_cuda.h).  Drop this directory into the engine's modules/ tree;
// register_types.cpp registers it as "BradotPhysicsCUDA" with PhysicsServer3DManager
// (servers/physics_server_3d.cpp:1170-1175), selectable through physics/3d/physics_engine.
//
// Implemented by hand: what a scene of RigidBody3D / StaticBody3D nodes needs -- spaces (create, activate, parameters, free),
// world-boundary / sphere / box / capsule / convex shapes, bodies with one shape (every setter also after the body is in a
// space; state-sync and force-integration callbacks; PhysicsDirectBodyState3D; impulses, forces, constant forces; axis
// locks; collision exceptions; free), pin, hinge, slider, cone-twist and 6-DOF joints with their parameters, flags and solver priority, step, stats,
// PhysicsDirectSpaceState3D queries.  Every other pure virtual of the interface gets a generated default override
// (engine_module/gen_stubs.py) that logs once and returns the type's default, which is how the reference answers an
// invalid RID.
#ifndef BRADOT_PHYSICS_SERVER_CUDA_H
#define BRADOT_PHYSICS_SERVER_CUDA_H

#include "core/templates/hash_map.h"
#include "core/templates/local_vector.h"
#include "core/templates/rid_owner.h"
#include "servers/physics_server_3d.h"

#include "bradot_physics_cuda.h"

#define BPC_UNSUPPORTED(m_name) WARN_PRINT_ONCE("BradotPhysicsCUDA: " m_name "() is not supported by this backend yet.")

class BradotPhysicsServerCUDA;

// PhysicsDirectSpaceState3D of one space (servers/physics_server_3d.h:123-214; reference implementation
// modules/bradot_physics_3d/bradot_space_3d.h:45-62).  intersect_ray and intersect_point run on the GPU (bpc_space_intersect_rays / _points,
// one query per call here; intersect_rays_batch answers many at once); the shape queries are "next" rows and answer "nothing".
class BradotPhysicsDirectSpaceStateCUDA : public PhysicsDirectSpaceState3D {
	BRCLASS(BradotPhysicsDirectSpaceStateCUDA, PhysicsDirectSpaceState3D);

public:
	BradotPhysicsServerCUDA *server = nullptr;
	int space_handle = -1;

	virtual bool intersect_ray(const RayParameters &p_parameters, RayResult &r_result) override;
	int intersect_rays_batch(const Vector<Vector3> &p_from_to, uint32_t p_collision_mask, bool p_hit_from_inside, LocalVector<RayResult> &r_results, LocalVector<bool> &r_hit, bool p_hit_back_faces = true, bool p_pick_ray = false, bool p_with_areas = false, bool p_with_bodies = true);
	virtual int intersect_point(const PointParameters &p_parameters, ShapeResult *r_results, int p_result_max) override;
	virtual int intersect_shape(const ShapeParameters &p_parameters, ShapeResult *r_results, int p_result_max) override;
	virtual bool cast_motion(const ShapeParameters &p_parameters, real_t &p_closest_safe, real_t &p_closest_unsafe, ShapeRestInfo *r_info = nullptr) override;
	virtual bool collide_shape(const ShapeParameters &p_parameters, Vector3 *r_results, int p_result_max, int &r_result_count) override;
	virtual bool rest_info(const ShapeParameters &p_parameters, ShapeRestInfo *r_info) override;
	void _fill_rest_info(const bpc_rest_info &p_info, ShapeRestInfo *r_info) const;
	virtual Vector3 get_closest_point_to_object_volume(RID p_object, const Vector3 p_point) const override;
};

// PhysicsDirectBodyState3D of one body (servers/physics_server_3d.h:46-115; reference implementation
// modules/bradot_physics_3d/bradot_body_direct_state_3d.cpp).  Getters answer from the state the server read back in bulk at
// the start of flush_queries(); setters go through the server's own entry points (body_set_state, body_apply_*).
class BradotPhysicsDirectBodyStateCUDA : public PhysicsDirectBodyState3D {
	BRCLASS(BradotPhysicsDirectBodyStateCUDA, PhysicsDirectBodyState3D);

public:
	BradotPhysicsServerCUDA *server = nullptr;
	RID body;

	virtual Vector3 get_total_gravity() const override;
	virtual real_t get_total_angular_damp() const override;
	virtual real_t get_total_linear_damp() const override;
	virtual Vector3 get_center_of_mass() const override;
	virtual Vector3 get_center_of_mass_local() const override;
	virtual Basis get_principal_inertia_axes() const override;
	virtual real_t get_inverse_mass() const override;
	virtual Vector3 get_inverse_inertia() const override;
	virtual Basis get_inverse_inertia_tensor() const override;
	virtual void set_linear_velocity(const Vector3 &p_velocity) override;
	virtual Vector3 get_linear_velocity() const override;
	virtual void set_angular_velocity(const Vector3 &p_velocity) override;
	virtual Vector3 get_angular_velocity() const override;
	virtual void set_transform(const Transform3D &p_transform) override;
	virtual Transform3D get_transform() const override;
	virtual Vector3 get_velocity_at_local_position(const Vector3 &p_position) const override;
	virtual void apply_central_impulse(const Vector3 &p_impulse) override;
	virtual void apply_impulse(const Vector3 &p_impulse, const Vector3 &p_position = Vector3()) override;
	virtual void apply_torque_impulse(const Vector3 &p_impulse) override;
	virtual void apply_central_force(const Vector3 &p_force) override;
	virtual void apply_force(const Vector3 &p_force, const Vector3 &p_position = Vector3()) override;
	virtual void apply_torque(const Vector3 &p_torque) override;
	virtual void add_constant_central_force(const Vector3 &p_force) override;
	virtual void add_constant_force(const Vector3 &p_force, const Vector3 &p_position = Vector3()) override;
	virtual void add_constant_torque(const Vector3 &p_torque) override;
	virtual void set_constant_force(const Vector3 &p_force) override;
	virtual Vector3 get_constant_force() const override;
	virtual void set_constant_torque(const Vector3 &p_torque) override;
	virtual Vector3 get_constant_torque() const override;
	virtual void set_sleep_state(bool p_sleep) override;
	virtual bool is_sleeping() const override;
	// the contacts BradotBody3D::add_contact kept for this body in the last step (bradot_body_direct_state_3d.cpp:205-262)
	virtual int get_contact_count() const override;
	virtual Vector3 get_contact_local_position(int p_contact_idx) const override;
	virtual Vector3 get_contact_local_normal(int p_contact_idx) const override;
	virtual Vector3 get_contact_impulse(int p_contact_idx) const override;
	virtual int get_contact_local_shape(int p_contact_idx) const override { return 0; } // one shape per body
	virtual Vector3 get_contact_local_velocity_at_position(int p_contact_idx) const override;
	virtual RID get_contact_collider(int p_contact_idx) const override;
	virtual Vector3 get_contact_collider_position(int p_contact_idx) const override;
	virtual ObjectID get_contact_collider_id(int p_contact_idx) const override;
	virtual int get_contact_collider_shape(int p_contact_idx) const override { return 0; }
	virtual Vector3 get_contact_collider_velocity_at_position(int p_contact_idx) const override;
	virtual real_t get_step() const override;
	virtual PhysicsDirectSpaceState3D *get_space_state() override;
};

class BradotPhysicsServerCUDA : public PhysicsServer3D {
	BRCLASS(BradotPhysicsServerCUDA, PhysicsServer3D);
	friend class BradotPhysicsDirectSpaceStateCUDA;
	friend class BradotPhysicsDirectBodyStateCUDA;

	struct Shape {
		ShapeType type = SHAPE_SPHERE;
		Variant data;
		real_t custom_bias = 0.0;
		int handle = -1; // bpc shape index, created lazily once the data is known
	};
	struct Space {
		int handle = -1;
		bool active = false;
		BradotPhysicsDirectSpaceStateCUDA *direct_state = nullptr;
		real_t params[8] = { 0.01, 0.05, 0.01, 0.8, 0.1, Math::deg_to_rad(8.0), 0.5, 16 };
		// the space's default area (P3D/bradot_space_3d.cpp:1220-1226; BradotArea3D defaults)
		real_t default_gravity = 9.80665, default_linear_damp = 0.1, default_angular_damp = 0.1;
		Vector3 default_gravity_vector = Vector3(0, -1, 0);
	};
	struct Body {
		RID self, space, shape;
		Transform3D shape_xf;
		bool shape_disabled = false;
		BodyMode mode = BODY_MODE_RIGID;
		Transform3D xf;
		Vector3 linvel, angvel;
		real_t mass = 1, friction = 1, bounce = 0, gravity_scale = 1, lin_damp = 0, ang_damp = 0;
		uint32_t layer = 1, mask = 1;
		bool can_sleep = true;
		bool ccd = false;
		bool ray_pickable = true;
		real_t collision_priority = 1.0;
		// body_set_param(BODY_PARAM_INERTIA / BODY_PARAM_CENTER_OF_MASS): handed over with bpc_bodies_set_mass_override before the
		// body is described (anew)
		Vector3 custom_inertia, custom_com;
		bool has_custom_inertia = false, has_custom_com = false, mass_override_dirty = false;
		bool omit_force_integration = false;
		bool lin_damp_replace = false, ang_damp_replace = false;
		uint32_t locked_axis = 0;
		int max_contacts_reported = 0;
		ObjectID instance_id;
		Vector3 applied_force, applied_torque; // accumulated by body_apply_*, handed over in one write before the next step
		Vector3 constant_force, constant_torque;
		bool pushed = false;      // in pushed_bodies
		bool dirty = false;       // described anew since the upload (mode, parameters, layers, shape ...): in dirty_bodies
		bool pending = false;     // in pending_bodies
		bool has_callbacks = false; // in callback_bodies
		Callable state_sync_callback;
		Callable fi_callable;
		Variant fi_udata;
		BradotPhysicsDirectBodyStateCUDA *direct_state = nullptr;
		LocalVector<RID> exceptions;
		// several shapes: a STATIC body's further shapes are bodies of their own on the GPU side (exactly what the reference's
		// step sees: one broadphase entry and one pair object per shape, P3D/bradot_collision_object_3d.cpp:156-184,
		// P3D/bradot_body_pair_3d.cpp) -- hidden records that take everything but shape and shape transform from `parent`
		Body *parent = nullptr;
		int shape_index = 0;
		LocalVector<Body *> parts;
		int slot = -1; // bpc body index once uploaded
	};
	struct Joint {
		JointType type = JOINT_TYPE_MAX;
		RID self, a, b;
		Transform3D fa, fb;
		bool disable_collisions = false;
		int priority = 1;
		real_t pin_params[3] = { 0.3, 1.0, 0.0 };
		// HingeJointParam order; defaults of modules/bradot_physics_3d/joints/bradot_hinge_joint_3d.h:73-88
		real_t hinge_params[HINGE_JOINT_MAX] = { 0.3, -Math_PI, Math_PI, 0.3, 0.9, 1.0, 0.0, 0.0 };
		bool hinge_flags[HINGE_JOINT_FLAG_MAX] = { false, false };
		// ConeTwistJointParam order, then the second swing span (bradot_cone_twist_joint_3d.h:80-86: tau / 8 and 0 until
		// SWING_SPAN is set, which writes both)
		real_t cone_params[CONE_TWIST_MAX + 1] = { Math_TAU / 8.0, 0.0, 0.3, 0.0, 1.0, 0.0 };
		// SliderJointParam order; the members of BradotSliderJoint3D as constructed (joints/bradot_slider_joint_3d.h:80-110)
		float slider_params[SLIDER_JOINT_MAX] = { -1.0f, 1.0f, 1.0f, 0.7f, 1.0f, 1.0f, 0.7f, 0.0f, 1.0f, 0.7f, 1.0f,
			0.0f, 0.0f, 1.0f, 0.7f, 1.0f, 1.0f, 0.7f, 0.0f, 1.0f, 0.7f, 1.0f };
		// generic 6-DOF: per axis the G6DOFJointAxisParam values then the flags, as the members of the reference's joint
		// stand after construction (joints/bradot_generic_6dof_joint_3d.h:67-112); filled by joint_make_generic_6dof
		float g6_params[BPC_G6DOF_PARAMS] = {};
		bool pending = false;
		int handle = -1;
	};

	struct Area { // BradotArea3D (P3D/bradot_area_3d.h:46-70); its further shapes are hidden records (`parts`) that take everything
		// but shape and shape transform from it and share its `group` on the GPU side (include/bradot_physics_cuda.h: bpc_area_desc.group)
		RID self, space, shape;
		Transform3D xf, shape_xf;
		bool shape_disabled = false;
		uint32_t layer = 1, mask = 1;
		bool monitorable = false, ray_pickable = false, gravity_is_point = false; // (BradotArea3D's constructor: not pickable, bradot_area_3d.cpp:343)
		Callable monitor_callback, area_monitor_callback;
		ObjectID instance_id;
		int gravity_mode = 0, linear_damp_mode = 0, angular_damp_mode = 0, priority = 0;
		real_t gravity = 9.80665, point_unit_distance = 0.0, linear_damp = 0.1, angular_damp = 0.1;
		Vector3 gravity_vector = Vector3(0, -1, 0);
		Variant wind[4]; // stored, used by soft bodies only
		Area *parent = nullptr;
		int shape_index = 0;
		uint32_t group = 0;
		LocalVector<Area *> parts;
		bool pending = false, dirty = false;
		int handle = -1; // bpc area index once uploaded
	};
	struct FreedBody { // what an exit event of a body freed since the last flush_queries() still has to name
		RID rid;
		ObjectID instance_id;
	};

	mutable RID_PtrOwner<Shape, true> shape_owner;
	mutable RID_PtrOwner<Area, true> area_owner;
	LocalVector<Area *> pending_areas, dirty_areas;
	LocalVector<Area *> area_of_handle;
	HashMap<int32_t, FreedBody> freed_bodies, freed_areas;
	void _area_changed(Area *p_area);
	Area *_area_shape_record(Area *p_area, int p_shape_idx) const;
	uint32_t area_group_counter = 0;
	void _area_leave_space(Area *p_area);
	void _fill_area_desc(const Area *p_area, bpc_area_desc &r_desc);
	void _flush_area_events();
	mutable RID_PtrOwner<Space, true> space_owner;
	mutable RID_PtrOwner<Body, true> body_owner;
	mutable RID_PtrOwner<Joint, true> joint_owner;

	bpc_context *ctx = nullptr;
	bool active = true;
	bool flushing_queries = false;
	bool doing_sync = false;
	bool using_threads = false;
	bool stepped_since_flush = false;
	real_t last_step = 0.0;
	LocalVector<Body *> pending_bodies;
	LocalVector<Body *> dirty_bodies;
	LocalVector<Body *> body_of_slot; // bpc body index -> body (what a query result is mapped back through); nullptr = free slot
	LocalVector<Joint *> pending_joints;
	LocalVector<Body *> pushed_bodies;   // bodies with a non-zero applied force / torque for the next step
	LocalVector<Body *> callback_bodies; // bodies with a state-sync or force-integration callback
	// body state of the last tick, read back in ONE copy at the start of flush_queries(): 19 floats per slot
	// (transform, linear velocity, angular velocity, active) -- what the callbacks and the direct state answer from
	LocalVector<float> state_cache;
	LocalVector<uint8_t> was_active; // active flag per slot one tick earlier
	bool state_cache_valid = false;
	// reported contacts of the last tick: all bodies in ONE copy at flush_queries() when anybody has a callback, else the
	// one body asked about (single_contacts)
	LocalVector<int32_t> contact_counts, contact_offsets;
	LocalVector<bpc_contact> contact_records, single_contacts;
	bool contact_cache_valid = false;
	int single_contacts_slot = -1, single_contacts_count = 0;
	int reporting_bodies = 0;
	bpc_stats last_stats = {};

	int _shape_handle(Shape *p_shape);
	void _flush_pending();
	void _fill_desc(const Body *p_body, bpc_body_desc &r_desc);
	void _fill_joint_desc(const Joint *p_joint, bpc_joint_desc &r_desc);
	void _body_changed(Body *p_body);
	Body *_body_shape_record(Body *p_body, int p_shape_idx) const;
	void _body_drop_part(Body *p_part);
	void _joint_changed(Joint *p_joint);
	void _body_leave_space(Body *p_body);
	void _body_update_callback_list(Body *p_body);
	void _body_call_queries(Body *p_body);
	void _body_push(Body *p_body);
	const float *_body_cached_state(const Body *p_body);

public:
	// body slot of a RID for the query exclude lists (-1: not a body, or not in a space yet)
	int32_t body_slot_of(const RID &p_rid) const {
		const Body *b = body_owner.get_or_null(p_rid);
		return b ? b->slot : -1;
	}
	void body_part_slots(const RID &p_rid, LocalVector<int32_t> &r_slots) const { // the slots of a static body's further shapes
		const Body *b = body_owner.get_or_null(p_rid);
		if (b) {
			for (const Body *part : b->parts) {
				if (part->slot >= 0) {
					r_slots.push_back(part->slot);
				}
			}
		}
	}

private:
	const bpc_contact *_body_contacts(const Body *p_body, int &r_count);
	bool _body_mass_world(const Body *p_body, Basis &r_iinv, Vector3 &r_com, real_t &r_inv_mass) const;

protected:
	static void _bind_methods() {}

public:
	// shapes
	virtual RID world_boundary_shape_create() override;
	virtual RID sphere_shape_create() override;
	virtual RID box_shape_create() override;
	virtual RID capsule_shape_create() override;
	virtual RID convex_polygon_shape_create() override;
	virtual RID concave_polygon_shape_create() override;
	virtual RID heightmap_shape_create() override;
	virtual RID custom_shape_create() override;
	virtual void shape_set_data(RID p_shape, const Variant &p_data) override;
	virtual void shape_set_custom_solver_bias(RID p_shape, real_t p_bias) override;
	virtual real_t shape_get_custom_solver_bias(RID p_shape) const override;
	virtual void shape_set_margin(RID p_shape, real_t p_margin) override;
	virtual real_t shape_get_margin(RID p_shape) const override;
	virtual void body_set_ray_pickable(RID p_body, bool p_enable) override;
	virtual void body_set_collision_priority(RID p_body, real_t p_priority) override;
	virtual real_t body_get_collision_priority(RID p_body) const override;
	virtual void body_set_shape_disabled(RID p_body, int p_shape_idx, bool p_disabled) override;
	virtual void body_set_user_flags(RID p_body, uint32_t p_flags) override;
	virtual uint32_t body_get_user_flags(RID p_body) const override;
	virtual void body_set_contacts_reported_depth_threshold(RID p_body, real_t p_threshold) override;
	virtual real_t body_get_contacts_reported_depth_threshold(RID p_body) const override;
	virtual void joint_make_hinge_simple(RID p_joint, RID p_body_A, const Vector3 &p_pivot_A, const Vector3 &p_axis_A, RID p_body_B, const Vector3 &p_pivot_B, const Vector3 &p_axis_B) override;
	virtual ShapeType shape_get_type(RID p_shape) const override;
	virtual Variant shape_get_data(RID p_shape) const override;
	// spaces
	virtual RID space_create() override;
	virtual void space_set_active(RID p_space, bool p_active) override;
	virtual bool space_is_active(RID p_space) const override;
	virtual void space_set_param(RID p_space, SpaceParameter p_param, real_t p_value) override;
	virtual real_t space_get_param(RID p_space, SpaceParameter p_param) const override;
	virtual PhysicsDirectSpaceState3D *space_get_direct_state(RID p_space) override;
	// areas
	virtual RID area_create() override;
	virtual void area_set_space(RID p_area, RID p_space) override;
	virtual RID area_get_space(RID p_area) const override;
	virtual void area_add_shape(RID p_area, RID p_shape, const Transform3D &p_transform = Transform3D(), bool p_disabled = false) override;
	virtual void area_set_shape(RID p_area, int p_shape_idx, RID p_shape) override;
	virtual void area_set_shape_transform(RID p_area, int p_shape_idx, const Transform3D &p_transform) override;
	virtual int area_get_shape_count(RID p_area) const override;
	virtual RID area_get_shape(RID p_area, int p_shape_idx) const override;
	virtual Transform3D area_get_shape_transform(RID p_area, int p_shape_idx) const override;
	virtual void area_remove_shape(RID p_area, int p_shape_idx) override;
	virtual void area_clear_shapes(RID p_area) override;
	virtual void area_set_shape_disabled(RID p_area, int p_shape_idx, bool p_disabled) override;
	virtual void area_attach_object_instance_id(RID p_area, ObjectID p_id) override;
	virtual ObjectID area_get_object_instance_id(RID p_area) const override;
	virtual void area_set_param(RID p_area, AreaParameter p_param, const Variant &p_value) override;
	virtual void area_set_transform(RID p_area, const Transform3D &p_transform) override;
	virtual Variant area_get_param(RID p_parea, AreaParameter p_param) const override;
	virtual Transform3D area_get_transform(RID p_area) const override;
	virtual void area_set_collision_layer(RID p_area, uint32_t p_layer) override;
	virtual uint32_t area_get_collision_layer(RID p_area) const override;
	virtual void area_set_collision_mask(RID p_area, uint32_t p_mask) override;
	virtual uint32_t area_get_collision_mask(RID p_area) const override;
	virtual void area_set_monitorable(RID p_area, bool p_monitorable) override;
	virtual void area_set_monitor_callback(RID p_area, const Callable &p_callback) override;
	virtual void area_set_area_monitor_callback(RID p_area, const Callable &p_callback) override;
	virtual void area_set_ray_pickable(RID p_area, bool p_enable) override;
	// bodies
	virtual RID body_create() override;
	virtual void body_set_space(RID p_body, RID p_space) override;
	virtual RID body_get_space(RID p_body) const override;
	virtual void body_set_mode(RID p_body, BodyMode p_mode) override;
	virtual BodyMode body_get_mode(RID p_body) const override;
	virtual void body_add_shape(RID p_body, RID p_shape, const Transform3D &p_transform = Transform3D(), bool p_disabled = false) override;
	virtual void body_set_shape(RID p_body, int p_shape_idx, RID p_shape) override;
	virtual void body_set_shape_transform(RID p_body, int p_shape_idx, const Transform3D &p_transform) override;
	virtual RID body_get_shape(RID p_body, int p_shape_idx) const override;
	virtual Transform3D body_get_shape_transform(RID p_body, int p_shape_idx) const override;
	virtual void body_remove_shape(RID p_body, int p_shape_idx) override;
	virtual void body_clear_shapes(RID p_body) override;
	virtual int body_get_shape_count(RID p_body) const override;
	virtual void body_attach_object_instance_id(RID p_body, ObjectID p_id) override;
	virtual ObjectID body_get_object_instance_id(RID p_body) const override;
	virtual void body_reset_mass_properties(RID p_body) override;
	void _push_mass_override(Body *p_body);
	virtual void body_set_collision_layer(RID p_body, uint32_t p_layer) override;
	virtual uint32_t body_get_collision_layer(RID p_body) const override;
	virtual void body_set_collision_mask(RID p_body, uint32_t p_mask) override;
	virtual uint32_t body_get_collision_mask(RID p_body) const override;
	virtual void body_set_param(RID p_body, BodyParameter p_param, const Variant &p_value) override;
	virtual Variant body_get_param(RID p_body, BodyParameter p_param) const override;
	virtual void body_set_state(RID p_body, BodyState p_state, const Variant &p_variant) override;
	virtual Variant body_get_state(RID p_body, BodyState p_state) const override;
	virtual void body_set_enable_continuous_collision_detection(RID p_body, bool p_enable) override;
	virtual bool body_is_continuous_collision_detection_enabled(RID p_body) const override;
	virtual void body_apply_central_impulse(RID p_body, const Vector3 &p_impulse) override;
	virtual void body_apply_impulse(RID p_body, const Vector3 &p_impulse, const Vector3 &p_position = Vector3()) override;
	virtual void body_apply_torque_impulse(RID p_body, const Vector3 &p_impulse) override;
	virtual void body_apply_central_force(RID p_body, const Vector3 &p_force) override;
	virtual void body_apply_force(RID p_body, const Vector3 &p_force, const Vector3 &p_position = Vector3()) override;
	virtual void body_apply_torque(RID p_body, const Vector3 &p_torque) override;
	virtual void body_add_constant_central_force(RID p_body, const Vector3 &p_force) override;
	virtual void body_add_constant_force(RID p_body, const Vector3 &p_force, const Vector3 &p_position = Vector3()) override;
	virtual void body_add_constant_torque(RID p_body, const Vector3 &p_torque) override;
	virtual void body_set_constant_force(RID p_body, const Vector3 &p_force) override;
	virtual Vector3 body_get_constant_force(RID p_body) const override;
	virtual void body_set_constant_torque(RID p_body, const Vector3 &p_torque) override;
	virtual Vector3 body_get_constant_torque(RID p_body) const override;
	virtual void body_set_axis_velocity(RID p_body, const Vector3 &p_axis_velocity) override;
	virtual void body_set_axis_lock(RID p_body, BodyAxis p_axis, bool p_lock) override;
	virtual bool body_is_axis_locked(RID p_body, BodyAxis p_axis) const override;
	virtual void body_add_collision_exception(RID p_body, RID p_body_b) override;
	virtual void body_remove_collision_exception(RID p_body, RID p_body_b) override;
	virtual void body_get_collision_exceptions(RID p_body, List<RID> *p_exceptions) override;
	virtual void body_set_max_contacts_reported(RID p_body, int p_contacts) override;
	virtual int body_get_max_contacts_reported(RID p_body) const override;
	virtual void body_set_omit_force_integration(RID p_body, bool p_omit) override;
	virtual bool body_is_omitting_force_integration(RID p_body) const override;
	virtual void body_set_state_sync_callback(RID p_body, const Callable &p_callable) override;
	virtual void body_set_force_integration_callback(RID p_body, const Callable &p_callable, const Variant &p_udata = Variant()) override;
	virtual PhysicsDirectBodyState3D *body_get_direct_state(RID p_body) override;
	virtual bool body_test_motion(RID p_body, const MotionParameters &p_parameters, MotionResult *r_result = nullptr) override;
	// joints
	virtual RID joint_create() override;
	virtual void joint_clear(RID p_joint) override;
	virtual void joint_make_pin(RID p_joint, RID p_body_A, const Vector3 &p_local_A, RID p_body_B, const Vector3 &p_local_B) override;
	virtual void joint_make_hinge(RID p_joint, RID p_body_A, const Transform3D &p_hinge_A, RID p_body_B, const Transform3D &p_hinge_B) override;
	virtual void joint_disable_collisions_between_bodies(RID p_joint, bool p_disable) override;
	virtual bool joint_is_disabled_collisions_between_bodies(RID p_joint) const override;
	virtual JointType joint_get_type(RID p_joint) const override;
	virtual void joint_set_solver_priority(RID p_joint, int p_priority) override;
	virtual int joint_get_solver_priority(RID p_joint) const override;
	virtual void pin_joint_set_param(RID p_joint, PinJointParam p_param, real_t p_value) override;
	virtual real_t pin_joint_get_param(RID p_joint, PinJointParam p_param) const override;
	virtual void pin_joint_set_local_a(RID p_joint, const Vector3 &p_A) override;
	virtual Vector3 pin_joint_get_local_a(RID p_joint) const override;
	virtual void pin_joint_set_local_b(RID p_joint, const Vector3 &p_B) override;
	virtual Vector3 pin_joint_get_local_b(RID p_joint) const override;
	virtual void hinge_joint_set_param(RID p_joint, HingeJointParam p_param, real_t p_value) override;
	virtual real_t hinge_joint_get_param(RID p_joint, HingeJointParam p_param) const override;
	virtual void hinge_joint_set_flag(RID p_joint, HingeJointFlag p_flag, bool p_enabled) override;
	virtual bool hinge_joint_get_flag(RID p_joint, HingeJointFlag p_flag) const override;
	virtual void joint_make_slider(RID p_joint, RID p_body_A, const Transform3D &p_local_frame_A, RID p_body_B, const Transform3D &p_local_frame_B) override;
	virtual void slider_joint_set_param(RID p_joint, SliderJointParam p_param, real_t p_value) override;
	virtual real_t slider_joint_get_param(RID p_joint, SliderJointParam p_param) const override;
	virtual void joint_make_generic_6dof(RID p_joint, RID p_body_A, const Transform3D &p_local_frame_A, RID p_body_B, const Transform3D &p_local_frame_B) override;
	virtual void generic_6dof_joint_set_param(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisParam p_param, real_t p_value) override;
	virtual real_t generic_6dof_joint_get_param(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisParam p_param) const override;
	virtual void generic_6dof_joint_set_flag(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisFlag p_flag, bool p_enable) override;
	virtual bool generic_6dof_joint_get_flag(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisFlag p_flag) const override;
	virtual void joint_make_cone_twist(RID p_joint, RID p_body_A, const Transform3D &p_local_frame_A, RID p_body_B, const Transform3D &p_local_frame_B) override;
	virtual void cone_twist_joint_set_param(RID p_joint, ConeTwistJointParam p_param, real_t p_value) override;
	virtual real_t cone_twist_joint_get_param(RID p_joint, ConeTwistJointParam p_param) const override;
	// lifecycle
	virtual void free(RID p_rid) override;
	virtual void set_active(bool p_active) override;
	virtual void init() override;
	virtual void step(real_t p_step) override;
	virtual void sync() override;
	virtual void flush_queries() override;
	virtual void end_sync() override;
	virtual void finish() override;
	virtual bool is_flushing_queries() const override;
	virtual int get_process_info(ProcessInfo p_info) override;

#include "bradot_physics_server_cuda_stubs.gen.inc"

	BradotPhysicsServerCUDA(bool p_using_threads = false);
	~BradotPhysicsServerCUDA();
};

#endif // BRADOT_PHYSICS_SERVER_CUDA_H
