// Engine-side backend: a PhysicsServer3D that forwards the rigid-body path to libbradot_physics_cuda.so
// through the C ABI (include/bradot_physics_cuda.h).  Drop this directory into the engine's modules/ tree;
// register_types.cpp registers it as "BradotPhysicsCUDA" with PhysicsServer3DManager
// (servers/physics_server_3d.cpp:1170-1175), selectable through physics/3d/physics_engine.
//
// Implemented by hand: what the step needs (spaces, world-boundary / sphere / box / capsule / convex shapes,
// rigid and static bodies with one shape, pin and hinge joints, state access, step, stats).  Every other pure
// virtual of the interface gets a generated default override (engine_module/gen_stubs.py) that logs once and
// returns the type's default, which is how the reference answers an invalid RID.
#ifndef BRADOT_PHYSICS_SERVER_CUDA_H
#define BRADOT_PHYSICS_SERVER_CUDA_H

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
	int intersect_rays_batch(const Vector<Vector3> &p_from_to, uint32_t p_collision_mask, bool p_hit_from_inside, LocalVector<RayResult> &r_results, LocalVector<bool> &r_hit);
	virtual int intersect_point(const PointParameters &p_parameters, ShapeResult *r_results, int p_result_max) override;
	virtual int intersect_shape(const ShapeParameters &p_parameters, ShapeResult *r_results, int p_result_max) override;
	virtual bool cast_motion(const ShapeParameters &p_parameters, real_t &p_closest_safe, real_t &p_closest_unsafe, ShapeRestInfo *r_info = nullptr) override { BPC_UNSUPPORTED("cast_motion"); return false; }
	virtual bool collide_shape(const ShapeParameters &p_parameters, Vector3 *r_results, int p_result_max, int &r_result_count) override;
	virtual bool rest_info(const ShapeParameters &p_parameters, ShapeRestInfo *r_info) override { BPC_UNSUPPORTED("rest_info"); return false; }
	virtual Vector3 get_closest_point_to_object_volume(RID p_object, const Vector3 p_point) const override { BPC_UNSUPPORTED("get_closest_point_to_object_volume"); return p_point; }
};

class BradotPhysicsServerCUDA : public PhysicsServer3D {
	BRCLASS(BradotPhysicsServerCUDA, PhysicsServer3D);
	friend class BradotPhysicsDirectSpaceStateCUDA;

	struct Shape {
		ShapeType type = SHAPE_SPHERE;
		Variant data;
		int handle = -1; // bpc shape index, created lazily once the data is known
	};
	struct Space {
		int handle = -1;
		bool active = false;
		BradotPhysicsDirectSpaceStateCUDA *direct_state = nullptr;
		real_t params[8] = { 0.01, 0.05, 0.01, 0.8, 0.1, Math::deg_to_rad(8.0), 0.5, 16 };
	};
	struct Body {
		RID self, space, shape;
		Transform3D shape_xf;
		BodyMode mode = BODY_MODE_RIGID;
		Transform3D xf;
		Vector3 linvel, angvel;
		real_t mass = 1, friction = 1, bounce = 0, gravity_scale = 1, lin_damp = 0, ang_damp = 0;
		uint32_t layer = 1, mask = 1;
		bool can_sleep = true;
		bool ccd = false;
		Vector3 applied_force, applied_torque; // accumulated by body_apply_*, handed over at the next step
		int slot = -1; // bpc body index once uploaded
	};
	struct Joint {
		JointType type = JOINT_TYPE_MAX;
		RID a, b;
		Transform3D fa, fb;
		bool disable_collisions = false;
		int handle = -1;
	};

	mutable RID_PtrOwner<Shape, true> shape_owner;
	mutable RID_PtrOwner<Space, true> space_owner;
	mutable RID_PtrOwner<Body, true> body_owner;
	mutable RID_PtrOwner<Joint, true> joint_owner;

	bpc_context *ctx = nullptr;
	bool active = true;
	bool flushing_queries = false;
	LocalVector<Body *> pending_bodies;
	LocalVector<Body *> body_of_slot; // bpc body index -> body (what a query result is mapped back through)
	LocalVector<Joint *> pending_joints;
	LocalVector<Body *> pushed_bodies; // bodies with a non-zero applied force / torque for the next step
	bpc_stats last_stats = {};

	int _shape_handle(Shape *p_shape);
	void _flush_pending();

protected:
	static void _bind_methods() {}

public:
	// shapes
	virtual RID world_boundary_shape_create() override;
	virtual RID sphere_shape_create() override;
	virtual RID box_shape_create() override;
	virtual RID capsule_shape_create() override;
	virtual RID convex_polygon_shape_create() override;
	virtual void shape_set_data(RID p_shape, const Variant &p_data) override;
	virtual ShapeType shape_get_type(RID p_shape) const override;
	virtual Variant shape_get_data(RID p_shape) const override;
	// spaces
	virtual RID space_create() override;
	virtual void space_set_active(RID p_space, bool p_active) override;
	virtual bool space_is_active(RID p_space) const override;
	virtual void space_set_param(RID p_space, SpaceParameter p_param, real_t p_value) override;
	virtual real_t space_get_param(RID p_space, SpaceParameter p_param) const override;
	virtual PhysicsDirectSpaceState3D *space_get_direct_state(RID p_space) override;
	// bodies
	virtual RID body_create() override;
	virtual void body_set_space(RID p_body, RID p_space) override;
	virtual RID body_get_space(RID p_body) const override;
	virtual void body_set_mode(RID p_body, BodyMode p_mode) override;
	virtual BodyMode body_get_mode(RID p_body) const override;
	virtual void body_add_shape(RID p_body, RID p_shape, const Transform3D &p_transform = Transform3D(), bool p_disabled = false) override;
	virtual int body_get_shape_count(RID p_body) const override;
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
	virtual void body_apply_central_force(RID p_body, const Vector3 &p_force) override;
	virtual void body_apply_torque(RID p_body, const Vector3 &p_torque) override;
	// joints
	virtual RID joint_create() override;
	virtual void joint_clear(RID p_joint) override;
	virtual void joint_make_pin(RID p_joint, RID p_body_A, const Vector3 &p_local_A, RID p_body_B, const Vector3 &p_local_B) override;
	virtual void joint_make_hinge(RID p_joint, RID p_body_A, const Transform3D &p_hinge_A, RID p_body_B, const Transform3D &p_hinge_B) override;
	virtual void joint_disable_collisions_between_bodies(RID p_joint, bool p_disable) override;
	virtual JointType joint_get_type(RID p_joint) const override;
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
