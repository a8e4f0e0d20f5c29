// See bradot_physics_server_cuda.h.  Object bookkeeping follows the reference backend
// (modules/bradot_physics_3d/bradot_physics_server_3d.cpp): RID owners, ERR_FAIL_NULL on unknown RIDs,
// nothing thrown.  Bodies are described on the host until the next step() and then created in ONE
// bpc_bodies_create call, because one driver round-trip per body is what limits large scenes.
#include "bradot_physics_server_cuda.h"

#include "core/math/convex_hull.h"
#include "core/math/geometry_3d.h"

static void _xf_to_floats(const Transform3D &p_xf, float *r_out) {
	for (int r = 0; r < 3; r++) {
		for (int c = 0; c < 3; c++) {
			r_out[r * 3 + c] = p_xf.basis[r][c];
		}
	}
	r_out[9] = p_xf.origin.x;
	r_out[10] = p_xf.origin.y;
	r_out[11] = p_xf.origin.z;
}

static Transform3D _floats_to_xf(const float *p_in) {
	Transform3D t;
	t.basis = Basis(p_in[0], p_in[1], p_in[2], p_in[3], p_in[4], p_in[5], p_in[6], p_in[7], p_in[8]);
	t.origin = Vector3(p_in[9], p_in[10], p_in[11]);
	return t;
}

/* SHAPES */

#define MAKE_SHAPE(m_type)                  \
	Shape *shape = memnew(Shape);           \
	shape->type = m_type;                   \
	return shape_owner.make_rid(shape);

RID BradotPhysicsServerCUDA::world_boundary_shape_create() { MAKE_SHAPE(SHAPE_WORLD_BOUNDARY) }
RID BradotPhysicsServerCUDA::sphere_shape_create() { MAKE_SHAPE(SHAPE_SPHERE) }
RID BradotPhysicsServerCUDA::box_shape_create() { MAKE_SHAPE(SHAPE_BOX) }
RID BradotPhysicsServerCUDA::capsule_shape_create() { MAKE_SHAPE(SHAPE_CAPSULE) }
RID BradotPhysicsServerCUDA::convex_polygon_shape_create() { MAKE_SHAPE(SHAPE_CONVEX_POLYGON) }

void BradotPhysicsServerCUDA::shape_set_data(RID p_shape, const Variant &p_data) {
	Shape *shape = shape_owner.get_or_null(p_shape);
	ERR_FAIL_NULL(shape);
	ERR_FAIL_COND_MSG(shape->handle >= 0, "BradotPhysicsCUDA: shape data cannot change once a body uses the shape.");
	shape->data = p_data;
}

PhysicsServer3D::ShapeType BradotPhysicsServerCUDA::shape_get_type(RID p_shape) const {
	const Shape *shape = shape_owner.get_or_null(p_shape);
	ERR_FAIL_NULL_V(shape, SHAPE_CUSTOM);
	return shape->type;
}

Variant BradotPhysicsServerCUDA::shape_get_data(RID p_shape) const {
	const Shape *shape = shape_owner.get_or_null(p_shape);
	ERR_FAIL_NULL_V(shape, Variant());
	return shape->data;
}

// Upload the shape on first use.  Data formats are the reference's (modules/bradot_physics_3d/bradot_shape_3d.cpp
// set_data: Plane / float radius / Vector3 half extents / Dictionary{radius,height} / PackedVector3Array).
int BradotPhysicsServerCUDA::_shape_handle(Shape *p_shape) {
	if (p_shape->handle >= 0) {
		return p_shape->handle;
	}
	float p[4] = { 0, 0, 0, 0 };
	switch (p_shape->type) {
		case SHAPE_WORLD_BOUNDARY: {
			Plane pl = p_shape->data;
			p[0] = pl.normal.x;
			p[1] = pl.normal.y;
			p[2] = pl.normal.z;
			p[3] = pl.d;
			p_shape->handle = bpc_shape_create(ctx, BPC_SHAPE_WORLD_BOUNDARY, p);
		} break;
		case SHAPE_SPHERE: {
			p[0] = (real_t)p_shape->data;
			p_shape->handle = bpc_shape_create(ctx, BPC_SHAPE_SPHERE, p);
		} break;
		case SHAPE_BOX: {
			Vector3 he = p_shape->data;
			p[0] = he.x;
			p[1] = he.y;
			p[2] = he.z;
			p_shape->handle = bpc_shape_create(ctx, BPC_SHAPE_BOX, p);
		} break;
		case SHAPE_CAPSULE: {
			Dictionary d = p_shape->data;
			p[0] = (real_t)d["radius"];
			p[1] = (real_t)d["height"];
			p_shape->handle = bpc_shape_create(ctx, BPC_SHAPE_CAPSULE, p);
		} break;
		case SHAPE_CONVEX_POLYGON: {
			// the hull is built once on the host, exactly as the reference does at shape set-up
			// (modules/bradot_physics_3d/bradot_shape_3d.cpp:1099)
			Vector<Vector3> points = p_shape->data;
			Geometry3D::MeshData mesh;
			Error err = ConvexHullComputer::convex_hull(points, mesh);
			ERR_FAIL_COND_V_MSG(err != OK, -1, "Failed to build convex hull");
			LocalVector<float> verts, planes;
			LocalVector<int32_t> faces, indices, edges;
			for (const Vector3 &v : mesh.vertices) {
				verts.push_back(v.x);
				verts.push_back(v.y);
				verts.push_back(v.z);
			}
			for (const Geometry3D::MeshData::Face &f : mesh.faces) {
				planes.push_back(f.plane.normal.x);
				planes.push_back(f.plane.normal.y);
				planes.push_back(f.plane.normal.z);
				planes.push_back(f.plane.d);
				faces.push_back(indices.size());
				faces.push_back(f.indices.size());
				for (int idx : f.indices) {
					indices.push_back(idx);
				}
			}
			for (const Geometry3D::MeshData::Edge &e : mesh.edges) {
				edges.push_back(e.vertex_a);
				edges.push_back(e.vertex_b);
				edges.push_back(e.face_a);
				edges.push_back(e.face_b);
			}
			p_shape->handle = bpc_shape_create_convex(ctx, verts.ptr(), mesh.vertices.size(), planes.ptr(), faces.ptr(), mesh.faces.size(),
					indices.ptr(), indices.size(), edges.ptr(), mesh.edges.size());
		} break;
		default:
			ERR_FAIL_V_MSG(-1, "BradotPhysicsCUDA: unsupported shape type.");
	}
	ERR_FAIL_COND_V_MSG(p_shape->handle < 0, -1, bpc_last_error(ctx));
	return p_shape->handle;
}

/* SPACES */

RID BradotPhysicsServerCUDA::space_create() {
	ERR_FAIL_NULL_V_MSG(ctx, RID(), "BradotPhysicsCUDA: init() failed, no CUDA device.");
	Space *space = memnew(Space);
	space->handle = bpc_space_create(ctx);
	return space_owner.make_rid(space);
}

PhysicsDirectSpaceState3D *BradotPhysicsServerCUDA::space_get_direct_state(RID p_space) {
	// same contract as the reference (bradot_physics_server_3d.cpp:189-197): one object per space, owned by the server
	Space *space = space_owner.get_or_null(p_space);
	ERR_FAIL_NULL_V(space, nullptr);
	if (!space->direct_state) {
		space->direct_state = memnew(BradotPhysicsDirectSpaceStateCUDA);
		space->direct_state->server = this;
		space->direct_state->space_handle = space->handle;
	}
	return space->direct_state;
}

int BradotPhysicsDirectSpaceStateCUDA::intersect_rays_batch(const Vector<Vector3> &p_from_to, uint32_t p_collision_mask, bool p_hit_from_inside, LocalVector<RayResult> &r_results, LocalVector<bool> &r_hit) {
	ERR_FAIL_NULL_V(server, 0);
	ERR_FAIL_COND_V(p_from_to.size() % 2, 0);
	server->_flush_pending();
	int n = p_from_to.size() / 2;
	LocalVector<float> in;
	in.resize(n * 6);
	for (int i = 0; i < n * 2; i++) {
		in[i * 3 + 0] = p_from_to[i].x;
		in[i * 3 + 1] = p_from_to[i].y;
		in[i * 3 + 2] = p_from_to[i].z;
	}
	LocalVector<bpc_ray_result> out;
	out.resize(n);
	ERR_FAIL_COND_V_MSG(bpc_space_intersect_rays(server->ctx, space_handle, n, in.ptr(), p_collision_mask, p_hit_from_inside ? BPC_RAY_HIT_FROM_INSIDE : 0, out.ptr()) < 0, 0, bpc_last_error(server->ctx));
	r_results.resize(n);
	r_hit.resize(n);
	int hits = 0;
	for (int i = 0; i < n; i++) {
		r_hit[i] = out[i].hit != 0 && out[i].body >= 0 && (uint32_t)out[i].body < server->body_of_slot.size();
		if (!r_hit[i]) {
			continue;
		}
		hits++;
		RayResult &r = r_results[i];
		r.position = Vector3(out[i].position[0], out[i].position[1], out[i].position[2]);
		r.normal = Vector3(out[i].normal[0], out[i].normal[1], out[i].normal[2]);
		r.rid = server->body_of_slot[out[i].body]->self;
		r.collider_id = ObjectID(); // body_attach_object_instance_id is not carried yet
		r.collider = nullptr;
		r.shape = 0;
		r.face_index = -1;
	}
	return hits;
}

int BradotPhysicsDirectSpaceStateCUDA::intersect_point(const PointParameters &p_parameters, ShapeResult *r_results, int p_result_max) {
	ERR_FAIL_NULL_V(server, 0);
	ERR_FAIL_COND_V_MSG(!p_parameters.exclude.is_empty() || p_parameters.collide_with_areas, 0, "BradotPhysicsCUDA: intersect_point with exclude lists or areas is not supported yet.");
	if (p_result_max <= 0 || !p_parameters.collide_with_bodies) {
		return 0;
	}
	server->_flush_pending();
	int max_r = p_result_max > 2048 ? 2048 : p_result_max;
	float p[3] = { (float)p_parameters.position.x, (float)p_parameters.position.y, (float)p_parameters.position.z };
	int32_t count = 0;
	LocalVector<int32_t> slots;
	slots.resize(max_r);
	ERR_FAIL_COND_V_MSG(bpc_space_intersect_points(server->ctx, space_handle, 1, p, p_parameters.collision_mask, max_r, &count, slots.ptr()) < 0, 0, bpc_last_error(server->ctx));
	int cc = 0;
	for (int i = 0; i < count; i++) {
		if (slots[i] < 0 || (uint32_t)slots[i] >= server->body_of_slot.size()) {
			continue;
		}
		r_results[cc].rid = server->body_of_slot[slots[i]]->self;
		r_results[cc].collider_id = ObjectID();
		r_results[cc].collider = nullptr;
		r_results[cc].shape = 0;
		cc++;
	}
	return cc;
}

int BradotPhysicsDirectSpaceStateCUDA::intersect_shape(const ShapeParameters &p_parameters, ShapeResult *r_results, int p_result_max) {
	ERR_FAIL_NULL_V(server, 0);
	ERR_FAIL_COND_V_MSG(!p_parameters.exclude.is_empty() || p_parameters.collide_with_areas || p_parameters.margin != 0.0, 0,
			"BradotPhysicsCUDA: intersect_shape with exclude lists, areas or a margin is not supported yet.");
	if (p_result_max <= 0 || !p_parameters.collide_with_bodies) {
		return 0;
	}
	BradotPhysicsServerCUDA::Shape *shape = server->shape_owner.get_or_null(p_parameters.shape_rid);
	ERR_FAIL_NULL_V(shape, 0);
	int32_t handle = server->_shape_handle(shape);
	ERR_FAIL_COND_V(handle < 0, 0);
	server->_flush_pending();
	int max_r = p_result_max > 2048 ? 2048 : p_result_max;
	float f[12];
	_xf_to_floats(p_parameters.transform, f);
	int32_t count = 0;
	LocalVector<int32_t> slots;
	slots.resize(max_r);
	ERR_FAIL_COND_V_MSG(bpc_space_intersect_shapes(server->ctx, space_handle, 1, &handle, f, p_parameters.collision_mask, max_r, &count, slots.ptr()) < 0, 0, bpc_last_error(server->ctx));
	int cc = 0;
	for (int i = 0; i < count; i++) {
		if (slots[i] < 0 || (uint32_t)slots[i] >= server->body_of_slot.size()) {
			continue;
		}
		if (r_results) {
			r_results[cc].rid = server->body_of_slot[slots[i]]->self;
			r_results[cc].collider_id = ObjectID();
			r_results[cc].collider = nullptr;
			r_results[cc].shape = 0;
		}
		cc++;
	}
	return cc;
}

bool BradotPhysicsDirectSpaceStateCUDA::collide_shape(const ShapeParameters &p_parameters, Vector3 *r_results, int p_result_max, int &r_result_count) {
	r_result_count = 0;
	ERR_FAIL_NULL_V(server, false);
	ERR_FAIL_COND_V_MSG(!p_parameters.exclude.is_empty() || p_parameters.collide_with_areas || p_parameters.margin != 0.0, false,
			"BradotPhysicsCUDA: collide_shape with exclude lists, areas or a margin is not supported yet.");
	if (p_result_max <= 0 || !r_results || !p_parameters.collide_with_bodies) {
		return false;
	}
	BradotPhysicsServerCUDA::Shape *shape = server->shape_owner.get_or_null(p_parameters.shape_rid);
	ERR_FAIL_NULL_V(shape, false);
	int32_t handle = server->_shape_handle(shape);
	ERR_FAIL_COND_V(handle < 0, false);
	server->_flush_pending();
	int max_r = p_result_max > 32 ? 32 : p_result_max;
	float f[12];
	_xf_to_floats(p_parameters.transform, f);
	int32_t count = 0;
	LocalVector<float> pts;
	pts.resize(max_r * 6);
	ERR_FAIL_COND_V_MSG(bpc_space_collide_shapes(server->ctx, space_handle, 1, &handle, f, p_parameters.collision_mask, max_r, &count, pts.ptr()) < 0, false, bpc_last_error(server->ctx));
	for (int i = 0; i < count * 2; i++) {
		r_results[i] = Vector3(pts[i * 3 + 0], pts[i * 3 + 1], pts[i * 3 + 2]);
	}
	r_result_count = count;
	return count > 0;
}

bool BradotPhysicsDirectSpaceStateCUDA::intersect_ray(const RayParameters &p_parameters, RayResult &r_result) {
	ERR_FAIL_COND_V_MSG(!p_parameters.exclude.is_empty() || p_parameters.collide_with_areas || p_parameters.pick_ray, false,
			"BradotPhysicsCUDA: intersect_ray with exclude lists, areas or pick_ray is not supported yet.");
	if (!p_parameters.collide_with_bodies) {
		return false;
	}
	Vector<Vector3> ft;
	ft.push_back(p_parameters.from);
	ft.push_back(p_parameters.to);
	LocalVector<RayResult> res;
	LocalVector<bool> hit;
	if (intersect_rays_batch(ft, p_parameters.collision_mask, p_parameters.hit_from_inside, res, hit) != 1) {
		return false;
	}
	r_result = res[0];
	return true;
}

void BradotPhysicsServerCUDA::space_set_active(RID p_space, bool p_active) {
	Space *space = space_owner.get_or_null(p_space);
	ERR_FAIL_NULL(space);
	space->active = p_active; // every created space is stepped by bpc_step; inactive spaces are a TODO of the C ABI
}

bool BradotPhysicsServerCUDA::space_is_active(RID p_space) const {
	const Space *space = space_owner.get_or_null(p_space);
	ERR_FAIL_NULL_V(space, false);
	return space->active;
}

void BradotPhysicsServerCUDA::space_set_param(RID p_space, SpaceParameter p_param, real_t p_value) {
	Space *space = space_owner.get_or_null(p_space);
	ERR_FAIL_NULL(space);
	ERR_FAIL_INDEX((int)p_param, 8);
	space->params[p_param] = p_value;
	// SpaceParameter values and BPC_SPACE_* values are the same numbers (servers/physics_server_3d.h:283-292)
	bpc_space_set_param(ctx, space->handle, (int)p_param, p_value);
}

real_t BradotPhysicsServerCUDA::space_get_param(RID p_space, SpaceParameter p_param) const {
	const Space *space = space_owner.get_or_null(p_space);
	ERR_FAIL_NULL_V(space, 0);
	ERR_FAIL_INDEX_V((int)p_param, 8, 0);
	return space->params[p_param];
}

/* BODIES */

RID BradotPhysicsServerCUDA::body_create() {
	Body *body = memnew(Body);
	RID rid = body_owner.make_rid(body);
	body->self = rid;
	return rid;
}

void BradotPhysicsServerCUDA::body_set_space(RID p_body, RID p_space) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_COND_MSG(body->slot >= 0, "BradotPhysicsCUDA: moving a body between spaces is not supported yet.");
	if (p_space.is_valid()) {
		ERR_FAIL_NULL(space_owner.get_or_null(p_space));
		body->space = p_space;
		pending_bodies.push_back(body);
	}
}

RID BradotPhysicsServerCUDA::body_get_space(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, RID());
	return body->space;
}

void BradotPhysicsServerCUDA::body_set_mode(RID p_body, BodyMode p_mode) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_COND_MSG(body->slot >= 0, "BradotPhysicsCUDA: the body mode is fixed once the body is in a space.");
	body->mode = p_mode;
}

PhysicsServer3D::BodyMode BradotPhysicsServerCUDA::body_get_mode(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, BODY_MODE_STATIC);
	return body->mode;
}

void BradotPhysicsServerCUDA::body_add_shape(RID p_body, RID p_shape, const Transform3D &p_transform, bool p_disabled) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_NULL(shape_owner.get_or_null(p_shape));
	ERR_FAIL_COND_MSG(body->shape.is_valid(), "BradotPhysicsCUDA: one shape per body in this version.");
	ERR_FAIL_COND_MSG(p_disabled, "BradotPhysicsCUDA: disabled shapes are not supported yet.");
	body->shape = p_shape;
	body->shape_xf = p_transform;
}

int BradotPhysicsServerCUDA::body_get_shape_count(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, -1);
	return body->shape.is_valid() ? 1 : 0;
}

void BradotPhysicsServerCUDA::body_set_collision_layer(RID p_body, uint32_t p_layer) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->layer = p_layer;
}
uint32_t BradotPhysicsServerCUDA::body_get_collision_layer(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, 0);
	return body->layer;
}
void BradotPhysicsServerCUDA::body_set_collision_mask(RID p_body, uint32_t p_mask) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->mask = p_mask;
}
uint32_t BradotPhysicsServerCUDA::body_get_collision_mask(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, 0);
	return body->mask;
}

void BradotPhysicsServerCUDA::body_set_param(RID p_body, BodyParameter p_param, const Variant &p_value) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_COND_MSG(body->slot >= 0, "BradotPhysicsCUDA: body parameters are fixed once the body is in a space (bulk upload).");
	switch (p_param) {
		case BODY_PARAM_BOUNCE: body->bounce = p_value; break;
		case BODY_PARAM_FRICTION: body->friction = p_value; break;
		case BODY_PARAM_MASS: {
			real_t m = p_value;
			ERR_FAIL_COND(m <= 0);
			body->mass = m;
		} break;
		case BODY_PARAM_GRAVITY_SCALE: body->gravity_scale = p_value; break;
		case BODY_PARAM_LINEAR_DAMP: body->lin_damp = p_value; break;
		case BODY_PARAM_ANGULAR_DAMP: body->ang_damp = p_value; break;
		default:
			BPC_UNSUPPORTED("body_set_param(inertia / centre of mass / damp modes)");
	}
}

Variant BradotPhysicsServerCUDA::body_get_param(RID p_body, BodyParameter p_param) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, Variant());
	switch (p_param) {
		case BODY_PARAM_BOUNCE: return body->bounce;
		case BODY_PARAM_FRICTION: return body->friction;
		case BODY_PARAM_MASS: return body->mass;
		case BODY_PARAM_GRAVITY_SCALE: return body->gravity_scale;
		case BODY_PARAM_LINEAR_DAMP: return body->lin_damp;
		case BODY_PARAM_ANGULAR_DAMP: return body->ang_damp;
		default: return Variant();
	}
}

void BradotPhysicsServerCUDA::body_set_state(RID p_body, BodyState p_state, const Variant &p_variant) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	float f[12];
	switch (p_state) {
		case BODY_STATE_TRANSFORM: {
			body->xf = p_variant;
			if (body->slot >= 0) {
				_xf_to_floats(body->xf, f);
				// a kinematic body is told where to be after the next step (new_transform, bradot_body_3d.cpp:355-364)
				bpc_bodies_write(ctx, body->mode == BODY_MODE_KINEMATIC ? BPC_FIELD_KINEMATIC_TARGET : BPC_FIELD_TRANSFORM, body->slot, 1, f);
			}
		} break;
		case BODY_STATE_LINEAR_VELOCITY:
		case BODY_STATE_ANGULAR_VELOCITY: {
			Vector3 v = p_variant;
			(p_state == BODY_STATE_LINEAR_VELOCITY ? body->linvel : body->angvel) = v;
			if (body->slot >= 0 && body->mode == BODY_MODE_KINEMATIC) {
				// constant_linear_velocity / constant_angular_velocity (bradot_body_3d.cpp:383-393)
				f[0] = body->linvel.x; f[1] = body->linvel.y; f[2] = body->linvel.z;
				f[3] = body->angvel.x; f[4] = body->angvel.y; f[5] = body->angvel.z;
				bpc_bodies_write(ctx, BPC_FIELD_KINEMATIC_VELOCITY, body->slot, 1, f);
			} else if (body->slot >= 0) {
				f[0] = v.x;
				f[1] = v.y;
				f[2] = v.z;
				bpc_bodies_write(ctx, p_state == BODY_STATE_LINEAR_VELOCITY ? BPC_FIELD_LINEAR_VELOCITY : BPC_FIELD_ANGULAR_VELOCITY, body->slot, 1, f);
			}
		} break;
		case BODY_STATE_SLEEPING: {
			if (body->slot >= 0) {
				f[0] = bool(p_variant) ? 0.f : 1.f;
				bpc_bodies_write(ctx, BPC_FIELD_ACTIVE, body->slot, 1, f);
			}
		} break;
		case BODY_STATE_CAN_SLEEP:
			body->can_sleep = p_variant;
			break;
	}
}

Variant BradotPhysicsServerCUDA::body_get_state(RID p_body, BodyState p_state) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, Variant());
	float f[12];
	switch (p_state) {
		case BODY_STATE_TRANSFORM:
			if (body->slot < 0) {
				return body->xf;
			}
			ERR_FAIL_COND_V_MSG(bpc_bodies_read(ctx, BPC_FIELD_TRANSFORM, body->slot, 1, f) < 0, Variant(), bpc_last_error(ctx));
			return _floats_to_xf(f);
		case BODY_STATE_LINEAR_VELOCITY:
		case BODY_STATE_ANGULAR_VELOCITY:
			if (body->slot < 0) {
				return p_state == BODY_STATE_LINEAR_VELOCITY ? body->linvel : body->angvel;
			}
			ERR_FAIL_COND_V_MSG(bpc_bodies_read(ctx, p_state == BODY_STATE_LINEAR_VELOCITY ? BPC_FIELD_LINEAR_VELOCITY : BPC_FIELD_ANGULAR_VELOCITY, body->slot, 1, f) < 0,
					Variant(), bpc_last_error(ctx));
			return Vector3(f[0], f[1], f[2]);
		case BODY_STATE_SLEEPING:
			if (body->slot < 0) {
				return false;
			}
			bpc_bodies_read(ctx, BPC_FIELD_ACTIVE, body->slot, 1, f);
			return f[0] == 0.f;
		case BODY_STATE_CAN_SLEEP:
			return body->can_sleep;
	}
	return Variant();
}

/* JOINTS */

RID BradotPhysicsServerCUDA::joint_create() {
	Joint *joint = memnew(Joint);
	return joint_owner.make_rid(joint);
}

void BradotPhysicsServerCUDA::joint_clear(RID p_joint) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND_MSG(joint->handle >= 0, "BradotPhysicsCUDA: joints cannot be cleared once uploaded.");
	joint->type = JOINT_TYPE_MAX;
}

void BradotPhysicsServerCUDA::joint_make_pin(RID p_joint, RID p_body_A, const Vector3 &p_local_A, RID p_body_B, const Vector3 &p_local_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	joint->type = JOINT_TYPE_PIN;
	joint->a = p_body_A;
	joint->b = p_body_B;
	joint->fa = Transform3D(Basis(), p_local_A);
	joint->fb = Transform3D(Basis(), p_local_B);
	pending_joints.push_back(joint);
}

void BradotPhysicsServerCUDA::joint_make_hinge(RID p_joint, RID p_body_A, const Transform3D &p_hinge_A, RID p_body_B, const Transform3D &p_hinge_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	joint->type = JOINT_TYPE_HINGE;
	joint->a = p_body_A;
	joint->b = p_body_B;
	joint->fa = p_hinge_A;
	joint->fb = p_hinge_B;
	pending_joints.push_back(joint);
}

void BradotPhysicsServerCUDA::joint_disable_collisions_between_bodies(RID p_joint, bool p_disable) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND_MSG(joint->handle >= 0, "BradotPhysicsCUDA: set before the first step after creating the joint.");
	joint->disable_collisions = p_disable;
}

PhysicsServer3D::JointType BradotPhysicsServerCUDA::joint_get_type(RID p_joint) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, JOINT_TYPE_PIN);
	return joint->type;
}

/* LIFECYCLE */

void BradotPhysicsServerCUDA::free(RID p_rid) {
	// GPU slots are not recycled in this version; host records are released like the reference does
	// (modules/bradot_physics_3d/bradot_physics_server_3d.cpp:1557-1622).
	if (shape_owner.owns(p_rid)) {
		Shape *s = shape_owner.get_or_null(p_rid);
		shape_owner.free(p_rid);
		memdelete(s);
	} else if (body_owner.owns(p_rid)) {
		Body *b = body_owner.get_or_null(p_rid);
		ERR_FAIL_COND_MSG(b->slot >= 0, "BradotPhysicsCUDA: removing an uploaded body is not supported yet.");
		pending_bodies.erase(b);
		pushed_bodies.erase(b);
		body_owner.free(p_rid);
		memdelete(b);
	} else if (space_owner.owns(p_rid)) {
		Space *s = space_owner.get_or_null(p_rid);
		if (s->direct_state) {
			memdelete(s->direct_state);
		}
		space_owner.free(p_rid);
		memdelete(s);
	} else if (joint_owner.owns(p_rid)) {
		Joint *j = joint_owner.get_or_null(p_rid);
		pending_joints.erase(j);
		joint_owner.free(p_rid);
		memdelete(j);
	} else {
		ERR_FAIL_MSG("Invalid ID.");
	}
}

void BradotPhysicsServerCUDA::set_active(bool p_active) {
	active = p_active;
}

void BradotPhysicsServerCUDA::init() {
	ctx = bpc_create(0);
	ERR_FAIL_NULL_MSG(ctx, "BradotPhysicsCUDA: no usable CUDA device (this backend has no CPU fallback).");
}

// Upload everything created since the last tick in bulk.
void BradotPhysicsServerCUDA::_flush_pending() {
	if (!pending_bodies.is_empty()) {
		LocalVector<bpc_body_desc> descs;
		LocalVector<Body *> uploaded;
		for (Body *b : pending_bodies) {
			Shape *shape = shape_owner.get_or_null(b->shape);
			Space *space = space_owner.get_or_null(b->space);
			if (!shape || !space) {
				continue; // shapeless bodies (the reference's static global body) take no part in the step
			}
			int sh = _shape_handle(shape);
			if (sh < 0) {
				continue;
			}
			bpc_body_desc d = {};
			d.space = space->handle;
			d.mode = (uint32_t)b->mode;
			d.shape = sh;
			d.flags = (b->can_sleep ? 1 : 0) | (b->ccd ? 2 : 0);
			_xf_to_floats(b->xf, d.xf);
			_xf_to_floats(b->shape_xf, d.shape_xf);
			d.linvel[0] = b->linvel.x; d.linvel[1] = b->linvel.y; d.linvel[2] = b->linvel.z;
			d.angvel[0] = b->angvel.x; d.angvel[1] = b->angvel.y; d.angvel[2] = b->angvel.z;
			d.mass = b->mass; d.friction = b->friction; d.bounce = b->bounce; d.gravity_scale = b->gravity_scale;
			d.lin_damp = b->lin_damp; d.ang_damp = b->ang_damp; d.layer = b->layer; d.mask = b->mask;
			descs.push_back(d);
			uploaded.push_back(b);
		}
		if (!descs.is_empty()) {
			int first = bpc_bodies_create(ctx, descs.ptr(), descs.size());
			ERR_FAIL_COND_MSG(first < 0, bpc_last_error(ctx));
			for (uint32_t i = 0; i < uploaded.size(); i++) {
				uploaded[i]->slot = first + i;
				if (body_of_slot.size() <= (uint32_t)(first + i)) {
					body_of_slot.resize(first + i + 1);
				}
				body_of_slot[first + i] = uploaded[i];
			}
		}
		pending_bodies.clear();
	}
	if (!pending_joints.is_empty()) {
		LocalVector<bpc_joint_desc> descs;
		LocalVector<Joint *> uploaded;
		for (Joint *j : pending_joints) {
			Body *a = body_owner.get_or_null(j->a);
			Body *b = body_owner.get_or_null(j->b);
			if (!a || !b || a->slot < 0 || b->slot < 0 || j->type == JOINT_TYPE_MAX) {
				continue;
			}
			bpc_joint_desc d = {};
			d.type = j->type == JOINT_TYPE_PIN ? BPC_JOINT_PIN : BPC_JOINT_HINGE;
			d.body_a = a->slot;
			d.body_b = b->slot;
			d.flags = j->disable_collisions ? 1 : 0;
			_xf_to_floats(j->fa, d.frame_a);
			_xf_to_floats(j->fb, d.frame_b);
			descs.push_back(d);
			uploaded.push_back(j);
		}
		if (!descs.is_empty()) {
			int first = bpc_joints_create(ctx, descs.ptr(), descs.size());
			ERR_FAIL_COND_MSG(first < 0, bpc_last_error(ctx));
			for (uint32_t i = 0; i < uploaded.size(); i++) {
				uploaded[i]->handle = first + i;
			}
		}
		pending_joints.clear();
	}
}

void BradotPhysicsServerCUDA::body_set_enable_continuous_collision_detection(RID p_body, bool p_enable) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_COND_MSG(body->slot >= 0, "BradotPhysicsCUDA: continuous collision detection is fixed once the body is in a space (bulk upload).");
	body->ccd = p_enable;
}

bool BradotPhysicsServerCUDA::body_is_continuous_collision_detection_enabled(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, false);
	return body->ccd;
}

// body_apply_central_force / body_apply_torque (bradot_physics_server_3d.cpp:859-897): accumulated on the host, handed to
// the GPU as one BPC_FIELD_APPLIED_FORCE row per pushed body before the step (a non-zero row wakes the body there)
void BradotPhysicsServerCUDA::body_apply_central_force(RID p_body, const Vector3 &p_force) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	if (body->applied_force == Vector3() && body->applied_torque == Vector3()) {
		pushed_bodies.push_back(body);
	}
	body->applied_force += p_force;
}

void BradotPhysicsServerCUDA::body_apply_torque(RID p_body, const Vector3 &p_torque) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	if (body->applied_force == Vector3() && body->applied_torque == Vector3()) {
		pushed_bodies.push_back(body);
	}
	body->applied_torque += p_torque;
}

void BradotPhysicsServerCUDA::step(real_t p_step) {
	if (!active || !ctx) {
		return;
	}
	_flush_pending();
	for (Body *b : pushed_bodies) {
		if (b->slot >= 0) {
			float f[6] = { (float)b->applied_force.x, (float)b->applied_force.y, (float)b->applied_force.z,
				(float)b->applied_torque.x, (float)b->applied_torque.y, (float)b->applied_torque.z };
			bpc_bodies_write(ctx, BPC_FIELD_APPLIED_FORCE, b->slot, 1, f);
		}
		b->applied_force = Vector3();
		b->applied_torque = Vector3();
	}
	pushed_bodies.clear();
	ERR_FAIL_COND_MSG(bpc_step(ctx, p_step) < 0, bpc_last_error(ctx));
}

void BradotPhysicsServerCUDA::sync() {
	if (ctx) {
		bpc_synchronize(ctx);
	}
}

void BradotPhysicsServerCUDA::flush_queries() {
	if (!active || !ctx) {
		return;
	}
	flushing_queries = true;
	bpc_get_stats(ctx, &last_stats);
	flushing_queries = false;
}

void BradotPhysicsServerCUDA::end_sync() {}

void BradotPhysicsServerCUDA::finish() {
	if (ctx) {
		bpc_destroy(ctx);
		ctx = nullptr;
	}
}

bool BradotPhysicsServerCUDA::is_flushing_queries() const {
	return flushing_queries;
}

int BradotPhysicsServerCUDA::get_process_info(ProcessInfo p_info) {
	if (!ctx) {
		return 0;
	}
	bpc_get_stats(ctx, &last_stats);
	switch (p_info) {
		case INFO_ACTIVE_OBJECTS: return last_stats.active_objects;
		case INFO_COLLISION_PAIRS: return last_stats.collision_pairs;
		case INFO_ISLAND_COUNT: return last_stats.island_count;
	}
	return 0;
}

BradotPhysicsServerCUDA::BradotPhysicsServerCUDA(bool p_using_threads) {}

BradotPhysicsServerCUDA::~BradotPhysicsServerCUDA() {
	finish();
}
