// See bradot_physics_server_cuda.h.  Object bookkeeping follows the reference backend
// (modules/bradot_physics_3d/bradot_physics_server_3d.cpp): RID owners, ERR_FAIL_NULL on unknown RIDs,
// nothing thrown.  Bodies are described on the host until the next step() and then created in ONE
// bpc_bodies_create call, because one driver round-trip per body is what limits large scenes.
#include "bradot_physics_server_cuda.h"

#include "core/math/convex_hull.h"
#include "core/math/geometry_3d.h"
#include "core/object/object.h"

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
RID BradotPhysicsServerCUDA::concave_polygon_shape_create() { MAKE_SHAPE(SHAPE_CONCAVE_POLYGON) }
RID BradotPhysicsServerCUDA::heightmap_shape_create() { MAKE_SHAPE(SHAPE_HEIGHTMAP) }

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
		case SHAPE_CONCAVE_POLYGON: {
			// static triangle mesh: {"faces": three vertices per face, "backface_collision"}; the tree over the faces is built by
			// the backend the way the reference builds its own (P3D/bradot_shape_3d.cpp:1614-1672)
			Dictionary d = p_shape->data;
			ERR_FAIL_COND_V(!d.has("faces"), -1);
			Vector<Vector3> faces = d["faces"];
			ERR_FAIL_COND_V(faces.size() == 0 || faces.size() % 3, -1);
			LocalVector<float> f;
			for (const Vector3 &v : faces) {
				f.push_back(v.x);
				f.push_back(v.y);
				f.push_back(v.z);
			}
			p_shape->handle = bpc_shape_create_concave(ctx, f.ptr(), faces.size() / 3, bool(d.get("backface_collision", false)) ? 1 : 0, nullptr, 0, nullptr);
		} break;
		case SHAPE_HEIGHTMAP: {
			// {"width", "depth", "heights" (PackedFloat32Array), optionally "min_height" / "max_height"} (P3D/bradot_shape_3d.cpp:2180-2250)
			Dictionary d = p_shape->data;
			ERR_FAIL_COND_V(!d.has("width") || !d.has("depth") || !d.has("heights"), -1);
			int w = d["width"], dp = d["depth"];
			PackedFloat32Array hs = d["heights"];
			ERR_FAIL_COND_V(w <= 0 || dp <= 0 || hs.size() != w * dp, -1);
			float mn = 0.f, mx = 0.f;
			if (d.has("min_height") && d.has("max_height")) {
				mn = d["min_height"];
				mx = d["max_height"];
			} else {
				for (int i = 0; i < hs.size(); i++) {
					mn = MIN(mn, hs[i]);
					mx = MAX(mx, hs[i]);
				}
			}
			p_shape->handle = bpc_shape_create_heightmap(ctx, hs.ptr(), w, dp, mn, mx);
		} break;
		default:
			ERR_FAIL_V_MSG(-1, "BradotPhysicsCUDA: unsupported shape type.");
	}
	ERR_FAIL_COND_V_MSG(p_shape->handle < 0, -1, bpc_last_error(ctx));
	if (p_shape->custom_bias != 0.0) {
		bpc_shape_set_custom_bias(ctx, p_shape->handle, (float)p_shape->custom_bias);
	}
	return p_shape->handle;
}

// the reference has no custom shapes either (bradot_physics_server_3d.cpp:102-104)
RID BradotPhysicsServerCUDA::custom_shape_create() {
	ERR_FAIL_V(RID());
}

// shape_set_custom_solver_bias (bradot_physics_server_3d.cpp:110-116): used by the pair setup in place of 0.8
void BradotPhysicsServerCUDA::shape_set_custom_solver_bias(RID p_shape, real_t p_bias) {
	Shape *shape = shape_owner.get_or_null(p_shape);
	ERR_FAIL_NULL(shape);
	shape->custom_bias = p_bias;
	if (shape->handle >= 0) {
		ERR_FAIL_COND_MSG(bpc_shape_set_custom_bias(ctx, shape->handle, (float)p_bias) < 0, bpc_last_error(ctx));
	}
}

real_t BradotPhysicsServerCUDA::shape_get_custom_solver_bias(RID p_shape) const {
	const Shape *shape = shape_owner.get_or_null(p_shape);
	ERR_FAIL_NULL_V(shape, 0);
	return shape->custom_bias;
}

// the following are accepted and ignored by the reference as well (bradot_physics_server_3d.cpp:131-136, 648-657, 859-868)
void BradotPhysicsServerCUDA::shape_set_margin(RID p_shape, real_t p_margin) {
}

real_t BradotPhysicsServerCUDA::shape_get_margin(RID p_shape) const {
	return 0.0;
}

void BradotPhysicsServerCUDA::body_set_user_flags(RID p_body, uint32_t p_flags) {
	ERR_FAIL_NULL(body_owner.get_or_null(p_body));
}

uint32_t BradotPhysicsServerCUDA::body_get_user_flags(RID p_body) const {
	ERR_FAIL_NULL_V(body_owner.get_or_null(p_body), 0);
	return 0;
}

void BradotPhysicsServerCUDA::body_set_contacts_reported_depth_threshold(RID p_body, real_t p_threshold) {
	ERR_FAIL_NULL(body_owner.get_or_null(p_body));
}

real_t BradotPhysicsServerCUDA::body_get_contacts_reported_depth_threshold(RID p_body) const {
	ERR_FAIL_NULL_V(body_owner.get_or_null(p_body), 0);
	return 0;
}

/* SPACES */

RID BradotPhysicsServerCUDA::space_create() {
	ERR_FAIL_NULL_V_MSG(ctx, RID(), "BradotPhysicsCUDA: init() failed, no CUDA device.");
	Space *space = memnew(Space);
	space->handle = bpc_space_create(ctx);
	// a new space is inactive until space_set_active(true) (bradot_physics_server_3d.cpp:144-170)
	bpc_space_set_active(ctx, space->handle, 0);
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

// The exclude list of a query (RayParameters / PointParameters / ShapeParameters ::exclude) for the duration of one call:
// RIDs mapped to body slots, handed to the backend, cleared again on the way out.
struct BpcExcludeScope {
	bpc_context *ctx = nullptr;
	bool set = false;
	BpcExcludeScope(BradotPhysicsServerCUDA *p_server, bpc_context *p_ctx, const HashSet<RID> &p_exclude, LocalVector<int32_t> &r_slots) {
		ctx = p_ctx;
		for (const RID &r : p_exclude) {
			int32_t slot = p_server->body_slot_of(r);
			if (slot >= 0) {
				r_slots.push_back(slot);
			}
			p_server->body_part_slots(r, r_slots); // a static body's further shapes
		}
		if (!r_slots.is_empty()) {
			set = bpc_space_set_query_exclude(ctx, r_slots.ptr(), (int)r_slots.size()) == 0;
		}
	}
	~BpcExcludeScope() {
		if (set) {
			bpc_space_set_query_exclude(ctx, nullptr, 0);
		}
	}
};

int BradotPhysicsDirectSpaceStateCUDA::intersect_rays_batch(const Vector<Vector3> &p_from_to, uint32_t p_collision_mask, bool p_hit_from_inside, LocalVector<RayResult> &r_results, LocalVector<bool> &r_hit, bool p_hit_back_faces, bool p_pick_ray, bool p_with_areas, bool p_with_bodies) {
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
	ERR_FAIL_COND_V_MSG(bpc_space_intersect_rays(server->ctx, space_handle, n, in.ptr(), p_collision_mask, (p_hit_from_inside ? BPC_RAY_HIT_FROM_INSIDE : 0) | (p_hit_back_faces ? BPC_RAY_HIT_BACK_FACES : 0) | (p_pick_ray ? BPC_RAY_PICK : 0) | (p_with_areas ? BPC_QUERY_AREAS : 0) | (p_with_bodies ? 0 : BPC_QUERY_NO_BODIES), out.ptr()) < 0, 0, bpc_last_error(server->ctx));
	r_results.resize(n);
	r_hit.resize(n);
	int hits = 0;
	for (int i = 0; i < n; i++) {
		const int area_handle = -2 - out[i].body; // an area answers as -2 - its index
		const bool is_area = out[i].hit != 0 && out[i].body <= -2 && (uint32_t)area_handle < server->area_of_handle.size() && server->area_of_handle[area_handle];
		r_hit[i] = is_area || (out[i].hit != 0 && out[i].body >= 0 && (uint32_t)out[i].body < server->body_of_slot.size() && server->body_of_slot[out[i].body]);
		if (!r_hit[i]) {
			continue;
		}
		hits++;
		RayResult &r = r_results[i];
		r.position = Vector3(out[i].position[0], out[i].position[1], out[i].position[2]);
		r.normal = Vector3(out[i].normal[0], out[i].normal[1], out[i].normal[2]);
		r.rid = is_area ? server->area_of_handle[area_handle]->self : server->body_of_slot[out[i].body]->self;
		r.collider_id = is_area ? server->area_of_handle[area_handle]->instance_id : server->body_of_slot[out[i].body]->instance_id;
		r.collider = r.collider_id.is_valid() ? ObjectDB::get_instance(r.collider_id) : nullptr;
		r.shape = is_area ? server->area_of_handle[area_handle]->shape_index : server->body_of_slot[out[i].body]->shape_index;
		r.face_index = out[i].face_index;
	}
	return hits;
}

// get_closest_point_to_object_volume (bradot_space_3d.cpp:573-611) of a body: every record of the body (its shapes) is asked in one
// bpc_bodies_closest_points, the nearest answer wins; a body without shapes answers its origin.  Areas are not bodies on
// the GPU side and are not answered here.
Vector3 BradotPhysicsDirectSpaceStateCUDA::get_closest_point_to_object_volume(RID p_object, const Vector3 p_point) const {
	ERR_FAIL_NULL_V(server, Vector3());
	BradotPhysicsServerCUDA::Body *body = server->body_owner.get_or_null(p_object);
	ERR_FAIL_NULL_V_MSG(body, Vector3(), "BradotPhysicsCUDA: get_closest_point_to_object_volume answers for bodies only.");
	server->_flush_pending();
	LocalVector<int32_t> slots;
	if (body->slot >= 0 && !body->shape_disabled) {
		slots.push_back(body->slot);
	}
	for (const BradotPhysicsServerCUDA::Body *part : body->parts) {
		if (part->slot >= 0 && !part->shape_disabled) {
			slots.push_back(part->slot);
		}
	}
	if (slots.is_empty()) {
		return Transform3D(server->body_get_state(p_object, PhysicsServer3D::BODY_STATE_TRANSFORM)).origin;
	}
	LocalVector<float> pts, out;
	pts.resize(slots.size() * 3);
	out.resize(slots.size() * 3);
	for (uint32_t i = 0; i < slots.size(); i++) {
		pts[i * 3 + 0] = p_point.x;
		pts[i * 3 + 1] = p_point.y;
		pts[i * 3 + 2] = p_point.z;
	}
	ERR_FAIL_COND_V_MSG(bpc_bodies_closest_points(server->ctx, (int)slots.size(), slots.ptr(), pts.ptr(), out.ptr()) < 0, Vector3(), bpc_last_error(server->ctx));
	real_t min_distance = 1e20;
	Vector3 min_point;
	for (uint32_t i = 0; i < slots.size(); i++) {
		Vector3 c(out[i * 3 + 0], out[i * 3 + 1], out[i * 3 + 2]);
		real_t d = c.distance_to(p_point);
		if (d < min_distance) {
			min_distance = d;
			min_point = c;
		}
	}
	return min_point;
}

int BradotPhysicsDirectSpaceStateCUDA::intersect_point(const PointParameters &p_parameters, ShapeResult *r_results, int p_result_max) {
	ERR_FAIL_NULL_V(server, 0);
	if (p_result_max <= 0 || (!p_parameters.collide_with_bodies && !p_parameters.collide_with_areas)) {
		return 0;
	}
	server->_flush_pending();
	LocalVector<int32_t> excluded;
	BpcExcludeScope exclude_scope(server, server->ctx, p_parameters.exclude, excluded);
	int max_r = p_result_max > 2048 ? 2048 : p_result_max;
	float p[3] = { (float)p_parameters.position.x, (float)p_parameters.position.y, (float)p_parameters.position.z };
	int32_t count = 0;
	LocalVector<int32_t> slots;
	slots.resize(max_r);
	ERR_FAIL_COND_V_MSG(bpc_space_intersect_points(server->ctx, space_handle, 1, p, p_parameters.collision_mask, (p_parameters.collide_with_areas ? BPC_QUERY_AREAS : 0) | (p_parameters.collide_with_bodies ? 0 : BPC_QUERY_NO_BODIES), max_r, &count, slots.ptr()) < 0, 0, bpc_last_error(server->ctx));
	int cc = 0;
	for (int i = 0; i < count; i++) {
		if (slots[i] <= -2) { // an area
			const int ah = -2 - slots[i];
			if ((uint32_t)ah < server->area_of_handle.size() && server->area_of_handle[ah]) {
				r_results[cc].rid = server->area_of_handle[ah]->self;
				r_results[cc].collider_id = server->area_of_handle[ah]->instance_id;
				r_results[cc].collider = r_results[cc].collider_id.is_valid() ? ObjectDB::get_instance(r_results[cc].collider_id) : nullptr;
				r_results[cc].shape = server->area_of_handle[ah]->shape_index;
				cc++;
			}
			continue;
		}
		if (slots[i] < 0 || (uint32_t)slots[i] >= server->body_of_slot.size() || !server->body_of_slot[slots[i]]) {
			continue;
		}
		r_results[cc].rid = server->body_of_slot[slots[i]]->self;
		r_results[cc].collider_id = server->body_of_slot[slots[i]]->instance_id;
		r_results[cc].collider = r_results[cc].collider_id.is_valid() ? ObjectDB::get_instance(r_results[cc].collider_id) : nullptr;
		r_results[cc].shape = server->body_of_slot[slots[i]]->shape_index;
		cc++;
	}
	return cc;
}

int BradotPhysicsDirectSpaceStateCUDA::intersect_shape(const ShapeParameters &p_parameters, ShapeResult *r_results, int p_result_max) {
	ERR_FAIL_NULL_V(server, 0);
	if (p_result_max <= 0 || (!p_parameters.collide_with_bodies && !p_parameters.collide_with_areas)) {
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
	LocalVector<int32_t> excluded;
	BpcExcludeScope exclude_scope(server, server->ctx, p_parameters.exclude, excluded);
	ERR_FAIL_COND_V_MSG(bpc_space_intersect_shapes(server->ctx, space_handle, 1, &handle, f, p_parameters.collision_mask, (p_parameters.collide_with_areas ? BPC_QUERY_AREAS : 0) | (p_parameters.collide_with_bodies ? 0 : BPC_QUERY_NO_BODIES), p_parameters.margin, max_r, &count, slots.ptr()) < 0, 0, bpc_last_error(server->ctx));
	int cc = 0;
	for (int i = 0; i < count; i++) {
		if (slots[i] <= -2) { // an area
			const int ah = -2 - slots[i];
			if ((uint32_t)ah < server->area_of_handle.size() && server->area_of_handle[ah]) {
				if (r_results) {
					r_results[cc].rid = server->area_of_handle[ah]->self;
					r_results[cc].collider_id = server->area_of_handle[ah]->instance_id;
					r_results[cc].collider = r_results[cc].collider_id.is_valid() ? ObjectDB::get_instance(r_results[cc].collider_id) : nullptr;
					r_results[cc].shape = server->area_of_handle[ah]->shape_index;
				}
				cc++;
			}
			continue;
		}
		if (slots[i] < 0 || (uint32_t)slots[i] >= server->body_of_slot.size() || !server->body_of_slot[slots[i]]) {
			continue;
		}
		if (r_results) {
			r_results[cc].rid = server->body_of_slot[slots[i]]->self;
			r_results[cc].collider_id = server->body_of_slot[slots[i]]->instance_id;
			r_results[cc].collider = r_results[cc].collider_id.is_valid() ? ObjectDB::get_instance(r_results[cc].collider_id) : nullptr;
			r_results[cc].shape = server->body_of_slot[slots[i]]->shape_index;
		}
		cc++;
	}
	return cc;
}

bool BradotPhysicsDirectSpaceStateCUDA::collide_shape(const ShapeParameters &p_parameters, Vector3 *r_results, int p_result_max, int &r_result_count) {
	r_result_count = 0;
	ERR_FAIL_NULL_V(server, false);
	ERR_FAIL_COND_V_MSG(p_parameters.collide_with_areas, false, "BradotPhysicsCUDA: collide_shape with areas is not supported yet.");
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
	LocalVector<int32_t> excluded;
	BpcExcludeScope exclude_scope(server, server->ctx, p_parameters.exclude, excluded);
	ERR_FAIL_COND_V_MSG(bpc_space_collide_shapes(server->ctx, space_handle, 1, &handle, f, p_parameters.collision_mask, p_parameters.margin, max_r, &count, pts.ptr()) < 0, false, bpc_last_error(server->ctx));
	for (int i = 0; i < count * 2; i++) {
		r_results[i] = Vector3(pts[i * 3 + 0], pts[i * 3 + 1], pts[i * 3 + 2]);
	}
	r_result_count = count;
	return count > 0;
}

// what bpc_space_rest_infos / bpc_space_cast_motions report about a body, as the engine's ShapeRestInfo
void BradotPhysicsDirectSpaceStateCUDA::_fill_rest_info(const bpc_rest_info &p_info, ShapeRestInfo *r_info) const {
	BradotPhysicsServerCUDA::Body *b = (p_info.body >= 0 && (uint32_t)p_info.body < server->body_of_slot.size()) ? server->body_of_slot[p_info.body] : nullptr;
	r_info->point = Vector3(p_info.point[0], p_info.point[1], p_info.point[2]);
	r_info->normal = Vector3(p_info.normal[0], p_info.normal[1], p_info.normal[2]);
	r_info->linear_velocity = Vector3(p_info.linear_velocity[0], p_info.linear_velocity[1], p_info.linear_velocity[2]);
	r_info->rid = b ? b->self : RID();
	r_info->collider_id = b ? b->instance_id : ObjectID();
	r_info->shape = b ? b->shape_index : 0;
}

bool BradotPhysicsDirectSpaceStateCUDA::cast_motion(const ShapeParameters &p_parameters, real_t &p_closest_safe, real_t &p_closest_unsafe, ShapeRestInfo *r_info) {
	p_closest_safe = 1;
	p_closest_unsafe = 1;
	ERR_FAIL_NULL_V(server, false);
	ERR_FAIL_COND_V_MSG(p_parameters.collide_with_areas, false, "BradotPhysicsCUDA: cast_motion with areas is not supported yet.");
	if (!p_parameters.collide_with_bodies) {
		return true;
	}
	BradotPhysicsServerCUDA::Shape *shape = server->shape_owner.get_or_null(p_parameters.shape_rid);
	ERR_FAIL_NULL_V(shape, false);
	int32_t handle = server->_shape_handle(shape);
	ERR_FAIL_COND_V(handle < 0, false);
	server->_flush_pending();
	float f[12];
	_xf_to_floats(p_parameters.transform, f);
	float motion[3] = { (float)p_parameters.motion.x, (float)p_parameters.motion.y, (float)p_parameters.motion.z };
	float fractions[2] = { 1.f, 1.f };
	bpc_rest_info info;
	LocalVector<int32_t> excluded;
	BpcExcludeScope exclude_scope(server, server->ctx, p_parameters.exclude, excluded);
	ERR_FAIL_COND_V_MSG(bpc_space_cast_motions(server->ctx, space_handle, 1, &handle, f, motion, p_parameters.margin, p_parameters.collision_mask, fractions, &info) < 0, false, bpc_last_error(server->ctx));
	p_closest_safe = fractions[0];
	p_closest_unsafe = fractions[1];
	if (r_info && info.hit) {
		_fill_rest_info(info, r_info);
	}
	return true;
}

bool BradotPhysicsDirectSpaceStateCUDA::rest_info(const ShapeParameters &p_parameters, ShapeRestInfo *r_info) {
	ERR_FAIL_NULL_V(server, false);
	ERR_FAIL_COND_V_MSG(p_parameters.collide_with_areas, false, "BradotPhysicsCUDA: rest_info with areas is not supported yet.");
	if (!p_parameters.collide_with_bodies) {
		return false;
	}
	BradotPhysicsServerCUDA::Shape *shape = server->shape_owner.get_or_null(p_parameters.shape_rid);
	ERR_FAIL_NULL_V(shape, false);
	int32_t handle = server->_shape_handle(shape);
	ERR_FAIL_COND_V(handle < 0, false);
	server->_flush_pending();
	float f[12];
	_xf_to_floats(p_parameters.transform, f);
	float motion[3] = { (float)p_parameters.motion.x, (float)p_parameters.motion.y, (float)p_parameters.motion.z };
	bpc_rest_info info;
	LocalVector<int32_t> excluded;
	BpcExcludeScope exclude_scope(server, server->ctx, p_parameters.exclude, excluded);
	ERR_FAIL_COND_V_MSG(bpc_space_rest_infos(server->ctx, space_handle, 1, &handle, f, motion, p_parameters.margin, p_parameters.collision_mask, &info) < 0, false, bpc_last_error(server->ctx));
	if (!info.hit) {
		return false;
	}
	if (r_info) {
		_fill_rest_info(info, r_info);
	}
	return true;
}

bool BradotPhysicsDirectSpaceStateCUDA::intersect_ray(const RayParameters &p_parameters, RayResult &r_result) {
	if (!p_parameters.collide_with_bodies && !p_parameters.collide_with_areas) {
		return false;
	}
	ERR_FAIL_NULL_V(server, false);
	server->_flush_pending();
	LocalVector<int32_t> excluded;
	BpcExcludeScope exclude_scope(server, server->ctx, p_parameters.exclude, excluded);
	Vector<Vector3> ft;
	ft.push_back(p_parameters.from);
	ft.push_back(p_parameters.to);
	LocalVector<RayResult> res;
	LocalVector<bool> hit;
	if (intersect_rays_batch(ft, p_parameters.collision_mask, p_parameters.hit_from_inside, res, hit, p_parameters.hit_back_faces, p_parameters.pick_ray, p_parameters.collide_with_areas, p_parameters.collide_with_bodies) != 1) {
		return false;
	}
	r_result = res[0];
	return true;
}

void BradotPhysicsServerCUDA::space_set_active(RID p_space, bool p_active) {
	Space *space = space_owner.get_or_null(p_space);
	ERR_FAIL_NULL(space);
	space->active = p_active;
	ERR_FAIL_COND_MSG(bpc_space_set_active(ctx, space->handle, p_active ? 1 : 0) < 0, bpc_last_error(ctx));
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
	// SpaceParameter values and BPC_SPACE_* values are the same numbers (servers/physics_server_3d.h:283-292)
	ERR_FAIL_COND_MSG(bpc_space_set_param(ctx, space->handle, (int)p_param, p_value) < 0, bpc_last_error(ctx));
	space->params[p_param] = p_value;
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

void BradotPhysicsServerCUDA::_fill_desc(const Body *p_body, bpc_body_desc &r_desc) {
	const Body *b = p_body->parent ? p_body->parent : p_body; // a further shape of a static body: the body's own values ...
	r_desc = {};
	const Space *space = space_owner.get_or_null(b->space);
	Shape *shape = shape_owner.get_or_null(p_body->shape); // ... but this record's shape
	r_desc.space = space ? space->handle : 0;
	r_desc.mode = (uint32_t)b->mode;
	r_desc.shape = shape ? _shape_handle(shape) : 0;
	r_desc.flags = (b->can_sleep ? 1 : 0) | (b->ccd ? 2 : 0) | (b->lin_damp_replace ? 16 : 0) | (b->ang_damp_replace ? 32 : 0) |
			(b->omit_force_integration ? 64 : 0) | (b->ray_pickable ? 0 : 128) | ((b->locked_axis & 63) << 8) | (p_body->shape_disabled ? 0x8000 : 0);
	_xf_to_floats(b->xf, r_desc.xf);
	_xf_to_floats(p_body->shape_xf, r_desc.shape_xf);
	r_desc.linvel[0] = b->linvel.x; r_desc.linvel[1] = b->linvel.y; r_desc.linvel[2] = b->linvel.z;
	r_desc.angvel[0] = b->angvel.x; r_desc.angvel[1] = b->angvel.y; r_desc.angvel[2] = b->angvel.z;
	r_desc.mass = b->mass; r_desc.friction = b->friction; r_desc.bounce = b->bounce; r_desc.gravity_scale = b->gravity_scale;
	r_desc.lin_damp = b->lin_damp; r_desc.ang_damp = b->ang_damp; r_desc.layer = b->layer; r_desc.mask = b->mask;
}

// A property that is part of the body's description changed.  Before the upload it simply sits in the host record; after it
// the body is re-described in bulk before the next step or query (bpc_bodies_update keeps its transform, velocities and
// sleep state).
void BradotPhysicsServerCUDA::_body_changed(Body *p_body) {
	if (p_body->slot >= 0 && !p_body->dirty) {
		p_body->dirty = true;
		dirty_bodies.push_back(p_body);
	}
	for (Body *part : p_body->parts) {
		_body_changed(part);
	}
}

// the record that holds shape `p_shape_idx` of a body: the body itself for 0, one of its parts beyond
BradotPhysicsServerCUDA::Body *BradotPhysicsServerCUDA::_body_shape_record(Body *p_body, int p_shape_idx) const {
	if (p_shape_idx == 0) {
		return p_body->shape.is_valid() ? p_body : nullptr;
	}
	return (p_shape_idx > 0 && (uint32_t)p_shape_idx <= p_body->parts.size()) ? p_body->parts[p_shape_idx - 1] : nullptr;
}

void BradotPhysicsServerCUDA::_body_drop_part(Body *p_part) {
	_body_leave_space(p_part);
	memdelete(p_part);
}

void BradotPhysicsServerCUDA::_body_leave_space(Body *p_body) {
	for (Body *part : p_body->parts) {
		_body_leave_space(part);
	}
	if (p_body->pending) {
		pending_bodies.erase(p_body);
		p_body->pending = false;
	}
	if (p_body->dirty) {
		dirty_bodies.erase(p_body);
		p_body->dirty = false;
	}
	if (p_body->slot >= 0) {
		int32_t slot = p_body->slot;
		// joints attached to the body go with its slot on the GPU side; their host records fall back to "not uploaded"
		List<RID> joints;
		joint_owner.get_owned_list(&joints);
		for (const RID &jr : joints) {
			Joint *j = joint_owner.get_or_null(jr);
			if (j && j->handle >= 0 && (j->a == p_body->self || j->b == p_body->self)) {
				j->handle = -1;
			}
		}
		ERR_FAIL_COND_MSG(bpc_bodies_destroy(ctx, &slot, 1) < 0, bpc_last_error(ctx));
		if (area_owner.get_rid_count() > 0) { // an area it stood in may still have to be told who left
			FreedBody fb;
			fb.rid = p_body->self;
			fb.instance_id = p_body->instance_id;
			freed_bodies[slot] = fb;
		}
		body_of_slot[slot] = nullptr;
		p_body->slot = -1;
		state_cache_valid = false;
		contact_cache_valid = false;
		single_contacts_slot = -1;
	}
}

void BradotPhysicsServerCUDA::body_set_space(RID p_body, RID p_space) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	if (p_space.is_valid()) {
		ERR_FAIL_NULL(space_owner.get_or_null(p_space));
	}
	if (body->space == p_space) {
		return; // pointless (bradot_physics_server_3d.cpp:461-463)
	}
	// leaving a space: the body's slot is freed and its pairs go with the next step
	if (body->slot >= 0) {
		const float *st = _body_cached_state(body);
		if (st) { // it takes its state along
			body->xf = _floats_to_xf(st);
			body->linvel = Vector3(st[12], st[13], st[14]);
			body->angvel = Vector3(st[15], st[16], st[17]);
		}
	}
	_body_leave_space(body);
	body->space = p_space;
	if (p_space.is_valid() && !body->pending) {
		body->pending = true;
		pending_bodies.push_back(body);
	}
	for (Body *part : body->parts) {
		part->space = p_space;
		if (p_space.is_valid() && !part->pending) {
			part->pending = true;
			pending_bodies.push_back(part);
		}
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
	if (body->mode == p_mode) {
		return;
	}
	ERR_FAIL_COND_MSG(!body->parts.is_empty() && p_mode > BODY_MODE_KINEMATIC, "BradotPhysicsCUDA: only static and kinematic bodies can have several shapes in this version.");
	body->mode = p_mode;
	_body_changed(body);
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
	if (body->shape.is_valid()) { // a further shape: bodies without dynamics only (static, kinematic), as a record of its own
		ERR_FAIL_COND_MSG(body->mode > BODY_MODE_KINEMATIC, "BradotPhysicsCUDA: only static and kinematic bodies can have several shapes in this version.");
		Body *part = memnew(Body);
		part->parent = body;
		part->self = body->self;
		part->instance_id = body->instance_id;
		part->space = body->space;
		part->shape = p_shape;
		part->shape_xf = p_transform;
		part->shape_disabled = p_disabled;
		part->shape_index = body->parts.size() + 1;
		body->parts.push_back(part);
		if (body->space.is_valid()) {
			part->pending = true;
			pending_bodies.push_back(part);
		}
		return;
	}
	body->shape = p_shape;
	body->shape_xf = p_transform;
	body->shape_disabled = p_disabled;
	_body_changed(body);
}

// body_set_shape_disabled (bradot_collision_object_3d.cpp:69-96): the shape leaves the broadphase and the mass data; the record stays
void BradotPhysicsServerCUDA::body_set_shape_disabled(RID p_body, int p_shape_idx, bool p_disabled) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	Body *rec = _body_shape_record(body, p_shape_idx);
	ERR_FAIL_NULL(rec);
	rec->shape_disabled = p_disabled;
	_body_changed(rec);
}

void BradotPhysicsServerCUDA::body_set_shape(RID p_body, int p_shape_idx, RID p_shape) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_NULL(shape_owner.get_or_null(p_shape));
	Body *rec = _body_shape_record(body, p_shape_idx);
	ERR_FAIL_NULL(rec);
	rec->shape = p_shape;
	_body_changed(rec);
}

void BradotPhysicsServerCUDA::body_set_shape_transform(RID p_body, int p_shape_idx, const Transform3D &p_transform) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	Body *rec = _body_shape_record(body, p_shape_idx);
	ERR_FAIL_NULL(rec);
	rec->shape_xf = p_transform;
	_body_changed(rec);
}

RID BradotPhysicsServerCUDA::body_get_shape(RID p_body, int p_shape_idx) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, RID());
	const Body *rec = _body_shape_record(const_cast<Body *>(body), p_shape_idx);
	ERR_FAIL_NULL_V(rec, RID());
	return rec->shape;
}

Transform3D BradotPhysicsServerCUDA::body_get_shape_transform(RID p_body, int p_shape_idx) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, Transform3D());
	const Body *rec = _body_shape_record(const_cast<Body *>(body), p_shape_idx);
	ERR_FAIL_NULL_V(rec, Transform3D());
	return rec->shape_xf;
}

// A body without its shape takes no part in the step: it leaves the GPU side until it is given a shape again.
void BradotPhysicsServerCUDA::body_remove_shape(RID p_body, int p_shape_idx) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_NULL(_body_shape_record(body, p_shape_idx));
	if (body->parts.is_empty()) {
		body_clear_shapes(p_body);
		return;
	}
	// the shapes behind it move down one index (bradot_collision_object_3d.cpp:98-115): shift them through the records
	// and drop the last one
	for (int i = p_shape_idx; i < (int)body->parts.size(); i++) {
		Body *dst = _body_shape_record(body, i), *src = body->parts[i];
		dst->shape = src->shape;
		dst->shape_xf = src->shape_xf;
		dst->shape_disabled = src->shape_disabled;
		_body_changed(dst);
	}
	_body_drop_part(body->parts[body->parts.size() - 1]);
	body->parts.resize(body->parts.size() - 1);
}

void BradotPhysicsServerCUDA::body_clear_shapes(RID p_body) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	if (!body->shape.is_valid()) {
		return;
	}
	for (Body *part : body->parts) {
		_body_drop_part(part);
	}
	body->parts.clear();
	bool in_space = body->slot >= 0 || body->pending;
	_body_leave_space(body);
	body->shape = RID();
	if (in_space && !body->pending) { // re-uploaded by _flush_pending once it has a shape again
		body->pending = true;
		pending_bodies.push_back(body);
	}
}

int BradotPhysicsServerCUDA::body_get_shape_count(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, -1);
	return body->shape.is_valid() ? 1 + (int)body->parts.size() : 0;
}

void BradotPhysicsServerCUDA::body_attach_object_instance_id(RID p_body, ObjectID p_id) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->instance_id = p_id; // host-side only: query results are mapped back through body_of_slot
	for (Body *part : body->parts) {
		part->instance_id = p_id;
	}
}

ObjectID BradotPhysicsServerCUDA::body_get_object_instance_id(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, ObjectID());
	return body->instance_id;
}

void BradotPhysicsServerCUDA::body_set_collision_layer(RID p_body, uint32_t p_layer) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->layer = p_layer;
	_body_changed(body);
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
	_body_changed(body);
}
uint32_t BradotPhysicsServerCUDA::body_get_collision_mask(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, 0);
	return body->mask;
}

void BradotPhysicsServerCUDA::body_set_param(RID p_body, BodyParameter p_param, const Variant &p_value) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
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
		case BODY_PARAM_LINEAR_DAMP_MODE: body->lin_damp_replace = (int)p_value == BODY_DAMP_MODE_REPLACE; break;
		case BODY_PARAM_ANGULAR_DAMP_MODE: body->ang_damp_replace = (int)p_value == BODY_DAMP_MODE_REPLACE; break;
		case BODY_PARAM_INERTIA:
			body->custom_inertia = p_value;
			body->has_custom_inertia = true;
			body->mass_override_dirty = true;
			break;
		case BODY_PARAM_CENTER_OF_MASS:
			body->custom_com = p_value;
			body->has_custom_com = true;
			body->mass_override_dirty = true;
			break;
		default:
			return;
	}
	_body_changed(body);
}

// body_reset_mass_properties (bradot_body_3d.cpp:152-160 reset_mass_properties): inertia and centre of mass come from the shape again
void BradotPhysicsServerCUDA::body_reset_mass_properties(RID p_body) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->has_custom_inertia = body->has_custom_com = false;
	body->mass_override_dirty = true;
	_body_changed(body);
}

// the body's own inertia / centre of mass go to its slot before it is described
void BradotPhysicsServerCUDA::_push_mass_override(Body *p_body) {
	if (!p_body->mass_override_dirty || p_body->slot < 0) {
		return;
	}
	p_body->mass_override_dirty = false;
	int32_t slot = p_body->slot;
	float in[3] = { (float)p_body->custom_inertia.x, (float)p_body->custom_inertia.y, (float)p_body->custom_inertia.z };
	float cm[3] = { (float)p_body->custom_com.x, (float)p_body->custom_com.y, (float)p_body->custom_com.z };
	bpc_bodies_set_mass_override(ctx, &slot, 1, nullptr, nullptr);
	if (p_body->has_custom_inertia || p_body->has_custom_com) {
		ERR_FAIL_COND_MSG(bpc_bodies_set_mass_override(ctx, &slot, 1, p_body->has_custom_inertia ? in : nullptr, p_body->has_custom_com ? cm : nullptr) < 0, bpc_last_error(ctx));
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
		case BODY_PARAM_LINEAR_DAMP_MODE: return body->lin_damp_replace ? BODY_DAMP_MODE_REPLACE : BODY_DAMP_MODE_COMBINE;
		case BODY_PARAM_ANGULAR_DAMP_MODE: return body->ang_damp_replace ? BODY_DAMP_MODE_REPLACE : BODY_DAMP_MODE_COMBINE;
		case BODY_PARAM_INERTIA:
		case BODY_PARAM_CENTER_OF_MASS: { // as the body stands on the GPU side (bradot_body_3d.cpp:266-275: _inv_inertia.inverse(), center_of_mass_local)
			float mp[16];
			if (body->slot < 0 || bpc_bodies_read(ctx, BPC_FIELD_MASS_PROPS, body->slot, 1, mp) < 0) {
				return p_param == BODY_PARAM_INERTIA ? (body->has_custom_inertia ? body->custom_inertia : Vector3()) : (body->has_custom_com ? body->custom_com : Vector3());
			}
			if (p_param == BODY_PARAM_CENTER_OF_MASS) {
				return Vector3(mp[4], mp[5], mp[6]);
			}
			return body->mode == BODY_MODE_RIGID ? Vector3(1.0f / mp[1], 1.0f / mp[2], 1.0f / mp[3]) : Vector3();
		}
		default: return Variant();
	}
}

// The body's row of the state read back at the last flush_queries(), or nullptr when there is none (then getters ask the GPU).
const float *BradotPhysicsServerCUDA::_body_cached_state(const Body *p_body) {
	if (!state_cache_valid || p_body->slot < 0 || (uint32_t)(p_body->slot + 1) * 19 > state_cache.size()) {
		return nullptr;
	}
	return &state_cache[p_body->slot * 19];
}

// body_set_state: on an uploaded body it runs on the GPU with the reference's side effects (bpc_bodies_set_state restates
// BradotBody3D::set_state, bradot_body_3d.cpp:352-417: wake-ups, orthonormalised transform, broadphase box, neighbours of a
// moved static body)
void BradotPhysicsServerCUDA::body_set_state(RID p_body, BodyState p_state, const Variant &p_variant) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	float f[12] = {};
	switch (p_state) {
		case BODY_STATE_TRANSFORM: {
			body->xf = p_variant;
			_xf_to_floats(body->xf, f);
		} break;
		case BODY_STATE_LINEAR_VELOCITY:
		case BODY_STATE_ANGULAR_VELOCITY: {
			Vector3 v = p_variant;
			(p_state == BODY_STATE_LINEAR_VELOCITY ? body->linvel : body->angvel) = v;
			f[0] = v.x;
			f[1] = v.y;
			f[2] = v.z;
		} break;
		case BODY_STATE_SLEEPING:
			f[0] = bool(p_variant) ? 1.f : 0.f;
			break;
		case BODY_STATE_CAN_SLEEP:
			body->can_sleep = p_variant;
			f[0] = body->can_sleep ? 1.f : 0.f;
			break;
	}
	if (body->slot < 0) {
		return;
	}
	int32_t slot = body->slot;
	ERR_FAIL_COND_MSG(bpc_bodies_set_state(ctx, (int)p_state, &slot, 1, f) < 0, bpc_last_error(ctx));
	for (Body *part : body->parts) { // the further shapes of a static or kinematic body are placed, and moved, with it
		if (part->slot >= 0 && (p_state == BODY_STATE_TRANSFORM || p_state == BODY_STATE_LINEAR_VELOCITY || p_state == BODY_STATE_ANGULAR_VELOCITY)) {
			int32_t ps = part->slot;
			ERR_FAIL_COND_MSG(bpc_bodies_set_state(ctx, (int)p_state, &ps, 1, f) < 0, bpc_last_error(ctx));
		}
	}
	// keep the read-back copy in step with what was just set (a callback may set and then get)
	float *st = state_cache_valid && (uint32_t)(slot + 1) * 19 <= state_cache.size() ? &state_cache[slot * 19] : nullptr;
	if (st) {
		switch (p_state) {
			case BODY_STATE_TRANSFORM: {
				Transform3D t = body->xf;
				if (body->mode >= BODY_MODE_RIGID) {
					t.orthonormalize();
				}
				if (body->mode != BODY_MODE_KINEMATIC) {
					_xf_to_floats(t, st);
				}
			} break;
			case BODY_STATE_LINEAR_VELOCITY: st[12] = f[0]; st[13] = f[1]; st[14] = f[2]; break;
			case BODY_STATE_ANGULAR_VELOCITY: st[15] = f[0]; st[16] = f[1]; st[17] = f[2]; break;
			case BODY_STATE_SLEEPING:
				if (body->mode >= BODY_MODE_RIGID) {
					st[18] = f[0] != 0.f ? 0.f : 1.f;
					if (f[0] != 0.f) {
						for (int k = 12; k < 18; k++) {
							st[k] = 0.f;
						}
					}
				}
				break;
			default:
				break;
		}
		if (p_state != BODY_STATE_SLEEPING && p_state != BODY_STATE_CAN_SLEEP && body->mode >= BODY_MODE_RIGID) {
			st[18] = 1.f;
		}
	}
}

Variant BradotPhysicsServerCUDA::body_get_state(RID p_body, BodyState p_state) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, Variant());
	if (p_state == BODY_STATE_CAN_SLEEP) {
		return body->can_sleep;
	}
	if (body->slot < 0) {
		switch (p_state) {
			case BODY_STATE_TRANSFORM: return body->xf;
			case BODY_STATE_LINEAR_VELOCITY: return body->linvel;
			case BODY_STATE_ANGULAR_VELOCITY: return body->angvel;
			default: return false;
		}
	}
	float row[19];
	const float *st = const_cast<BradotPhysicsServerCUDA *>(this)->_body_cached_state(body);
	if (!st) {
		ERR_FAIL_COND_V_MSG(bpc_bodies_read_state(ctx, body->slot, 1, row) < 0, Variant(), bpc_last_error(ctx));
		st = row;
	}
	switch (p_state) {
		case BODY_STATE_TRANSFORM: return _floats_to_xf(st);
		case BODY_STATE_LINEAR_VELOCITY: return Vector3(st[12], st[13], st[14]);
		case BODY_STATE_ANGULAR_VELOCITY: return Vector3(st[15], st[16], st[17]);
		case BODY_STATE_SLEEPING: return st[18] == 0.f;
		default: return Variant();
	}
}

// body_set_collision_priority (bradot_physics_server_3d.cpp:611-623): what body_test_motion's recovery weighs its obstacles by
void BradotPhysicsServerCUDA::body_set_collision_priority(RID p_body, real_t p_priority) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_COND_MSG(p_priority <= 0, "Priority must be greater than 0.");
	body->collision_priority = p_priority;
	if (body->slot >= 0) {
		int32_t slot = body->slot;
		float v = (float)p_priority;
		ERR_FAIL_COND_MSG(bpc_bodies_set_collision_priority(ctx, &slot, 1, &v) < 0, bpc_last_error(ctx));
	}
}

real_t BradotPhysicsServerCUDA::body_get_collision_priority(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, 0);
	return body->collision_priority;
}

// body_set_ray_pickable (bradot_physics_server_3d.cpp:907-911): only rays cast with RayParameters::pick_ray pass such a body by
void BradotPhysicsServerCUDA::body_set_ray_pickable(RID p_body, bool p_enable) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->ray_pickable = p_enable;
	_body_changed(body);
}

void BradotPhysicsServerCUDA::body_set_enable_continuous_collision_detection(RID p_body, bool p_enable) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->ccd = p_enable;
	_body_changed(body);
}

bool BradotPhysicsServerCUDA::body_is_continuous_collision_detection_enabled(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, false);
	return body->ccd;
}

void BradotPhysicsServerCUDA::body_set_axis_lock(RID p_body, BodyAxis p_axis, bool p_lock) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->locked_axis = p_lock ? (body->locked_axis | (uint32_t)p_axis) : (body->locked_axis & ~(uint32_t)p_axis);
	_body_changed(body);
	if (body->slot >= 0) { // body->wakeup() (bradot_physics_server_3d.cpp:820-826)
		body_set_state(p_body, BODY_STATE_SLEEPING, false);
	}
}

bool BradotPhysicsServerCUDA::body_is_axis_locked(RID p_body, BodyAxis p_axis) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, false);
	return body->locked_axis & (uint32_t)p_axis;
}

// body_set_max_contacts_reported (bradot_physics_server_3d.cpp:819-824; BradotBody3D::set_max_contacts_reported resizes the
// body's contact list): the body's table of the N deepest contacts on the GPU
void BradotPhysicsServerCUDA::body_set_max_contacts_reported(RID p_body, int p_contacts) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_COND(p_contacts < 0);
	reporting_bodies += (p_contacts > 0 ? 1 : 0) - (body->max_contacts_reported > 0 ? 1 : 0);
	body->max_contacts_reported = p_contacts;
	if (body->slot >= 0) {
		int32_t slot = body->slot, cap = p_contacts;
		ERR_FAIL_COND_MSG(bpc_bodies_set_max_contacts_reported(ctx, &slot, 1, &cap) < 0, bpc_last_error(ctx));
		contact_cache_valid = false;
		single_contacts_slot = -1;
	}
}

int BradotPhysicsServerCUDA::body_get_max_contacts_reported(RID p_body) const {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, -1);
	return body->max_contacts_reported;
}

// The contacts reported to one body in the last step: from the bulk copy of flush_queries() when there is one, else read
// for this body alone (and kept until the next step or the next body asked about).
const bpc_contact *BradotPhysicsServerCUDA::_body_contacts(const Body *p_body, int &r_count) {
	r_count = 0;
	if (p_body->slot < 0 || p_body->max_contacts_reported <= 0) {
		return nullptr;
	}
	if (contact_cache_valid && (uint32_t)p_body->slot < contact_counts.size()) {
		r_count = contact_counts[p_body->slot];
		return contact_records.ptr() + contact_offsets[p_body->slot];
	}
	if (single_contacts_slot != p_body->slot) {
		single_contacts.resize(p_body->max_contacts_reported);
		int n = bpc_body_read_contacts(ctx, p_body->slot, p_body->max_contacts_reported, single_contacts.ptr());
		ERR_FAIL_COND_V_MSG(n < 0, nullptr, bpc_last_error(ctx));
		single_contacts_slot = p_body->slot;
		single_contacts_count = MIN(n, p_body->max_contacts_reported);
	}
	r_count = single_contacts_count;
	return single_contacts.ptr();
}

void BradotPhysicsServerCUDA::body_set_omit_force_integration(RID p_body, bool p_omit) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->omit_force_integration = p_omit;
	_body_changed(body);
}

bool BradotPhysicsServerCUDA::body_is_omitting_force_integration(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, false);
	return body->omit_force_integration;
}

// body_add_collision_exception (bradot_physics_server_3d.cpp:690-720): one-sided in the reference's bookkeeping, symmetric
// in effect (BradotBodyPair3D::setup skips the pair if either body excepts the other, bradot_body_pair_3d.cpp:267-270)
void BradotPhysicsServerCUDA::body_add_collision_exception(RID p_body, RID p_body_b) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_b));
	if (!body->exceptions.has(p_body_b)) {
		body->exceptions.push_back(p_body_b);
	}
	_flush_pending();
	Body *other = body_owner.get_or_null(p_body_b);
	if (body->slot >= 0 && other->slot >= 0) {
		bpc_bodies_set_exception(ctx, body->slot, other->slot, 1);
	}
	if (body->slot >= 0) {
		body_set_state(p_body, BODY_STATE_SLEEPING, false); // body->wakeup()
	}
}

void BradotPhysicsServerCUDA::body_remove_collision_exception(RID p_body, RID p_body_b) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->exceptions.erase(p_body_b);
	Body *other = body_owner.get_or_null(p_body_b);
	if (other && body->slot >= 0 && other->slot >= 0 && !other->exceptions.has(p_body)) {
		bpc_bodies_set_exception(ctx, body->slot, other->slot, 0);
	}
	if (body->slot >= 0) {
		body_set_state(p_body, BODY_STATE_SLEEPING, false);
	}
}

void BradotPhysicsServerCUDA::body_get_collision_exceptions(RID p_body, List<RID> *p_exceptions) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	for (const RID &r : body->exceptions) {
		p_exceptions->push_back(r);
	}
}

/* forces and impulses (bradot_physics_server_3d.cpp:827-940, bradot_body_3d.h:228-281) */

void BradotPhysicsServerCUDA::_body_push(Body *p_body) {
	if (!p_body->pushed) {
		p_body->pushed = true;
		pushed_bodies.push_back(p_body);
	}
}

// world inverse inertia tensor, world-oriented centre of mass and inverse mass as the GPU holds them
bool BradotPhysicsServerCUDA::_body_mass_world(const Body *p_body, Basis &r_iinv, Vector3 &r_com, real_t &r_inv_mass) const {
	if (p_body->slot < 0) {
		return false;
	}
	float m[13];
	ERR_FAIL_COND_V_MSG(bpc_bodies_read(ctx, BPC_FIELD_MASS_WORLD, p_body->slot, 1, m) < 0, false, bpc_last_error(ctx));
	r_iinv = Basis(m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8]);
	r_com = Vector3(m[9], m[10], m[11]);
	r_inv_mass = m[12];
	return true;
}

void BradotPhysicsServerCUDA::body_apply_central_impulse(RID p_body, const Vector3 &p_impulse) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	_flush_pending();
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	ERR_FAIL_COND_MSG(!_body_mass_world(body, iinv, com, im), "BradotPhysicsCUDA: impulses need a body that is in a space.");
	Vector3 v = body_get_state(p_body, BODY_STATE_LINEAR_VELOCITY);
	body_set_state(p_body, BODY_STATE_LINEAR_VELOCITY, v + p_impulse * im); // wakes the body, as body->wakeup() does
}

void BradotPhysicsServerCUDA::body_apply_impulse(RID p_body, const Vector3 &p_impulse, const Vector3 &p_position) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	_flush_pending();
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	ERR_FAIL_COND_MSG(!_body_mass_world(body, iinv, com, im), "BradotPhysicsCUDA: impulses need a body that is in a space.");
	Vector3 v = body_get_state(p_body, BODY_STATE_LINEAR_VELOCITY), w = body_get_state(p_body, BODY_STATE_ANGULAR_VELOCITY);
	body_set_state(p_body, BODY_STATE_LINEAR_VELOCITY, v + p_impulse * im);
	body_set_state(p_body, BODY_STATE_ANGULAR_VELOCITY, w + iinv.xform((p_position - com).cross(p_impulse)));
}

void BradotPhysicsServerCUDA::body_apply_torque_impulse(RID p_body, const Vector3 &p_impulse) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	_flush_pending();
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	ERR_FAIL_COND_MSG(!_body_mass_world(body, iinv, com, im), "BradotPhysicsCUDA: impulses need a body that is in a space.");
	Vector3 w = body_get_state(p_body, BODY_STATE_ANGULAR_VELOCITY);
	body_set_state(p_body, BODY_STATE_ANGULAR_VELOCITY, w + iinv.xform(p_impulse));
}

// body_apply_central_force / body_apply_force / body_apply_torque: accumulated on the host, handed to the GPU in ONE
// scattered write for all pushed bodies before the step (a non-zero row wakes its body there)
void BradotPhysicsServerCUDA::body_apply_central_force(RID p_body, const Vector3 &p_force) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	_body_push(body);
	body->applied_force += p_force;
}

void BradotPhysicsServerCUDA::body_apply_force(RID p_body, const Vector3 &p_force, const Vector3 &p_position) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	_flush_pending();
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	_body_mass_world(body, iinv, com, im);
	_body_push(body);
	body->applied_force += p_force;
	body->applied_torque += (p_position - com).cross(p_force);
}

void BradotPhysicsServerCUDA::body_apply_torque(RID p_body, const Vector3 &p_torque) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	_body_push(body);
	body->applied_torque += p_torque;
}

static void _write_constant(bpc_context *ctx, int p_slot, const Vector3 &p_f, const Vector3 &p_t) {
	float f[6] = { (float)p_f.x, (float)p_f.y, (float)p_f.z, (float)p_t.x, (float)p_t.y, (float)p_t.z };
	bpc_bodies_write(ctx, BPC_FIELD_CONSTANT_FORCE, p_slot, 1, f);
}

void BradotPhysicsServerCUDA::body_add_constant_central_force(RID p_body, const Vector3 &p_force) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body_set_constant_force(p_body, body->constant_force + p_force);
}

void BradotPhysicsServerCUDA::body_add_constant_force(RID p_body, const Vector3 &p_force, const Vector3 &p_position) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	_flush_pending();
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	_body_mass_world(body, iinv, com, im);
	body->constant_torque += (p_position - com).cross(p_force);
	body_set_constant_force(p_body, body->constant_force + p_force);
}

void BradotPhysicsServerCUDA::body_add_constant_torque(RID p_body, const Vector3 &p_torque) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body_set_constant_torque(p_body, body->constant_torque + p_torque);
}

void BradotPhysicsServerCUDA::body_set_constant_force(RID p_body, const Vector3 &p_force) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->constant_force = p_force;
	_flush_pending();
	if (body->slot >= 0) {
		_write_constant(ctx, body->slot, body->constant_force, body->constant_torque);
	}
}

Vector3 BradotPhysicsServerCUDA::body_get_constant_force(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, Vector3());
	return body->constant_force;
}

void BradotPhysicsServerCUDA::body_set_constant_torque(RID p_body, const Vector3 &p_torque) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->constant_torque = p_torque;
	_flush_pending();
	if (body->slot >= 0) {
		_write_constant(ctx, body->slot, body->constant_force, body->constant_torque);
	}
}

Vector3 BradotPhysicsServerCUDA::body_get_constant_torque(RID p_body) const {
	const Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, Vector3());
	return body->constant_torque;
}

// BradotBody3D::set_axis_velocity (bradot_body_3d.h:221-226): the velocity component along the axis is replaced
void BradotPhysicsServerCUDA::body_set_axis_velocity(RID p_body, const Vector3 &p_axis_velocity) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	Vector3 axis = p_axis_velocity.normalized();
	Vector3 v = body_get_state(p_body, BODY_STATE_LINEAR_VELOCITY);
	v -= axis * axis.dot(v);
	v += p_axis_velocity;
	body_set_state(p_body, BODY_STATE_LINEAR_VELOCITY, v);
}

/* callbacks and direct state */

void BradotPhysicsServerCUDA::_body_update_callback_list(Body *p_body) {
	bool want = p_body->state_sync_callback.is_valid() || p_body->fi_callable.is_valid();
	if (want && !p_body->has_callbacks) {
		callback_bodies.push_back(p_body);
	} else if (!want && p_body->has_callbacks) {
		callback_bodies.erase(p_body);
	}
	p_body->has_callbacks = want;
}

void BradotPhysicsServerCUDA::body_set_state_sync_callback(RID p_body, const Callable &p_callable) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->state_sync_callback = p_callable;
	_body_update_callback_list(body);
}

void BradotPhysicsServerCUDA::body_set_force_integration_callback(RID p_body, const Callable &p_callable, const Variant &p_udata) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL(body);
	body->fi_callable = p_callable;
	body->fi_udata = p_udata;
	_body_update_callback_list(body);
}

PhysicsDirectBodyState3D *BradotPhysicsServerCUDA::body_get_direct_state(RID p_body) {
	// the reference allows this during sync / flush_queries only (bradot_physics_server_3d.cpp:924-940)
	ERR_FAIL_COND_V_MSG((using_threads && !doing_sync), nullptr, "Body state is inaccessible right now, wait for iteration or physics process notification.");
	if (!body_owner.owns(p_body)) {
		return nullptr;
	}
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, nullptr);
	if (body->slot < 0 && !body->pending) {
		return nullptr; // not in a space
	}
	if (!body->direct_state) {
		body->direct_state = memnew(BradotPhysicsDirectBodyStateCUDA);
		body->direct_state->server = this;
		body->direct_state->body = p_body;
	}
	return body->direct_state;
}


// body_test_motion (P3D/bradot_physics_server_3d.cpp:903-916 -> BradotSpace3D::test_body_motion): one request through
// bpc_bodies_test_motion; exclude_bodies and exclude_objects become the query exclude list for the duration of the call.
bool BradotPhysicsServerCUDA::body_test_motion(RID p_body, const MotionParameters &p_parameters, MotionResult *r_result) {
	Body *body = body_owner.get_or_null(p_body);
	ERR_FAIL_NULL_V(body, false);
	ERR_FAIL_COND_V(!body->space.is_valid(), false);
	ERR_FAIL_COND_V(p_parameters.max_collisions < 0 || p_parameters.max_collisions > MotionResult::MAX_COLLISIONS, false);
	_flush_pending();
	if (r_result) {
		*r_result = MotionResult();
	}
	if (body->slot < 0) { // no shape: nothing can stop it (test_body_motion's "no shapes found" exit)
		if (r_result) {
			r_result->travel = p_parameters.motion;
		}
		return false;
	}
	LocalVector<int32_t> excluded;
	for (const RID &r : p_parameters.exclude_bodies) {
		int32_t s = body_slot_of(r);
		if (s >= 0) {
			excluded.push_back(s);
		}
		body_part_slots(r, excluded);
	}
	if (!p_parameters.exclude_objects.is_empty()) {
		for (uint32_t s = 0; s < body_of_slot.size(); s++) {
			if (body_of_slot[s] && p_parameters.exclude_objects.has(body_of_slot[s]->instance_id)) {
				excluded.push_back((int32_t)s);
			}
		}
	}
	if (!excluded.is_empty()) {
		ERR_FAIL_COND_V_MSG(bpc_space_set_query_exclude(ctx, excluded.ptr(), (int)excluded.size()) < 0, false, bpc_last_error(ctx));
	}
	bpc_motion_params prm;
	memset(&prm, 0, sizeof(prm));
	prm.body = body->slot;
	_xf_to_floats(p_parameters.from, prm.from);
	prm.motion[0] = p_parameters.motion.x;
	prm.motion[1] = p_parameters.motion.y;
	prm.motion[2] = p_parameters.motion.z;
	prm.margin = p_parameters.margin;
	prm.max_collisions = p_parameters.max_collisions;
	prm.flags = p_parameters.recovery_as_collision ? 2u : 0u;
	bpc_motion_result res;
	int rc = bpc_bodies_test_motion(ctx, 1, &prm, &res);
	if (!excluded.is_empty()) {
		bpc_space_set_query_exclude(ctx, nullptr, 0);
	}
	ERR_FAIL_COND_V_MSG(rc < 0, false, bpc_last_error(ctx));
	if (r_result) {
		r_result->travel = Vector3(res.travel[0], res.travel[1], res.travel[2]);
		r_result->remainder = Vector3(res.remainder[0], res.remainder[1], res.remainder[2]);
		r_result->collision_depth = res.collision_depth;
		r_result->collision_safe_fraction = res.collision_safe_fraction;
		r_result->collision_unsafe_fraction = res.collision_unsafe_fraction;
		r_result->collision_count = res.collided ? res.collision_count : 0;
		for (int i = 0; i < r_result->collision_count; i++) {
			const bpc_motion_collision &c = res.collisions[i];
			MotionCollision &o = r_result->collisions[i];
			o.position = Vector3(c.position[0], c.position[1], c.position[2]);
			o.normal = Vector3(c.normal[0], c.normal[1], c.normal[2]);
			o.collider_velocity = Vector3(c.collider_velocity[0], c.collider_velocity[1], c.collider_velocity[2]);
			o.collider_angular_velocity = Vector3(c.collider_angular_velocity[0], c.collider_angular_velocity[1], c.collider_angular_velocity[2]);
			o.depth = c.depth;
			Body *other = (c.collider >= 0 && (uint32_t)c.collider < body_of_slot.size()) ? body_of_slot[c.collider] : nullptr;
			o.local_shape = 0;
			o.collider_shape = other ? other->shape_index : 0;
			o.collider = other ? other->self : RID();
			o.collider_id = other ? other->instance_id : ObjectID();
		}
	}
	return res.collided != 0;
}

// BradotBody3D::call_queries (bradot_body_3d.cpp:762-781): the force-integration callback, then the state-sync callback,
// both handed the body's direct state
void BradotPhysicsServerCUDA::_body_call_queries(Body *p_body) {
	Variant direct_state_variant = body_get_direct_state(p_body->self);
	if (p_body->fi_callable.is_valid()) {
		const Variant *vp[2] = { &direct_state_variant, &p_body->fi_udata };
		Callable::CallError ce;
		int argc = (p_body->fi_udata.get_type() == Variant::NIL) ? 1 : 2;
		Variant rv;
		p_body->fi_callable.callp(vp, argc, rv, ce);
	}
	if (p_body->state_sync_callback.is_valid()) {
		p_body->state_sync_callback.call(direct_state_variant);
	}
}

#define DIRECT_BODY(m_ret)                                                \
	BradotPhysicsServerCUDA::Body *b = server->body_owner.get_or_null(body); \
	ERR_FAIL_NULL_V(b, m_ret);

#define DIRECT_CONTACT(m_ret)                                   \
	DIRECT_BODY(m_ret);                                         \
	int contact_count = 0;                                      \
	const bpc_contact *c = server->_body_contacts(b, contact_count); \
	ERR_FAIL_INDEX_V(p_contact_idx, contact_count, m_ret);      \
	c += p_contact_idx;

int BradotPhysicsDirectBodyStateCUDA::get_contact_count() const {
	DIRECT_BODY(0);
	int contact_count = 0;
	server->_body_contacts(b, contact_count);
	return contact_count;
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_contact_local_position(int p_contact_idx) const {
	DIRECT_CONTACT(Vector3());
	return Vector3(c->local_pos[0], c->local_pos[1], c->local_pos[2]);
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_contact_local_normal(int p_contact_idx) const {
	DIRECT_CONTACT(Vector3());
	return Vector3(c->local_normal[0], c->local_normal[1], c->local_normal[2]);
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_contact_impulse(int p_contact_idx) const {
	DIRECT_CONTACT(Vector3());
	return Vector3(c->impulse[0], c->impulse[1], c->impulse[2]);
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_contact_local_velocity_at_position(int p_contact_idx) const {
	DIRECT_CONTACT(Vector3());
	return Vector3(c->local_velocity[0], c->local_velocity[1], c->local_velocity[2]);
}
RID BradotPhysicsDirectBodyStateCUDA::get_contact_collider(int p_contact_idx) const {
	DIRECT_CONTACT(RID());
	BradotPhysicsServerCUDA::Body *o = (uint32_t)c->collider < server->body_of_slot.size() ? server->body_of_slot[c->collider] : nullptr;
	return o ? o->self : RID();
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_contact_collider_position(int p_contact_idx) const {
	DIRECT_CONTACT(Vector3());
	return Vector3(c->collider_pos[0], c->collider_pos[1], c->collider_pos[2]);
}
ObjectID BradotPhysicsDirectBodyStateCUDA::get_contact_collider_id(int p_contact_idx) const {
	DIRECT_CONTACT(ObjectID());
	BradotPhysicsServerCUDA::Body *o = (uint32_t)c->collider < server->body_of_slot.size() ? server->body_of_slot[c->collider] : nullptr;
	return o ? o->instance_id : ObjectID();
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_contact_collider_velocity_at_position(int p_contact_idx) const {
	DIRECT_CONTACT(Vector3());
	return Vector3(c->collider_velocity[0], c->collider_velocity[1], c->collider_velocity[2]);
}

Vector3 BradotPhysicsDirectBodyStateCUDA::get_total_gravity() const {
	DIRECT_BODY(Vector3());
	// default area of the body's space only (areas are a "next" row): gravity vector * gravity * gravity_scale
	return Vector3(0, -9.80665, 0) * b->gravity_scale;
}
real_t BradotPhysicsDirectBodyStateCUDA::get_total_angular_damp() const {
	DIRECT_BODY(0);
	return b->ang_damp_replace ? b->ang_damp : 0.1 + b->ang_damp;
}
real_t BradotPhysicsDirectBodyStateCUDA::get_total_linear_damp() const {
	DIRECT_BODY(0);
	return b->lin_damp_replace ? b->lin_damp : 0.1 + b->lin_damp;
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_center_of_mass() const {
	DIRECT_BODY(Vector3());
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	server->_body_mass_world(b, iinv, com, im);
	return com;
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_center_of_mass_local() const {
	DIRECT_BODY(Vector3());
	float m[16];
	ERR_FAIL_COND_V(b->slot < 0 || bpc_bodies_read(server->ctx, BPC_FIELD_MASS_PROPS, b->slot, 1, m) < 0, Vector3());
	return Vector3(m[4], m[5], m[6]);
}
Basis BradotPhysicsDirectBodyStateCUDA::get_principal_inertia_axes() const {
	DIRECT_BODY(Basis());
	float m[16];
	ERR_FAIL_COND_V(b->slot < 0 || bpc_bodies_read(server->ctx, BPC_FIELD_MASS_PROPS, b->slot, 1, m) < 0, Basis());
	return get_transform().basis * Basis(m[7], m[8], m[9], m[10], m[11], m[12], m[13], m[14], m[15]);
}
real_t BradotPhysicsDirectBodyStateCUDA::get_inverse_mass() const {
	DIRECT_BODY(0);
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	server->_body_mass_world(b, iinv, com, im);
	return im;
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_inverse_inertia() const {
	DIRECT_BODY(Vector3());
	float m[16];
	ERR_FAIL_COND_V(b->slot < 0 || bpc_bodies_read(server->ctx, BPC_FIELD_MASS_PROPS, b->slot, 1, m) < 0, Vector3());
	return Vector3(m[1], m[2], m[3]);
}
Basis BradotPhysicsDirectBodyStateCUDA::get_inverse_inertia_tensor() const {
	DIRECT_BODY(Basis());
	Basis iinv;
	Vector3 com;
	real_t im = 0;
	server->_body_mass_world(b, iinv, com, im);
	return iinv;
}
void BradotPhysicsDirectBodyStateCUDA::set_linear_velocity(const Vector3 &p_velocity) {
	server->body_set_state(body, PhysicsServer3D::BODY_STATE_LINEAR_VELOCITY, p_velocity);
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_linear_velocity() const {
	return server->body_get_state(body, PhysicsServer3D::BODY_STATE_LINEAR_VELOCITY);
}
void BradotPhysicsDirectBodyStateCUDA::set_angular_velocity(const Vector3 &p_velocity) {
	server->body_set_state(body, PhysicsServer3D::BODY_STATE_ANGULAR_VELOCITY, p_velocity);
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_angular_velocity() const {
	return server->body_get_state(body, PhysicsServer3D::BODY_STATE_ANGULAR_VELOCITY);
}
void BradotPhysicsDirectBodyStateCUDA::set_transform(const Transform3D &p_transform) {
	server->body_set_state(body, PhysicsServer3D::BODY_STATE_TRANSFORM, p_transform);
}
Transform3D BradotPhysicsDirectBodyStateCUDA::get_transform() const {
	return server->body_get_state(body, PhysicsServer3D::BODY_STATE_TRANSFORM);
}
Vector3 BradotPhysicsDirectBodyStateCUDA::get_velocity_at_local_position(const Vector3 &p_position) const {
	// BradotBody3D::get_velocity_in_local_point (bradot_body_3d.h:217-219)
	return get_linear_velocity() + get_angular_velocity().cross(p_position - get_center_of_mass());
}
void BradotPhysicsDirectBodyStateCUDA::apply_central_impulse(const Vector3 &p_impulse) { server->body_apply_central_impulse(body, p_impulse); }
void BradotPhysicsDirectBodyStateCUDA::apply_impulse(const Vector3 &p_impulse, const Vector3 &p_position) { server->body_apply_impulse(body, p_impulse, p_position); }
void BradotPhysicsDirectBodyStateCUDA::apply_torque_impulse(const Vector3 &p_impulse) { server->body_apply_torque_impulse(body, p_impulse); }
void BradotPhysicsDirectBodyStateCUDA::apply_central_force(const Vector3 &p_force) { server->body_apply_central_force(body, p_force); }
void BradotPhysicsDirectBodyStateCUDA::apply_force(const Vector3 &p_force, const Vector3 &p_position) { server->body_apply_force(body, p_force, p_position); }
void BradotPhysicsDirectBodyStateCUDA::apply_torque(const Vector3 &p_torque) { server->body_apply_torque(body, p_torque); }
void BradotPhysicsDirectBodyStateCUDA::add_constant_central_force(const Vector3 &p_force) { server->body_add_constant_central_force(body, p_force); }
void BradotPhysicsDirectBodyStateCUDA::add_constant_force(const Vector3 &p_force, const Vector3 &p_position) { server->body_add_constant_force(body, p_force, p_position); }
void BradotPhysicsDirectBodyStateCUDA::add_constant_torque(const Vector3 &p_torque) { server->body_add_constant_torque(body, p_torque); }
void BradotPhysicsDirectBodyStateCUDA::set_constant_force(const Vector3 &p_force) { server->body_set_constant_force(body, p_force); }
Vector3 BradotPhysicsDirectBodyStateCUDA::get_constant_force() const { return server->body_get_constant_force(body); }
void BradotPhysicsDirectBodyStateCUDA::set_constant_torque(const Vector3 &p_torque) { server->body_set_constant_torque(body, p_torque); }
Vector3 BradotPhysicsDirectBodyStateCUDA::get_constant_torque() const { return server->body_get_constant_torque(body); }
void BradotPhysicsDirectBodyStateCUDA::set_sleep_state(bool p_sleep) { server->body_set_state(body, PhysicsServer3D::BODY_STATE_SLEEPING, p_sleep); }
bool BradotPhysicsDirectBodyStateCUDA::is_sleeping() const { return server->body_get_state(body, PhysicsServer3D::BODY_STATE_SLEEPING); }
real_t BradotPhysicsDirectBodyStateCUDA::get_step() const { return server->last_step; }
PhysicsDirectSpaceState3D *BradotPhysicsDirectBodyStateCUDA::get_space_state() {
	BradotPhysicsServerCUDA::Body *b = server->body_owner.get_or_null(body);
	ERR_FAIL_NULL_V(b, nullptr);
	return server->space_get_direct_state(b->space);
}

/* JOINTS */

RID BradotPhysicsServerCUDA::joint_create() {
	Joint *joint = memnew(Joint);
	RID rid = joint_owner.make_rid(joint);
	joint->self = rid;
	return rid;
}

void BradotPhysicsServerCUDA::_fill_joint_desc(const Joint *p_joint, bpc_joint_desc &r_desc) {
	const Joint *j = p_joint;
	const Body *a = body_owner.get_or_null(j->a), *b = body_owner.get_or_null(j->b);
	r_desc = {};
	r_desc.type = j->type == JOINT_TYPE_PIN ? BPC_JOINT_PIN : (j->type == JOINT_TYPE_HINGE ? BPC_JOINT_HINGE : (j->type == JOINT_TYPE_SLIDER ? BPC_JOINT_SLIDER : (j->type == JOINT_TYPE_6DOF ? BPC_JOINT_6DOF : BPC_JOINT_CONE_TWIST)));
	r_desc.body_a = a ? a->slot : 0;
	r_desc.body_b = b ? b->slot : 0;
	r_desc.flags = (j->disable_collisions ? 1 : 0) | 8; // parameters are always given
	_xf_to_floats(j->fa, r_desc.frame_a);
	_xf_to_floats(j->fb, r_desc.frame_b);
	if (j->type == JOINT_TYPE_PIN) {
		for (int k = 0; k < 3; k++) {
			r_desc.params[k] = j->pin_params[k];
		}
	} else if (j->type == JOINT_TYPE_SLIDER || j->type == JOINT_TYPE_6DOF) {
		// its values go through bpc_joint_params_write
	} else if (j->type == JOINT_TYPE_CONE_TWIST) {
		for (int k = 0; k < CONE_TWIST_MAX + 1; k++) {
			r_desc.params[k] = j->cone_params[k];
		}
	} else {
		for (int k = 0; k < HINGE_JOINT_MAX; k++) {
			r_desc.params[k] = j->hinge_params[k];
		}
		r_desc.flags |= (j->hinge_flags[HINGE_JOINT_FLAG_USE_LIMIT] ? 2 : 0) | (j->hinge_flags[HINGE_JOINT_FLAG_ENABLE_MOTOR] ? 4 : 0);
	}
	r_desc.priority = j->priority;
}

// a parameter, flag or the priority of an uploaded joint changed: rewrite its record
void BradotPhysicsServerCUDA::_joint_changed(Joint *p_joint) {
	if (p_joint->handle < 0) {
		return;
	}
	bpc_joint_desc d;
	_fill_joint_desc(p_joint, d);
	ERR_FAIL_COND_MSG(bpc_joints_write(ctx, p_joint->handle, 1, &d) < 0, bpc_last_error(ctx));
}

// joint_clear (bradot_physics_server_3d.cpp:1229-1246): the joint becomes an empty one of no type
void BradotPhysicsServerCUDA::joint_clear(RID p_joint) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	if (joint->handle >= 0) {
		bpc_joints_destroy(ctx, joint->handle, 1);
		joint->handle = -1;
	}
	if (joint->pending) {
		pending_joints.erase(joint);
		joint->pending = false;
	}
	joint->type = JOINT_TYPE_MAX;
}

void BradotPhysicsServerCUDA::joint_make_pin(RID p_joint, RID p_body_A, const Vector3 &p_local_A, RID p_body_B, const Vector3 &p_local_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	joint_clear(p_joint);
	joint->type = JOINT_TYPE_PIN;
	joint->a = p_body_A;
	joint->b = p_body_B;
	joint->fa = Transform3D(Basis(), p_local_A);
	joint->fb = Transform3D(Basis(), p_local_B);
	joint->pending = true;
	pending_joints.push_back(joint);
}

void BradotPhysicsServerCUDA::joint_make_hinge(RID p_joint, RID p_body_A, const Transform3D &p_hinge_A, RID p_body_B, const Transform3D &p_hinge_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	joint_clear(p_joint);
	joint->type = JOINT_TYPE_HINGE;
	joint->a = p_body_A;
	joint->b = p_body_B;
	joint->fa = p_hinge_A;
	joint->fb = p_hinge_B;
	joint->pending = true;
	pending_joints.push_back(joint);
}

// joint_make_hinge_simple (bradot_physics_server_3d.cpp:1320-1342): pivots and axes instead of frames; the frames are built as
// the reference's second hinge constructor builds them (joints/bradot_hinge_joint_3d.cpp:69-108) from body A's transform
// at this moment.  (That constructor stores -axis_B in frame B; joint_make_hinge's flips the z column of the frame it is
// given, so +axis_B is handed over here.)
void BradotPhysicsServerCUDA::joint_make_hinge_simple(RID p_joint, RID p_body_A, const Vector3 &p_pivot_A, const Vector3 &p_axis_A, RID p_body_B, const Vector3 &p_pivot_B, const Vector3 &p_axis_B) {
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	const Basis ba = Transform3D(body_get_state(p_body_A, BODY_STATE_TRANSFORM)).basis;
	Vector3 a1 = ba.get_column(0), a2;
	real_t projection = p_axis_A.dot(a1);
	if (projection >= 1.0f - CMP_EPSILON) {
		a1 = -ba.get_column(2);
		a2 = ba.get_column(1);
	} else if (projection <= -1.0f + CMP_EPSILON) {
		a1 = ba.get_column(2);
		a2 = ba.get_column(1);
	} else {
		a2 = p_axis_A.cross(a1);
		a1 = a2.cross(p_axis_A);
	}
	Transform3D fa, fb;
	fa.origin = p_pivot_A;
	fa.basis = Basis(a1.x, a2.x, p_axis_A.x, a1.y, a2.y, p_axis_A.y, a1.z, a2.z, p_axis_A.z);
	Quaternion arc(p_axis_A, p_axis_B);
	Vector3 b1 = arc.xform(a1);
	Vector3 b2 = p_axis_B.cross(b1);
	fb.origin = p_pivot_B;
	fb.basis = Basis(b1.x, b2.x, p_axis_B.x, b1.y, b2.y, p_axis_B.y, b1.z, b2.z, p_axis_B.z);
	joint_make_hinge(p_joint, p_body_A, fa, p_body_B, fb);
}

void BradotPhysicsServerCUDA::joint_disable_collisions_between_bodies(RID p_joint, bool p_disable) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	joint->disable_collisions = p_disable;
	_joint_changed(joint);
}

bool BradotPhysicsServerCUDA::joint_is_disabled_collisions_between_bodies(RID p_joint) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, true);
	return joint->disable_collisions;
}

PhysicsServer3D::JointType BradotPhysicsServerCUDA::joint_get_type(RID p_joint) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, JOINT_TYPE_PIN);
	return joint->type;
}

void BradotPhysicsServerCUDA::joint_set_solver_priority(RID p_joint, int p_priority) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND_MSG(p_priority < 1 || p_priority > 15, "BradotPhysicsCUDA: joint solver priority must be 1..15.");
	joint->priority = p_priority;
	_joint_changed(joint);
}

int BradotPhysicsServerCUDA::joint_get_solver_priority(RID p_joint) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, 0);
	return joint->priority;
}

void BradotPhysicsServerCUDA::pin_joint_set_param(RID p_joint, PinJointParam p_param, real_t p_value) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_PIN);
	ERR_FAIL_INDEX((int)p_param, 3);
	joint->pin_params[p_param] = p_value;
	_joint_changed(joint);
}

real_t BradotPhysicsServerCUDA::pin_joint_get_param(RID p_joint, PinJointParam p_param) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, 0);
	ERR_FAIL_COND_V(joint->type != JOINT_TYPE_PIN, 0);
	ERR_FAIL_INDEX_V((int)p_param, 3, 0);
	return joint->pin_params[p_param];
}

void BradotPhysicsServerCUDA::pin_joint_set_local_a(RID p_joint, const Vector3 &p_A) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_PIN);
	joint->fa.origin = p_A;
	_joint_changed(joint);
}
Vector3 BradotPhysicsServerCUDA::pin_joint_get_local_a(RID p_joint) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, Vector3());
	return joint->fa.origin;
}
void BradotPhysicsServerCUDA::pin_joint_set_local_b(RID p_joint, const Vector3 &p_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_PIN);
	joint->fb.origin = p_B;
	_joint_changed(joint);
}
Vector3 BradotPhysicsServerCUDA::pin_joint_get_local_b(RID p_joint) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, Vector3());
	return joint->fb.origin;
}

void BradotPhysicsServerCUDA::hinge_joint_set_param(RID p_joint, HingeJointParam p_param, real_t p_value) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_HINGE);
	ERR_FAIL_INDEX((int)p_param, (int)HINGE_JOINT_MAX);
	joint->hinge_params[p_param] = p_value;
	_joint_changed(joint);
}

real_t BradotPhysicsServerCUDA::hinge_joint_get_param(RID p_joint, HingeJointParam p_param) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, 0);
	ERR_FAIL_COND_V(joint->type != JOINT_TYPE_HINGE, 0);
	ERR_FAIL_INDEX_V((int)p_param, (int)HINGE_JOINT_MAX, 0);
	return joint->hinge_params[p_param];
}

void BradotPhysicsServerCUDA::hinge_joint_set_flag(RID p_joint, HingeJointFlag p_flag, bool p_enabled) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_HINGE);
	ERR_FAIL_INDEX((int)p_flag, (int)HINGE_JOINT_FLAG_MAX);
	joint->hinge_flags[p_flag] = p_enabled;
	_joint_changed(joint);
}

bool BradotPhysicsServerCUDA::hinge_joint_get_flag(RID p_joint, HingeJointFlag p_flag) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, false);
	ERR_FAIL_COND_V(joint->type != JOINT_TYPE_HINGE, false);
	ERR_FAIL_INDEX_V((int)p_flag, (int)HINGE_JOINT_FLAG_MAX, false);
	return joint->hinge_flags[p_flag];
}

// joint_make_slider / slider_joint_*_param (bradot_physics_server_3d.cpp:1421-1459; the joint itself:
// joints/bradot_slider_joint_3d.cpp)
void BradotPhysicsServerCUDA::joint_make_slider(RID p_joint, RID p_body_A, const Transform3D &p_local_frame_A, RID p_body_B, const Transform3D &p_local_frame_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	joint_clear(p_joint);
	joint->type = JOINT_TYPE_SLIDER;
	joint->a = p_body_A;
	joint->b = p_body_B;
	joint->fa = p_local_frame_A;
	joint->fb = p_local_frame_B;
	joint->pending = true;
	pending_joints.push_back(joint);
}

void BradotPhysicsServerCUDA::slider_joint_set_param(RID p_joint, SliderJointParam p_param, real_t p_value) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_SLIDER);
	ERR_FAIL_INDEX((int)p_param, (int)SLIDER_JOINT_MAX);
	joint->slider_params[p_param] = p_value;
	if (joint->handle >= 0) {
		ERR_FAIL_COND_MSG(bpc_joint_params_write(ctx, joint->handle, (int)p_param, 1, &joint->slider_params[p_param]) < 0, bpc_last_error(ctx));
	}
}

real_t BradotPhysicsServerCUDA::slider_joint_get_param(RID p_joint, SliderJointParam p_param) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, 0);
	ERR_FAIL_COND_V(joint->type != JOINT_TYPE_SLIDER, 0);
	ERR_FAIL_INDEX_V((int)p_param, (int)SLIDER_JOINT_MAX, 0);
	return joint->slider_params[p_param];
}

// joint_make_generic_6dof / generic_6dof_joint_* (bradot_physics_server_3d.cpp:1501-1555; the joint itself:
// joints/bradot_generic_6dof_joint_3d.cpp).  The table holds the joint's members; set_param / get_param map the server's
// parameters on them exactly as the reference's switch does (:451-601) -- its quirks included: ANGULAR_MOTOR_FORCE_LIMIT
// writes the LIMIT force and reads the motor force, linear motors and all springs are ignored and read as 0.
void BradotPhysicsServerCUDA::joint_make_generic_6dof(RID p_joint, RID p_body_A, const Transform3D &p_local_frame_A, RID p_body_B, const Transform3D &p_local_frame_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	joint_clear(p_joint);
	joint->type = JOINT_TYPE_6DOF;
	joint->a = p_body_A;
	joint->b = p_body_B;
	joint->fa = p_local_frame_A;
	joint->fb = p_local_frame_B;
	for (int ax = 0; ax < 3; ax++) {
		float *q = joint->g6_params + ax * BPC_G6DOF_AXIS_STRIDE;
		for (int k = 0; k < BPC_G6DOF_AXIS_STRIDE; k++) {
			q[k] = 0.0f;
		}
		q[G6DOF_JOINT_LINEAR_LIMIT_SOFTNESS] = 0.7f;
		q[G6DOF_JOINT_LINEAR_RESTITUTION] = 0.5f;
		q[G6DOF_JOINT_LINEAR_DAMPING] = 1.0f;
		q[G6DOF_JOINT_ANGULAR_LOWER_LIMIT] = -1e30f;
		q[G6DOF_JOINT_ANGULAR_UPPER_LIMIT] = 1e30f;
		q[G6DOF_JOINT_ANGULAR_LIMIT_SOFTNESS] = 0.5f;
		q[G6DOF_JOINT_ANGULAR_DAMPING] = 1.0f;
		q[G6DOF_JOINT_ANGULAR_FORCE_LIMIT] = 300.0f;
		q[G6DOF_JOINT_ANGULAR_ERP] = 0.5f;
		q[G6DOF_JOINT_ANGULAR_MOTOR_FORCE_LIMIT] = 0.1f;
		q[G6DOF_JOINT_MAX + G6DOF_JOINT_FLAG_ENABLE_LINEAR_LIMIT] = 1.0f;
	}
	joint->pending = true;
	pending_joints.push_back(joint);
}

static int _g6_slot(PhysicsServer3D::G6DOFJointAxisParam p_param, bool p_write) {
	switch (p_param) {
		case PhysicsServer3D::G6DOF_JOINT_LINEAR_LOWER_LIMIT:
		case PhysicsServer3D::G6DOF_JOINT_LINEAR_UPPER_LIMIT:
		case PhysicsServer3D::G6DOF_JOINT_LINEAR_LIMIT_SOFTNESS:
		case PhysicsServer3D::G6DOF_JOINT_LINEAR_RESTITUTION:
		case PhysicsServer3D::G6DOF_JOINT_LINEAR_DAMPING:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_LOWER_LIMIT:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_UPPER_LIMIT:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_LIMIT_SOFTNESS:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_DAMPING:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_RESTITUTION:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_FORCE_LIMIT:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_ERP:
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_MOTOR_TARGET_VELOCITY:
			return (int)p_param;
		case PhysicsServer3D::G6DOF_JOINT_ANGULAR_MOTOR_FORCE_LIMIT:
			return p_write ? (int)PhysicsServer3D::G6DOF_JOINT_ANGULAR_FORCE_LIMIT : (int)p_param;
		default:
			return -1; // not implemented in the reference backend
	}
}

void BradotPhysicsServerCUDA::generic_6dof_joint_set_param(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisParam p_param, real_t p_value) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_6DOF);
	ERR_FAIL_INDEX((int)p_axis, 3);
	int slot = _g6_slot(p_param, true);
	if (slot < 0) {
		return;
	}
	int idx = (int)p_axis * BPC_G6DOF_AXIS_STRIDE + slot;
	joint->g6_params[idx] = p_value;
	if (joint->handle >= 0) {
		ERR_FAIL_COND_MSG(bpc_joint_params_write(ctx, joint->handle, idx, 1, &joint->g6_params[idx]) < 0, bpc_last_error(ctx));
	}
}

real_t BradotPhysicsServerCUDA::generic_6dof_joint_get_param(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisParam p_param) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, 0);
	ERR_FAIL_COND_V(joint->type != JOINT_TYPE_6DOF, 0);
	ERR_FAIL_INDEX_V((int)p_axis, 3, 0);
	int slot = _g6_slot(p_param, false);
	return slot < 0 ? real_t(0) : real_t(joint->g6_params[(int)p_axis * BPC_G6DOF_AXIS_STRIDE + slot]);
}

void BradotPhysicsServerCUDA::generic_6dof_joint_set_flag(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisFlag p_flag, bool p_enable) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_6DOF);
	ERR_FAIL_INDEX((int)p_axis, 3);
	if (p_flag != G6DOF_JOINT_FLAG_ENABLE_LINEAR_LIMIT && p_flag != G6DOF_JOINT_FLAG_ENABLE_ANGULAR_LIMIT && p_flag != G6DOF_JOINT_FLAG_ENABLE_MOTOR) {
		return; // not implemented in the reference backend
	}
	int idx = (int)p_axis * BPC_G6DOF_AXIS_STRIDE + (int)G6DOF_JOINT_MAX + (int)p_flag;
	joint->g6_params[idx] = p_enable ? 1.0f : 0.0f;
	if (joint->handle >= 0) {
		ERR_FAIL_COND_MSG(bpc_joint_params_write(ctx, joint->handle, idx, 1, &joint->g6_params[idx]) < 0, bpc_last_error(ctx));
	}
}

bool BradotPhysicsServerCUDA::generic_6dof_joint_get_flag(RID p_joint, Vector3::Axis p_axis, G6DOFJointAxisFlag p_flag) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, false);
	ERR_FAIL_COND_V(joint->type != JOINT_TYPE_6DOF, false);
	ERR_FAIL_INDEX_V((int)p_axis, 3, false);
	if (p_flag != G6DOF_JOINT_FLAG_ENABLE_LINEAR_LIMIT && p_flag != G6DOF_JOINT_FLAG_ENABLE_ANGULAR_LIMIT && p_flag != G6DOF_JOINT_FLAG_ENABLE_MOTOR) {
		return false;
	}
	return joint->g6_params[(int)p_axis * BPC_G6DOF_AXIS_STRIDE + (int)G6DOF_JOINT_MAX + (int)p_flag] != 0.0f;
}

// joint_make_cone_twist / cone_twist_joint_*_param (bradot_physics_server_3d.cpp:1461-1499; the joint itself:
// joints/bradot_cone_twist_joint_3d.cpp)
void BradotPhysicsServerCUDA::joint_make_cone_twist(RID p_joint, RID p_body_A, const Transform3D &p_local_frame_A, RID p_body_B, const Transform3D &p_local_frame_B) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_A));
	ERR_FAIL_NULL(body_owner.get_or_null(p_body_B));
	joint_clear(p_joint);
	joint->type = JOINT_TYPE_CONE_TWIST;
	joint->a = p_body_A;
	joint->b = p_body_B;
	joint->fa = p_local_frame_A;
	joint->fb = p_local_frame_B;
	joint->pending = true;
	pending_joints.push_back(joint);
}

void BradotPhysicsServerCUDA::cone_twist_joint_set_param(RID p_joint, ConeTwistJointParam p_param, real_t p_value) {
	Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL(joint);
	ERR_FAIL_COND(joint->type != JOINT_TYPE_CONE_TWIST);
	ERR_FAIL_INDEX((int)p_param, (int)CONE_TWIST_MAX);
	joint->cone_params[p_param] = p_value;
	if (p_param == CONE_TWIST_JOINT_SWING_SPAN) { // both spans (bradot_cone_twist_joint_3d.cpp:276-279)
		joint->cone_params[CONE_TWIST_MAX] = p_value;
	}
	_joint_changed(joint);
}

real_t BradotPhysicsServerCUDA::cone_twist_joint_get_param(RID p_joint, ConeTwistJointParam p_param) const {
	const Joint *joint = joint_owner.get_or_null(p_joint);
	ERR_FAIL_NULL_V(joint, 0);
	ERR_FAIL_COND_V(joint->type != JOINT_TYPE_CONE_TWIST, 0);
	ERR_FAIL_INDEX_V((int)p_param, (int)CONE_TWIST_MAX, 0);
	return joint->cone_params[p_param];
}

/* LIFECYCLE */

// free(RID) of any kind (bradot_physics_server_3d.cpp:1557-1622).  A body leaves the GPU side with its slot (reused by the
// next upload); freeing a space frees what is in it.

// ---------------------------------------------------------------------------------------------- areas
// PhysicsServer3D::area_* (servers/physics_server_3d.h:335-381; reference P3D/bradot_physics_server_3d.cpp:215-436).  An area is
// a host record like a body; it goes to the backend once it has a space and a shape (bpc_areas_create), every later setter
// marks it and the whole record is written again before the next step (bpc_areas_write, which also counts as "the area moved").
// A space RID in place of an area RID addresses the space's default area, as in the reference (:291-295).
RID BradotPhysicsServerCUDA::area_create() {
	Area *area = memnew(Area);
	RID rid = area_owner.make_rid(area);
	area->self = rid;
	return rid;
}

void BradotPhysicsServerCUDA::_fill_area_desc(const Area *p_part, bpc_area_desc &r_desc) {
	memset(&r_desc, 0, sizeof(r_desc));
	const Area *p_area = p_part->parent ? p_part->parent : p_part; // a further shape: the area's own values, this record's shape
	const Space *space = space_owner.get_or_null(p_area->space);
	Shape *shape = shape_owner.get_or_null(p_part->shape);
	r_desc.group = p_area->group;
	r_desc.space = space ? space->handle : -1;
	r_desc.shape = shape ? _shape_handle(shape) : -1;
	r_desc.flags = (p_area->ray_pickable ? 16u : 0u) | (p_area->monitorable ? 1u : 0u) | (p_area->monitor_callback.is_valid() ? 2u : 0u) | (p_area->gravity_is_point ? 4u : 0u) |
			(p_area->area_monitor_callback.is_valid() ? 8u : 0u);
	r_desc.priority = p_area->priority;
	_xf_to_floats(p_area->xf, r_desc.xf);
	_xf_to_floats(p_part->shape_xf, r_desc.shape_xf);
	r_desc.layer = p_area->layer;
	r_desc.mask = p_area->mask;
	r_desc.gravity_mode = p_area->gravity_mode;
	r_desc.linear_damp_mode = p_area->linear_damp_mode;
	r_desc.angular_damp_mode = p_area->angular_damp_mode;
	r_desc.gravity = p_area->gravity;
	r_desc.gravity_vector[0] = p_area->gravity_vector.x;
	r_desc.gravity_vector[1] = p_area->gravity_vector.y;
	r_desc.gravity_vector[2] = p_area->gravity_vector.z;
	r_desc.point_unit_distance = p_area->point_unit_distance;
	r_desc.linear_damp = p_area->linear_damp;
	r_desc.angular_damp = p_area->angular_damp;
}

// a setter touched the area: upload it if it can be now, else write it again before the next step
BradotPhysicsServerCUDA::Area *BradotPhysicsServerCUDA::_area_shape_record(Area *p_area, int p_shape_idx) const {
	if (p_shape_idx == 0) {
		return p_area->shape.is_valid() ? p_area : nullptr;
	}
	return (p_shape_idx > 0 && (uint32_t)p_shape_idx <= p_area->parts.size()) ? p_area->parts[p_shape_idx - 1] : nullptr;
}

void BradotPhysicsServerCUDA::_area_changed(Area *p_area) {
	for (Area *part : p_area->parts) {
		_area_changed(part);
	}
	const bool wanted = (p_area->parent ? p_area->parent->space : p_area->space).is_valid() && p_area->shape.is_valid() && !p_area->shape_disabled;
	if (p_area->handle >= 0) {
		if (!wanted) {
			_area_leave_space(p_area);
		} else if (!p_area->dirty) {
			p_area->dirty = true;
			dirty_areas.push_back(p_area);
		}
	} else if (wanted && !p_area->pending) {
		p_area->pending = true;
		pending_areas.push_back(p_area);
	}
}

void BradotPhysicsServerCUDA::_area_leave_space(Area *p_area) {
	for (Area *part : p_area->parts) {
		_area_leave_space(part);
	}
	if (p_area->pending) {
		pending_areas.erase(p_area);
		p_area->pending = false;
	}
	if (p_area->dirty) {
		dirty_areas.erase(p_area);
		p_area->dirty = false;
	}
	if (p_area->handle >= 0) {
		int32_t h = p_area->handle;
		ERR_FAIL_COND_MSG(bpc_areas_destroy(ctx, &h, 1) < 0, bpc_last_error(ctx));
		FreedBody fa; // an area that watched it still has to be told who left
		fa.rid = p_area->self;
		fa.instance_id = p_area->instance_id;
		freed_areas[h] = fa;
		area_of_handle[h] = nullptr;
		p_area->handle = -1;
	}
}

void BradotPhysicsServerCUDA::area_set_space(RID p_area, RID p_space) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	if (p_space.is_valid()) {
		ERR_FAIL_NULL(space_owner.get_or_null(p_space));
	}
	if (area->space == p_space) {
		return;
	}
	_area_leave_space(area); // an area does not change space on the backend: it leaves and is created anew
	area->space = p_space;
	_area_changed(area);
}

RID BradotPhysicsServerCUDA::area_get_space(RID p_area) const {
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, RID());
	return area->space;
}

void BradotPhysicsServerCUDA::area_add_shape(RID p_area, RID p_shape, const Transform3D &p_transform, bool p_disabled) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	ERR_FAIL_NULL(shape_owner.get_or_null(p_shape));
	if (area->shape.is_valid()) { // a further shape: a record of its own that shares the area's group
		Area *part = memnew(Area);
		part->parent = area;
		part->self = area->self;
		part->instance_id = area->instance_id;
		part->shape = p_shape;
		part->shape_xf = p_transform;
		part->shape_disabled = p_disabled;
		part->shape_index = area->parts.size() + 1;
		if (area->group == 0) {
			area->group = ++area_group_counter;
		}
		area->parts.push_back(part);
		_area_changed(area); // (the area itself is described anew with its group, the part is uploaded)
		return;
	}
	area->shape = p_shape;
	area->shape_xf = p_transform;
	area->shape_disabled = p_disabled;
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_set_shape(RID p_area, int p_shape_idx, RID p_shape) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	Area *rec = _area_shape_record(area, p_shape_idx);
	ERR_FAIL_NULL(rec);
	ERR_FAIL_NULL(shape_owner.get_or_null(p_shape));
	rec->shape = p_shape;
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_set_shape_transform(RID p_area, int p_shape_idx, const Transform3D &p_transform) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	Area *rec = _area_shape_record(area, p_shape_idx);
	ERR_FAIL_NULL(rec);
	rec->shape_xf = p_transform;
	_area_changed(area);
}

int BradotPhysicsServerCUDA::area_get_shape_count(RID p_area) const {
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, -1);
	return area->shape.is_valid() ? 1 + (int)area->parts.size() : 0;
}

RID BradotPhysicsServerCUDA::area_get_shape(RID p_area, int p_shape_idx) const {
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, RID());
	const Area *rec = _area_shape_record(const_cast<Area *>(area), p_shape_idx);
	ERR_FAIL_NULL_V(rec, RID());
	return rec->shape;
}

Transform3D BradotPhysicsServerCUDA::area_get_shape_transform(RID p_area, int p_shape_idx) const {
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, Transform3D());
	const Area *rec = _area_shape_record(const_cast<Area *>(area), p_shape_idx);
	ERR_FAIL_NULL_V(rec, Transform3D());
	return rec->shape_xf;
}

void BradotPhysicsServerCUDA::area_remove_shape(RID p_area, int p_shape_idx) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	ERR_FAIL_NULL(_area_shape_record(area, p_shape_idx));
	if (area->parts.is_empty()) {
		area->shape = RID();
		_area_changed(area);
		return;
	}
	// the shapes behind it move down one index: shift them through the records, drop the last one
	for (int i = p_shape_idx; i < (int)area->parts.size(); i++) {
		Area *dst = _area_shape_record(area, i), *src = area->parts[i];
		dst->shape = src->shape;
		dst->shape_xf = src->shape_xf;
		dst->shape_disabled = src->shape_disabled;
	}
	Area *last = area->parts[area->parts.size() - 1];
	_area_leave_space(last);
	memdelete(last);
	area->parts.resize(area->parts.size() - 1);
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_clear_shapes(RID p_area) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	for (Area *part : area->parts) {
		_area_leave_space(part);
		memdelete(part);
	}
	area->parts.clear();
	area->shape = RID();
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_set_shape_disabled(RID p_area, int p_shape_idx, bool p_disabled) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	Area *rec = _area_shape_record(area, p_shape_idx);
	ERR_FAIL_NULL(rec);
	rec->shape_disabled = p_disabled;
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_attach_object_instance_id(RID p_area, ObjectID p_id) {
	if (space_owner.owns(p_area)) {
		return; // the default area of a space is nobody's node
	}
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	area->instance_id = p_id;
	for (Area *part : area->parts) {
		part->instance_id = p_id;
	}
}

ObjectID BradotPhysicsServerCUDA::area_get_object_instance_id(RID p_area) const {
	if (space_owner.owns(p_area)) {
		return ObjectID();
	}
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, ObjectID());
	return area->instance_id;
}


void BradotPhysicsServerCUDA::area_set_param(RID p_area, AreaParameter p_param, const Variant &p_value) {
	if (space_owner.owns(p_area)) { // the default area: what bodies get where no area settles gravity / damping
		Space *space = space_owner.get_or_null(p_area);
		switch (p_param) {
			case AREA_PARAM_GRAVITY: space->default_gravity = p_value; break;
			case AREA_PARAM_GRAVITY_VECTOR: space->default_gravity_vector = p_value; break;
			case AREA_PARAM_LINEAR_DAMP: space->default_linear_damp = p_value; break;
			case AREA_PARAM_ANGULAR_DAMP: space->default_angular_damp = p_value; break;
			default: return; // override modes, point gravity, priority and wind mean nothing on the default area of this backend
		}
		// compute_gravity of a non-point area: vector * magnitude in fp32 (P3D/bradot_area_3d.cpp:330)
		Vector3 g = space->default_gravity_vector * space->default_gravity;
		bpc_space_set_param(ctx, space->handle, BPC_SPACE_GRAVITY_X, g.x);
		bpc_space_set_param(ctx, space->handle, BPC_SPACE_GRAVITY_Y, g.y);
		bpc_space_set_param(ctx, space->handle, BPC_SPACE_GRAVITY_Z, g.z);
		bpc_space_set_param(ctx, space->handle, BPC_SPACE_LINEAR_DAMP, space->default_linear_damp);
		bpc_space_set_param(ctx, space->handle, BPC_SPACE_ANGULAR_DAMP, space->default_angular_damp);
		return;
	}
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	switch (p_param) {
		case AREA_PARAM_GRAVITY_OVERRIDE_MODE: area->gravity_mode = p_value; break;
		case AREA_PARAM_GRAVITY: area->gravity = p_value; break;
		case AREA_PARAM_GRAVITY_VECTOR: area->gravity_vector = p_value; break;
		case AREA_PARAM_GRAVITY_IS_POINT: area->gravity_is_point = p_value; break;
		case AREA_PARAM_GRAVITY_POINT_UNIT_DISTANCE: area->point_unit_distance = p_value; break;
		case AREA_PARAM_LINEAR_DAMP_OVERRIDE_MODE: area->linear_damp_mode = p_value; break;
		case AREA_PARAM_LINEAR_DAMP: area->linear_damp = p_value; break;
		case AREA_PARAM_ANGULAR_DAMP_OVERRIDE_MODE: area->angular_damp_mode = p_value; break;
		case AREA_PARAM_ANGULAR_DAMP: area->angular_damp = p_value; break;
		case AREA_PARAM_PRIORITY: area->priority = p_value; break;
		case AREA_PARAM_WIND_FORCE_MAGNITUDE: area->wind[0] = p_value; return;
		case AREA_PARAM_WIND_SOURCE: area->wind[1] = p_value; return;
		case AREA_PARAM_WIND_DIRECTION: area->wind[2] = p_value; return;
		case AREA_PARAM_WIND_ATTENUATION_FACTOR: area->wind[3] = p_value; return;
	}
	_area_changed(area);
}

Variant BradotPhysicsServerCUDA::area_get_param(RID p_area, AreaParameter p_param) const {
	if (space_owner.owns(p_area)) {
		const Space *space = space_owner.get_or_null(p_area);
		switch (p_param) {
			case AREA_PARAM_GRAVITY: return space->default_gravity;
			case AREA_PARAM_GRAVITY_VECTOR: return space->default_gravity_vector;
			case AREA_PARAM_LINEAR_DAMP: return space->default_linear_damp;
			case AREA_PARAM_ANGULAR_DAMP: return space->default_angular_damp;
			case AREA_PARAM_GRAVITY_IS_POINT: return false;
			default: return 0;
		}
	}
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, Variant());
	switch (p_param) {
		case AREA_PARAM_GRAVITY_OVERRIDE_MODE: return area->gravity_mode;
		case AREA_PARAM_GRAVITY: return area->gravity;
		case AREA_PARAM_GRAVITY_VECTOR: return area->gravity_vector;
		case AREA_PARAM_GRAVITY_IS_POINT: return area->gravity_is_point;
		case AREA_PARAM_GRAVITY_POINT_UNIT_DISTANCE: return area->point_unit_distance;
		case AREA_PARAM_LINEAR_DAMP_OVERRIDE_MODE: return area->linear_damp_mode;
		case AREA_PARAM_LINEAR_DAMP: return area->linear_damp;
		case AREA_PARAM_ANGULAR_DAMP_OVERRIDE_MODE: return area->angular_damp_mode;
		case AREA_PARAM_ANGULAR_DAMP: return area->angular_damp;
		case AREA_PARAM_PRIORITY: return area->priority;
		case AREA_PARAM_WIND_FORCE_MAGNITUDE: return area->wind[0];
		case AREA_PARAM_WIND_SOURCE: return area->wind[1];
		case AREA_PARAM_WIND_DIRECTION: return area->wind[2];
		case AREA_PARAM_WIND_ATTENUATION_FACTOR: return area->wind[3];
	}
	return Variant();
}

void BradotPhysicsServerCUDA::area_set_transform(RID p_area, const Transform3D &p_transform) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	area->xf = p_transform;
	_area_changed(area);
}

Transform3D BradotPhysicsServerCUDA::area_get_transform(RID p_area) const {
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, Transform3D());
	return area->xf;
}

void BradotPhysicsServerCUDA::area_set_collision_layer(RID p_area, uint32_t p_layer) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	area->layer = p_layer;
	_area_changed(area);
}

uint32_t BradotPhysicsServerCUDA::area_get_collision_layer(RID p_area) const {
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, 0);
	return area->layer;
}

void BradotPhysicsServerCUDA::area_set_collision_mask(RID p_area, uint32_t p_mask) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	area->mask = p_mask;
	_area_changed(area);
}

uint32_t BradotPhysicsServerCUDA::area_get_collision_mask(RID p_area) const {
	const Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL_V(area, 0);
	return area->mask;
}

void BradotPhysicsServerCUDA::area_set_monitorable(RID p_area, bool p_monitorable) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	if (area->monitorable == p_monitorable) {
		return;
	}
	area->monitorable = p_monitorable;
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_set_monitor_callback(RID p_area, const Callable &p_callback) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	area->monitor_callback = p_callback.is_valid() ? p_callback : Callable();
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_set_area_monitor_callback(RID p_area, const Callable &p_callback) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	area->area_monitor_callback = p_callback.is_valid() ? p_callback : Callable();
	_area_changed(area);
}

void BradotPhysicsServerCUDA::area_set_ray_pickable(RID p_area, bool p_enable) {
	Area *area = area_owner.get_or_null(p_area);
	ERR_FAIL_NULL(area);
	area->ray_pickable = p_enable;
	_area_changed(area);
}

// BradotArea3D::call_queries (P3D/bradot_area_3d.cpp:232-271): what happened to each (area, body) since the last flush, summed
// (entered and left again = nothing), handed to the area's monitor callback as (status, body RID, object id, body shape, area shape)
void BradotPhysicsServerCUDA::_flush_area_events() {
	if (area_owner.get_rid_count() == 0 && freed_bodies.is_empty()) {
		return;
	}
	LocalVector<bpc_area_event> ev;
	ev.resize(4096);
	int n = bpc_area_events_read(ctx, ev.ptr(), (int)ev.size());
	if (n > (int)ev.size()) { // more than fitted: the rest waited in the backend
		uint32_t had = ev.size();
		ev.resize(n);
		int more = bpc_area_events_read(ctx, ev.ptr() + had, n - (int)had);
		n = (int)had + MAX(more, 0);
	}
	if (n <= 0) {
		freed_bodies.clear();
		freed_areas.clear();
		return;
	}
	HashMap<uint64_t, int> state; // (area << 32 | body or other area, bit 31 = other area) -> entered minus left
	LocalVector<uint64_t> order;
	for (int i = 0; i < n && i < (int)ev.size(); i++) {
		uint64_t key = ((uint64_t)(uint32_t)ev[i].area << 32) | (uint32_t)ev[i].body | (ev[i].status >= 2 ? 0x80000000u : 0u);
		if (!state.has(key)) {
			state[key] = 0;
			order.push_back(key);
		}
		state[key] += (ev[i].status & 1) == 0 ? 1 : -1;
	}
	for (uint64_t key : order) {
		int s = state[key];
		if (s == 0) {
			continue;
		}
		uint32_t ai = (uint32_t)(key >> 32);
		Area *record = ai < area_of_handle.size() ? area_of_handle[ai] : nullptr;
		if (!record) {
			continue;
		}
		Area *area = record->parent ? record->parent : record; // a further shape reports to the area it belongs to, with its index
		const int area_shape = record->shape_index;
		if (key & 0x80000000u) { // an area entered / left: the area monitor callback (P3D/bradot_area_3d.cpp:273-310)
			uint32_t oi = (uint32_t)key & 0x7fffffffu;
			if (!area->area_monitor_callback.is_valid()) {
				continue;
			}
			RID other_rid;
			ObjectID other_id;
			if (oi < area_of_handle.size() && area_of_handle[oi]) {
				other_rid = area_of_handle[oi]->self;
				other_id = area_of_handle[oi]->instance_id;
			} else if (freed_areas.has((int32_t)oi)) {
				other_rid = freed_areas[(int32_t)oi].rid;
				other_id = freed_areas[(int32_t)oi].instance_id;
			} else {
				continue;
			}
			const int other_shape = (oi < area_of_handle.size() && area_of_handle[oi]) ? area_of_handle[oi]->shape_index : 0;
			Variant res[5] = { s > 0 ? AREA_BODY_ADDED : AREA_BODY_REMOVED, other_rid, other_id, other_shape, area_shape };
			const Variant *resptr[5] = { &res[0], &res[1], &res[2], &res[3], &res[4] };
			Callable::CallError ce;
			Variant ret;
			area->area_monitor_callback.callp(resptr, 5, ret, ce);
			continue;
		}
		int32_t slot = (int32_t)(uint32_t)key;
		if (!area->monitor_callback.is_valid()) {
			continue;
		}
		RID body_rid;
		ObjectID body_id;
		Body *b = (slot >= 0 && (uint32_t)slot < body_of_slot.size()) ? body_of_slot[slot] : nullptr;
		if (freed_bodies.has(slot) && s < 0) { // it left by being freed: the event still names it
			body_rid = freed_bodies[slot].rid;
			body_id = freed_bodies[slot].instance_id;
		} else if (b) {
			body_rid = b->self;
			body_id = b->instance_id;
		} else {
			continue;
		}
		Variant res[5] = { s > 0 ? AREA_BODY_ADDED : AREA_BODY_REMOVED, body_rid, body_id, b ? b->shape_index : 0, area_shape };
		const Variant *resptr[5] = { &res[0], &res[1], &res[2], &res[3], &res[4] };
		Callable::CallError ce;
		Variant ret;
		area->monitor_callback.callp(resptr, 5, ret, ce);
		if (ce.error != Callable::CallError::CALL_OK) {
			ERR_PRINT_ONCE("Error calling monitor callback method " + Variant::get_callable_error_text(area->monitor_callback, resptr, 5, ce));
		}
	}
	freed_bodies.clear();
	freed_areas.clear();
}

void BradotPhysicsServerCUDA::free(RID p_rid) {
	if (shape_owner.owns(p_rid)) {
		Shape *s = shape_owner.get_or_null(p_rid);
		shape_owner.free(p_rid);
		memdelete(s);
	} else if (body_owner.owns(p_rid)) {
		Body *b = body_owner.get_or_null(p_rid);
		reporting_bodies -= b->max_contacts_reported > 0 ? 1 : 0;
		_body_leave_space(b);
		for (Body *part : b->parts) {
			memdelete(part);
		}
		b->parts.clear();
		if (b->pushed) {
			pushed_bodies.erase(b);
		}
		if (b->has_callbacks) {
			callback_bodies.erase(b);
		}
		// joints of the body stay valid RIDs of no effect, as in the reference until they are freed themselves
		List<RID> joints;
		joint_owner.get_owned_list(&joints);
		for (const RID &jr : joints) {
			Joint *j = joint_owner.get_or_null(jr);
			if (j && (j->a == p_rid || j->b == p_rid)) {
				joint_clear(jr);
			}
		}
		if (b->direct_state) {
			memdelete(b->direct_state);
		}
		body_owner.free(p_rid);
		memdelete(b);
	} else if (space_owner.owns(p_rid)) {
		Space *s = space_owner.get_or_null(p_rid);
		List<RID> bodies;
		body_owner.get_owned_list(&bodies);
		for (const RID &br : bodies) {
			Body *b = body_owner.get_or_null(br);
			if (b && b->space == p_rid) { // the body outlives its space, outside of any (co->set_space(nullptr))
				_body_leave_space(b);
				b->space = RID();
			}
		}
		List<RID> areas;
		area_owner.get_owned_list(&areas);
		for (const RID &ar : areas) {
			Area *a = area_owner.get_or_null(ar);
			if (a && a->space == p_rid) {
				_area_leave_space(a);
				a->space = RID();
			}
		}
		bpc_space_destroy(ctx, s->handle);
		if (s->direct_state) {
			memdelete(s->direct_state);
		}
		space_owner.free(p_rid);
		memdelete(s);
	} else if (joint_owner.owns(p_rid)) {
		Joint *j = joint_owner.get_or_null(p_rid);
		joint_clear(p_rid);
		joint_owner.free(p_rid);
		memdelete(j);
	} else if (area_owner.owns(p_rid)) {
		Area *a = area_owner.get_or_null(p_rid);
		_area_leave_space(a);
		for (Area *part : a->parts) {
			memdelete(part);
		}
		a->parts.clear();
		area_owner.free(p_rid);
		memdelete(a);
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

// Everything created or re-described since the last tick goes to the GPU in bulk: new bodies in one bpc_bodies_create_slots
// (freed slots first), changed ones in one bpc_bodies_update, new joints in one bpc_joints_create.
void BradotPhysicsServerCUDA::_flush_pending() {
	if (!pending_bodies.is_empty()) {
		LocalVector<bpc_body_desc> descs;
		LocalVector<Body *> uploaded, waiting;
		for (Body *b : pending_bodies) {
			Shape *shape = shape_owner.get_or_null(b->shape);
			Space *space = space_owner.get_or_null(b->space);
			if (!space) {
				b->pending = false;
				continue;
			}
			if (!shape || _shape_handle(shape) < 0) {
				waiting.push_back(b); // shapeless bodies (the reference's static global body) take no part in the step
				continue;
			}
			bpc_body_desc d;
			_fill_desc(b, d);
			descs.push_back(d);
			uploaded.push_back(b);
		}
		pending_bodies.clear();
		for (Body *b : waiting) {
			pending_bodies.push_back(b);
		}
		if (!descs.is_empty()) {
			LocalVector<int32_t> slots;
			slots.resize(descs.size());
			ERR_FAIL_COND_MSG(bpc_bodies_create_slots(ctx, descs.ptr(), descs.size(), slots.ptr()) < 0, bpc_last_error(ctx));
			for (uint32_t i = 0; i < uploaded.size(); i++) {
				Body *b = uploaded[i];
				b->slot = slots[i];
				b->pending = false;
				if (body_of_slot.size() <= (uint32_t)slots[i]) {
					body_of_slot.resize(slots[i] + 1);
				}
				body_of_slot[slots[i]] = b;
				if (b->constant_force != Vector3() || b->constant_torque != Vector3()) {
					_write_constant(ctx, b->slot, b->constant_force, b->constant_torque);
				}
				if (b->max_contacts_reported > 0) {
					int32_t slot = b->slot, cap = b->max_contacts_reported;
					bpc_bodies_set_max_contacts_reported(ctx, &slot, 1, &cap);
				}
				const Body *owner = b->parent ? b->parent : b;
				if (owner->collision_priority != 1.0) {
					int32_t slot = b->slot;
					float v = (float)owner->collision_priority;
					bpc_bodies_set_collision_priority(ctx, &slot, 1, &v);
				}
				if (b->has_custom_inertia || b->has_custom_com) {
					b->mass_override_dirty = true;
					_body_changed(b); // described once more below, now that the slot its own mass data go with is known
				}
			}
			// the records of one body never pair with each other (static ones would not anyway: this is for kinematic bodies)
			for (Body *b : uploaded) {
				Body *owner = b->parent ? b->parent : b;
				if (owner->parts.is_empty()) {
					continue;
				}
				if (owner != b && owner->slot >= 0) {
					bpc_bodies_set_same_object(ctx, b->slot, owner->slot, 1);
				}
				for (Body *other : owner->parts) {
					if (other != b && other->slot >= 0) {
						bpc_bodies_set_same_object(ctx, b->slot, other->slot, 1);
					}
				}
			}
			// exceptions named before both bodies were in a space
			for (Body *b : uploaded) {
				for (const RID &r : b->exceptions) {
					Body *o = body_owner.get_or_null(r);
					if (o && o->slot >= 0) {
						bpc_bodies_set_exception(ctx, b->slot, o->slot, 1);
					}
				}
			}
			state_cache_valid = false;
		}
	}
	if (!dirty_bodies.is_empty()) {
		LocalVector<bpc_body_desc> descs;
		LocalVector<int32_t> slots;
		for (Body *b : dirty_bodies) {
			b->dirty = false;
			if (b->slot < 0 || !shape_owner.get_or_null(b->shape)) {
				continue;
			}
			_push_mass_override(b);
			bpc_body_desc d;
			_fill_desc(b, d);
			descs.push_back(d);
			slots.push_back(b->slot);
		}
		dirty_bodies.clear();
		if (!descs.is_empty()) {
			ERR_FAIL_COND_MSG(bpc_bodies_update(ctx, slots.ptr(), slots.size(), descs.ptr(), 1) < 0, bpc_last_error(ctx));
		}
	}
	if (!pending_joints.is_empty()) {
		LocalVector<bpc_joint_desc> descs;
		LocalVector<Joint *> uploaded, waiting;
		for (Joint *j : pending_joints) {
			Body *a = body_owner.get_or_null(j->a);
			Body *b = body_owner.get_or_null(j->b);
			if (!a || !b || j->type == JOINT_TYPE_MAX) {
				j->pending = false;
				continue;
			}
			if (a->slot < 0 || b->slot < 0) {
				waiting.push_back(j); // one of its bodies is not in a space yet
				continue;
			}
			bpc_joint_desc d;
			_fill_joint_desc(j, d);
			descs.push_back(d);
			uploaded.push_back(j);
		}
		pending_joints.clear();
		for (Joint *j : waiting) {
			pending_joints.push_back(j);
		}
		if (!descs.is_empty()) {
			int first = bpc_joints_create(ctx, descs.ptr(), descs.size());
			ERR_FAIL_COND_MSG(first < 0, bpc_last_error(ctx));
			for (uint32_t i = 0; i < uploaded.size(); i++) {
				uploaded[i]->handle = first + i;
				uploaded[i]->pending = false;
				if (uploaded[i]->type == JOINT_TYPE_SLIDER) {
					ERR_FAIL_COND_MSG(bpc_joint_params_write(ctx, first + i, 0, SLIDER_JOINT_MAX, uploaded[i]->slider_params) < 0, bpc_last_error(ctx));
				} else if (uploaded[i]->type == JOINT_TYPE_6DOF) {
					ERR_FAIL_COND_MSG(bpc_joint_params_write(ctx, first + i, 0, BPC_G6DOF_PARAMS, uploaded[i]->g6_params) < 0, bpc_last_error(ctx));
				}
			}
		}
	}
	LocalVector<Area *> waiting_areas;
	for (Area *a : pending_areas) {
		bpc_area_desc d;
		_fill_area_desc(a, d);
		if (d.space < 0 || d.shape < 0) {
			waiting_areas.push_back(a); // its shape has no data yet
			continue;
		}
		a->pending = false;
		int h = bpc_areas_create(ctx, &d, 1);
		ERR_CONTINUE_MSG(h < 0, bpc_last_error(ctx));
		a->handle = h;
		if (area_of_handle.size() <= (uint32_t)h) {
			area_of_handle.resize(h + 1);
		}
		area_of_handle[h] = a;
	}
	pending_areas.clear();
	for (Area *a : waiting_areas) {
		pending_areas.push_back(a);
	}
	for (Area *a : dirty_areas) {
		a->dirty = false;
		if (a->handle < 0) {
			continue;
		}
		bpc_area_desc d;
		_fill_area_desc(a, d);
		ERR_CONTINUE_MSG(bpc_areas_write(ctx, a->handle, 1, &d) < 0, bpc_last_error(ctx));
	}
	dirty_areas.clear();
}

void BradotPhysicsServerCUDA::step(real_t p_step) {
	if (!active || !ctx) {
		return;
	}
	_flush_pending();
	if (!pushed_bodies.is_empty()) { // ONE scattered write for every pushed body
		LocalVector<int32_t> slots;
		LocalVector<float> rows;
		for (Body *b : pushed_bodies) {
			if (b->slot >= 0) {
				slots.push_back(b->slot);
				rows.push_back(b->applied_force.x); rows.push_back(b->applied_force.y); rows.push_back(b->applied_force.z);
				rows.push_back(b->applied_torque.x); rows.push_back(b->applied_torque.y); rows.push_back(b->applied_torque.z);
			}
			b->applied_force = Vector3();
			b->applied_torque = Vector3();
			b->pushed = false;
		}
		pushed_bodies.clear();
		if (!slots.is_empty()) {
			ERR_FAIL_COND_MSG(bpc_bodies_write_slots(ctx, BPC_FIELD_APPLIED_FORCE, slots.ptr(), slots.size(), rows.ptr()) < 0, bpc_last_error(ctx));
		}
	}
	last_step = p_step;
	stepped_since_flush = true;
	state_cache_valid = false;
	contact_cache_valid = false;
	single_contacts_slot = -1;
	ERR_FAIL_COND_MSG(bpc_step(ctx, p_step) < 0, bpc_last_error(ctx));
}

void BradotPhysicsServerCUDA::sync() {
	doing_sync = true;
	if (ctx) {
		bpc_synchronize(ctx);
	}
}

// BradotSpace3D::call_queries (bradot_space_3d.cpp:1164-1176): every body that was integrated in the last step and has a
// state-sync or force-integration callback gets it, with its direct state.  The whole body state comes back in ONE copy
// (19 floats per slot); "was integrated" = active before or after the step (a body that fell asleep in this very step was
// still moved by it and reports its last transform, as in the reference, where integrate_velocities queues the callback).
void BradotPhysicsServerCUDA::flush_queries() {
	if (!active || !ctx) {
		return;
	}
	flushing_queries = true;
	_flush_pending();
	bpc_get_stats(ctx, &last_stats);
	int n = bpc_body_count(ctx);
	if (n > 0 && !callback_bodies.is_empty() && stepped_since_flush) { // no step, nothing was integrated, nobody is told
		stepped_since_flush = false;
		state_cache.resize((uint32_t)n * 19);
		if (bpc_bodies_read_state(ctx, 0, n, state_cache.ptr()) < 0) {
			flushing_queries = false;
			ERR_FAIL_MSG(bpc_last_error(ctx));
		}
		state_cache_valid = true;
		if (reporting_bodies > 0) {
			contact_counts.resize(n);
			contact_offsets.resize(n);
			contact_records.resize(MAX(1, bpc_contacts_capacity(ctx)));
			contact_cache_valid = bpc_contacts_read(ctx, contact_counts.ptr(), contact_offsets.ptr(), contact_records.ptr()) >= 0;
		}
		uint32_t old = was_active.size();
		was_active.resize(n);
		for (uint32_t i = old; i < (uint32_t)n; i++) {
			was_active[i] = 1; // a new body counts as active before its first step
		}
		// the callbacks may create, free or re-describe bodies: walk a copy
		LocalVector<Body *> todo;
		for (Body *b : callback_bodies) {
			// (a static body is never integrated: integrate_velocities returns before queueing it, bradot_body_3d.cpp:675-683)
			if (b->slot >= 0 && b->slot < n && b->mode != BODY_MODE_STATIC && (state_cache[b->slot * 19 + 18] != 0.f || was_active[b->slot])) {
				todo.push_back(b);
			}
		}
		for (int i = 0; i < n; i++) {
			was_active[i] = state_cache[i * 19 + 18] != 0.f;
		}
		for (Body *b : todo) {
			if (callback_bodies.has(b)) {
				_body_call_queries(b);
			}
		}
	}
	_flush_area_events(); // after the bodies, as BradotSpace3D::call_queries does (P3D/bradot_space_3d.cpp:1164-1176)
	flushing_queries = false;
}

void BradotPhysicsServerCUDA::end_sync() {
	doing_sync = false;
}

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

BradotPhysicsServerCUDA::BradotPhysicsServerCUDA(bool p_using_threads) :
		using_threads(p_using_threads) {}

BradotPhysicsServerCUDA::~BradotPhysicsServerCUDA() {
	finish();
}
