// Registers the CUDA backend with the engine's physics-server registry, next to "BradotPhysics3D"
// (pattern: modules/bradot_physics_3d/register_types.cpp:38-56).
#include "register_types.h"

#include "bradot_physics_server_cuda.h"
#include "core/config/project_settings.h"
#include "servers/physics_server_3d.h"
#include "servers/physics_server_3d_wrap_mt.h"

static PhysicsServer3D *_createBradotPhysicsCUDACallback() {
	bool using_threads = false;
	if (ProjectSettings::get_singleton()->has_setting("physics/3d/run_on_separate_thread")) {
		using_threads = GLOBAL_GET("physics/3d/run_on_separate_thread");
	}
	PhysicsServer3D *physics_server_3d = memnew(BradotPhysicsServerCUDA(using_threads));
	return memnew(PhysicsServer3DWrapMT(physics_server_3d, using_threads));
}

void initialize_bradot_physics_cuda_module(ModuleInitializationLevel p_level) {
	if (p_level != MODULE_INITIALIZATION_LEVEL_SERVERS) {
		return;
	}
	PhysicsServer3DManager::get_singleton()->register_server("BradotPhysicsCUDA", callable_mp_static(_createBradotPhysicsCUDACallback));
}

void uninitialize_bradot_physics_cuda_module(ModuleInitializationLevel p_level) {
}
