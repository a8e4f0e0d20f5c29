// Link stubs for symbols the reference's physics/core closure references but that live in
// parts of the engine we deliberately do not build (rendering, display, input mappings).
// TEST INFRASTRUCTURE ONLY (see oracle/build_ref.py).  Each one was named by the linker; none of
// them is reachable from the rigid-body step (soft bodies / FPS cap / joypads are not exercised).
#include "core/input/default_controller_mappings.h"
#include "servers/display_server.h"
#include "servers/rendering/rendering_device.h"
#include "servers/rendering_server.h"

RenderingServer *RenderingServer::get_singleton() { return nullptr; }
Array RenderingServer::mesh_surface_get_arrays(RID, int) const { return Array(); }
RenderingDevice *RenderingDevice::get_singleton() { return nullptr; }
void RenderingDevice::_set_max_fps(int) {}
DisplayServer *DisplayServer::singleton = nullptr;
const char *DefaultControllerMappings::mappings[] = { nullptr };
