def can_build(env, platform):
    # needs libbradot_physics_cuda.so (built with nvcc from bradot_b200/csrc) at link time
    return not env["disable_3d"]


def configure(env):
    pass
