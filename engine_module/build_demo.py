#!/usr/bin/env python
"""Compile the engine-side module against the engine headers and link engine_demo.

Needs the engine tree (/root/reference) and the engine objects built by oracle/build_ref.py
(oracle/_ref/libbradot_ref.a): the demo IS the engine core plus our module.  Output goes to
engine_module/_build/ (git-ignored; it travels to the GPU box like the other built files)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import build_ref  # noqa: E402  (compiler flags / include paths of the engine build)

MINE = ["world_boundary_shape_create", "sphere_shape_create", "box_shape_create", "capsule_shape_create", "convex_polygon_shape_create",
        "shape_set_data", "shape_get_type", "shape_get_data", "space_create", "space_set_active", "space_is_active", "space_set_param",
        "space_get_param", "space_get_direct_state", "body_create", "body_set_space", "body_get_space", "body_set_mode", "body_get_mode", "body_add_shape",
        "body_get_shape_count", "body_set_collision_layer", "body_get_collision_layer", "body_set_collision_mask", "body_get_collision_mask",
        "body_set_param", "body_get_param", "body_set_state", "body_get_state", "body_set_enable_continuous_collision_detection",
        "body_is_continuous_collision_detection_enabled", "body_apply_central_force", "body_apply_torque", "joint_create", "joint_clear", "joint_make_pin",
        "joint_make_hinge", "joint_disable_collisions_between_bodies", "joint_get_type", "free", "set_active", "init", "step", "sync",
        "flush_queries", "end_sync", "finish", "is_flushing_queries", "get_process_info"]


def build():
    if not os.path.isdir(build_ref.REF):
        print("engine tree not present: keeping the prebuilt engine_demo as is")
        return os.path.exists(os.path.join(HERE, "_build", "engine_demo"))
    out = os.path.join(HERE, "_build")
    gen = os.path.join(out, "gen")
    os.makedirs(gen, exist_ok=True)
    subprocess.run([sys.executable, os.path.join(HERE, "gen_stubs.py"), build_ref.REF, os.path.join(gen, "bradot_physics_server_cuda_stubs.gen.inc")] + MINE, check=True)
    lib_dir = os.path.join(ROOT, "bradot_b200")
    srcs = [os.path.join(HERE, "bradot_physics_cuda", "bradot_physics_server_cuda.cpp"), os.path.join(HERE, "bradot_physics_cuda", "register_types.cpp"),
            os.path.join(HERE, "engine_demo.cpp"), os.path.join(ROOT, "oracle", "ref_stubs.cpp")]
    base = (["g++"] + build_ref.CXXFLAGS + ["-D" + d for d in build_ref.DEFINES] + ["-I" + i for i in build_ref.INCLUDES] +
            ["-I" + HERE, "-I" + gen, "-I" + os.path.join(ROOT, "include")])
    objs = []
    for src in srcs:
        obj = os.path.join(out, os.path.basename(src).replace(".cpp", ".o"))
        # only the demo's main() needs access to engine bring-up calls that are private to Main; the module is compiled strictly
        extra = ["-fno-access-control"] if src.endswith("engine_demo.cpp") else []
        subprocess.run(base + extra + ["-c", src, "-o", obj], check=True)
        objs.append(obj)
    cmd = (["g++"] + objs +
           ["-o", os.path.join(out, "engine_demo"), "-Wl,--start-group", os.path.join(build_ref.OUT, "libbradot_ref.a"), "-Wl,--end-group",
            "-L" + lib_dir, "-lbradot_physics_cuda", "-Wl,-rpath,$ORIGIN/../../bradot_b200", "-lpthread", "-ldl", "-lm"])
    subprocess.run(cmd, check=True)
    print("built", os.path.join(out, "engine_demo"))
    return True


if __name__ == "__main__":
    build()
