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

def hand_written():
    """Names of the PhysicsServer3D virtuals the module implements by hand: every `virtual ... name(...) override` declared
    in the server class of bradot_physics_server_cuda.h (the rest get generated defaults)."""
    import re
    text = open(os.path.join(HERE, "bradot_physics_cuda", "bradot_physics_server_cuda.h")).read()
    cls = text[text.index("class BradotPhysicsServerCUDA : public PhysicsServer3D"):]
    cls = cls[:cls.index('#include "bradot_physics_server_cuda_stubs.gen.inc"')]
    return sorted(set(re.findall(r"virtual\s+[^;{}]*?\b(\w+)\s*\([^;{}]*\)\s*(?:const\s*)?override", cls)))


def build():
    if not os.path.isdir(build_ref.REF):
        print("engine tree not present: keeping the prebuilt engine_demo as is")
        return os.path.exists(os.path.join(HERE, "_build", "engine_demo"))
    out = os.path.join(HERE, "_build")
    gen = os.path.join(out, "gen")
    os.makedirs(gen, exist_ok=True)
    subprocess.run([sys.executable, os.path.join(HERE, "gen_stubs.py"), build_ref.REF, os.path.join(gen, "bradot_physics_server_cuda_stubs.gen.inc")] + hand_written(), check=True)
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
