#!/usr/bin/env python
"""Build the reference's own 3D physics (BradotPhysics3D) into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing under bradot_b200/ may import, link or
execute anything produced here; only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs use it, as the checker and the
timed CPU baseline.

What it does (plan A):
  * runs the reference's own header generators standalone
    (core/object/make_virtuals.py, core/extension/make_wrappers.py,
    core/core_builders.py) and writes the remaining trivial generated headers
    from the templates in core/SCsub -- all into oracle/_ref/gen/, never into
    /root/reference;
  * compiles core/**, drivers/unix/**, the thirdparty C files core/SCsub
    lists, servers/physics_server_3d*.cpp and modules/bradot_physics_3d/**
    FROM WHERE THEY LIE under /root/reference with the reference's release
    flags (-O3 -ffp-contract=off, real_t = float, THREADS_ENABLED);
  * archives them (so the linker pulls only what the harness needs) and links
    oracle/ref_harness.cpp + oracle/ref_stubs.cpp into
    oracle/_ref/oracle_step.

No reference source is copied into this repository.  SCons is absent from the
image, which is why this is a hand-rolled recipe.
"""
import concurrent.futures as cf
import glob
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("BRADOT_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT, "gen")
OBJ = os.path.join(OUT, "obj")

DEFINES = [
    "UNIX_ENABLED", "DEBUG_ENABLED", "THREADS_ENABLED", "ZSTD_STATIC_LINKING_ONLY",
    "BROTLI_ENABLED", "CLIPPER2_ENABLED", "MINIZIP_ENABLED", "NDEBUG",
    "_FILE_OFFSET_BITS=64", "MBEDTLS_CONFIG_FILE=\"thirdparty/mbedtls/include/bradot_core_mbedtls_config.h\"",
]
INCLUDES = [
    GEN, REF, os.path.join(REF, "platform/linuxbsd"),
    os.path.join(REF, "thirdparty/zlib"), os.path.join(REF, "thirdparty/zstd"),
    os.path.join(REF, "thirdparty/zstd/common"),
    os.path.join(REF, "thirdparty/mbedtls/include"),
    os.path.join(REF, "thirdparty/brotli/include"),
    os.path.join(REF, "thirdparty/clipper2/include"),
]
# Reference release build: SConstruct "optimize=speed" -> -O3; platform/linuxbsd/detect.py:177 -ffp-contract=off
CXXFLAGS = ["-std=gnu++17", "-O3", "-ffp-contract=off", "-fno-exceptions", "-w", "-pipe", "-fPIC", "-pthread"]
CFLAGS = ["-O2", "-w", "-pipe", "-fPIC"]


def sh(cmd, **kw):
    return subprocess.run(cmd, check=True, **kw)


def generate_headers():
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "core"))
    sys.path.insert(0, os.path.join(REF, "core/object"))
    sys.path.insert(0, os.path.join(REF, "core/extension"))
    for d in ("core/object", "core/extension", "core/io", "modules"):
        os.makedirs(os.path.join(GEN, d), exist_ok=True)
    import importlib
    mv = importlib.import_module("make_virtuals")
    mv.run([os.path.join(GEN, "core/object/gdvirtual.gen.inc")], None, None)
    mw = importlib.import_module("make_wrappers")
    mw.run([os.path.join(GEN, "core/extension/ext_wrappers.gen.inc")], None, None)
    cb = importlib.import_module("core_builders")
    cb.make_authors_header([os.path.join(GEN, "core/authors.gen.h")], [os.path.join(REF, "AUTHORS.md")], None)
    cb.make_donors_header([os.path.join(GEN, "core/donors.gen.h")], [os.path.join(REF, "DONORS.md")], None)
    cb.make_license_header([os.path.join(GEN, "core/license.gen.h")],
                           [os.path.join(REF, "COPYRIGHT.txt"), os.path.join(REF, "LICENSE.txt")], None)
    with open(os.path.join(GEN, "core/io/certs_compressed.gen.h"), "w") as f:
        f.write("#ifndef CERTS_COMPRESSED_GEN_H\n#define CERTS_COMPRESSED_GEN_H\n#define _SYSTEM_CERTS_PATH \"\"\n#endif\n")
    with open(os.path.join(GEN, "core/disabled_classes.gen.h"), "w") as f:
        f.write("#pragma once\n")
    with open(os.path.join(GEN, "modules/modules_enabled.gen.h"), "w") as f:
        f.write("#pragma once\n#define MODULE_BRADOT_PHYSICS_3D_ENABLED\n")
    ver = {}
    exec(open(os.path.join(REF, "version.py")).read(), ver)
    with open(os.path.join(GEN, "core/version_generated.gen.h"), "w") as f:
        f.write("#pragma once\n")
        f.write('#define VERSION_SHORT_NAME "%s"\n#define VERSION_NAME "%s"\n' % (ver["short_name"], ver["name"]))
        f.write("#define VERSION_MAJOR %d\n#define VERSION_MINOR %d\n#define VERSION_PATCH %d\n" %
                (ver["major"], ver["minor"], ver["patch"]))
        f.write('#define VERSION_STATUS "%s"\n#define VERSION_BUILD "oracle"\n#define VERSION_MODULE_CONFIG "%s"\n' %
                (ver["status"], ver.get("module_config", "")))
        f.write('#define VERSION_WEBSITE "%s"\n#define VERSION_DOCS_BRANCH "latest"\n' % ver.get("website", ""))
        f.write('#define VERSION_DOCS_URL "https://docs.bradotengine.org/en/" VERSION_DOCS_BRANCH\n')
    with open(os.path.join(GEN, "core/version_hash.gen.cpp"), "w") as f:
        f.write('#include "core/version.h"\nconst char *const VERSION_HASH = "";\nconst uint64_t VERSION_TIMESTAMP = 0;\n')
    with open(os.path.join(GEN, "core/script_encryption_key.gen.cpp"), "w") as f:
        f.write('#include "core/config/project_settings.h"\nuint8_t script_encryption_key[32] = {0};\n')


def reference_sources():
    cpp = []
    for pat in ("core/**/*.cpp", "drivers/unix/**/*.cpp", "modules/bradot_physics_3d/**/*.cpp"):
        cpp += glob.glob(os.path.join(REF, pat), recursive=True)
    cpp += [os.path.join(REF, "servers/physics_server_3d.cpp"),
            os.path.join(REF, "servers/physics_server_3d_wrap_mt.cpp")]
    cpp += [os.path.join(GEN, "core/version_hash.gen.cpp"), os.path.join(GEN, "core/script_encryption_key.gen.cpp")]
    # module registration needs the scene-side ClassDB of the whole engine; the harness creates the server directly
    cpp = [c for c in cpp if not c.endswith("bradot_physics_3d/register_types.cpp")]
    tp = "thirdparty/"
    c = [tp + "misc/" + f for f in ("fastlz.c", "r128.c", "smaz.c")]
    cpp += [os.path.join(REF, tp + "misc/" + f) for f in ("pcg.cpp", "polypartition.cpp", "smolv.cpp")]
    cpp += [os.path.join(REF, tp + "clipper2/src/" + f) for f in
            ("clipper.engine.cpp", "clipper.offset.cpp", "clipper.rectclip.cpp")]
    c += [tp + "brotli/" + f for f in (
        "common/constants.c", "common/context.c", "common/dictionary.c", "common/platform.c",
        "common/shared_dictionary.c", "common/transform.c", "dec/bit_reader.c", "dec/decode.c",
        "dec/huffman.c", "dec/state.c")]
    c += [tp + "zlib/" + f for f in ("adler32.c", "compress.c", "crc32.c", "deflate.c", "inffast.c", "inflate.c",
                                      "inftrees.c", "trees.c", "uncompr.c", "zutil.c")]
    c += [tp + "minizip/" + f for f in ("ioapi.c", "unzip.c", "zip.c")]
    c += [os.path.relpath(p, REF) for p in glob.glob(os.path.join(REF, tp, "zstd/common/*.c"))]
    c += [os.path.relpath(p, REF) for p in glob.glob(os.path.join(REF, tp, "zstd/compress/*.c"))]
    c += [os.path.relpath(p, REF) for p in glob.glob(os.path.join(REF, tp, "zstd/decompress/*.c"))]
    c += [tp + "zstd/decompress/huf_decompress_amd64.S"]
    c += [tp + "mbedtls/library/" + f for f in (  # the "light" list of core/crypto/SCsub:32-43
        "aes.c", "base64.c", "constant_time.c", "ctr_drbg.c", "entropy.c", "md.c", "md5.c", "sha1.c", "sha256.c",
        "bradot_core_mbedtls_platform.c")]
    c = [os.path.join(REF, x) for x in c]
    return sorted(set(cpp)), sorted(set(c))


def obj_path(src):
    h = hashlib.md5(src.encode()).hexdigest()[:8]
    return os.path.join(OBJ, os.path.basename(src).rsplit(".", 1)[0] + "_" + h + ".o")


def compile_one(src):
    o = obj_path(src)
    if os.path.exists(o) and os.path.getmtime(o) >= os.path.getmtime(src):
        return o, None
    flags = CXXFLAGS if src.endswith(".cpp") else CFLAGS
    cc = "g++" if src.endswith(".cpp") else "gcc"
    cmd = [cc] + flags + ["-D" + d for d in DEFINES] + ["-I" + i for i in INCLUDES] + ["-c", src, "-o", o]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        return o, r.stderr[-3000:]
    return o, None


def build(jobs=None):
    if not os.path.isdir(REF):
        print("reference tree not present at %s: keeping prebuilt oracle/_ref as is" % REF)
        return os.path.exists(os.path.join(OUT, "oracle_step"))
    os.makedirs(OBJ, exist_ok=True)
    generate_headers()
    cpp, c = reference_sources()
    jobs = jobs or os.cpu_count()
    objs, failed = [], []
    with cf.ThreadPoolExecutor(jobs) as ex:
        for o, err in ex.map(compile_one, cpp + c):
            if err:
                failed.append((o, err))
            else:
                objs.append(o)
    for o, err in failed:
        print("FAILED", o, "\n", err)
    if failed:
        raise SystemExit("reference compile failed for %d TUs" % len(failed))
    lib = os.path.join(OUT, "libbradot_ref.a")
    if os.path.exists(lib):
        os.remove(lib)
    # ar with a response file: the object list is long
    rsp = os.path.join(OUT, "objs.rsp")
    with open(rsp, "w") as f:
        f.write("\n".join(objs))
    sh(["ar", "rcs", lib, "@" + rsp])
    link_harness()
    return True


def link_harness():
    lib = os.path.join(OUT, "libbradot_ref.a")
    exe = os.path.join(OUT, "oracle_step")
    srcs = [os.path.join(HERE, "ref_harness.cpp"), os.path.join(HERE, "ref_stubs.cpp")]
    cmd = (["g++"] + CXXFLAGS + ["-fno-access-control"] + ["-D" + d for d in DEFINES] + ["-I" + i for i in INCLUDES] +
           srcs + ["-o", exe, "-Wl,--start-group", lib, "-Wl,--end-group", "-lpthread", "-ldl", "-lm"])
    sh(cmd)
    print("built", exe)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "link":
        link_harness()
    else:
        build()
