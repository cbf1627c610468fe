#!/usr/bin/env python3
"""Build recipe for oracle/_ref: the reference's own CPU path-tracing path, compiled in place.

TEST INFRASTRUCTURE ONLY.  Compiles the reference's libaten / libatenscene sources *where they lie*
under /root/reference (nothing is copied into this repository) together with our harness
(oracle/ref_harness.cpp) into oracle/_ref/ (git-ignored, travels to the GPU box with gpurun).

The only change to the reference is one generated header: g++ 13 rejects
`this->operator float()` inside `PolymorphicValue::getAs` (src/libaten/misc/value.h:109-127; the
reference's CI compiles with clang-17).  The recipe writes a sed-patched copy of that one header to
oracle/_ref/shadow/libaten/misc/value.h (a build artefact, never committed) inside a tree of
symlinks to every other libaten file, and compiles the reference through that tree.  GL calls are never executed (they are guarded by ctxt.IsWindowInitialized(),
src/libaten/geometry/TriangleGroupMesh.cpp:36), so unresolved gl*/glew symbols are ignored at link.

Usage: python oracle/build_ref.py [--ref /root/reference] [-j N]
"""
import argparse
import concurrent.futures as cf
import hashlib
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

# libaten sources on (or needed to link) the CPU path-tracing path; GL-only TUs are left out.
LIBATEN_EXCLUDE = re.compile(
    r"^(visualizer/(?!GeomDataBuffer|shader|fbo)|display/|filter/|os/windows/|"
    r"accelerator/(ThreadedBvhFrustum|stackless_qbvh|qbvh)|cuda/|ui/)")

ASSETS = [  # copied (assets, not sources) so that oracle/_ref can run on the GPU box
    "cornellbox/orig.obj", "cornellbox/orig.mtl",
    "sponza/sponza_lod.obj", "sponza/sponza_lod.sbvh", "sponza/sponza.mtl",
    "teapot/teapot.obj", "suzanne/suzanne.obj", "plane/plane.obj", "dragon/dragon.obj",
]


def sh(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return r.returncode, r.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("-j", type=int, default=os.cpu_count() or 4)
    ap.add_argument("--force", action="store_true")
    a = ap.parse_args()
    ref = a.ref
    if not os.path.isdir(os.path.join(ref, "src", "libaten")):
        print(f"[oracle/_ref] reference not present at {ref}; keeping prebuilt files", file=sys.stderr)
        return 0 if os.path.exists(os.path.join(OUT, "ref_harness")) else 2
    src = os.path.join(ref, "src")
    tp = os.path.join(ref, "3rdparty")
    os.makedirs(os.path.join(OUT, "obj"), exist_ok=True)
    
    # --- shadow tree: symlinks to every libaten file, plus the one generated header ------------
    # (quote-includes are resolved relative to the includer's directory first, so only a tree in
    # which *every* includer sits next to the patched header shadows the original reliably)
    la_real = os.path.join(src, "libaten")
    la = os.path.join(OUT, "shadow", "libaten")
    for root, dirs, files in os.walk(la_real):
        rel = os.path.relpath(root, la_real)
        os.makedirs(os.path.join(la, rel), exist_ok=True)
        for f in files:
            d = os.path.join(la, rel, f)
            if os.path.normpath(os.path.join(rel, f)) == "misc/value.h":
                continue
            if not os.path.islink(d):
                os.symlink(os.path.join(root, f), d)
    txt = open(os.path.join(la_real, "misc/value.h")).read()
    for t, rep in (("float", "std::get<float>(val_)"), ("bool", "std::get<bool>(val_)"),
                   ("vec4", "std::get<vec4>(val_)"), ("int32_t", "std::get<int32_t>(val_)"),
                   ("vec3", "static_cast<vec3>(std::get<vec4>(val_))")):
        txt = txt.replace("return this->operator %s();" % t, "return %s;" % rep)
    txt = txt.replace("static_cast<vec3>(this->operator vec4())", "static_cast<vec3>(std::get<vec4>(val_))")
    pv = os.path.join(la, "misc", "value.h")
    if not os.path.exists(pv) or open(pv).read() != txt:
        open(pv, "w").write(txt)

    inc = [la, os.path.join(src, "libatenscene"),
           os.path.join(src, "common"), os.path.join(tp, "glm"), os.path.join(tp, "glew/include"),
           os.path.join(tp, "stb"), os.path.join(tp, "openvdb/nanovdb"), os.path.join(tp, "imgui"),
           os.path.join(tp, "tinyobjloader"), os.path.join(tp, "tinyxml2"), os.path.join(tp, "glfw/include")]
    cxx = ["g++", "-std=c++17", "-O2", "-fopenmp", "-fPIC", "-w", "-DGLEW_NO_GLU", "-DNDEBUG"]
    for i in inc:
        cxx += ["-I", i]

    tus = []
    for root, _, files in os.walk(la):
        for f in files:
            if not f.endswith(".cpp"):
                continue
            rel = os.path.relpath(os.path.join(root, f), la)
            if LIBATEN_EXCLUDE.search(rel):
                continue
            tus.append(os.path.join(root, f))
    for f in ("ObjLoader.cpp", "ImageLoader.cpp", "MaterialLoader.cpp"):
        tus.append(os.path.join(src, "libatenscene", f))
    tus.append(os.path.join(tp, "tinyobjloader", "tiny_obj_loader.cc"))
    tus.append(os.path.join(tp, "tinyxml2", "tinyxml2.cpp"))
    tus.append(os.path.join(HERE, "ref_stb_impl.cpp"))
    tus.append(os.path.join(HERE, "ref_harness.cpp"))

    def obj_of(tu):
        h = hashlib.md5(tu.encode()).hexdigest()[:8]
        return os.path.join(OUT, "obj", os.path.splitext(os.path.basename(tu))[0] + "_" + h + ".o")

    def compile_one(tu):
        o = obj_of(tu)
        if not a.force and os.path.exists(o) and os.path.getmtime(o) > os.path.getmtime(tu) \
                and os.path.getmtime(o) > os.path.getmtime(__file__):
            return tu, 0, ""
        rc, out = sh(cxx + ["-c", tu, "-o", o])
        return tu, rc, out

    failed = []
    optional_fail = []
    with cf.ThreadPoolExecutor(a.j) as ex:
        for tu, rc, out in ex.map(compile_one, tus):
            if rc != 0:
                if tu.startswith(la) and "pathtracing" not in tu:
                    optional_fail.append((tu, out))
                else:
                    failed.append((tu, out))
    for tu, out in optional_fail:
        print(f"[oracle/_ref] skipped (does not compile, not on the path): {tu}\n{out[-1500:]}", file=sys.stderr)
    if failed:
        for tu, out in failed:
            print(f"[oracle/_ref] FAILED {tu}\n{out[-3000:]}", file=sys.stderr)
        return 1
    skip = {obj_of(tu) for tu, _ in optional_fail}
    objs = [obj_of(tu) for tu in tus if obj_of(tu) not in skip]
    exe = os.path.join(OUT, "ref_harness")
    rc, out = sh(["g++", "-fopenmp", "-o", exe] + objs + ["-Wl,--unresolved-symbols=ignore-all", "-lpthread", "-ldl"])
    if rc != 0:
        print(out[-4000:], file=sys.stderr)
        return 1
    # assets (not source) needed at run time on a box without /root/reference
    sponza_maps = [os.path.join("sponza", f) for f in sorted(os.listdir(os.path.join(ref, "asset", "sponza")))
                   if f.lower().endswith((".jpg", ".png"))]
    for rel in ASSETS + sponza_maps:
        s = os.path.join(ref, "asset", rel)
        d = os.path.join(OUT, "asset", rel)
        if os.path.exists(s):
            os.makedirs(os.path.dirname(d), exist_ok=True)
            if not os.path.exists(d) or os.path.getsize(d) != os.path.getsize(s):
                shutil.copyfile(s, d)
    print(f"[oracle/_ref] built {exe}")
    # device_host: the same harness compiled with the `device` command and linked against the product library -- the
    # reference's loaders + the C++ shim class + libaten_b200.so in one C++ program (tests/test_gpu_cpp_host.py).
    lib_dir = os.path.join(os.path.dirname(HERE), "aten_b200")
    if os.path.exists(os.path.join(lib_dir, "libaten_b200.so")):
        hsrc = os.path.join(HERE, "ref_harness.cpp")
        dobj = os.path.join(OUT, "obj", "ref_harness_device.o")
        if a.force or not os.path.exists(dobj) or os.path.getmtime(dobj) < max(os.path.getmtime(hsrc), os.path.getmtime(__file__)) \
                or os.path.getmtime(dobj) < os.path.getmtime(os.path.join(os.path.dirname(HERE), "include", "idaten_b200_shim.h")):
            rc, out = sh(cxx + ["-DATN_WITH_DEVICE", "-c", hsrc, "-o", dobj])
            if rc != 0:
                print(out[-3000:], file=sys.stderr)
                return 1
        dexe = os.path.join(OUT, "device_host")
        rc, out = sh(["g++", "-fopenmp", "-o", dexe] + [o for o in objs if o != obj_of(hsrc)] + [dobj] +
                     ["-L", lib_dir, "-laten_b200", "-Wl,-rpath,$ORIGIN/../../aten_b200",
                      "-Wl,--unresolved-symbols=ignore-all", "-lpthread", "-ldl"])
        if rc != 0:
            print(out[-4000:], file=sys.stderr)
            return 1
        print(f"[oracle/_ref] built {dexe}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
