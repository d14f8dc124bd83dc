"""In-tree build of the native libraries (nvcc, sm_100a only).

    python -m spark_druid_olap_b200.build [--force]

Outputs (git-ignored, shipped to the GPU box by gpurun):
    spark_druid_olap_b200/libsdo_engine.so    engine + C ABI (include/sdo_engine.h)
    spark_druid_olap_b200/libsdo_tpchgen.so   synthetic flattened-TPC-H generator (tests / bench input)
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]

ENGINE_SRCS = ["kernels.cu", "select_kernels.cu", "noalloc.cu", "jit.cpp", "comm.cpp", "minijs.cpp", "spec.cpp", "store.cpp", "plan.cpp", "engine.cpp", "capi.cpp"]
ENGINE_HDRS = ["scan_kernel.cuh", "jit.hpp", "common.hpp", "json.hpp", "timeutil.hpp", "spec.hpp", "store.hpp", "plan.hpp", "kernels.hpp",
               "engine.hpp", "xxh64.hpp", os.path.join("..", "..", "include", "sdo_engine.h")]
ENGINE_LIB = os.path.join(HERE, "libsdo_engine.so")
GEN_LIB = os.path.join(HERE, "libsdo_tpchgen.so")
OBJ_DIR = os.path.join(HERE, "build")


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + "\n")
        raise RuntimeError("build failed: " + " ".join(cmd[:3]) + " ...")
    return r.stdout


def build_engine(force=False, verbose=False):
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in ENGINE_HDRS]
    objs = []
    procs = []
    for src in ENGINE_SRCS:
        sp = os.path.join(CSRC, src)
        obj = os.path.join(OBJ_DIR, os.path.splitext(src)[0] + ".o")
        objs.append(obj)
        if force or _newer(obj, [sp] + hdrs):
            cmd = [NVCC, "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function", "-x", "cu"] + ARCH + \
                  ["-Xptxas", "-v" if verbose else "-O3", "-c", sp, "-o", obj]
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + out + "\n")
            raise RuntimeError("build failed: " + cmd[-3])
        if verbose and out.strip():
            print(out)
    if force or procs or _newer(ENGINE_LIB, objs):
        _run([NVCC, "-shared", "-o", ENGINE_LIB] + objs + ARCH + ["-cudart", "static", "-Xlinker", "--no-undefined", "-lpthread", "-ldl"])
    return ENGINE_LIB


def build_generator(force=False):
    src = os.path.join(CSRC, "tpch_flat_gen.cpp")
    if force or _newer(GEN_LIB, [src]):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", GEN_LIB, src])
    return GEN_LIB


def build_all(force=False, verbose=False):
    return build_engine(force, verbose), build_generator(force)


if __name__ == "__main__":
    libs = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    for l in libs:
        print(l)
