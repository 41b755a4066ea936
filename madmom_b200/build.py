"""In-tree build of libmadmom_b200.so (explicit nvcc, sm_100a only).

    python -m madmom_b200.build            # build if sources are newer than the library
    python -m madmom_b200.build --force
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIBPATH = os.path.join(LIBDIR, "libmadmom_b200.so")
# (source, object name, extra flags): the front-end kernel is compiled once per frame size and window path (table /
# cosine sum evaluated in the kernel), in parallel
SOURCES = [("frontend_kernel.cu", "frontend_kernel_%d.o" % n, ["-DMMB_FRONTEND_N=%d" % n]) for n in (2048, 4096, 1024)]
SOURCES += [("frontend_kernel.cu", "frontend_kernel_%d_wcos.o" % n, ["-DMMB_FRONTEND_N=%d" % n, "-DMMB_FRONTEND_WCOS=1"])
            for n in (2048, 4096)]
SOURCES += [("frontend_dispatch.cu", "frontend_dispatch.o", ["-DMMB_SPLIT_FRONTEND"]), ("capi.cu", "capi.o", []),
            ("calibration.cu", "calibration.o", []), ("postproc.cu", "postproc.o", []),
            ("stft64.cu", "stft64.o", [])]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]


if os.environ.get("MMB_NVCC_EXTRA"):              # experiments only, e.g. "-Xptxas -O2"
    NVCC_FLAGS = NVCC_FLAGS + os.environ["MMB_NVCC_EXTRA"].split()


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _stale():
    if not os.path.exists(LIBPATH):
        return True
    t = os.path.getmtime(LIBPATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "madmom_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    """Compile csrc/*.cu into lib/libmadmom_b200.so; returns the library path."""
    if not force and not _stale():
        return LIBPATH
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(os.path.dirname(HERE), "build")
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    objs = []
    procs = []
    for src, objname, extra in SOURCES:
        obj = os.path.join(objdir, objname)
        cmd = ([nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) +
               ["-c", os.path.join(CSRC, src), "-o", obj])
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, pr in procs:
        out, _ = pr.communicate()
        if verbose and out:
            print(out)
        if pr.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), out))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIBPATH] + objs
    subprocess.check_call(cmd)
    return LIBPATH


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
