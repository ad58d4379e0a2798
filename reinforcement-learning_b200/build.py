"""Builds libgigaenv_cuda.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libgigaenv_cuda.so")
OBJ = os.path.join(HERE, "build")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"] + ARCH

# per-file extra flags: the parity-critical TUs are built without FMA contraction so that scalar float
# expressions round like the reference's x86-64 SSE2 build (no FMA there either).
SOURCES = {
    "capi.cu": [],
    "gae.cu": ["-fmad=false"],
    "envset.cu": ["-fmad=false"],
    "expbuf.cu": [],
    "traj.cu": [],
    "obsstat.cu": ["-fmad=false"],
}


def _stale(src, obj, extra_deps):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [src] + extra_deps)


def build_prof():
    """Profiling variant (per-phase cycle counters in the tick, tools/phase_profile.py): libgigaenv_cuda_prof.so."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    out = os.path.join(HERE, "libgigaenv_cuda_prof.so")
    os.makedirs(OBJ, exist_ok=True)
    objs = []
    for name, extra in SOURCES.items():
        obj = os.path.join(OBJ, name.replace(".cu", ".prof.o"))
        subprocess.check_call([nvcc] + COMMON + extra + ["-DGGL_PHASE_TIMING", "-c", os.path.join(CSRC, name), "-o", obj])
        objs.append(obj)
    subprocess.check_call([nvcc, "-shared", "-o", out] + ARCH + objs + ["-lcudart"])
    return out


def build(verbose=False, force=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ, exist_ok=True)
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(HERE, "..", "include", "gigaenv.h"))
    objs = []
    relink = force or not os.path.exists(OUT)
    procs = []
    for name, extra in SOURCES.items():
        src = os.path.join(CSRC, name)
        if not os.path.exists(src):
            continue
        obj = os.path.join(OBJ, name.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(src, obj, hdrs):
            cmd = [nvcc] + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
            relink = True
    for name, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError("nvcc failed on %s" % name)
        if verbose:
            print(out)
    if relink:
        cmd = [nvcc, "-shared", "-o", OUT] + ARCH + objs + ["-lcudart"]
        subprocess.check_call(cmd)
    return OUT


def build_variant(tag, defines):
    """Experiment variant: libgigaenv_cuda_<tag>.so with extra -D flags (selected at run time through GGL_LIB_PATH)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    out = os.path.join(HERE, "libgigaenv_cuda_%s.so" % tag)
    os.makedirs(OBJ, exist_ok=True)
    objs, procs = [], []
    for name, extra in SOURCES.items():
        obj = os.path.join(OBJ, name.replace(".cu", ".%s.o" % tag))
        objs.append(obj)
        procs.append(subprocess.Popen([nvcc] + COMMON + extra + list(defines) + ["-c", os.path.join(CSRC, name), "-o", obj]))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed (variant %s)" % tag)
    subprocess.check_call([nvcc, "-shared", "-o", out] + ARCH + objs + ["-lcudart"])
    return out


if __name__ == "__main__":
    if "--variant" in sys.argv:
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], sys.argv[i + 2:]))
    elif "--prof" in sys.argv:
        print(build_prof())
    else:
        print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
