"""In-tree nvcc build of libsdmm_b200.so (sm_100a only).  Used by __graft_entry__.build(); the
built library travels to the GPU box with the repo snapshot (it is git-ignored, not
gpurun-ignored).  The product never falls back to anything else when the library is missing."""
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_DIR = os.path.join(_HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libsdmm_b200.so")
SOURCES = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps():
    """Every file a translation unit can include: stale() compares all of them with the library."""
    inc = os.path.join(_HERE, "..", "include")
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h", ".hpp", ".inc"))]
    out += [os.path.join(inc, f) for f in os.listdir(inc) if f.endswith(".h")]
    return out


NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(d) > t for d in _deps())


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link the C-ABI shared library in-tree."""
    if not force and not stale():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    objs = []
    procs = []
    for s in SOURCES:
        o = os.path.join(LIB_DIR, s.replace(".cu", ".o"))
        objs.append(o)
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", os.path.join(CSRC, s), "-o", o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for s, p in procs:
        out, _ = p.communicate()
        log.append(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (s, out))
    cmd = [_nvcc(), "-Wno-deprecated-gpu-targets", "-shared", "-cudart", "shared", "-o", LIB_PATH] + objs + ["-ldl"]
    subprocess.check_call(cmd)
    with open(os.path.join(LIB_DIR, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return LIB_PATH


if __name__ == "__main__":
    import sys
    build(force="--force" in sys.argv, verbose=True)
    print(LIB_PATH)
