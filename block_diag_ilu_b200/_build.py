"""In-tree nvcc build of libbdilu_b200.so (sm_100a only; no JIT cache, the .so travels with the tree)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libbdilu_b200.so")
SOURCES = ["bdilu_factor.cu", "bdilu_solve.cu", "bdilu_solve_pf.cu", "bdilu_solve_skew.cu", "bdilu_solve_wave.cu", "bdilu_solve_wave32.cu", "bdilu_solve_quad32.cu", "bdilu_solve_half32.cu", "bdilu_solve_mrhs.cu", "bdilu_solve_mrhs_wave.cu", "bdilu_matvec.cu", "bdilu_f32.cu", "bdilu_capi.cu", "bdilu_host.cu"]
HEADERS = ["bdilu_common.cuh", "bdilu_kernels.h", "bdilu_wave_common.cuh", "bdilu_mrhs_common.cuh", os.path.join("..", "..", "include", "bdilu.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--cudart", "shared"]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every .cu for sm_100a and link the shared library.  Returns its path."""
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("BDILU_NVCC_EXTRA", "").split()  # e.g. -DBDILU_W32_PROF for the phase-profile build
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc] + NVCC_FLAGS + extra + ["-Xptxas", "-v", "-c", os.path.join(CSRC, src), "-o", obj]
        log = open(obj + ".log", "w")
        procs.append((src, obj, subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), log))
    objs = []
    for src, obj, p, log in procs:
        rc = p.wait()
        log.close()
        if rc != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src, open(obj + ".log").read()))
        if verbose:
            print(open(obj + ".log").read())
        objs.append(obj)
    cmd = [nvcc, "-shared", "--cudart", "shared", "-o", LIB] + objs
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose=True))
