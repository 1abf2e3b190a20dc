"""Build libminterpy_b200.so in-tree with nvcc for sm_100a (no torch involved)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libminterpy_b200.so")
SOURCES = ["mpx_plan.cu", "mpx_eval.cu", "mpx_eval_deriv.cu", "mpx_eval_1d.cu", "mpx_eval_mma.cu", "mpx_eval_tmma.cu", "mpx_dds.cu", "mpx_monomials.cu", "mpx_dense.cu",
           "mpx_api.cu", "mpx_noalloc.cu", "mpx_index.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--fmad=true"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, tuning=False):
    """tuning=True (`python -m minterpy_b200.build --tuning`) compiles the lab-bench knobs and the
    measured-and-dropped kernel variants in (-DMPX_TUNING); the shipped library has neither."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, "mpx_common.h"), os.path.join(CSRC, "mpx_ptx.h"),
               os.path.join(CSRC, "mpx_eval_kernel.cuh"),
               os.path.join(HERE, "..", "include", "minterpy_b200.h")]
    jobs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        if force or _stale(o, [s] + headers):
            cmd = ([nvcc] + NVCC_FLAGS + (["-DMPX_TUNING"] if tuning else []) +
                   (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o])
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0 or verbose:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))

    with ThreadPoolExecutor(max_workers=6) as ex:
        list(ex.map(run, jobs))
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if force or jobs or _stale(LIB, objs):
        run([nvcc, "-shared", "-o", LIB] + objs + ["-lcudart"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv or "--tuning" in sys.argv, verbose="-v" in sys.argv,
                tuning="--tuning" in sys.argv))
