"""Build libindy7_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

The fused kernels are instantiated per (arithmetic type, CTA size) in separate translation units that compile in
parallel; the objects are linked into one shared library that exports the C ABI of include/indy7_b200.h."""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libindy7_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]

# (object name, source, extra defines)
UNITS = [("i7_api", "i7_api.cu", []), ("i7_admm", "i7_admm_inst.cu", [])]
for tname, t in (("f64", "double"), ("f32", "float")):
    for nt in (128, 256, 384):
        for sp in ("false", "true"):
            UNITS.append((f"i7_sqp_{tname}_{nt}_{sp[0]}", "i7_sqp_inst.cu", [f"-DI7_T={t}", f"-DI7_NT={nt}", f"-DI7_SPILL={sp}"]))


for _cl in (8, 4, 2):
    UNITS.append((f"i7_cluster_f64_{_cl}", "i7_cluster_inst.cu",
                  ["-DI7_T=double", f"-DI7_CL={_cl}"] + (["-DI7_CLUSTER_DEFINE_LAYOUT"] if _cl == 8 else [])))
    UNITS.append((f"i7_cluster_f32_{_cl}", "i7_cluster_inst.cu", ["-DI7_T=float", f"-DI7_CL={_cl}"]))


def nvcc_path():
    for p in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc"):
        if p and os.path.exists(p):
            return p
    return "nvcc"


def _sources():
    """Everything the library is built from: every file under csrc/, the public header, this script."""
    srcs = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")))
    return [os.path.join(CSRC, s) for s in srcs] + [os.path.join(HERE, "..", "include", "indy7_b200.h"),
                                                     os.path.abspath(__file__)]


def _stamp():
    h = hashlib.sha256()
    for f in _sources():
        h.update(open(f, "rb").read())
    return h.hexdigest()


def needs_build():
    st = os.path.join(OBJ, "stamp")
    return not (os.path.exists(LIB) and os.path.exists(st) and open(st).read() == _stamp())


def build(force=False, verbose=False, extra=()):
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    nvcc = nvcc_path()

    def compile_unit(u):
        name, src, defs = u
        obj = os.path.join(OBJ, name + ".o")
        cmd = [nvcc] + NVCC_FLAGS + list(extra) + defs + ["-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"nvcc failed for {name}:\n{r.stdout}\n{r.stderr}")
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(len(UNITS), os.cpu_count() or 4)) as ex:
        res = list(ex.map(compile_unit, UNITS))
    if verbose:
        for _, err in res:
            if err.strip():
                print(err, flush=True)
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + [o for o, _ in res]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)
    open(os.path.join(OBJ, "stamp"), "w").write(_stamp())
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True, extra=["-Xptxas", "-v"] if "--ptxas" in sys.argv else ())
    print("built", LIB)
