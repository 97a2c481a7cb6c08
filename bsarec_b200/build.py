"""Build libbsarec_sm100a.so in-tree with nvcc (sm_100a only; cross-compiles without a GPU).

    python -m bsarec_b200.build [--force]
"""
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libbsarec_sm100a.so")
SOURCES = ["capi.cu", "gemm_simt.cu", "rowwise.cu", "block.cu", "ce.cu", "topk.cu", "umma_engine.cu", "tmap.cu", "attn.cu", "attn_umma.cu", "attn_kv_umma.cu", "adam.cu", "lastrow.cu", "dataprep.cu", "rowwise_vec.cu", "bce.cu", "fmlp.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-I", INCLUDE,
         "--expt-relaxed-constexpr", "--expt-extended-lambda", "-Xptxas", "-v",
         *os.environ.get("BSA_EXTRA_NVCC_FLAGS", "").split()]      # e.g. -DBSA_ATTN_STAMPS for profiles/attn_phases.py


def _digest():
    h = hashlib.sha256(" ".join(FLAGS).encode())
    for root in (CSRC, INCLUDE):
        for fn in sorted(os.listdir(root)):
            with open(os.path.join(root, fn), "rb") as f:
                h.update(fn.encode())
                h.update(f.read())
    return h.hexdigest()


def _compile(src):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(obj + ".log", "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force=False, verbose=True):
    os.makedirs(OBJ, exist_ok=True)
    stamp = os.path.join(OBJ, "digest.txt")
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return LIB
    if not os.path.exists(NVCC):
        if os.path.exists(LIB):        # GPU box without a toolkit: use the library that travelled with the tree
            return LIB
        raise RuntimeError(f"{NVCC} not found and {LIB} is not built")
    with cf.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(_compile, SOURCES))
    cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as f:
        f.write(dig)
    if verbose:
        print(f"built {LIB}", file=sys.stderr)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
