"""Build libb200envs.so (the C-ABI library of sm_100a kernels) in-tree with nvcc.

`python -m madrona_rl_envs_playground_b200.build` or `build()`; nvcc cross-compiles without a GPU.  The built library
sits next to this file so that it travels with the repository snapshot to the GPU box.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb200envs.so")
SOURCES = ["b200envs.cu", "overcooked.cu", "hanabi.cu", "simple_envs.cu", "random_policy.cu", "sm_rates.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # Cartpole parity: no FMA contraction of the double-intermediate Euler step; nothing else here uses floats
    "--fmad=false",
    "-Xcompiler", "-fPIC,-O2,-Wall", "-shared", "-cudart", "static",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "b200envs.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """One nvcc -c per source (in parallel, only the sources that changed), then one link."""
    if not force and not needs_build():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not f.endswith(".cu")] + \
              [os.path.join(HERE, "..", "include", "b200envs.h"), os.path.abspath(__file__)]
    newest_header = max(os.path.getmtime(h) for h in headers)
    compile_flags = [f for f in NVCC_FLAGS if f != "-shared"]

    def compile_one(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        path = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(path), newest_header):
            return obj, None
        cmd = [_nvcc()] + compile_flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
        return obj, subprocess.run(cmd, capture_output=True, text=True)

    with ThreadPoolExecutor(len(SOURCES)) as pool:
        results = list(pool.map(compile_one, SOURCES))
    for obj, res in results:
        if res is not None and (verbose or res.returncode != 0):
            sys.stderr.write(res.stdout + res.stderr)
        if res is not None and res.returncode != 0:
            raise RuntimeError(f"nvcc failed compiling {os.path.basename(obj)}")
    link = [_nvcc(), "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a",
            "-Xcompiler", "-fPIC"] + [obj for obj, _ in results] + ["-o", LIB]
    res = subprocess.run(link, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed linking libb200envs.so")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
