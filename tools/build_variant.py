#!/usr/bin/env python
"""Build a variant of libb200envs.so for A/B runs on one GPU box: tmp_variants/lib_<name>.so from the working tree, or
from the csrc/ + include/ of a git ref (--ref), optionally with extra -D flags.  Select it at run time with
B200ENVS_LIB=tmp_variants/lib_<name>.so (madrona_rl_envs_playground_b200/_lib.py).

    python tools/build_variant.py base --ref HEAD
    python tools/build_variant.py roll3 -DHB_ROLL_BUFS=3
"""
import os
import shutil
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from madrona_rl_envs_playground_b200.build import NVCC_FLAGS, SOURCES, _nvcc   # noqa: E402


def main():
    name, ref, defs = sys.argv[1], None, []
    args = sys.argv[2:]
    while args:
        a = args.pop(0)
        if a == "--ref":
            ref = args.pop(0)
        else:
            defs.append(a)
    out_dir = os.path.join(ROOT, "tmp_variants")
    os.makedirs(out_dir, exist_ok=True)
    work = tempfile.mkdtemp(prefix="b200variant_")
    pkg = "madrona_rl_envs_playground_b200"
    if ref:
        for sub in (f"{pkg}/csrc", "include"):
            files = subprocess.run(["git", "ls-tree", "-r", "--name-only", ref, sub], cwd=ROOT, capture_output=True,
                                   text=True, check=True).stdout.split()
            for f in files:
                dst = os.path.join(work, f)
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                with open(dst, "wb") as fh:
                    fh.write(subprocess.run(["git", "show", f"{ref}:{f}"], cwd=ROOT, capture_output=True, check=True).stdout)
    else:
        shutil.copytree(os.path.join(ROOT, pkg, "csrc"), os.path.join(work, pkg, "csrc"))
        shutil.copytree(os.path.join(ROOT, "include"), os.path.join(work, "include"))
    csrc = os.path.join(work, pkg, "csrc")
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + defs

    def compile_one(src):
        obj = os.path.join(work, src.replace(".cu", ".o"))
        r = subprocess.run([_nvcc()] + flags + ["-c", os.path.join(csrc, src), "-o", obj], capture_output=True, text=True)
        if r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
            raise SystemExit(f"nvcc failed on {src}")
        return obj

    with ThreadPoolExecutor(len(SOURCES)) as pool:
        objs = list(pool.map(compile_one, SOURCES))
    lib = os.path.join(out_dir, f"lib_{name}.so")
    subprocess.run([_nvcc(), "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler",
                    "-fPIC"] + objs + ["-o", lib], check=True)
    shutil.rmtree(work, ignore_errors=True)
    print(lib)


if __name__ == "__main__":
    main()
