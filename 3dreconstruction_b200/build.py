"""In-tree build of libb200match.so: explicit nvcc for sm_100a only (no torch, no JIT cache).

    python 3dreconstruction_b200/build.py [--force] [--verbose]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libb200match.so")
SOURCES = ["capi.cu", "hamming.cu", "hamming_tc.cu", "l2_i8.cu", "l2_f32.cu", "tail.cu", "prep.cu", "adapter.cu", "ransac.cu", "essential.cu"]
# ransac.cu restates cv2's float32/fp64 arithmetic of findHomography: a*b+c must round twice there
PER_SOURCE_FLAGS = {"ransac.cu": ["-fmad=false"], "essential.cu": ["-fmad=false"]}
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "b200match.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, defs=(), so: str = SO, tag: str = "") -> str:
    """defs/so/tag: experiment builds (extra -D flags into another .so); the product build uses none."""
    if not force and not defs and not needs_build():
        return SO
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for src in SOURCES:
        obj = os.path.join(HERE, "build", src.replace(".cu", tag + ".o"))
        cmd = [_nvcc(), *NVCC_FLAGS, *PER_SOURCE_FLAGS.get(src, []), *defs, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {src}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    # -cudart shared would need libcudart on the box; static cudart keeps the .so self-contained.
    link = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", so, *objs,
            "-Xcompiler", "-fPIC", "-cudart", "static"]
    subprocess.run(link, check=True)
    return so


if __name__ == "__main__":
    _defs = [a for a in sys.argv[1:] if a.startswith("-D")]
    _tag = next((a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--tag=")), "")
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, defs=_defs,
                so=SO.replace(".so", _tag + ".so") if _tag else SO, tag=_tag))
