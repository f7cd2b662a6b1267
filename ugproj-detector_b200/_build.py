"""In-tree build of libugdet.so: nvcc, sm_100a only, one shared library with a C ABI."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libugdet.so")
SOURCES = ["api.cu", "k_image.cu", "k_cascade.cu", "k_group.cu", "k_area.cu", "k_probe.cu", "k_legacy.cu", "k_noalloc.cu", "cascade_xml.cc", "config_ini.cc", "plan.cc"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--fmad=false",                      # every float op rounds exactly as written (bit-exact with the oracle)
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-Wall",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _compile_all(objdir: str, defines: list[str]):
    """Compiles every source to objdir (the translation units in parallel: k_cascade.cu alone takes a minute)."""
    from concurrent.futures import ThreadPoolExecutor

    def one(s):
        o = os.path.join(objdir, s + ".o")
        cmd = [_nvcc(), *NVCC_FLAGS, *[f"-D{d}" for d in defines], "-c", os.path.join(CSRC, s), "-o", o]
        return s, o, subprocess.run(cmd, capture_output=True, text=True)

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as ex:
        results = list(ex.map(one, SOURCES))
    objs, log = [], []
    for s, o, r in results:
        log.append(r.stderr)
        if r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError(f"nvcc failed on {s}")
        objs.append(o)
    return objs, log


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "ugdet.h")]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs, log = _compile_all(objdir, [])
    cmd = [_nvcc(), "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    with open(os.path.join(objdir, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        sys.stderr.write("\n".join(log))
    return LIB


def build_variant(name: str, defines: list[str]) -> str:
    """A differently compiled copy of the library (e.g. -DUGD_CHECKED=1: device-side bounds checks on every access to the
    scan tiles) as build/libugdet_<name>.so; load it with UGD_LIB=<path>.  The product build is untouched."""
    objdir = os.path.join(HERE, "build", name)
    os.makedirs(objdir, exist_ok=True)
    out = os.path.join(HERE, "build", f"libugdet_{name}.so")
    objs, _ = _compile_all(objdir, defines)
    r = subprocess.run([_nvcc(), "-shared", "-o", out, *objs, "-gencode", "arch=compute_100a,code=sm_100a"], capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    return out


if __name__ == "__main__":
    if "--checked" in sys.argv:
        print(build_variant("checked", ["UGD_CHECKED=1"]))
        sys.exit(0)

    print(build(force="--force" in sys.argv, verbose=True))
