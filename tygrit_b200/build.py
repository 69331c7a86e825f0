"""Build ``tygrit_b200/lib/libtygrit_b200.so`` (and the oracle's C checker) in-tree.

sm_100a only: ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``.  nvcc
cross-compiles without a GPU, so this runs in the build container; the built
``.so`` files travel to the GPU box with the repo snapshot.
"""

from __future__ import annotations

import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "tygrit_b200", "csrc")
LIBDIR = os.path.join(ROOT, "tygrit_b200", "lib")
LIB = os.path.join(LIBDIR, "libtygrit_b200.so")
SOURCES = ["env.cu", "query.cu", "fkapi.cu", "filter.cu", "host.cu", "synth.cu", "ingest.cu", "occlusion.cu", "c5.cu", "planner.cu"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
]


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps if os.path.exists(d))


def nvcc_path() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the tygrit_b200 CUDA library cannot be built")


def build_cuda(force: bool = False, verbose: bool = False, stats: bool = False) -> str:
    """``stats=True`` builds ``libtygrit_b200_stats.so`` with -DTG_STATS (narrow-phase work counters,
    a profiling aid selected with TYGRIT_B200_LIB; never the product library)."""
    os.makedirs(LIBDIR, exist_ok=True)
    if stats:
        return _build(os.path.join(LIBDIR, "libtygrit_b200_stats.so"), os.path.join(LIBDIR, "stats"), ["-DTG_STATS"], force, verbose)
    return _build(LIB, LIBDIR, [], force, verbose)


def _build(LIB: str, OBJDIR: str, extra: list[str], force: bool, verbose: bool) -> str:
    os.makedirs(OBJDIR, exist_ok=True)
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(ROOT, "include", "tygrit_b200.h"))
    if not force and _newer(LIB, deps):
        return LIB
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(OBJDIR, os.path.basename(s).replace(".cu", ".o"))
        objs.append(o)
        cmd = [nvcc_path(), *NVCC_FLAGS, *extra, "-c", s, "-o", o]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            print(" ".join(cmd), file=sys.stderr)
            print(out, file=sys.stderr)
        if p.returncode:
            raise RuntimeError("nvcc failed for " + cmd[-3])
    # shared cudart (the toolkit's libcudart.so.12, the same soname torch loads): the library then holds only its own code
    link = [nvcc_path(), "-shared", "--cudart", "shared", "-Xlinker", "-rpath=/usr/local/cuda/lib64", "-gencode",
            "arch=compute_100a,code=sm_100a", "-o", LIB, *objs]
    subprocess.run(link, check=True)
    return LIB


def build_oracle(force: bool = False) -> str | None:
    """Compile the C restatement (test infrastructure); building the checker is not using it."""
    src = os.path.join(ROOT, "oracle", "cpu_checker.c")
    if not os.path.exists(src):
        return None
    out = os.path.join(ROOT, "oracle", "libtg_oracle.so")
    if not force and _newer(out, [src]):
        return out
    subprocess.run(
        ["gcc", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-shared", "-o", out, src, "-lm"], check=True
    )
    return out


if __name__ == "__main__":
    print(build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_cuda(force="--force" in sys.argv, stats=True))  # the counting build must never lag behind the product library
    if "--variant" in sys.argv:  # tuning aid: python -m tygrit_b200.build --variant NAME -DFOO=1 ... -> lib/libtygrit_b200_NAME.so
        i = sys.argv.index("--variant")
        name, flags = sys.argv[i + 1], [a for a in sys.argv[i + 2:] if a.startswith("-D")]
        print(_build(os.path.join(LIBDIR, f"libtygrit_b200_{name}.so"), os.path.join(LIBDIR, name), flags, True, False))
    print(build_oracle(force="--force" in sys.argv))
