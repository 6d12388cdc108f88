"""Builds libpyrapose_b200.so in-tree with nvcc for sm_100a.

The shared library is the C-ABI drop-in boundary (include/pyrapose_b200.h). It is compiled
straight with nvcc (no torch extension machinery: the signatures carry no torch types) and
lands in pyrapose_b200/_lib/ so that it travels to the GPU box with the source snapshot.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "_lib")
LIBNAME = "libpyrapose_b200.so"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]

# Translation units. `exact=True` disables FMA contraction: those kernels must reproduce the
# reference's separately-rounded float32/float64 arithmetic bit for bit.
SOURCES = [
    ("api.cu", False),
    ("conv_igemm.cu", False),
]


def _optional_sources():
    extra = [
        ("overlap.cu", True),
        ("anchors.cu", True),
        ("decode.cu", True),
        ("image.cu", True),
        ("filter.cu", True),
        ("elementwise.cu", False),
        ("split_ops.cu", False),
        ("engine.cu", False),
        ("conv_wgrad.cu", False),
        ("pnp.cu", False),
        ("train_kernels.cu", False),
        ("train.cu", False),
        ("exchange.cu", True),
    ]
    return [(s, e) for s, e in extra if os.path.exists(os.path.join(CSRC, s))]


def lib_path():
    return os.path.join(LIBDIR, LIBNAME)


def _digest(paths):
    h = hashlib.sha256()
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _file_sha(path):
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link the shared library. Returns its path."""
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    sources = SOURCES + _optional_sources()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "pyrapose_b200.h"))
    stamp = os.path.join(objdir, "stamp.txt")
    digest = _digest([os.path.join(CSRC, s) for s, _ in sources] + headers + [os.path.abspath(__file__)])
    if not force and os.path.exists(lib_path()) and os.path.exists(stamp) and open(stamp).read() == digest:
        return lib_path()

    def compile_one(item):
        src, exact = item
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = ["nvcc"] + ARCH + COMMON + (["-fmad=false"] if exact else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(sources))) as ex:
        objs = list(ex.map(compile_one, sources))
    cmd = ["nvcc"] + ARCH + ["-shared", "-o", lib_path()] + objs + ["-Xcompiler", "-fPIC", "-cudart", "static"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    with open(stamp, "w") as f:
        f.write(digest)
    # what was built, from what, with which compiler: travels with the .so (bench.py prints it as `build`)
    import json
    import time
    ver = subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout.strip().splitlines()
    with open(os.path.join(LIBDIR, "build_info.json"), "w") as f:
        json.dump({"sources_sha256": digest, "sources": [s for s, _ in sources], "nvcc": ver[-2:] if ver else None,
                   "arch_flags": ARCH, "flags": COMMON, "exact_flags": ["-fmad=false"],
                   "built_utc": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()),
                   "lib_sha256": _file_sha(lib_path())}, f, indent=1)
    return lib_path()


def build_info():
    """The record build() left beside the library (None when the library was not built by it)."""
    import json
    p = os.path.join(LIBDIR, "build_info.json")
    if not os.path.exists(p):
        return None
    info = json.load(open(p))
    info["lib_sha256_now"] = _file_sha(lib_path()) if os.path.exists(lib_path()) else None
    return info


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
