"""Builds the oracle's native pieces — TEST INFRASTRUCTURE ONLY.

  port : oracle/compute_overlap_port.c  -> oracle/_port/liboracle_port.so   (always)
  ref  : the REFERENCE's own Cython extension, compiled from the source where it lies
         (/root/reference/PyraPose/utils/compute_overlap.pyx) -> oracle/_ref/  (only where /root/reference exists;
         the GPU box uses the prebuilt file that travels with the snapshot). No reference source is copied
         into the repo: Cython writes its generated C into oracle/_ref/ (git-ignored).
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_DIR = os.path.join(HERE, "_port")
REF_DIR = os.path.join(HERE, "_ref")
REF_PYX = "/root/reference/PyraPose/utils/compute_overlap.pyx"


def port_lib_path():
    return os.path.join(PORT_DIR, "liboracle_port.so")


def build_port(force=False):
    os.makedirs(PORT_DIR, exist_ok=True)
    src = os.path.join(HERE, "compute_overlap_port.c")
    out = port_lib_path()
    if not force and os.path.exists(out) and os.path.getmtime(out) >= os.path.getmtime(src):
        return out
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", src, "-o", out])
    return out


def ref_ext_path():
    if not os.path.isdir(REF_DIR):
        return None
    for f in os.listdir(REF_DIR):
        if f.startswith("compute_overlap") and f.endswith(".so"):
            return os.path.join(REF_DIR, f)
    return None


def build_ref(force=False):
    """Cythonize + compile the reference .pyx in place (outputs only under oracle/_ref/). Returns the
    extension path, or the prebuilt one / None when /root/reference is absent."""
    existing = ref_ext_path()
    if not os.path.exists(REF_PYX):
        return existing
    if existing and not force:
        return existing
    import numpy as np
    os.makedirs(REF_DIR, exist_ok=True)
    c_file = os.path.join(REF_DIR, "compute_overlap.c")
    subprocess.check_call([sys.executable, "-m", "cython", "-3", REF_PYX, "-o", c_file])
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    out = os.path.join(REF_DIR, "compute_overlap" + ext)
    subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-fwrapv", "-fno-strict-aliasing",
                           "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
                           "-I", sysconfig.get_paths()["include"], "-I", np.get_include(), c_file, "-o", out])
    return out


def load_ref_compute_overlap():
    """Returns the reference's compute_overlap function (from oracle/_ref) or None."""
    path = ref_ext_path()
    if path is None:
        return None
    import importlib.util
    spec = importlib.util.spec_from_file_location("compute_overlap", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.compute_overlap


if __name__ == "__main__":
    print(build_port(force=True))
    print(build_ref(force="--force" in sys.argv))
