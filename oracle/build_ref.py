"""Build the UNMODIFIED reference hot path into oracle/_ref/ (test infrastructure only).

Compiles the reference's own Cython sources where they lie under /root/reference/src
(wfpt.pyx + pdf.pxi + integrate.pxi -> module `wfpt`; cdfdif_wrapper.pyx + cdfdif.c ->
module `cdfdif_wrapper`; see /root/reference/setup.py:6-7) with the stock flags
(-O2, no -ffast-math, no -march=native).  Outputs go ONLY into oracle/_ref/ (git-ignored,
travels to the GPU box with the snapshot).  No reference source is kept in the repo:
the scratch copy lives under oracle/_ref/_build and is deleted after the build.

The only edit applied to the scratch copy is commenting out the unused `import hddm`
at src/wfpt.pyx:16 (the package itself needs pymc/kabuki which are not installed);
`language_level=2` keeps split_cdf's integer division (src/wfpt.pyx:420).

Usage:  python oracle/build_ref.py [--reference /root/reference]
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may load what this builds.
"""
import argparse
import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

SETUP = r'''
from setuptools import setup, Extension
from Cython.Build import cythonize
import numpy as np
ext = cythonize([Extension('wfpt', ['wfpt.pyx'], language='c++', extra_compile_args=['-O2']),
                 Extension('cdfdif_wrapper', ['cdfdif_wrapper.pyx', 'cdfdif.c'], extra_compile_args=['-O2'])],
                compiler_directives={'language_level': 2})
setup(name='hddm_ref_hotpath', ext_modules=ext, include_dirs=[np.get_include()])
'''


def built():
    return (glob.glob(os.path.join(OUT, "wfpt*.so")) and
            glob.glob(os.path.join(OUT, "cdfdif_wrapper*.so")))


def build(reference="/root/reference", force=False):
    if built() and not force:
        return True
    src = os.path.join(reference, "src")
    if not os.path.isdir(src):
        return False
    scratch = os.path.join(OUT, "_build")
    shutil.rmtree(scratch, ignore_errors=True)
    os.makedirs(scratch)
    for f in ("wfpt.pyx", "pdf.pxi", "integrate.pxi", "cdfdif.c", "cdfdif.h", "cdfdif_wrapper.pyx"):
        shutil.copy(os.path.join(src, f), scratch)
    p = os.path.join(scratch, "wfpt.pyx")
    lines = open(p).read().split("\n")
    assert lines[15].strip() == "import hddm", lines[15]
    lines[15] = "# import hddm  (unused in the module body; package needs pymc/kabuki)"
    open(p, "w").write("\n".join(lines))
    open(os.path.join(scratch, "setup.py"), "w").write(SETUP)
    env = dict(os.environ, CFLAGS="-O2", CXXFLAGS="-O2")
    r = subprocess.run([sys.executable, "setup.py", "-q", "build_ext", "--inplace"], cwd=scratch,
                       env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout[-4000:])
        raise RuntimeError("reference hot path failed to build")
    for so in glob.glob(os.path.join(scratch, "*.so")):
        shutil.copy(so, OUT)
    shutil.rmtree(scratch, ignore_errors=True)
    return True


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--force", action="store_true")
    a = ap.parse_args()
    ok = build(a.reference, a.force)
    print("oracle/_ref:", "built" if ok else "reference sources not present; nothing built",
          sorted(os.path.basename(s) for s in glob.glob(os.path.join(OUT, "*.so"))))
