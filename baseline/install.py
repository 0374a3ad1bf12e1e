#!/usr/bin/env python
"""Builds the UNMODIFIED reference into baseline/_ref (git-ignored; travels to the GPU box).

The reference is not a Python package (`pip install /root/reference` fails: neither
setup.py nor pyproject.toml), so "installing" it means copying its five files
byte-for-byte and compiling the three programs with the commands its Makefile
implies (Makefile:1-5: `$(MPI) $(CFLAGS) -o riemann riemann.cpp`, CFLAGS
undefined => no optimisation flags) plus the two rules the Makefile lacks:

  riemann     g++ -I<mpi shim> -o riemann riemann.cpp          (mpigxx -> g++ + shim)
  4main       gcc -I<mpi shim> -o 4main 4main.c
  cintegrate  nvcc -arch=sm_100 -o cintegrate cintegrate.cu

No MPI exists in this image, so mpi.h / mpirun come from baseline/mpi_shim
(numbers are labelled "MPI shim").  One extra, clearly labelled, MODIFIED
binary is built for context only: cintegrate_sin = cintegrate.cu with the
commented-out `cuda_function` launch (cintegrate.cu:128) enabled instead of
`cuda_test` — the reference's own GPU sin kernel, which it never runs.
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("BQ_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")
SHIM = os.path.join(HERE, "mpi_shim")
FILES = ["Makefile", "riemann.cpp", "4main.c", "cintegrate.cu", "ex4vel.h"]


def sh(cmd, cwd=None):
    p = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return p.returncode, p.stdout


def main():
    if not os.path.isdir(REF):
        sys.exit("reference sources not found at %s" % REF)
    src = os.path.join(DST, "src")
    binr = os.path.join(DST, "bin")
    os.makedirs(src, exist_ok=True)
    os.makedirs(binr, exist_ok=True)
    manifest = {"files": {}, "builds": {}}
    for f in FILES:
        shutil.copyfile(os.path.join(REF, f), os.path.join(src, f))
        manifest["files"][f] = hashlib.sha256(open(os.path.join(src, f), "rb").read()).hexdigest()
    builds = {
        "riemann": ["g++", "-I" + SHIM, "-o", os.path.join(binr, "riemann"), "riemann.cpp"],
        "riemann_O2": ["g++", "-O2", "-I" + SHIM, "-o", os.path.join(binr, "riemann_O2"), "riemann.cpp"],
        "4main": ["gcc", "-I" + SHIM, "-o", os.path.join(binr, "4main"), "4main.c"],
        "4main_O2": ["gcc", "-O2", "-I" + SHIM, "-o", os.path.join(binr, "4main_O2"), "4main.c"],
        "cintegrate": ["nvcc", "-arch=sm_100", "-o", os.path.join(binr, "cintegrate"), "cintegrate.cu"],
    }
    for name, cmd in builds.items():
        rc, out = sh(cmd, cwd=src)
        manifest["builds"][name] = {"cmd": " ".join(cmd), "rc": rc, "log": out[-2000:]}
        print("%-12s rc=%d" % (name, rc))
    # labelled, modified variant: enable the reference's own sin kernel
    mod = os.path.join(DST, "src_sin")
    os.makedirs(mod, exist_ok=True)
    text = open(os.path.join(src, "cintegrate.cu")).read()
    a = "\tcuda_test<<<grid_size,block_size>>>(d_InterpProfile, d_DefaultProfile, d_sums, 0, 1800, SP, SM); //calls kernel"
    b = "    // cuda_function<<<grid_size,block_size>>>(d_sums, 0, M_PI, SP, SM);"
    if a in text and b in text:
        text = text.replace(a, "\t// " + a.strip()).replace(b, "    cuda_function<<<grid_size,block_size>>>(d_sums, 0, M_PI, SP, SM);")
        open(os.path.join(mod, "cintegrate.cu"), "w").write(text)
        shutil.copyfile(os.path.join(src, "ex4vel.h"), os.path.join(mod, "ex4vel.h"))
        cmd = ["nvcc", "-arch=sm_100", "-o", os.path.join(binr, "cintegrate_sin"), "cintegrate.cu"]
        rc, out = sh(cmd, cwd=mod)
        manifest["builds"]["cintegrate_sin (MODIFIED: launch line swapped)"] = {
            "cmd": " ".join(cmd), "rc": rc, "log": out[-2000:]}
        print("%-12s rc=%d" % ("cintegrate_sin", rc))
    json.dump(manifest, open(os.path.join(DST, "manifest.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
