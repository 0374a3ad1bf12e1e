#!/usr/bin/env python
"""In-tree build of every native artefact (no JIT cache, no site-packages install).

  b200quad/lib/libb200quad.so        core: CUDA kernels (sm_100a) + runtime + CPU backend, C ABI
  b200quad/lib/libb200quad_cpu.so    same C ABI, CPU backend only (no libcudart needed)
  b200quad/lib/libb200quad_torch.so  TORCH_LIBRARY(b200quad, ...) ops on top of the core
  bin/riemann, bin/cintegrate, bin/train4   reference-compatible CLIs

The reference's Makefile (Makefile:1-9) builds only riemann with an undefined
$(CFLAGS); this script is the whole build.
"""
import argparse
import hashlib
import os
import shutil
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "csrc")
LIBDIR = os.path.join(ROOT, "b200quad", "lib")
OBJDIR = os.path.join(ROOT, "build", "obj")
BINDIR = os.path.join(ROOT, "bin")

NVCC = os.environ.get("NVCC", shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = os.environ.get("BQ_NVCC_EXTRA", "").split() + ["-std=c++17", "-O3", "-lineinfo", "-fmad=false", "-Xcompiler", "-fPIC", "-Xptxas", "-v"] + ARCH
CXX_FLAGS = ["-std=c++17", "-O2", "-fPIC", "-ffp-contract=off", "-Wall", "-Wno-unknown-pragmas",
             "-Wno-unused-function"]

CUDA_SOURCES = ["quad_kernel.cu", "scan_kernel.cu", "ctx.cu", "scan_host.cu", "noalloc.cu", "drv.cu", "symm.cu"]
HOST_SOURCES = ["host_common.cpp", "cpu_backend.cpp", "cpu_scan.cpp"]
CPU_ONLY_SOURCES = ["cuda_stubs.cpp"]     # the CUDA entry points as stubs: only in libb200quad_cpu.so
SAN_FLAGS = ["-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-fno-sanitize-recover=undefined", "-g"]


def _run(cmd, log=None):
    t0 = time.time()
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.append((" ".join(cmd), p.stdout))
    if p.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + p.stdout + "\n")
        raise RuntimeError("build step failed: " + cmd[0])
    return p.stdout, time.time() - t0


def _digest(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _headers():
    out = [os.path.join(ROOT, "include", "b200quad.h")]
    for f in os.listdir(CSRC):
        if f.endswith((".h", ".cuh", ".inc")):
            out.append(os.path.join(CSRC, f))
    return out


def _compile(src, out, cmd_prefix, stamp_extra, force, log):
    """Compile src -> out unless a stamp says inputs are unchanged."""
    stamp = out + ".stamp"
    want = _digest([src] + _headers(), stamp_extra + os.environ.get("BQ_NVCC_EXTRA", ""))
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == want:
        return False
    _run(cmd_prefix + ["-c", src, "-o", out], log)
    with open(stamp, "w") as f:
        f.write(want)
    return True


def build_core(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    log = []
    objs, host_objs = [], []
    for s in HOST_SOURCES:
        src = os.path.join(CSRC, s)
        if not os.path.exists(src):
            continue
        o = os.path.join(OBJDIR, s + ".o")
        _compile(src, o, [CXX] + CXX_FLAGS + ["-I" + os.path.join(ROOT, "include")], "cxx", force, log)
        host_objs.append(o)
    for s in CUDA_SOURCES:
        src = os.path.join(CSRC, s)
        if not os.path.exists(src):
            continue
        o = os.path.join(OBJDIR, s + ".o")
        _compile(src, o, [NVCC] + NVCC_FLAGS + ["-I" + os.path.join(ROOT, "include")], "nvcc", force, log)
        objs.append(o)
    core = os.path.join(LIBDIR, "libb200quad.so")
    _run([NVCC, "-shared", "-o", core] + objs + host_objs + ["-cudart", "shared"] + ARCH, log)
    cpu_only_objs = []
    for src_name in CPU_ONLY_SOURCES:
        src = os.path.join(CSRC, src_name)
        o = os.path.join(OBJDIR, src_name + ".o")
        _compile(src, o, [CXX] + CXX_FLAGS + ["-I" + os.path.join(ROOT, "include")], "cxx", force, log)
        cpu_only_objs.append(o)
    cpu = os.path.join(LIBDIR, "libb200quad_cpu.so")
    _run([CXX, "-shared", "-o", cpu] + host_objs + cpu_only_objs + ["-lpthread", "-lm"], log)
    if verbose:
        for cmd, out in log:
            print(cmd)
            print(out)
    with open(os.path.join(ROOT, "build", "ptxas.log"), "w") as f:
        for cmd, out in log:
            f.write(cmd + "\n" + out + "\n")
    return core, cpu


def build_torch(force=False):
    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(CSRC, "torch_ops.cpp")
    if not os.path.exists(src):
        return None
    out = os.path.join(LIBDIR, "libb200quad_torch.so")
    stamp = out + ".stamp"
    want = _digest([src] + _headers(), torch.__version__)
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == want:
        return out
    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    import sysconfig
    inc += ["-isystem", sysconfig.get_paths()["include"]]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = [CXX, "-std=c++17", "-O2", "-fPIC", "-shared", abi, "-I" + os.path.join(ROOT, "include"),
           "-I/usr/local/cuda/include"] + inc + [src, "-o", out,
           "-L" + LIBDIR, "-lb200quad", "-Wl,-rpath,$ORIGIN",
           "-L" + tlib, "-ltorch", "-ltorch_cpu", "-lc10", "-Wl,-rpath," + tlib,
           "-Wl,--no-as-needed"]
    _run(cmd)
    with open(stamp, "w") as f:
        f.write(want)
    return out


def build_cli(force=False):
    os.makedirs(BINDIR, exist_ok=True)
    built = []
    clidir = os.path.join(ROOT, "cli")
    for name in ("riemann", "cintegrate", "train4"):
        src = os.path.join(clidir, name + ".cpp")
        if not os.path.exists(src):
            continue
        out = os.path.join(BINDIR, name)
        cmd = [CXX, "-std=c++17", "-O2", "-I" + os.path.join(ROOT, "include"), src, "-o", out,
               "-L" + LIBDIR, "-lb200quad", "-Wl,-rpath,$ORIGIN/../b200quad/lib", "-lpthread"]
        _run(cmd)
        built.append(out)
    return built


def build_sanitize():
    """Host sanitizer tier: the CPU library and the three CLIs built with AddressSanitizer +
    UndefinedBehaviorSanitizer into build/sanitize/ (tools/sanitize_host.sh runs the CPU test tier
    and the CLI golden runs under them).  The reference's undefined behaviour — uninitialised
    accumulators cintegrate.cu:86,135, strlen of an unwritten buffer 4main.c:110, the one-past-the-end
    table read cintegrate.cu:117 vs :39 — is the class of bug this tier exists to catch."""
    out_dir = os.path.join(ROOT, "build", "sanitize")
    os.makedirs(out_dir, exist_ok=True)
    # the distribution's own g++ knows where its sanitizer runtimes live (a relocated toolchain may not)
    CXX = os.environ.get("CXX_SANITIZE") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else globals()["CXX"])
    flags = [f for f in CXX_FLAGS if f != "-O2"] + ["-O1"] + SAN_FLAGS
    inc = ["-I" + os.path.join(ROOT, "include")]
    objs = []
    for src_name in HOST_SOURCES + CPU_ONLY_SOURCES:
        o = os.path.join(out_dir, src_name + ".o")
        _run([CXX] + flags + inc + ["-c", os.path.join(CSRC, src_name), "-o", o])
        objs.append(o)
    lib = os.path.join(out_dir, "libb200quad_cpu.so")
    _run([CXX, "-shared", "-o", lib] + objs + SAN_FLAGS + ["-lpthread", "-lm"])
    built = [lib]
    for name in ("riemann", "cintegrate", "train4"):
        exe = os.path.join(out_dir, name)
        _run([CXX, "-std=c++17", "-O1"] + SAN_FLAGS + inc + [os.path.join(ROOT, "cli", name + ".cpp")] + objs +
             ["-o", exe, "-lpthread", "-lm"])
        built.append(exe)
    return built


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sanitize", action="store_true", help="build the ASan+UBSan host tier into build/sanitize/")
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--no-torch", action="store_true")
    a = ap.parse_args()
    t0 = time.time()
    if a.sanitize:
        print("sanitize:", build_sanitize())
        print("done in %.1fs" % (time.time() - t0))
        return
    print("core :", build_core(a.force, a.verbose))
    if not a.no_torch:
        print("torch:", build_torch(a.force))
    print("cli  :", build_cli(a.force))
    print("done in %.1fs" % (time.time() - t0))


if __name__ == "__main__":
    main()
