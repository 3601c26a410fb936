"""Builds every native artefact in-tree (no JIT caches):

  avrecode_b200/libavrecode_b200.so   CUDA kernels (sm_100a) + C ABI + host indexer/container
  avrecode_b200/recode                drop-in CLI linked against the .so
  oracle/_build/libavr_oracle.so      CPU restatement (test infrastructure)
  build/libavr_emu.so                 host emulation of the device code (test infrastructure)
  build/libavr_emu_file.so            avr_file.cpp + h264_index.cpp on top of that emulation (test infrastructure)
  build/h264gen                       synthetic stream generator (test infrastructure)
  oracle/_ref/*                       the unmodified reference, only where /root/reference exists

Usage: python -m avrecode_b200.build [--force]
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "avrecode_b200", "csrc")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
LIB_SOURCES = ("avr_lib.cu", "avr_probe.cu", "avr_file.cpp", "h264_index.cpp")


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd, cwd=ROOT)


def build(force=False, with_ref=True):
    os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
    os.makedirs(os.path.join(ROOT, "oracle", "_build"), exist_ok=True)
    hdrs = [os.path.join(ROOT, "include", "avrecode_b200.h")] + [
        os.path.join(CSRC, f) for f in ("avr_core.cuh", "avr_walk.cuh", "avr_job.cuh", "avr_tables.h")]
    lib = os.path.join(ROOT, "avrecode_b200", "libavrecode_b200.so")
    lib_src = [os.path.join(CSRC, f) for f in LIB_SOURCES]
    # one object per source, compiled in parallel (nvcc spends ~40 s on each kernel file), then one link
    objdir = os.path.join(ROOT, "build", "obj")
    os.makedirs(objdir, exist_ok=True)
    objs, jobs = [], []
    for src in lib_src:
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + hdrs):
            if src.endswith(".cu"):
                cmd = [NVCC, "-std=c++17", *ARCH, "-lineinfo", "-O3", "-Xcompiler", "-fPIC", "-c", "-o", obj, src]
            else:
                cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-pthread", "-c", "-o", obj, src]
            print("+", " ".join(cmd), flush=True)
            jobs.append((cmd, subprocess.Popen(cmd, cwd=ROOT)))
    for cmd, p in jobs:
        if p.wait() != 0:
            raise subprocess.CalledProcessError(p.returncode, cmd)
    if force or jobs or _newer(lib, objs):
        _run([NVCC, *ARCH, "-shared", "-cudart", "static", "-Xcompiler", "-pthread", "-o", lib, *objs])
    cli = os.path.join(ROOT, "avrecode_b200", "recode")
    cli_src = [os.path.join(CSRC, "recode_main.cpp")]
    if force or _newer(cli, cli_src + [lib]):
        _run(["g++", "-std=c++17", "-O2", "-o", cli, *cli_src, "-L" + os.path.join(ROOT, "avrecode_b200"),
              "-lavrecode_b200", "-Wl,-rpath,$ORIGIN"])
    ora = os.path.join(ROOT, "oracle", "_build", "libavr_oracle.so")
    ora_src = [os.path.join(ROOT, "oracle", f) for f in ("avr_oracle_lib.cpp", "avr_oracle.hpp")]
    if force or _newer(ora, ora_src + hdrs):
        _run(["g++", "-std=c++14", "-O2", "-shared", "-fPIC", "-o", ora, ora_src[0]])
    emu = os.path.join(ROOT, "build", "libavr_emu.so")
    emu_src = [os.path.join(ROOT, "tests", "emu", "avr_emu.cpp")]
    if force or _newer(emu, emu_src + hdrs):
        _run(["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-o", emu, emu_src[0]])
    # the product's host file layer on top of the emulation (multi-device sharding, container codec: CPU tests)
    emuf = os.path.join(ROOT, "build", "libavr_emu_file.so")
    emuf_src = [os.path.join(ROOT, "tests", "emu", "avr_emu_file.cpp"), emu_src[0], os.path.join(CSRC, "avr_file.cpp"),
                os.path.join(CSRC, "h264_index.cpp")]
    if force or _newer(emuf, emuf_src + hdrs):
        _run(["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-pthread", "-o", emuf, emuf_src[0], emuf_src[2], emuf_src[3]])
    gen = os.path.join(ROOT, "build", "h264gen")
    gen_src = [os.path.join(ROOT, "tools", "h264gen.cpp"), ora_src[1]]
    if force or _newer(gen, gen_src + hdrs):
        _run(["g++", "-std=c++14", "-O2", "-o", gen, gen_src[0]])
    if with_ref and os.path.isdir("/root/reference/ffmpeg/libavcodec"):
        _run(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")])


if __name__ == "__main__":
    build(force="--force" in sys.argv)
