"""In-tree build recipes (no JIT cache: the .so files must travel with the repo snapshot).

  csrc/librtc_cuda.so   nvcc, sm_100a, --fmad=false (bit parity with the reference's f64: Rust never
                        fuses a*b+c), -lineinfo for ncu source pages
  host/librth_host.so   g++ -ffp-contract=off, links librtc_cuda.so ($ORIGIN rpath)
Test infrastructure (built here too, never loaded by the package):
  oracle/liboracle.so, oracle/kat        the CPU restatement of the reference + its known-answer tests
  tests/cpu_emul/libemul.so              host compile of the kernel source for the CPU-only test tier
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "--fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden", "-shared"]
GXX_FLAGS = ["-O2", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-Wall", "-Wextra", "-pthread"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def _glob(d, exts):
    return [os.path.join(d, f) for f in sorted(os.listdir(d)) if f.endswith(exts)]


def _run(cmd, cwd=None):
    print("+", " ".join(cmd), flush=True)
    subprocess.run(cmd, cwd=cwd, check=True)


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found")
    return p


def build_cuda(force=False):
    csrc = os.path.join(HERE, "csrc")
    out = os.path.join(csrc, "librtc_cuda.so")
    srcs = _glob(csrc, (".cu", ".cuh", ".hpp")) + [os.path.join(ROOT, "include", "rtc.h")]
    if force or _newer(out, srcs):
        _run([nvcc_path()] + NVCC_FLAGS + ["-o", out, os.path.join(csrc, "rtc_api.cu")])
    return out


def build_host(force=False):
    host = os.path.join(HERE, "host")
    out = os.path.join(host, "librth_host.so")
    srcs = _glob(host, (".cpp", ".hpp")) + [os.path.join(ROOT, "include", "rtc.h")]
    cuda = build_cuda()
    if force or _newer(out, srcs + [cuda]):
        _run(["g++"] + GXX_FLAGS + ["-shared", "-o", out, os.path.join(host, "rth_capi.cpp"),
                                    "-L" + os.path.dirname(cuda), "-lrtc_cuda", "-Wl,-rpath,$ORIGIN/../csrc"])
    # the reference's `render_scene <scene-file> <output-file>` binary over the same library
    exe = os.path.join(host, "render_scene")
    if force or _newer(exe, srcs + [out]):
        _run(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-o", exe, os.path.join(host, "render_scene.cpp"),
              "-L" + host, "-lrth_host", "-L" + os.path.dirname(cuda), "-lrtc_cuda",
              "-Wl,-rpath,$ORIGIN", "-Wl,-rpath,$ORIGIN/../csrc"])
    return out


def build_oracle(force=False):
    d = os.path.join(ROOT, "oracle")
    if force:
        _run(["make", "-C", d, "clean"])
    _run(["make", "-C", d, "-j4", "all"])
    return os.path.join(d, "liboracle.so")


def build_emul(force=False):
    d = os.path.join(ROOT, "tests", "cpu_emul")
    out = os.path.join(d, "libemul.so")
    srcs = [os.path.join(d, "emul.cpp")] + _glob(os.path.join(HERE, "csrc"), (".cuh", ".hpp")) + [os.path.join(ROOT, "include", "rtc.h")]
    if force or _newer(out, srcs):
        # hidden visibility + -Bsymbolic: the inline rtk:: helpers it shares with librtc_cuda.so must not interpose
        _run(["g++"] + GXX_FLAGS + ["-fvisibility=hidden", "-Wl,-Bsymbolic", "-shared", "-o", out, os.path.join(d, "emul.cpp")])
    return out


def build_all(force=False):
    return {"cuda": build_cuda(force), "host": build_host(force), "oracle": build_oracle(force), "emul": build_emul(force)}


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
