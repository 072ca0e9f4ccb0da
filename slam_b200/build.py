"""In-tree build of every native piece (no JIT cache: the .so files travel with
the repo snapshot to the GPU box).

  slam_b200/lib/libccicp2d_b200.so   product: sm_100a kernels + C ABI (nvcc)
  slam_b200/lib/libccicp2d.so        product: header-compatible C++ facade (g++), links the above
  slam_b200/lib/libscangen.so        workload generator (gcc)
  tests/host/libkdreplay_host.so     host build of the tie-replay code, for CPU tests (g++)
  oracle/libccicp_port.so            TEST INFRASTRUCTURE: C restatement (gcc)
  oracle/_ref/libccicp_ref.so        TEST INFRASTRUCTURE: unmodified reference sources, only when
                                     /root/reference is present (prebuilt file is used otherwise)
"""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "slam_b200", "csrc")
LIB = os.path.join(ROOT, "slam_b200", "lib")
MIN_BLOCKS = os.environ.get("CCICP_MIN_BLOCKS", "4")   # 64 registers, 4 CTAs/SM: +6 % over 3 CTAs at 80 registers

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",                       # parity: no FMA contraction anywhere (see ccicp_kernels.cuh)
    "-DCCICP_MIN_BLOCKS=" + MIN_BLOCKS,
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def _run(cmd, verbose):
    if verbose:
        print("+", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build step failed: %s\n%s" % (" ".join(cmd), r.stdout))
    if verbose and r.stdout.strip():
        print(r.stdout)
    return r.stdout


def source_sha256():
    """SHA-256 over the CUDA sources of the product library and its compiler flags.  nvcc does not produce bit-identical
    libraries from identical sources (two builds in a row differ), so profiles/traffic.json is stamped with this as well
    as with the SHA-256 of the binary the captures were taken from."""
    import hashlib
    h = hashlib.sha256()
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".h"))]
    files.append(os.path.join(ROOT, "include", "ccicp2d_b200.h"))
    for f in files:
        h.update(os.path.basename(f).encode() + b"\0")
        h.update(open(f, "rb").read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found")
    return p


def build_cuda(force=False, verbose=False, extra=()):
    os.makedirs(LIB, exist_ok=True)
    out = os.path.join(LIB, "libccicp2d_b200.so")
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "ccicp2d_b200.h"), __file__]
    if force or _newer(out, srcs):
        tmp = out + ".tmp"              # the library is replaced in one rename: a snapshot never sees a half-written file
        _run([nvcc_path()] + NVCC_FLAGS + list(extra) + ["-o", tmp, os.path.join(CSRC, "ccicp_api.cu"),
                                                       os.path.join(CSRC, "ccicp_multi.cu"), "-ldl"], verbose)
        os.replace(tmp, out)
    return out


def build_facade(force=False, verbose=False):
    out = os.path.join(LIB, "libccicp2d.so")
    src = os.path.join(CSRC, "facade", "icp_facade.cpp")
    if not os.path.exists(src):
        return None
    hdrs = [os.path.join(ROOT, "include", "ccicp2d", f) for f in os.listdir(os.path.join(ROOT, "include", "ccicp2d"))]
    hdrs.append(os.path.join(ROOT, "include", "ccicp2d_b200.h"))       # struct layouts of the C ABI
    if force or _newer(out, [src] + hdrs):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "include"), "-o", out, src,
              "-L" + LIB, "-lccicp2d_b200", "-Wl,-rpath,$ORIGIN"], verbose)
    return out


def build_scangen(force=False, verbose=False):
    os.makedirs(LIB, exist_ok=True)
    out = os.path.join(LIB, "libscangen.so")
    srcs = [os.path.join(CSRC, "scangen.c"), os.path.join(CSRC, "scangen.h")]
    if force or _newer(out, srcs):
        _run(["gcc", "-O2", "-std=c11", "-fopenmp", "-fPIC", "-shared", "-o", out, srcs[0], "-lm"], verbose)
    return out


def build_host_tests(force=False, verbose=False):
    d = os.path.join(ROOT, "tests", "host")
    out = os.path.join(d, "libkdreplay_host.so")
    srcs = [os.path.join(d, "kdreplay_host.cpp"), os.path.join(d, "kdreplay_ref.h"), os.path.join(CSRC, "kdreplay.h")]
    if force or _newer(out, srcs):
        _run(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", out, srcs[0]], verbose)
    # host build of the grid search (nn_grid.h) for the CPU exactness test
    out2 = os.path.join(d, "libnn_grid_host.so")
    srcs2 = [os.path.join(d, "nn_grid_host.cpp"), os.path.join(CSRC, "nn_grid.h")]
    if force or _newer(out2, srcs2):
        _run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", out2, srcs2[0]], verbose)
    return out


def build_callsite(force=False, verbose=False):
    """The reference call site compiled against the drop-in headers/library (tests/host/callsite_b200)."""
    d = os.path.join(ROOT, "tests", "host")
    out = os.path.join(d, "callsite_b200")
    src = os.path.join(d, "callsite_demo.cpp")
    if not os.path.exists(os.path.join(LIB, "libccicp2d.so")):
        return None
    if force or _newer(out, [src, os.path.join(LIB, "libccicp2d.so")]):
        _run(["g++", "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-o", out, src, "-L" + LIB,
              "-lccicp2d", "-lccicp2d_b200", "-Wl,-rpath,$ORIGIN/../../slam_b200/lib"], verbose)
    # the batched doICPMatch facade (include/ccicp2d/icpToolsBatch.h) driven with Eigen/ROS-shaped stand-in types
    out2, src2 = os.path.join(d, "batch_demo_b200"), os.path.join(d, "batch_demo.cpp")
    hdr2 = os.path.join(ROOT, "include", "ccicp2d", "icpToolsBatch.h")
    if force or _newer(out2, [src2, hdr2, os.path.join(ROOT, "include", "ccicp2d_b200.h"), os.path.join(LIB, "libccicp2d_b200.so")]):
        _run(["g++", "-O2", "-std=c++17", "-Wall", "-I" + os.path.join(ROOT, "include"), "-o", out2, src2, "-L" + LIB,
              "-lccicp2d_b200", "-Wl,-rpath,$ORIGIN/../../slam_b200/lib"], verbose)
    # several host threads on the reference's class surface (SURVEY.md 8(b) threading)
    out4, src4 = os.path.join(d, "facade_threads_b200"), os.path.join(d, "facade_threads.cpp")
    if os.path.exists(src4) and (force or _newer(out4, [src4, os.path.join(LIB, "libccicp2d.so"), os.path.join(LIB, "libscangen.so")])):
        _run(["g++", "-O2", "-std=c++17", "-pthread", "-I" + os.path.join(ROOT, "include"), "-o", out4, src4, "-L" + LIB,
              "-lccicp2d", "-lccicp2d_b200", "-lscangen", "-Wl,-rpath,$ORIGIN/../../slam_b200/lib"], verbose)
    # a C++ host driving several GPUs through the multi-device entry points (no Python, no torch)
    out3, src3 = os.path.join(d, "multi_demo_b200"), os.path.join(d, "multi_demo.cu")
    if os.path.exists(src3) and (force or _newer(out3, [src3, os.path.join(ROOT, "include", "ccicp2d_b200.h"),
                                                        os.path.join(LIB, "libccicp2d_b200.so"), os.path.join(LIB, "libscangen.so")])):
        _run([nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-o", out3, src3, "-L" + LIB,
              "-lccicp2d_b200", "-lscangen", "-Xlinker", "-rpath=$ORIGIN/../../slam_b200/lib"], verbose)
    return out


def build_oracle(force=False, verbose=False):
    """Builds the checker (TEST INFRASTRUCTURE); building it is not using it."""
    od = os.path.join(ROOT, "oracle")
    if os.path.exists(os.path.join(od, "ccicp_port.c")):
        _run(["make", "-C", od, "port"] + (["-B"] if force else []), verbose)
    if os.path.isdir("/root/reference/ccicp2d/src"):
        _run(["make", "-C", od, "ref"] + (["-B"] if force else []), verbose)


def build_tools(force=False, verbose=False):
    """profiles/fp32_peak: micro-benchmark of the un-fused FP32 issue peak (denominator of bench.py's roofline_useful)."""
    src = os.path.join(ROOT, "profiles", "fp32_peak.cu")
    out = os.path.join(ROOT, "profiles", "fp32_peak")
    if os.path.exists(src) and (force or _newer(out, [src])):
        _run([nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-fmad=false", "-o", out, src], verbose)
    return out


def build_all(force=False, verbose=False):
    build_cuda(force, verbose)
    build_tools(force, verbose)
    build_scangen(force, verbose)
    build_facade(force, verbose)
    build_host_tests(force, verbose)
    build_callsite(force, verbose)
    build_oracle(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
