"""Build libbamcl_cuda.so (sm_100a) in-tree with nvcc. No JIT cache: the .so travels with the repo snapshot."""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libbamcl_cuda.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
         "-Xcompiler", "-fPIC,-O2", "-shared", "-cudart", "static"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "bamcl.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """one object per .cu (compiled in parallel, rebuilt only when it or a header changed), then one link"""
    from concurrent.futures import ThreadPoolExecutor
    extra = os.environ.get("BAMCL_EXTRA_NVCC", "").split()  # kernel tuning sweeps
    only = os.environ.get("BAMCL_EXTRA_ONLY", "").split()    # ... applied to these sources only (e.g. "score_cloud")
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "bamcl.h")]
    hdr_t = max(os.path.getmtime(h) for h in hdrs)
    compile_flags = [f for f in FLAGS if f not in ("-shared", "-cudart", "static")]
    jobs = []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        name = os.path.basename(src)[:-3]
        mine = not only or name in only
        stale = ((force or extra) and mine) or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_t)
        if stale:
            jobs.append([NVCC] + compile_flags + (extra if mine else []) + (["-Xptxas", "-v"] if verbose else []) +
                        ["-c", "-o", obj, src])
    if not jobs and os.path.exists(OUT) and not force:
        return OUT
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        list(ex.map(subprocess.check_call, jobs))
    objs = [os.path.join(objdir, os.path.basename(src)[:-3] + ".o") for src in sources()]
    subprocess.check_call([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static", "-o", OUT] + objs +
                          ["-ldl"])
    return OUT


HOST = os.path.join(HERE, "host")
HOST_LIB = os.path.join(HERE, "libbadger_amcl.so")
HOST_TEST = os.path.join(HOST, "test", "test_host_api")
CXX = os.environ.get("CXX", "g++")


def build_host(force=False):
    """libbadger_amcl.so: the C++ mirror of the reference's class API (host/include/amcl/...) over the C ABI, and
    the test driver that runs the reference's own unit tests against it. Plain g++: no CUDA in these sources."""
    srcs = sorted(glob.glob(os.path.join(HOST, "src", "*.cpp")))
    hdrs = glob.glob(os.path.join(HOST, "include", "amcl", "*", "*.h")) + glob.glob(os.path.join(HOST, "include", "amcl", "*.h"))
    inc = ["-I", os.path.join(HOST, "include"), "-I", os.path.join(HERE, "..", "include")]
    common = [CXX, "-std=c++17", "-O2", "-fPIC", "-ffp-contract=off", "-Wall"] + inc
    def stale(out, deps):
        return force or not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps)
    if stale(HOST_LIB, srcs + hdrs + [OUT]):
        subprocess.check_call(common + ["-shared", "-o", HOST_LIB] + srcs + ["-L", HERE, "-lbamcl_cuda", "-Wl,-rpath,$ORIGIN"])
    test_src = os.path.join(HOST, "test", "test_host_api.cpp")
    if stale(HOST_TEST, [test_src, HOST_LIB] + hdrs):
        subprocess.check_call(common + ["-o", HOST_TEST, test_src, "-L", HERE, "-lbadger_amcl", "-lbamcl_cuda",
                                        "-Wl,-rpath,$ORIGIN/../.."])
    # the standalone C++ host of the distributed path (C ABI only)
    dist_src, dist_exe = os.path.join(HOST, "test", "bench_dist.cpp"), os.path.join(HOST, "test", "bench_dist")
    if stale(dist_exe, [dist_src, OUT]):
        subprocess.check_call([CXX, "-std=c++17", "-O2", "-Wall", "-I", os.path.join(HERE, "..", "include"), "-o", dist_exe,
                               dist_src, "-L", HERE, "-lbamcl_cuda", "-Wl,-rpath,$ORIGIN/../.."])
    return HOST_LIB


if __name__ == "__main__":
    build(force=True, verbose="-v" in sys.argv)
    build_host(force=True)
    print(OUT)
