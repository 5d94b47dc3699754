"""Build the native parts in-tree (no JIT cache):

  openmm_b200/lib/libb200nb.so      CUDA kernels + C ABI  (nvcc, sm_100a only)
  openmm_b200/lib/libOpenMMB200.so  OpenMM Platform plugin (g++), needs the OpenMM headers of the reference tree
                                    and oracle/_ref/libOpenMM.so at build time -> only built when /root/reference exists
  oracle/liboracle_nb.so            plain-C restatement (the checker), oracle/_ref/* (the compiled reference, same condition)

Usage: python -m openmm_b200.build [--force]
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "openmm_b200", "csrc")
PLUGIN = os.path.join(ROOT, "openmm_b200", "plugin")
LIBDIR = os.path.join(ROOT, "openmm_b200", "lib")
OBJDIR = os.path.join(ROOT, "build", "obj")
REF = os.environ.get("OPENMM_REFERENCE", "/root/reference")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, log=None, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, **kw)
    if log:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("command failed: " + " ".join(cmd))
    return r.stdout


def build_cuda(force=False):
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "b200_nonbonded.h"))
    cus = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    objs = []

    def compile_one(f):
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJDIR, f[:-3] + ".o")
        if force or _newer(obj, [src] + headers):
            _run([NVCC] + NVCC_FLAGS + ["-c", src, "-o", obj], log=obj + ".log")
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(compile_one, cus))
    lib = os.path.join(LIBDIR, "libb200nb.so")
    if force or _newer(lib, objs):
        _run([NVCC, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ldl"])
    return lib


def build_oracle(force=False):
    odir = os.path.join(ROOT, "oracle")
    _run(["make", "-C", odir, "restatement"])
    if os.path.isdir(REF):
        _run(["make", "-C", odir, "-j8", "reference", "REF=" + REF])
        try:   # the reference's CUDA platform (bench.py's "reference_gpu" row); its absence only removes that row
            _run(["make", "-C", odir, "-j8", "reference_cuda", "REF=" + REF], log=os.path.join(odir, "_ref", "reference_cuda_build.log"))
        except RuntimeError as e:
            sys.stderr.write("reference CUDA platform not built: %s\n" % e)
    return os.path.join(odir, "liboracle_nb.so")


def build_plugin(force=False):
    """OpenMM `B200` Platform plugin + its C++ parity/bench driver; needs the reference headers."""
    if not os.path.isdir(PLUGIN):
        return None
    lib = os.path.join(LIBDIR, "libOpenMMB200.so")
    ref_lib = os.path.join(ROOT, "oracle", "_ref", "libOpenMM.so")
    if not os.path.isdir(REF) or not os.path.exists(ref_lib):
        if os.path.exists(lib):
            return lib          # prebuilt copy travelled with the snapshot
        return None
    inc = ["-I" + os.path.join(REF, d, "include") for d in
           ("openmmapi", "olla", "serialization", "libraries/lepton", "libraries/sfmt", "libraries/vecmath", "platforms/reference")]
    inc += ["-I" + os.path.join(REF, "platforms/reference/src"), "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(REF, "tests")]
    srcs = sorted(os.path.join(PLUGIN, f) for f in os.listdir(PLUGIN) if f.endswith(".cpp") and not f.startswith("driver_"))
    hdrs = [os.path.join(PLUGIN, f) for f in os.listdir(PLUGIN) if f.endswith(".h")] + [os.path.join(ROOT, "include", "b200_nonbonded.h")]
    if srcs and (force or _newer(lib, srcs + hdrs)):
        _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", lib] + srcs + inc +
             ["-L" + LIBDIR, "-lb200nb", "-L" + os.path.dirname(ref_lib), "-lOpenMM",
              "-Wl,-rpath,$ORIGIN", "-Wl,-rpath,$ORIGIN/../../oracle/_ref"])
    # drivers (link against the reference's libOpenMM; they load the plugin at run time)
    for f in sorted(os.listdir(PLUGIN)):
        if f.startswith("driver_") and f.endswith(".cpp"):
            exe = os.path.join(ROOT, "oracle", "_ref", f[len("driver_"):-4])
            src = os.path.join(PLUGIN, f)
            if force or _newer(exe, [src] + hdrs):
                _run(["g++", "-O2", "-std=c++17", "-o", exe, src] + inc +
                     ["-L" + os.path.dirname(ref_lib), "-lOpenMM", "-ldl", "-pthread", "-Wl,-rpath,$ORIGIN"])
    return lib


def build_all(force=False):
    out = {"cuda": build_cuda(force), "oracle": build_oracle(force), "plugin": build_plugin(force)}
    return out


if __name__ == "__main__":
    print(build_all("--force" in sys.argv))
