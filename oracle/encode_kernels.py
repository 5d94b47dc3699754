"""oracle/encode_kernels.py -- TEST / BENCH INFRASTRUCTURE, not product code.

Build-time step of oracle/Makefile for the reference's own CUDA platform (the "reference on this GPU" row of bench.py).
The reference's cmake build turns every kernel file under <dir>/src/kernels into one `static const std::string` member of a
class <Class>KernelSources (cmake_modules/EncodeKernelFiles.cmake, platforms/cuda/sharedTarget/CMakeLists.txt:7-12,
platforms/common/CMakeLists.txt:3-14).  This script does the same job without cmake: it reads the kernel files and the two
templates <Class>.h.in / <Class>.cpp.in WHERE THEY LIE in the reference tree and writes the generated .h/.cpp into the
output directory given (oracle/_ref/gen, git-ignored).  Nothing from the reference is stored in this repository.

usage: encode_kernels.py <src dir holding kernels/ and the templates> <ClassName> <extension> <output dir>
"""
import glob
import os
import sys


def main():
    src, cls, ext, out = sys.argv[1:5]
    os.makedirs(out, exist_ok=True)
    decls, defs = [], []
    for path in sorted(glob.glob(os.path.join(src, "kernels", "*." + ext))):
        name = os.path.basename(path)[:-(len(ext) + 1)]
        text = open(path, encoding="utf-8", errors="replace").read()
        lines = text.split("\n")
        body = "\n".join('"' + l.replace("\\", "\\\\").replace('"', '\\"') + '\\n"' for l in lines)
        decls.append(f"static const std::string {name};")
        defs.append(f"const string {cls}::{name} = {body};")
    header = open(os.path.join(src, cls + ".h.in")).read().replace("@KERNEL_FILE_DECLARATIONS@", "\n".join(decls))
    source = open(os.path.join(src, cls + ".cpp.in")).read() + "\n" + "\n".join(defs) + "\n"
    for fname, content in ((cls + ".h", header), (cls + ".cpp", source)):
        p = os.path.join(out, fname)
        if not os.path.exists(p) or open(p).read() != content:
            open(p, "w").write(content)


if __name__ == "__main__":
    main()
