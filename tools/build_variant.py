"""Builds a variant of the library with extra -D flags for ONE source file into cvpce_b200/variants/<name>.so (same-box A/B
runs select it with CVPCE_B200_SO).  usage: python tools/build_variant.py name source.cu -DX=1 -DY=2"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import build
name, source, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
build.build_library()
vdir = os.path.join(build.HERE, "variants"); os.makedirs(vdir, exist_ok=True)
nvcc = build._nvcc()
objs = []
for src in build.SOURCES:
    o = os.path.join(build.HERE, "build", src.replace(".cu", ".o"))
    if src == source:
        o = os.path.join(vdir, name + "_" + src.replace(".cu", ".o"))
        subprocess.check_call([nvcc] + build.NVCC_FLAGS + flags + ["-c", os.path.join(build.CSRC, src), "-o", o])
    objs.append(o)
out = os.path.join(vdir, "lib%s.so" % name)
subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"])
print(out)
