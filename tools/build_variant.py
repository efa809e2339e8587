#!/usr/bin/env python
"""Build a variant of libhifigan_b200.so for same-box A/B runs: `build_variant.py NAME [file.cu=replacement.cu ...] [-DFLAG ...]`
-> build/variants/NAME/libhifigan_b200.so (load it with HFG_LIB=<path>).  Box-to-box spread of kernel times is +-15 %, so
two builds are only comparable inside one gpurun call."""
import os, subprocess, sys, concurrent.futures
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from neuraltexttospeech_b200 import build as B

name = sys.argv[1]
repl = dict(a.split("=", 1) for a in sys.argv[2:] if "=" in a and not a.startswith("-"))
flags = [a for a in sys.argv[2:] if a.startswith("-")]
out = os.path.join(ROOT, "build", "variants", name)
os.makedirs(out, exist_ok=True)


def one(src):
    path = repl.get(src, os.path.join(B.CSRC, src))
    obj = os.path.join(out, src.replace(".cu", ".o"))
    r = subprocess.run([B._nvcc(), *B.NVCC_FLAGS, *flags, "-I", B.CSRC, "-c", path, "-o", obj], capture_output=True, text=True)
    if r.returncode:
        raise SystemExit(r.stdout + r.stderr)
    return obj


with concurrent.futures.ThreadPoolExecutor(max_workers=len(B.SOURCES)) as ex:
    objs = list(ex.map(one, B.SOURCES))
lib = os.path.join(out, "libhifigan_b200.so")
r = subprocess.run([B._nvcc(), "-shared", "-o", lib, *objs, "-cudart", "shared"], capture_output=True, text=True)
if r.returncode:
    raise SystemExit(r.stdout + r.stderr)
print(lib)
