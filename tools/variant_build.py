#!/usr/bin/env python
"""Build an experimental variant of libfdb200.so with extra -D flags on one source file.

    python tools/variant_build.py NAME kshap_eval_tc.cu -DFDB_K2_SERVICE_SLEEP_NS=200 ...

writes fraud_detection_b200/csrc/build/variants/libfdb200_NAME.so (git-ignored, travels to the GPU box); select it
with FDB_LIB_PATH=<that path>.  The other objects are reused from the regular build (run build.py first).
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fraud_detection_b200"))
import build as fb  # noqa: E402

name, src, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
fb.build()
vdir = os.path.join(fb.BUILD_DIR, "variants")
os.makedirs(vdir, exist_ok=True)
obj = os.path.join(vdir, f"{src[:-3]}_{name}.o")
subprocess.run([fb._nvcc(), *fb.NVCC_FLAGS, *flags, "-c", os.path.join(fb.CSRC, src), "-o", obj], check=True)
objs = [obj if os.path.basename(s) == src else os.path.join(fb.BUILD_DIR, os.path.basename(s)[:-3] + ".o") for s in fb.sources()]
out = os.path.join(vdir, f"libfdb200_{name}.so")
subprocess.run([fb._nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out, *objs, "-lcuda"], check=True)
print(out)
