"""Compile a variant of the CUDA library with extra -D flags into _variants/<name>.so (git-ignored, travels with gpurun):

    python tools/build_variant.py split2 -DK5B_SPLIT_MAX=2 -DK5B_SPLIT_MINB=4

tools/variant_probe.py <lib> then times it on the M1 frame."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from shape_based_matching_b200 import build as b  # noqa: E402

name, flags = sys.argv[1], sys.argv[2:]
out = os.path.join(ROOT, "_variants", name + ".so")
os.makedirs(os.path.dirname(out), exist_ok=True)
flags_base = [f for f in b.NVCC_FLAGS if f not in ("-Xptxas", "-v")]
cmd = [b.NVCC] + flags_base + flags + ["-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(b.PKG, "csrc"),
                                        os.path.join(b.PKG, "csrc", "sbm_capi.cu"), "-o", out]
print("+", " ".join(cmd))
subprocess.check_call(cmd)
print(out)
