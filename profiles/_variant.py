"""Build a variant of the library for an A/B run on one GPU box (scratch tool):
    python profiles/_variant.py NAME [conv_umma source] [-DFLAG ...]   ->  build/variants/libNAME.so
The other translation units are taken from build/obj (run tmdiff_b200.build first).  Select it with TMDIFF_B200_LIB."""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tmdiff_b200 import build as B
name = sys.argv[1]
src = os.path.join(B.CSRC, "conv_umma.cu")
flags = []
for a in sys.argv[2:]:
    if a.startswith("-D"): flags.append(a)
    else: src = a
B.build()
root = os.path.join(B.HERE, "..")
out_dir = os.path.join(root, "build", "variants"); os.makedirs(out_dir, exist_ok=True)
obj = os.path.join(out_dir, name + ".o")
cmd = [B._nvcc()] + B.NVCC_FLAGS + ["-I", B.CSRC] + flags + ["-c", src, "-o", obj]
r = subprocess.run(cmd, capture_output=True, text=True)
if r.returncode: sys.exit(r.stderr)
for line in r.stderr.splitlines():
    if "spill" in line and "0 bytes spill stores" not in line: print(line.strip())
objs = [os.path.join(root, "build", "obj", s.replace(".cu", ".o")) for s in B.SOURCES if s != "conv_umma.cu"] + [obj]
r = subprocess.run([B._nvcc(), "-shared", "-o", os.path.join(out_dir, "lib%s.so" % name)] + objs, capture_output=True, text=True)
if r.returncode: sys.exit(r.stderr)
print("built", name)
