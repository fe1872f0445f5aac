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
from concurrent.futures import ThreadPoolExecutor
def one(part):          # conv_umma.cu is built in five parts (tmdiff_b200/build.py)
    obj = os.path.join(out_dir, "%s.p%s.o" % (name, part))
    cmd = [B._nvcc()] + B.NVCC_FLAGS + ["-I", B.CSRC] + flags + ["-DTMD_UMMA_PART=" + part, "-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode: sys.exit(r.stderr)
    return obj
with ThreadPoolExecutor(6) as ex: objs = list(ex.map(one, ["0", "1", "2", "3", "4", "5"]))
objs += [os.path.join(root, "build", "obj", s.replace(".cu", ".o")) for s in B.SOURCES if not s.startswith("conv_umma.cu")]
r = subprocess.run([B._nvcc(), "-shared", "-o", os.path.join(out_dir, "lib%s.so" % name)] + objs, capture_output=True, text=True)
if r.returncode: sys.exit(r.stderr)
print("built", name)
