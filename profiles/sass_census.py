"""Opcode census of the built library (cuobjdump -sass): per kernel, the instructions that prove which hardware paths it uses.
    python profiles/sass_census.py > profiles/sass_census_r02.md
UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA tensor load / store, UBLKCP = cp.async.bulk, LDTM / STTM = tcgen05.ld / st,
UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, MUFU.TANH = tanh.approx, RED = red.global (fused 32->1 projection), HMMA = mma.sync."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "tmdiff_b200", "lib", "libtmdiff_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
keys = ["UTCHMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "SYNCS", "MUFU.TANH", "RED.", "HMMA", "STL", "LDL"]
per = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::", "", cur)
        cur = re.sub(r"\(.*", "", cur)
        per[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        per[cur]["_total"] += 1
        for k in keys:
            if op.startswith(k):
                per[cur][k] += 1
tot = collections.Counter()
for c in per.values():
    tot.update(c)
print("# SASS opcode census of libtmdiff_b200.so (sm_100a), `profiles/sass_census.py`\n")
print("Whole library: " + ", ".join("%s %d" % (k, tot[k]) for k in keys) + ", %d instructions in %d kernels.\n" % (tot["_total"], len(per)))
print("| kernel | instr | " + " | ".join(keys) + " |")
print("|---|---:|" + "---:|" * len(keys))
groups = collections.OrderedDict()
for name, c in per.items():
    g = re.sub(r"conv_umma_kernel<(\d+), (\d+), (\d+), \(bool\)(\d), (-?\d+)>", lambda m: "conv_umma_kernel<%s,%s,%s,%s,EF=%s>" % m.groups(), name)
    groups[g] = c
for name, c in groups.items():
    if any(c[k] for k in keys[:7]) or "umma" in name or "th_" in name:
        print("| `%s` | %d | " % (name[:90], c["_total"]) + " | ".join(str(c[k]) if c[k] else "" for k in keys) + " |")
print("\nKernels without tensor-core / TMA / mbarrier instructions (plain SIMT): " + ", ".join("`%s`" % n.split("<")[0].split("::")[-1] for n, c in groups.items()
      if not any(c[k] for k in keys[:7]) and "umma" not in n and "th_" not in n))
