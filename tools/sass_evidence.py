"""Per-kernel SASS evidence for profiles/: which Blackwell-specific instructions each kernel of librt1b200.so contains.
   python tools/sass_evidence.py > profiles/sass_evidence.txt"""
import collections, re, subprocess, sys, os
LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rt1b200", "lib", "librt1b200.so")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "LDGMC", "MULTIMEM", "HMMA", "LDGSTS", "MUFU.TANH",
        "ATOMS", "RED.", "REDG", "ATOMG", "FFMA2", "ACQBULK", "STG.E.128", "LDG.E.128", "LDS.128", "STS.128"]
cur, counts, size = None, collections.OrderedDict(), {}
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); counts[cur] = collections.Counter(); size[cur] = 0; continue
    if cur and re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", line):
        size[cur] += 1
        for k in KEYS:
            if (re.search(r"(?<![A-Z])HMMA", line) if k == "HMMA" else k in line):
                counts[cur][k] += 1
dem = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
print(f"# cuobjdump -sass {os.path.relpath(LIB)}  (sm_100a) — instruction counts per kernel")
print("# UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA load/store (cp.async.bulk.tensor),")
print("# SYNCS = mbarrier ops, LDGMC/MULTIMEM = multimem.ld_reduce / multimem.st (NVLS), HMMA would be the legacy mma.sync path (none)")
print("# FFMA2 = fma.rn.f32x2 (packed fp32), REDG = red.global (incl. .F32x4 vector reductions), ACQBULK = griddepcontrol.wait")
for (name, c), d in zip(counts.items(), dem):
    short = re.sub(r"\(.*", "", d.replace("(anonymous namespace)::", ""))[:70]
    inter = " ".join(f"{k}={v}" for k, v in c.items())
    print(f"{short:72s} {size[name]:6d} instr  {inter}")
