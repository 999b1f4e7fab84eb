"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA use (cuobjdump -sass on the in-tree library).
usage: python tools/sass_summary.py > profiles/sass_summary.txt"""
import re
import subprocess
import sys
from collections import OrderedDict

LIB = "hcache_deepspeed_b200/ops/lib/libhds_b200.so"
COLS = ["UTCHMMA", "UTCQMMA", ".2CTA", "LDTM", "UTMALDG", "UTCBAR", "UTCATOMSWS", "SYNCS", "REDG", "MUFU.EX2", "LDGMC", "LD.SYS", "ST.SYS", "HMMA"]

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
demangle = lambda names: subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
kernels, cur = OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kernels[cur] = dict.fromkeys(COLS, 0)
        continue
    if cur is None:
        continue
    for c in COLS:
        if c == "HMMA":
            kernels[cur][c] += len(re.findall(r"(?<![A-Z])HMMA", line))
        elif c == "LD.SYS":     # system-scope loads: flag acquires and peer-memory reads
            kernels[cur][c] += len(re.findall(r"LDG\.E\S*\.SYS", line))
        elif c == "ST.SYS":     # system-scope stores: flag releases over NVLink
            kernels[cur][c] += len(re.findall(r"STG\.E\S*\.SYS", line))
        elif c in line:
            kernels[cur][c] += 1
names = demangle(list(kernels))
print("# SASS evidence: tcgen05 / TMEM / TMA mnemonics per kernel (cuobjdump -sass libhds_b200.so; tools/sass_summary.py)")
print("# UTCHMMA = tcgen05.mma kind::f16, UTCQMMA = tcgen05.mma kind::f8f6f4, .2CTA = cta_group::2 forms (MMA, TMA load, commit),")
print("# LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA), UTCBAR = tcgen05.commit, UTCATOMSWS = tcgen05.alloc/dealloc,")
print("# SYNCS = mbarrier ops, REDG = red.global (incl. .SYS peer reductions), LDGMC = multimem.ld_reduce (in-switch reduction),")
print("# LD.SYS / ST.SYS = system-scope loads / stores: peer-memory reads, the flag acquire / release of the tensor-parallel")
print("# protocol inside the GEMM and norm kernels, and multimem.st (it assembles to STG.E.128.STRONG.SYS on the multicast")
print("# address), HMMA = legacy mma.sync (must stay 0)")
print(f"# {len(kernels)} kernels\n")
print(f"{'kernel':<112}" + "".join(f"{c:>11}" for c in COLS))
for (mangled, counts), name in zip(kernels.items(), names):
    short = re.sub(r"\(.*", "", name)[:110]
    print(f"{short:<112}" + "".join(f"{counts[c]:>11}" for c in COLS))
tot = {c: sum(k[c] for k in kernels.values()) for c in COLS}
print(f"\n{'TOTAL':<112}" + "".join(f"{tot[c]:>11}" for c in COLS))
