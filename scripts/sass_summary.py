"""Per-kernel SASS opcode counts of libaha_b200.so (what proves the Blackwell-native path: tcgen05.mma -> UTC*MMA,
tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UBLKCP/UBLKPF, mbarrier -> SYNCS, cta_group::2 -> the .2CTA forms).
    python scripts/sass_summary.py > profiles/r02_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                          "aha_b200", "lib", "libaha_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
ops = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UBLKCP", "UBLKPF", "SYNCS", "MUFU.EX2", "HMMA", "ATOM", "RED"]
kern, counts, two_cta = None, collections.OrderedDict(), collections.Counter()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        continue
    if kern is None:
        continue
    for o in ops:
        if re.search(r"\b" + re.escape(o), line):
            counts[kern][o] += 1
    if "2CTA" in line:
        two_cta[kern] += 1
dem = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
print(f"cuobjdump -sass {os.path.basename(lib)} (sm_100a): opcode counts per kernel")
print(f"{'kernel':80s} " + " ".join(f"{o:>8s}" for o in ops) + "    .2CTA")
tot = collections.Counter()
for (k, c), d in zip(counts.items(), dem):
    name = re.sub(r"\(anonymous namespace\)::", "", d)
    name = re.sub(r"\(.*", "", name)[:80]
    if sum(c.values()) == 0:
        continue
    print(f"{name:80s} " + " ".join(f"{c[o]:8d}" for o in ops) + f" {two_cta[k]:8d}")
    tot.update(c)
print(f"{'TOTAL':80s} " + " ".join(f"{tot[o]:8d}" for o in ops) + f" {sum(two_cta.values()):8d}")
deps = subprocess.run(["ldd", lib], capture_output=True, text=True).stdout
print("\nldd:", ", ".join(sorted(set(l.split()[0] for l in deps.splitlines() if l.strip()))))
