"""Per-kernel census of the Blackwell-specific SASS in libi64b.so (cuobjdump -sass): tcgen05 MMAs (UTC*MMA), TMEM traffic
(LDTM / STTM), TMA (UTMALDG / UTMASTG / UBLKCP), mbarrier / cluster ops, and the legacy tensor-core mnemonics that must NOT
appear (HMMA / IMMA / QMMA warp-level MMAs).  Usage: python tools/sass_census.py [lib] > profiles/<round>_sass_census.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vllm-i64_b200", "lib", "libi64b.so")
PATTERNS = [("UTCHMMA", r"\bUTCHMMA"), ("UTCIMMA", r"\bUTCIMMA"), ("UTCQMMA", r"\bUTCQMMA"), (".2CTA", r"\.2CTA"),
            ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"),
            ("GATHER4", r"GATHER4"), ("UBLKCP", r"\bUBLKCP"), ("UTCBAR", r"\bUTCBAR"), ("MULTICAST", r"MULTICAST"),
            ("SYNCS", r"\bSYNCS"), ("UCGABAR", r"\bUCGABAR"), ("legacy HMMA/IMMA/QMMA", r"(?<![A-Z])(HMMA|IMMA|QMMA)\b")]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
per, total, cur = collections.OrderedDict(), collections.Counter(), None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        per[cur] = collections.Counter()
        continue
    if cur is None or not re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
        continue
    per[cur]["instructions"] += 1
    for name, pat in PATTERNS:
        if re.search(pat, line):
            per[cur][name] += 1
            total[name] += 1
arch = sorted(set(re.findall(r"arch = (sm_\w+)", txt)))
print(f"library: {os.path.relpath(lib, ROOT)}   arch: {', '.join(arch)}   kernels: {len(per)}")
print("whole library: " + ", ".join(f"{n} {total[n]}" for n, _ in PATTERNS))
print()
for fn, c in per.items():
    hits = [f"{n} {c[n]}" for n, _ in PATTERNS if c[n]]
    name = demangle(fn)
    name = (name[:name.find(">(") + 1] if ">(" in name else name.split("(")[0]).replace("(anonymous namespace)::", "").replace("<unnamed>::", "")[:110]
    print(f"{name:<112} {c['instructions']:>6} instr   " + (", ".join(hits) if hits else "-"))
