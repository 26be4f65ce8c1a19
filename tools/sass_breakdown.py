"""Instruction count of a kernel per source region (nvdisasm -g line info): which code makes a cold-running kernel big."""
import collections, re, subprocess, sys, os, tempfile
obj = sys.argv[1]
bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 20
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, check=True, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
cur, cnt = None, collections.Counter()
only = sys.argv[4] if len(sys.argv) > 4 else None      # substring of the function (section) name to keep
keep = True
for line in txt.splitlines():
    ms = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if ms:
        keep = only is None or only in ms.group(1)
        continue
    if not keep:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
    elif re.match(r"\s+/\*[0-9a-f]{4,}\*/", line) and cur:
        cnt[cur] += 1
print("total instructions", sum(cnt.values()))
b = collections.Counter()
for (f, l), c in cnt.items():
    b[(f, l // bucket * bucket)] += c
for k, c in sorted(b.items(), key=lambda kv: -kv[1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 30]:
    print(k, c)
