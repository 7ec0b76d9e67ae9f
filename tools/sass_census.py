"""SASS opcode census of libnerf_b200.so (no GPU needed): per kernel, how many tcgen05 / TMA / TMEM instructions the
sm_100a cubins contain -- the evidence that the tensor-core path is tcgen05 (UTCHMMA), its accumulators live in TMEM
(LDTM = tcgen05.ld), operands arrive by bulk async copies (UBLKCP) and tensor-map TMA (UTMALDG), and completion is
signalled through tcgen05.commit (UTCBAR).      python tools/sass_census.py > profiles/r2_sass_census.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "nerf_shared_b200", "libnerf_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
OPS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "UTCBAR", "UTCBAR.2CTA", "SYNCS", "HMMA", "FFMA", "MUFU"]
counts, arch, cur = collections.OrderedDict(), set(), None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    m = re.search(r"arch = (sm_\w+)", line)
    if m:
        arch.add(m.group(1))
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        c = counts[cur]
        c["_total"] += 1
        base = op.split(".")[0]
        if base in OPS:
            c[base] += 1
        if base in ("UTCHMMA", "UTCBAR") and ".2CTA" in op:
            c[base + ".2CTA"] += 1


def demangle(n):
    try:
        return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(")[0]
    except Exception:
        return n


print("# cuobjdump -sass nerf_shared_b200/libnerf_b200.so ; cubin architectures: %s ; %d kernels" % (", ".join(sorted(arch)), len(counts)))
print("%-58s %7s " % ("kernel", "instrs") + " ".join("%12s" % o for o in OPS))
tot = collections.Counter()
for k, c in counts.items():
    print("%-58s %7d " % (demangle(k)[:58], c["_total"]) + " ".join("%12d" % c[o] for o in OPS))
    tot.update(c)
print("%-58s %7d " % ("TOTAL", tot["_total"]) + " ".join("%12d" % tot[o] for o in OPS))
