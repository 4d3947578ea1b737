"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA / mbarrier / multimem use
(cuobjdump -sass on the in-tree library; runs on the CPU box). Output: profiles/r2/sass_evidence_r2.txt (or argv[1])."""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "distributed_tensorflow_framework_b200", "lib", "libdtf_b200_cuda.so")
KEEP = re.compile(r"^(LDGMC|STGMC|REDGMC|UTCHMMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|SYNCS|LDGSTS|MEMBAR|ELECT|RED\.|ATOM|MULTIMEM|LD\.E\..*SYS|ST\.E\..*SYS|UBLKCP)")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
counts = collections.OrderedDict()
fn = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        counts[fn] = collections.Counter()
        continue
    if fn:
        for op in re.findall(r"\b((?:LDGMC|STGMC|REDGMC|UTCHMMA|UTCQMMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|"
                             r"SYNCS|LDGSTS|UBLKCP|ELECT|LDS|STS|RED|ATOMG?)(?:\.[A-Za-z0-9_]+)*)\b", line):
            op = re.sub(r"^(LDTM)\..*", r"\1.", op)
            counts[fn][op] += 1
with open(sys.argv[1] if len(sys.argv) > 1 else os.path.join(root, "profiles", "r2", "sass_evidence_r2.txt"), "w") as f:
    for fn, c in counts.items():
        for op, n in sorted(c.items()):
            f.write("%-100s %-34s %d\n" % (fn, op, n))
print("kernels:", len(counts))
