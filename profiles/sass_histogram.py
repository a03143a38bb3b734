"""SASS opcode histogram per kernel of libb200cnn.so (what proves the Blackwell-native paths, B200_PROFILING.md):
UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (1-D TMA),
UTMALDG = cp.async.bulk.tensor (tensor-map TMA), HMMA = mma.sync, SYNCS = mbarrier ops.
  python profiles/sass_histogram.py > profiles/r2_sass_histogram.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "distributed-machine-learning-using-cnn-cifar-10-dataset-_b200", "libb200cnn.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
OPS = ["UTCHMMA", "LDTM", "UTCBAR", "UBLKCP", "UTMALDG", "HMMA", "SYNCS", "MULTIMEM", "ATOMG", "RED", "ACQBULK", "ELECT"]
cur, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"b200::\(anonymous namespace\)::", "", cur)
        hist[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        hist[cur]["_total"] += 1
        for o in OPS:
            if op.startswith(o):
                hist[cur][o] += 1
print("# cuobjdump -sass libb200cnn.so (sm_100a): opcode counts per kernel; columns:", " ".join(OPS), "| total instructions")
for k, c in hist.items():
    if not any(c[o] for o in OPS):
        continue
    print("%-110s %s | %d" % (k[:110], " ".join("%5d" % c[o] for o in OPS), c["_total"]))
