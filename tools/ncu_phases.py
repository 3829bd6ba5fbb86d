"""Attribute executed instructions / stall samples of an .ncu-rep to source REGIONS of one kernel
source file using the full inline chain of every SASS instruction (nvdisasm line info of the
liblslam.so the report was taken with).
usage: ncu_phases.py <report.ncu-rep> <kernel-substring> <source-file> <name:first-last> ..."""
import csv, io, os, re, subprocess, sys, tempfile, collections
rep, ksub, SRC = sys.argv[1:4]
regions = []
for spec in sys.argv[4:]:
    name, rng = spec.split(":")
    a, b = rng.split("-")
    regions.append((name, int(a), int(b)))
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "lidar_slam_b200/csrc/liblslam.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info-inline", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
if not dis.strip():
    dis = subprocess.run(["nvdisasm", "--print-line-info", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
lines = dis.split("\n")
start = [i for i, l in enumerate(lines) if l.startswith("//---") and ksub in l and ".text." in l][0]
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith("//---")), len(lines))
cur = []; amap = {}; fresh = False
for ln in lines[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        if not fresh: cur = []; fresh = True          # a new chain starts (innermost frame first)
        cur.append((m.group(1).split('/')[-1], int(m.group(2))))
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S+)', ln)
    if m:
        amap[int(m.group(1), 16)] = (list(cur), m.group(2)); fresh = False
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]; i_ex = hdr.index('Instructions Executed'); i_smp = hdr.index('# Samples')
base = None; agg = collections.defaultdict(lambda: [0, 0]); tot = ts = 0
for r in rows[2:]:
    try: addr = int(r[0], 16) if r[0].startswith('0x') else int(r[0])
    except Exception: continue
    if base is None: base = addr
    ex = int(r[i_ex] or 0); smp = int(r[i_smp] or 0); tot += ex; ts += smp
    fr, op = amap.get(addr - base, ([], '?'))
    name = "(unattributed)"
    for f, l in fr:
        if f == SRC:
            hit = [n for n, a, b in regions if a <= l <= b]
            if hit: name = hit[0]; break
    if 'SYNCS' in op or op in ('YIELD',): pass
    a = agg[name]; a[0] += ex; a[1] += smp
print("total warp-instructions %d, samples %d" % (tot, ts))
for name, v in sorted(agg.items(), key=lambda x: -x[1][0]):
    print("  %5.1f%% inst %5.1f%% smp  %s" % (100 * v[0] / tot, 100 * v[1] / ts, name))
