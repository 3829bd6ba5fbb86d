"""Attribute executed instructions and stall samples of an .ncu-rep to source lines of
csrc/ransac.cuh (needs the liblslam.so the report was taken with, for line info).
usage: ncu_lines.py <report.ncu-rep> <kernel-mangled-substring> [topN]"""
import csv, io, re, subprocess, sys, os, tempfile
rep, ksub = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
SRC = sys.argv[4] if len(sys.argv) > 4 else 'ransac.cuh'
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "lidar_slam_b200/csrc/liblslam.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
# isolate the kernel's section
lines = dis.split("\n")
start = [i for i, l in enumerate(lines) if l.startswith("//---") and ksub in l and ".text." in l][0]
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith("//---")), len(lines))
cur = None; amap = {}
for ln in lines[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        frames = [(m.group(1).split('/')[-1], int(m.group(2)))] + [(a.split('/')[-1], int(b)) for a, b in re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))]
        cur = frames; continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/', ln)
    if m: amap[int(m.group(1), 16)] = cur
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]; i_ex = hdr.index('Instructions Executed'); i_smp = hdr.index('# Samples')
base = None; agg = {}; tot = ts = 0
for r in rows[2:]:
    try: addr = int(r[0], 16) if r[0].startswith('0x') else int(r[0])
    except Exception: continue
    if base is None: base = addr
    ex = int(r[i_ex] or 0); smp = int(r[i_smp] or 0); tot += ex; ts += smp
    fr = amap.get(addr - base)
    key = ('?', 0)
    if fr:
        key = fr[-1]
        for f, l in fr[::-1]:
            if f == SRC: key = (f, l); break
    a = agg.setdefault(key, [0, 0]); a[0] += ex; a[1] += smp
src = open(os.path.join(root, "lidar_slam_b200/csrc", SRC)).read().split("\n")
print("total warp-instructions %d, samples %d" % (tot, ts))
# phase totals by line ranges (markers in the source)
marks = [(i + 1, l.strip()) for i, l in enumerate(src) if "=================" in l]
marks = [(1, "prologue/staging")] + marks + [(len(src) + 1, "end")]
for (a, name), (b, _) in zip(marks[:-1], marks[1:]):
    e = sum(v[0] for k, v in agg.items() if k[0] == SRC and a <= k[1] < b)
    s = sum(v[1] for k, v in agg.items() if k[0] == SRC and a <= k[1] < b)
    print("  %5.1f%% inst %5.1f%% smp  lines %d-%d %s" % (100 * e / tot, 100 * s / ts, a, b - 1, name[:70]))
e = sum(v[0] for k, v in agg.items() if k[0] != SRC); s = sum(v[1] for k, v in agg.items() if k[0] != SRC)
print("  %5.1f%% inst %5.1f%% smp  (other files / unmapped)" % (100 * e / tot, 100 * s / ts))
for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    txt = src[k[1] - 1].strip()[:80] if k[0] == SRC and k[1] > 0 else ''
    print('%5.1f%% inst %5.1f%% smp  %s:%d  %s' % (100 * v[0] / tot, 100 * v[1] / ts, k[0], k[1], txt))
