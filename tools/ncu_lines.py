"""Analysis aid: join an ncu source-page CSV (SASS rows, in address order) with nvdisasm line info of the
same build -> executed warp instructions / stall samples per source line.
   python tools/ncu_lines.py <report.ncu-rep> <kernel-substring> [top] [--sass]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 and sys.argv[3].isdigit() else 40
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.environ.get("RT_SO") or os.path.join(root, "real_time_ray_tracer_b200", "librt_b200.so")  # RT_SO: the build the report was captured with
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=d, capture_output=True)
ins = []   # (loc, text)
for f in sorted(os.listdir(d)):
    if not f.endswith(".cubin") or f.startswith("rt_kernels-"):
        continue
    out = subprocess.run(["nvdisasm", "-gi" if "--outer" in sys.argv else "-g", "-c", os.path.join(d, f)], capture_output=True, text=True).stdout
    fn, loc = "", ""
    for line in out.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
        if m: fn = m.group(1); continue
        m = re.search(r'//## File "([^"]*)", line (\d+)', line)
        if m:
            loc = "%s:%s" % (os.path.basename(m.group(1)), m.group(2)); continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and pat in fn:
            ins.append((loc, m.group(2).strip()))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# a report may hold several kernels: take the section whose kernel name matches
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
sec = next(i for i in starts if os.environ.get("RT_KPAT", pat) in rows[i][1])   # RT_KPAT: the demangled name when it differs (templates)
hi = sec + 1
h = rows[hi]
end = next((i for i in starts if i > sec), len(rows))
body = [r for r in rows[hi + 1:end] if r]
ci, cs, ct = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
assert len(body) == len(ins), (len(body), len(ins))
by_line = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for (loc, text), r in zip(ins, body):
    v = (int(r[ci]), int(r[cs]), int(r[ct]))
    for k in range(3):
        by_line[loc][k] += v[k]; tot[k] += v[k]
print("total warp-inst %d samples %d thread-inst %d (lanes/inst %.2f)" % (tot[0], tot[1], tot[2], tot[2] / max(1, tot[0])))
print("%-28s %10s %6s %8s %6s %6s" % ("line", "warp-inst", "%", "samples", "%", "lanes"))
for loc, v in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-28s %10d %6.2f %8d %6.2f %6.1f" % (loc, v[0], 100 * v[0] / tot[0], v[1], 100 * v[1] / max(1, tot[1]), v[2] / max(1, v[0])))
if "--sass" in sys.argv:
    for (loc, text), r in zip(ins, body):
        print("%-26s %9s %6s %5.1f  %s" % (loc, r[ci], r[cs], int(r[ct]) / max(1, int(r[ci])), text))
