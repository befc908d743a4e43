"""Analysis aid: SASS instructions per source line of one kernel (needs -lineinfo).
   python tools/sass_lines.py <kernel-substring> [top]"""
import collections, os, re, subprocess, sys, tempfile
so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "real_time_ray_tracer_b200", "librt_b200.so")
pat, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=d, capture_output=True)
cnt, total = collections.Counter(), 0
for f in os.listdir(d):
    if not f.endswith(".cubin") or f.startswith("rt_kernels-"):
        continue
    out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, f)], capture_output=True, text=True).stdout
    fn, loc = "", ""
    for line in out.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
        if m: fn = m.group(1); continue
        m = re.search(r'//## File "([^"]*)", line (\d+)', line)
        if m: loc = "%s:%s" % (os.path.basename(m.group(1)), m.group(2)); continue
        if pat in fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
            cnt[loc] += 1; total += 1
print("total", total)
for l, c in cnt.most_common(top):
    print(c, l)
