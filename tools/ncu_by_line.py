"""Top source lines (innermost inlined frame) of an ncu capture of the step kernel.
   usage: python tools/ncu_by_line.py <report.ncu-rep> <kernel-substring> [file-filter] [top]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, kname = sys.argv[1], sys.argv[2]
filt = sys.argv[3] if len(sys.argv) > 3 else ""
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "flybody_b200", "libflyb.so")], cwd=tmp, stdout=subprocess.DEVNULL)
dis = []
for cubin in sorted(f for f in os.listdir(tmp) if f.endswith(".cubin")):
    syms = subprocess.run(["cuobjdump", "-elf", cubin], cwd=tmp, capture_output=True, text=True).stdout
    if kname in syms:
        dis = subprocess.run(["nvdisasm", "--print-line-info-inline", cubin], cwd=tmp, capture_output=True, text=True).stdout.splitlines()
        break
addr2line, inside, cur = {}, False, None
for ln in dis:
    if ln.startswith(".text."):
        inside = kname in ln; cur = None; continue
    if not inside: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)( inlined at)?', ln)
    if m:
        if cur is None or cur[2]: cur = (os.path.basename(m.group(1)), int(m.group(2)), False)
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m and cur:
        addr2line[int(m.group(1), 16)] = cur[:2]; cur = (cur[0], cur[1], True)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
keys = ("# Samples", "Instructions Executed", "Thread Instructions Executed", "stall_long_sb", "stall_barrier", "stall_wait", "stall_short_sb")
ci = {k: H.index(k) for k in ("Address",) + keys}
agg = collections.defaultdict(lambda: [0] * 7); base = None
for r in rows[hdr + 1:]:
    if len(r) <= max(ci.values()): continue
    a = int(r[ci["Address"]], 16); base = a if base is None else base
    f, line = addr2line.get(a - base, ("?", 0))
    for k, key in enumerate(keys): agg[(f, line)][k] += int(float(r[ci[key]] or 0))
ts = sum(v[0] for v in agg.values()); ti = sum(v[1] for v in agg.values())
src = {}
for (f, line), v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    if filt and filt not in f: continue
    if top <= 0: break
    top -= 1
    if f not in src:
        p = os.path.join(ROOT, "flybody_b200", "csrc", f)
        src[f] = open(p).read().splitlines() if os.path.exists(p) else []
    text = src[f][line - 1].strip()[:90] if 0 < line <= len(src[f]) else ""
    s = max(v[0], 1)
    print(f"{f}:{line:<4d} {100*v[0]/ts:5.2f}% ins {100*v[1]/ti:5.2f}% ln {v[2]/max(v[1],1):4.1f} lsb {100*v[3]/s:3.0f} bar {100*v[4]/s:3.0f} | {text}")
