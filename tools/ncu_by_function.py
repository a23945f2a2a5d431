"""Per-function breakdown of an ncu capture of the fully inlined step kernel.
   usage: python tools/ncu_by_function.py <report.ncu-rep> [kernel-substring]
Joins `ncu --page source --csv` (SASS view: samples and executed instructions per address) with the line table of the
in-tree cubin (`nvdisasm --print-line-info-inline`): every address is charged to the innermost source line it was inlined
from, and a line to the function whose definition precedes it in its file."""
import collections, csv, io, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = sys.argv[1]
kname = sys.argv[2] if len(sys.argv) > 2 else "fly_step_kernelILi16ELi1E"
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "flybody_b200", "libflyb.so")], cwd=tmp, stdout=subprocess.DEVNULL)
dis = []   # every kernel instantiation sits in a cubin of its own: take the one that holds the kernel asked for
for cubin in sorted(f for f in os.listdir(tmp) if f.endswith(".cubin")):
    syms = subprocess.run(["cuobjdump", "-elf", cubin], cwd=tmp, capture_output=True, text=True).stdout
    if kname in syms:
        dis = subprocess.run(["nvdisasm", "--print-line-info-inline", cubin], cwd=tmp, capture_output=True, text=True).stdout.splitlines()
        break
addr2line, inside, cur = {}, False, None
for ln in dis:
    if ln.startswith(".text."):
        inside = kname in ln
        cur = None
        continue
    if not inside:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)( inlined at)?', ln)
    if m:
        if cur is None or cur[2]:       # first line of a group = innermost frame
            cur = (os.path.basename(m.group(1)), int(m.group(2)), False)
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m and cur:
        addr2line[int(m.group(1), 16)] = cur[:2]
        cur = (cur[0], cur[1], True)    # next //## group starts a new mapping
# function table per source file
funcs = {}
for f in os.listdir(os.path.join(ROOT, "flybody_b200", "csrc")):
    tab = []
    for i, ln in enumerate(open(os.path.join(ROOT, "flybody_b200", "csrc", f)), 1):
        m = re.match(r"\s*(?:template\s*<[^>]*>\s*)?(?:FLY_FN|FLY_INL|FLY_MEM|FLY_NOINL|__global__|static|FLY_HD)\b[^;(]*?\b(\w+)\s*\(", ln)
        if m and not ln.strip().startswith("//"):
            tab.append((i, m.group(1)))
    funcs[f] = tab
def func_of(f, line):
    name = "?"
    for i, n in funcs.get(f, []):
        if i <= line:
            name = n
        else:
            break
    return f"{f}:{name}"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
ci = {k: H.index(k) for k in ("Address", "# Samples", "Instructions Executed", "Thread Instructions Executed", "stall_long_sb", "stall_barrier", "stall_wait", "stall_short_sb")}
agg = collections.defaultdict(lambda: [0] * 7)
base = None
for r in rows[hdr + 1:]:
    if len(r) <= max(ci.values()):
        continue
    a = int(r[ci["Address"]], 16)
    base = a if base is None else base
    f, line = addr2line.get(a - base, ("?", 0))
    key = func_of(f, line)
    vals = [int(float(r[ci[k]] or 0)) for k in ("# Samples", "Instructions Executed", "Thread Instructions Executed", "stall_long_sb", "stall_barrier", "stall_wait", "stall_short_sb")]
    for k, v in enumerate(vals):
        agg[key][k] += v
ts, ti = sum(v[0] for v in agg.values()), sum(v[1] for v in agg.values())
print(f"{'function':44s} {'samples%':>8s} {'instr%':>7s} {'lanes':>6s} {'long_sb%':>8s} {'barrier%':>8s} {'wait%':>6s} {'short%':>6s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
    s = max(v[0], 1)
    print(f"{k:44s} {100*v[0]/ts:8.2f} {100*v[1]/ti:7.2f} {v[2]/max(v[1],1):6.1f} {100*v[3]/s:8.1f} {100*v[4]/s:8.1f} {100*v[5]/s:6.1f} {100*v[6]/s:6.1f}")
print(f"total samples {ts}, warp instructions {ti}")
