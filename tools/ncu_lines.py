#!/usr/bin/env python3
"""Per-source-line instruction counts and stall samples of one kernel from an ncu report.
Joins `ncu --page source --csv` (SASS rows in program order) with `nvdisasm -g` line info of the
cubin extracted from libsoar_b200.so. Usage:
  python tools/ncu_lines.py <report.ncu-rep> <kernel-name> <object (e.g. blend_bwd)> [top] [invocation]"""
import csv, io, os, re, subprocess, sys, tempfile, collections

rep, kname, obj = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
invocation = sys.argv[5] if len(sys.argv) > 5 else None
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(root, "soar_b200", "libsoar_b200.so")],
               cwd=tmp, check=True, stdout=subprocess.DEVNULL)
cubin = os.path.join(tmp, f"{obj}.sm_100a.cubin")
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True, check=True).stdout
# walk the disassembly of the wanted function
lines, cur_line, in_fn = [], None, False
for l in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m:
        in_fn = kname in m.group(1)
        continue
    if not in_fn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur_line = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
        lines.append(cur_line)
nk = os.environ.get("NCU_KNAME", kname)  # template instantiations: mangled name for the cubin, plain for ncu
sel = ["--kernel-id", f"::{nk}:{invocation}"] if invocation else ["--kernel-name", nk]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + sel,
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
ci, cs = H.index("Instructions Executed"), H.index("# Samples")
body = []
for r in rows[hdr + 1:]:
    if r and r[0] == "Address":   # the next captured launch of the same kernel: one launch is enough
        break
    if len(r) == len(H):
        body.append(r)
if len(body) != len(lines):
    print(f"warning: {len(body)} SASS rows in the report vs {len(lines)} in the cubin", file=sys.stderr)
agg = collections.defaultdict(lambda: [0, 0])
for r, ln in zip(body, lines):
    agg[ln][0] += int(r[ci]); agg[ln][1] += int(r[cs])
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print(f"# {kname}: {ti} warp instructions, {ts} stall samples")
src_cache = {}
def src(ln):
    if ln is None: return ""
    f, n = ln
    if f not in src_cache:
        for d in ("soar_b200/csrc",):
            p = os.path.join(root, d, f)
            src_cache[f] = open(p).read().splitlines() if os.path.exists(p) else []
    L = src_cache[f]
    return L[n - 1].strip()[:90] if 0 < n <= len(L) else ""
for ln, (i, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    tag = f"{ln[0]}:{ln[1]}" if ln else "?"
    print(f"{100*i/ti:5.1f}% inst {100*s/max(ts,1):5.1f}% samp  {tag:24s} {src(ln)}")
