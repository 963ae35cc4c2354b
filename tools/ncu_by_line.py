"""Join an `ncu --page source --csv` dump of one kernel with nvdisasm line info: per source line, the warp
instructions executed, average active threads and stall samples.
    python tools/ncu_by_line.py SRC.csv CUBIN MANGLED_SUBSTR [top]"""
import collections, csv, re, subprocess, sys

src_csv, cubin, fn_sub = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dis = subprocess.run(["nvdisasm", "-c", "-g", cubin], capture_output=True, text=True).stdout
line_of, cur_fn, cur, inl = {}, None, None, ""
for l in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m:
        cur_fn = m.group(1); cur = None; continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
    if m and cur_fn and fn_sub in cur_fn:
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ia, ie, it, isamp = h.index("Address"), h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
base = None
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for r in rows[hi + 1:]:
    if len(r) <= it or not r[ia] or r[ia] == "Address":
        base = None if (r and r[0] in ("Address", "Kernel Name")) else base
        continue
    a = int(r[ia], 16) if not r[ia].isdigit() else int(r[ia])
    if base is None:
        base = a
    key = line_of.get(a - base)
    v = [int(r[ie] or 0), int(r[it] or 0), int(r[isamp] or 0)]
    for k in range(3):
        agg[key][k] += v[k]; tot[k] += v[k]
print(f"total inst {tot[0]:,} thread-inst {tot[1]:,} avg threads {tot[1] / max(1, tot[0]):.2f} samples {tot[2]}")
byfile = collections.defaultdict(lambda: [0, 0, 0])
for k, v in agg.items():
    f = k[0] if k else None
    for i in range(3):
        byfile[f][i] += v[i]
for f, v in sorted(byfile.items(), key=lambda x: -x[1][0]):
    print(f"  {str(f):28s} inst {100 * v[0] / tot[0]:5.1f}%  thr/inst {v[1] / max(1, v[0]):5.1f}  samples {100 * v[2] / max(1, tot[2]):5.1f}%")
for k, v in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"  {str(k):34s} inst {100 * v[0] / tot[0]:5.1f}%  thr/inst {v[1] / max(1, v[0]):5.1f}  samples {100 * v[2] / max(1, tot[2]):5.1f}%")
