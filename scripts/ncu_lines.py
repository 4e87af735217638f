"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump by CUDA source line (stall samples)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None
per_line = collections.Counter(); per_file = collections.Counter(); text = {}
stalls = collections.Counter(); line_stall = collections.defaultdict(collections.Counter)
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; si = hdr.index("# Samples"); sc = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]; continue
    if hdr is None or len(r) <= si or r[2] != "-": continue   # only the CUDA-line rows (Address == "-")
    try: n = int(r[si])
    except ValueError: continue
    key = (cur, int(r[0])); per_line[key] += n; per_file[cur] += n; text[key] = r[1].strip()
    for i in sc:
        try: v = int(r[i])
        except ValueError: v = 0
        stalls[hdr[i]] += v; line_stall[key][hdr[i]] += v
tot = sum(per_line.values())
print("total samples", tot, dict(per_file))
print("stalls", [(k, round(v / max(1, sum(stalls.values())), 3)) for k, v in stalls.most_common(8)])
for k, n in per_line.most_common(top):
    print(f"{n / tot:6.3f} {k[0]}:{k[1]:4d} {text[k][:110]}  | {[(a[6:], b) for a, b in line_stall[k].most_common(2)]}")
