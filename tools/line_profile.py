"""Attributes executed warp instructions / stall samples of one kernel in an .ncu-rep to source lines (via nvdisasm -g)."""
import collections, csv, os, re, subprocess, sys, tempfile
rep, kern, obj = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
for cub in sorted(f for f in os.listdir(tmp) if f.endswith(".cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.split("\n")
    hit = [i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l]
    if hit:
        break
start = hit[0]
cur, off2line = None, {}
for l in dis[start + 1:]:
    if l.startswith("//-----"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/", l)
    if m:
        off2line[int(m.group(1), 16)] = cur
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.split("\n")))
kre = os.environ.get("KNAME", "")   # substring of the demangled kernel name (first match)
kn = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and kre in r[1]][0]
hi = [i for i, r in enumerate(rows) if i > kn and r and r[0] == "Address"][0]
hdr = rows[hi]
ie, ns = hdr.index("Instructions Executed"), hdr.index("# Samples")
data = []
for r in rows[hi + 1:]:
    if len(r) >= len(hdr) and r[0].startswith("0x"): data.append(r)
    elif data: break
base = int(data[0][0], 16)
agg, samp = collections.Counter(), collections.Counter()
for r in data:
    k = off2line.get(int(r[0], 16) - base) or ("?", 0)
    agg[k] += int(r[ie]); samp[k] += int(r[ns])
tot, ts = sum(agg.values()), sum(samp.values())
print("kernel", kern, "warp instructions", tot, "sass", len(data))
byf = collections.Counter()
for k, v in agg.items():
    byf[k[0]] += v
print({k: round(100 * v / tot, 1) for k, v in byf.items()})
for k, v in agg.most_common(top):
    print("%-24s %5d  inst %5.2f%%  samples %5.2f%%" % (k[0], k[1], 100 * v / tot, 100 * samp[k] / ts))
if len(sys.argv) > 5:
    want = (sys.argv[5].split(":")[0], int(sys.argv[5].split(":")[1]))
    ops = collections.Counter()
    si = hdr.index("Source")
    for r in data:
        if (off2line.get(int(r[0], 16) - base) or ("?", 0)) == want:
            t = r[si].split()
            op = t[1] if t[0].startswith("@") else t[0]
            ops[op] += int(r[ie])
    print(want, [(k, round(100 * v / tot, 2)) for k, v in ops.most_common(20)])
if len(sys.argv) > 6:
    col = sys.argv[6]
    ci = hdr.index(col)
    st = collections.Counter()
    for r in data:
        try:
            st[off2line.get(int(r[0], 16) - base) or ("?", 0)] += int(r[ci])
        except ValueError:
            pass
    tt = sum(st.values()) or 1
    print(col, "total", tt, "of", ts)
    for k, v in st.most_common(14):
        print("   %-24s %5d  %5.1f%%" % (k[0], k[1], 100 * v / tt))
