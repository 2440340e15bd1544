"""Aggregates executed warp instructions / stall samples of one kernel in an .ncu-rep by source region.
usage: agg_profile.py rep mangled-kernel-substring lib.so [ncu kernel regex]"""
import collections, csv, os, re, subprocess, sys, tempfile
rep, kern, obj = sys.argv[1], sys.argv[2], sys.argv[3]
kre = sys.argv[4] if len(sys.argv) > 4 else kern
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
for cub in sorted(f for f in os.listdir(tmp) if f.endswith(".cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.split("\n")
    hit = [i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l]
    if hit: break
cur, off2line, off2op = None, {}, {}
for l in dis[hit[0] + 1:]:
    if l.startswith("//-----"): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*)", l)
    if m:
        off2line[int(m.group(1), 16)] = cur
        t = m.group(2).split()
        op = t[1] if t[0].startswith("@") else t[0]
        off2op[int(m.group(1), 16)] = op.split(".")[0].rstrip(";")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.split("\n")))
kn = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and kre in r[1]][0]   # kre: substring of the demangled name
hi = [i for i, r in enumerate(rows) if i > kn and r and r[0] == "Address"][0]
hdr = rows[hi]
ie, ns = hdr.index("Instructions Executed"), hdr.index("# Samples")
data = [r for r in rows[hi + 1:] if len(r) >= len(hdr) and r[0].startswith("0x")]
first = next(i for i, r in enumerate(rows[hi + 1:]) if len(r) >= len(hdr) and r[0].startswith("0x"))
# only the first kernel instance in the page
seen, d2 = set(), []
for r in data:
    if r[0] in seen: break
    seen.add(r[0]); d2.append(r)
data = d2
base = int(data[0][0], 16)
R = "thcm_rows.cuh"; K = "thcm_kernels.cu"; C = "thcm_cell.cuh"; ST = "thcm_stage.cuh"
def region(f, ln):
    if f == R:
        for name, a, b in [("div", 184, 188), ("rho", 193, 195), ("tanh", 196, 241), ("face/mix", 242, 320), ("jac rows", 322, 467), ("rhs rows", 468, 557),
                           ("boundaries", 558, 648), ("rowops", 145, 170), ("filter", 650, 655), ("row_dot", 700, 713), ("emit/seam", 714, 800), ("sigbits", 40, 46)]:
            if a <= ln <= b: return name
    return f
agg, samp, ops = collections.Counter(), collections.Counter(), collections.Counter()
for r in data:
    o = int(r[0], 16) - base
    k = off2line.get(o) or ("?", 0)
    g = region(*k)
    agg[g] += int(r[ie]); samp[g] += int(r[ns]); ops[off2op.get(o, "?")] += int(r[ie])
tot, ts = sum(agg.values()), sum(samp.values())
print("kernel", kern, "warp instructions", tot, "sass", len(data))
for k, v in agg.most_common():
    print("%-24s inst %5.2f%%  samples %5.2f%%" % (k, 100 * v / tot, 100 * samp[k] / ts))
print("opcodes:", ", ".join("%s %.1f%%" % (k, 100 * v / tot) for k, v in ops.most_common(24)))
if len(sys.argv) > 5:
    agg2, s2 = collections.Counter(), collections.Counter()
    for r in data:
        k = off2line.get(int(r[0], 16) - base) or ("?", 0)
        if k[0] == sys.argv[5]: agg2[k[1]] += int(r[ie]); s2[k[1]] += int(r[ns])
    for k, v in agg2.most_common(30): print("   %s:%d inst %.2f%% samples %.2f%%" % (sys.argv[5], k, 100 * v / tot, 100 * s2[k] / ts))
