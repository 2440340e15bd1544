"""Warp-stall sample breakdown per kernel of an .ncu-rep (source page), optionally the top source lines of one stall reason."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.split("\n")))
i = 0
done = set()
while i < len(rows):
    r = rows[i]
    if r and r[0] == "Kernel Name":
        name = r[1]; hdr = rows[i + 1]; i += 2
        cols = [(j, h) for j, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        tot = collections.Counter(); n = 0
        while i < len(rows) and rows[i] and rows[i][0].startswith("0x"):
            for j, h in cols:
                try: tot[h] += int(rows[i][j])
                except ValueError: pass
            i += 1
        if name not in done:
            done.add(name)
            s = sum(tot.values()) or 1
            print(name[:70], " ".join("%s %.1f%%" % (k[6:], 100 * v / s) for k, v in tot.most_common(9)))
    else:
        i += 1
