"""Summarise an ncu report's source page: samples by region (wait loops / softmax tile / MMA issue / exit)."""
import csv, subprocess, sys
from collections import Counter
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, data = rows[1], rows[2:]
iS, isrc, iex = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
tot = sum(int(r[iS]) for r in data)
print("total samples", tot, "instructions", len(data))
top = sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for i in sorted(top):
    r = data[i]
    print(f"{i:5d} {int(r[iS]):6d} {100*int(r[iS])/tot:5.1f}% ex={r[iex]:>9s}  {r[isrc].strip()[:80]}")
by_ex = Counter()
n_ex = Counter()
for r in data:
    by_ex[r[iex]] += int(r[iS]); n_ex[r[iex]] += 1
print("samples by execution count (count, n_instr, samples):")
for k, v in by_ex.most_common(10):
    print(f"   ex={k:>9s} n={n_ex[k]:5d} samples={v:6d} {100*v/tot:5.1f}%")
