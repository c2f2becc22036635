"""Compare two --dump-shapes tables from bench.py: per-shape time deltas (run in the build container)."""
import json
import sys

a = {(r["kernel"], r["shape"]): r for r in json.load(open(sys.argv[1]))}
b = {(r["kernel"], r["shape"]): r for r in json.load(open(sys.argv[2]))}
ta, tb = sum(r["ms"] for r in a.values()), sum(r["ms"] for r in b.values())
print(f"total {ta:.2f} -> {tb:.2f} ms")
fam = {}
for k, r in b.items():
    fam.setdefault(k[0], [0.0, 0.0])[1] += r["ms"]
for k, r in a.items():
    fam.setdefault(k[0], [0.0, 0.0])[0] += r["ms"]
for k, v in sorted(fam.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:24s} {v[0]:8.3f} -> {v[1]:8.3f}")
rows = []
for k in set(a) | set(b):
    ma, mb = a.get(k, {"ms": 0})["ms"], b.get(k, {"ms": 0})["ms"]
    rows.append((mb - ma, k, ma, mb))
rows.sort()
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
for d, k, ma, mb in rows[:n] + rows[-8:]:
    r = b.get(k) or a.get(k)
    tf = r["flops"] / max(mb, 1e-9) / 1e9 if r["flops"] else 0
    print(f"{d:+7.3f}  {ma:7.3f} -> {mb:7.3f}  n={r['launches']:3d} tf={tf:5.0f}  {k[0]} {k[1]}")
