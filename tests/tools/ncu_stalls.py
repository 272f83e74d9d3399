"""Summarise an ncu report's source page: stall reasons overall and for address ranges.
Usage: python tests/tools/ncu_stalls.py report.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[hi], [r for r in rows[hi + 1:] if len(r) >= len(rows[hi])]
idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
S = lambda r, h: int(r[idx[h]] or 0)
total = sum(S(r, "# Samples") for r in data)
print("kernel:", rows[0][1] if rows and len(rows[0]) > 1 else "?", " samples:", total)
agg = {s: sum(S(r, s) for r in data) for s in stalls}
for s, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]:
    print("  %-26s %9d %5.1f%%" % (s, v, 100.0 * v / max(total, 1)))
# segments between backward branches = loops; print per-loop instruction counts and samples
base = int(data[0][idx["Address"]], 16)
print("\naddr     samples  inst_exec  source  {dominant stalls}")
for r in sorted(data, key=lambda r: -S(r, "# Samples"))[:topn]:
    a = int(r[idx["Address"]], 16) - base
    dom = {s[6:]: S(r, s) for s in stalls if S(r, s) * 5 > S(r, "# Samples")}
    print("%06x %8d %10s  %-60s %s" % (a, S(r, "# Samples"), r[idx["Instructions Executed"]], r[idx["Source"]].strip()[:60], dom))
# cumulative samples by 0x400 bucket
print("\nsamples by code region (0x800 bytes):")
buckets = {}
for r in data:
    a = (int(r[idx["Address"]], 16) - base) // 0x800
    buckets[a] = buckets.get(a, 0) + S(r, "# Samples")
for a in sorted(buckets):
    if buckets[a] * 200 > total:
        print("  %06x  %8d %5.1f%%" % (a * 0x800, buckets[a], 100.0 * buckets[a] / total))
