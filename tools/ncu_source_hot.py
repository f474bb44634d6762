"""Top stall-sampled SASS instructions of an `ncu --page source --csv` dump, plus stall-reason totals."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) >= len(hdr) - 2]
H = {h: i for i, h in enumerate(hdr)}
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
def f(r, k):
    try: return float(r[H[k]].replace(',', ''))
    except Exception: return 0.0
tot = sum(f(r, "# Samples") for r in data)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(f(r, s) for r in data) for s in stalls}
print("total samples", tot)
print("stall totals:", ", ".join(f"{s[6:]}={v / max(tot,1) * 100:.1f}%" for s, v in sorted(agg.items(), key=lambda x: -x[1])[:10]))
cum = 0
for idx, r in sorted(enumerate(data), key=lambda x: -f(x[1], "# Samples"))[:top]:
    s = f(r, "# Samples")
    why = sorted(((f(r, k), k[6:]) for k in stalls), reverse=True)[:2]
    print(f"{idx:5d} {s / max(tot,1) * 100:5.1f}%  exec={f(r, 'Instructions Executed'):12.0f}  {r[H['Source']][:70]:70s} {why[0][1]}:{why[0][0]:.0f} {why[1][1]}:{why[1][0]:.0f}")
