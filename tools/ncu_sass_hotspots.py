"""Top SASS instructions by stall samples from `ncu --page source --csv` (sass view).  usage: ncu_sass_hotspots.py file.csv [N]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if len(r) > 6 and r[0] == "Address")
data = [r for r in rows if len(r) >= len(hdr) - 2 and r[0].startswith("0x")]
half = len(data) // 2 if len(data) > 1 and data[0][0] == data[len(data) // 2][0] else len(data)
data = data[:half]
iS, iI = hdr.index("# Samples"), hdr.index("Instructions Executed")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(d[iS]) for d in data)
st = collections.Counter()
for d in data:
    for i in stalls:
        st[hdr[i][6:]] += int(d[i])
print("samples", tot, "warp-instructions", sum(int(d[iI]) for d in data))
print(st.most_common(12))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
top = sorted(range(len(data)), key=lambda k: -int(data[k][iS]))[:n]
for k in sorted(top):
    d = data[k]
    s = {hdr[i][6:]: int(d[i]) for i in stalls if int(d[i]) > 0}
    print(k, d[1].strip()[:72].ljust(72), d[iS], d[iI], sorted(s.items(), key=lambda kv: -kv[1])[:2])
