"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: mean time per kernel name."""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, agg = None, collections.OrderedDict()
for r in rows:
    if len(r) > 5 and r[0] == "ID":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d["Metric Name"] != "gpu__time_duration.sum":
            continue
        k = re.sub(r"\(.*", "", d["Kernel Name"]).replace("void ", "").replace("jbk::", "").replace("<unnamed>::", "").replace("unnamed>::", "")[:48]
        v = float(d["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}[d["Metric Unit"]]
        agg.setdefault((k, d["Grid Size"], d["Block Size"]), []).append(v)
for (k, g, b), v in agg.items():
    print(f"{k:50s} grid={g:>16s} block={b:>12s} n={len(v):3d} mean={sum(v)/len(v):9.4f} ms")
