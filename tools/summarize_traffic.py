"""ncu csv (metrics dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum; -k regex:k_hist) ->
profiles/hist_traffic.json, which bench.py reports as roofline.traffic."""
import collections, csv, json, sys
src, dst = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(src)))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]
mi, ui, vi = h.index("Metric Name"), h.index("Metric Unit"), h.index("Metric Value")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}
per = collections.defaultdict(dict)
for r in rows[hdr + 1:]:
    if len(r) > vi:
        per[r[0]][r[mi]] = float(r[vi].replace(",", "")) * scale.get(r[ui], 1)
ids = sorted(per, key=int)
rd = sum(per[i]["dram__bytes_read.sum"] for i in ids)
wr = sum(per[i]["dram__bytes_write.sum"] for i in ids)
ns = sum(per[i]["gpu__time_duration.sum"] for i in ids)
root = per[ids[0]]
out = {"source": src, "launches": len(ids), "dram_bytes_read": rd, "dram_bytes_write": wr,
       "dram_bytes_per_launch": (rd + wr) / len(ids), "ncu_time_ms": ns / 1e6,
       "root_launch": {"dram_bytes": root["dram__bytes_read.sum"] + root["dram__bytes_write.sum"],
                       "ncu_time_us": root["gpu__time_duration.sum"] / 1e3},
       "note": "the rounds of one tree of bench.py workload"}
json.dump(out, open(dst, "w"), indent=1)
print(json.dumps(out, indent=1))
