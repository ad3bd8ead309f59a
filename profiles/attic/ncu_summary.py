"""ncu raw csv (ncu -i rep --page raw --csv) -> profiles/ncu_traffic.json + a text summary."""
import csv, json, sys
from pathlib import Path
raw, workload, out_txt = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.reader(open(raw)))
hdr, units, rows = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed.avg.per_cycle_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "lts__t_sector_hit_rate.pct",
        "smsp__inst_executed.sum"]
traffic = {}
lines = []
for r in rows:
    name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").split("<")[0]
    rd = float(r[col["dram__bytes_read.sum"]]) * scale[units[col["dram__bytes_read.sum"]]]
    wr = float(r[col["dram__bytes_write.sum"]]) * scale[units[col["dram__bytes_write.sum"]]]
    traffic[name] = int(rd + wr)
    lines.append(name)
    for w in want:
        if w in col:
            lines.append(f"    {w:58s} {r[col[w]]:>16s} {units[col[w]]}")
    lines.append(f"    {'dram bytes per launch (read+write)':58s} {int(rd + wr):>16d} byte")
p = Path("profiles/ncu_traffic.json")
d = json.loads(p.read_text()) if p.exists() else {}
d.setdefault(workload, {}).update(traffic)
p.write_text(json.dumps(d, indent=1, sort_keys=True) + "\n")
Path(out_txt).write_text("\n".join(lines) + "\n")
print(json.dumps(traffic, indent=1))
