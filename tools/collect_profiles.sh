#!/bin/bash
# After tools/gpu_round.sh TAG has come back: turn gpurun_out/TAG into the committed evidence under profiles/.
# usage: tools/collect_profiles.sh TAG   (e.g. r01f)
set -e
TAG=$1; SRC=gpurun_out/$TAG; DST=profiles
cp $SRC/bench.json $DST/${TAG}_bench.json
cp $SRC/launches.csv $DST/${TAG}_launches_bench.csv
ncu -i $SRC/full.ncu-rep --page raw --csv > $DST/${TAG}_ncu_full_raw.csv
python tools/ncu_lines.py $SRC/full.ncu-rep k_geom_march 45 > $DST/${TAG}_k_geom_march_lines.txt
python tools/ncu_regions.py $SRC/full.ncu-rep k_geom_march >> $DST/${TAG}_k_geom_march_lines.txt
python tools/ncu_lines.py $SRC/full.ncu-rep k_raster 30 > $DST/${TAG}_k_raster_lines.txt
python - $DST/${TAG}_ncu_full_raw.csv $TAG <<'PY'
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1]))); h = rows[0]
out = {"source": f"profiles/{sys.argv[2]}_ncu_full_raw.csv (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum and smsp__inst_executed.sum per launch, c3 frame)"}
def num(r, name, unit_row=rows[1]):
    i = h.index(name); v = float(r[i].replace(',', '')); u = unit_row[i]
    return v * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}.get(u, 1)
for key, pat in (("k_geom", "k_geom_march"), ("k_raster", "k_raster")):
    r = next(r for r in rows[2:] if pat in r[h.index("Kernel Name")])
    out[key] = int(num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum"))
    out[key + "_warp_inst"] = int(num(r, "smsp__inst_executed.sum"))
json.dump(out, open("profiles/dram_traffic.json", "w"), indent=1)
print(out)
PY
