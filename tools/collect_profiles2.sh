#!/bin/bash
# After tools/gpu_round2.sh SRC has come back: turn gpurun_out/SRC into the committed evidence profiles/TAG_*.
# usage: tools/collect_profiles2.sh SRC TAG   (e.g. r02u r02)
set -e
SRC=gpurun_out/$1; TAG=$2; DST=profiles
cp $SRC/bench.json $DST/${TAG}_bench.json
cp $SRC/bench_reference.json $DST/${TAG}_bench_reference.json
cp $SRC/launches.csv $DST/${TAG}_launches_bench.csv
for cfg in c3 c3d c3s c4 c5; do [ -f $SRC/profile_$cfg.log ] && cat $SRC/profile_$cfg.log; done > $DST/${TAG}_kernel_times.txt
ncu -i $SRC/full.ncu-rep --page raw --csv > $DST/${TAG}_ncu_full_raw.csv
ncu -i $SRC/full_c3d.ncu-rep --page raw --csv > $DST/${TAG}_ncu_full_c3depth_raw.csv
ncu -i $SRC/full_c5.ncu-rep --page raw --csv > $DST/${TAG}_ncu_full_c5_raw.csv
ncu -i $SRC/raster_nocacheflush.ncu-rep --page raw --csv > $DST/${TAG}_ncu_k_raster_nocacheflush_raw.csv
python tools/ncu_lines.py $SRC/full.ncu-rep k_geom_march 45 > $DST/${TAG}_k_geom_march_lines.txt
python tools/ncu_regions.py $SRC/full.ncu-rep k_geom_march >> $DST/${TAG}_k_geom_march_lines.txt
python tools/ncu_lines.py $SRC/full.ncu-rep k_shade_list 30 > $DST/${TAG}_k_raster_lines.txt
python tools/ncu_lines.py $SRC/full_c5.ncu-rep k_geom_march 30 > $DST/${TAG}_k_geom_march_c5_lines.txt
python tools/ncu_regions.py $SRC/full_c5.ncu-rep k_geom_march >> $DST/${TAG}_k_geom_march_c5_lines.txt
python tools/ncu_lines.py $SRC/full_c3d.ncu-rep k_geom_march 30 > $DST/${TAG}_k_geom_march_c3depth_lines.txt
python tools/ncu_regions.py $SRC/full_c3d.ncu-rep k_geom_march >> $DST/${TAG}_k_geom_march_c3depth_lines.txt
python - $DST/${TAG}_ncu_full_raw.csv $TAG <<'PY'
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1]))); h = rows[0]
out = {"source": f"profiles/{sys.argv[2]}_ncu_full_raw.csv (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum and smsp__inst_executed.sum per launch, c3 frame)"}
def num(r, name, unit_row=rows[1]):
    i = h.index(name); v = float(r[i].replace(',', '')); u = unit_row[i]
    return v * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}.get(u, 1)
for key, pat in (("k_geom", "k_geom_march"), ("k_raster", "k_shade_list"), ("k_resolve", "k_resolve_lean")):
    r = next(r for r in rows[2:] if pat in r[h.index("Kernel Name")])
    out[key] = int(num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum"))
    out[key + "_warp_inst"] = int(num(r, "smsp__inst_executed.sum"))
json.dump(out, open("profiles/dram_traffic.json", "w"), indent=1)
print(out)
PY
# SASS evidence from the library the captures ran: TMA / mbarrier / packed-f32 instructions of k_geom_march, the PNG kernels
LIB=vulkan-forge_b200/python/vulkan_forge/libvfcuda.so
{
  echo "# cuobjdump -sass $LIB : instruction counts per kernel (sm_100a cubin)"
  cuobjdump -sass $LIB 2>/dev/null | awk '
    /Function :/ { fn=$3 }
    /UTMALDG/ {c[fn,"UTMALDG"]++} /SYNCS/ {c[fn,"SYNCS"]++} /FFMA2/ {c[fn,"FFMA2"]++} /FMUL2/ {c[fn,"FMUL2"]++} /FADD2/ {c[fn,"FADD2"]++}
    /PRMT/ {c[fn,"PRMT"]++} /STS.128/ {c[fn,"STS.128"]++} /RED[G.]/ {c[fn,"RED"]++} /ATOMG/ {c[fn,"ATOMG"]++} /STL/ {c[fn,"STL"]++} /LDL/ {c[fn,"LDL"]++} /VOTE/ {c[fn,"VOTE"]++} /SHFL/ {c[fn,"SHFL"]++} /STG.E.128/ {c[fn,"STG.128"]++}
    { if (fn != "") n[fn]++ }
    END { for (f in n) { printf "%s lines=%d", f, n[f]; split("UTMALDG SYNCS FFMA2 FMUL2 FADD2 PRMT STS.128 RED ATOMG VOTE SHFL STG.128 STL LDL", k, " "); for (i=1;i<=14;i++) if (c[f,k[i]]) printf " %s=%d", k[i], c[f,k[i]]; printf "\n" } }' | sort
  echo
  echo "# k_geom_march<TMA=1, DEPTH=0, SHIPPED=0>: the lines themselves"
  cuobjdump -sass -fun '_Z12k_geom_marchILb1ELb0ELb0EEv7VfFrame14CUtensorMap_stPK6VfAxisPKfPyPjjS8_PK5uint2jjjPK6float2jS7_' $LIB 2>/dev/null | grep -E "UTMALDG|SYNCS|FFMA2|FMUL2|FADD2" | sed -e 's#/\* 0x[0-9a-f]* \*/##' -e 's/  */ /g' | head -120
} > $DST/${TAG}_sass_evidence.txt
wc -l $DST/${TAG}_sass_evidence.txt
