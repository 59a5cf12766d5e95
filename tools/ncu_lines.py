"""Summarise an ncu report: per-kernel headline metrics, and for one kernel the instruction split
between barriers plus the hottest source lines.  usage: ncu_lines.py REPORT.ncu-rep KERNEL_REGEX [top]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
def ncu(*a):
    return subprocess.run(["ncu", "-i", rep, *a], capture_output=True, text=True).stdout
raw = list(csv.reader(io.StringIO(ncu("--page", "raw", "--csv"))))
hdr, units = raw[0], raw[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'dram__throughput.avg.pct_of_peak_sustained_elapsed']
for r in raw[2:]:
    print('---', r[hdr.index('Kernel Name')][:50])
    for w in want:
        if w in hdr:
            print(f"   {w:62} {units[hdr.index(w)]:14} {r[hdr.index(w)][:20]}")
src = list(csv.reader(io.StringIO(ncu("--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                                      "regex:" + kre, "--launch-skip", "0", "--launch-count", "1"))))
agg, seg, segs, cur, h = {}, 0, {}, None, None
for r in src:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if r[0] == "Line No": h = r; iI = h.index('Instructions Executed'); iSm = h.index('# Samples'); continue
    if r[0] == "Function Name" or h is None or len(r) <= iI: continue
    try: v = float(r[iI] or 0); sm = float(r[iSm] or 0)
    except ValueError: continue
    if r[0] != '':
        k = (cur, int(r[0]), r[1][:100]); a = agg.get(k, (0, 0)); agg[k] = (a[0] + v, a[1] + sm)
tot = sum(v for v, _ in agg.values()) or 1; ts = sum(s for _, s in agg.values()) or 1
print(f"total warp-inst {tot/1e6:.1f}M, samples {ts:.0f}")
sass = list(csv.reader(io.StringIO(ncu("--page", "source", "--csv", "--kernel-name", "regex:" + kre, "--launch-skip", "0", "--launch-count", "1"))))
h = sass[1]; iI = h.index('Instructions Executed'); iS = h.index('Source'); iSm = h.index('# Samples')
for r in sass[2:]:
    if len(r) <= iI: continue
    try: v = float(r[iI] or 0); sm = float(r[iSm] or 0)
    except ValueError: break
    a = segs.get(seg, (0, 0)); segs[seg] = (a[0] + v, a[1] + sm)
    if 'BAR.SYNC' in r[iS]: seg += 1
for k, (v, sm) in segs.items():
    print(f"segment {k} (between barriers): {v/1e6:8.1f}M inst {v/tot*100:5.1f}%   samples {sm/ts*100:5.1f}%")
for k, (v, sm) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{v/tot*100:5.1f}% inst {sm/ts*100:5.1f}% smp | {k[0][:22]}:{k[1]:>4} | {k[2]}")
