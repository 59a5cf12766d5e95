"""Per-region summary of k_geom_march from an ncu report (source page): warp instructions, stall samples
and average active threads per code region.  usage: ncu_regions.py REPORT.ncu-rep [kernel-regex]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; kre = sys.argv[2] if len(sys.argv) > 2 else "k_geom_march"
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + kre, "--launch-skip", "0", "--launch-count", "1"], capture_output=True, text=True).stdout
out, cur, h = {}, None, None
for r in csv.reader(io.StringIO(txt)):
    if not r: continue
    if r[0] in ("File Path", "File Name"): cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": h = r; iI = h.index("Instructions Executed"); iT = h.index("Thread Instructions Executed"); iS = h.index("# Samples"); continue
    if h is None or len(r) <= iT or r[0] == "": continue
    try: ln = int(r[0]); v = (float(r[iI] or 0), float(r[iT] or 0), float(r[iS] or 0))
    except ValueError: continue
    a = out.get((cur, ln), (0, 0, 0)); out[(cur, ln)] = tuple(x + y for x, y in zip(a, v))
src = {}
def region(f, n):
    import re, os
    if f not in src:
        p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vulkan-forge_b200", "csrc", f)
        src[f] = open(p).read().split("\n") if os.path.exists(p) else []
    name = "?"
    for k in range(min(n, len(src[f])) - 1, -1, -1):     # nearest enclosing definition above the line
        m = re.match(r"\s*(?:template.*>\s*)?(?:static\s+)?(?:__device__|__global__).*?\b(\w+)\s*\(", src[f][k]) or \
            re.match(r"^(k_\w+)\(", src[f][k]) or re.match(r"\s+auto (\w+) = \[&\]", src[f][k])
        if m: name = m.group(1); break
    return f.split(".")[0][-12:] + ":" + name
agg = {}
for (f, n), v in out.items():
    k = region(f, n) if f and f.startswith("vf_") else (f or "?")
    a = agg.get(k, (0, 0, 0)); agg[k] = tuple(x + y for x, y in zip(a, v))
tot = sum(v[0] for v in agg.values()) or 1; ts = sum(v[2] for v in agg.values()) or 1
print(f"total warp-inst {tot/1e6:.1f} M, samples {ts:.0f}")
for k, v in sorted(agg.items(), key=lambda x: -x[1][2]):
    if v[2] / ts >= 0.003: print(f"{k:42s} inst {v[0]/1e6:6.2f} M {100*v[0]/tot:5.1f} %   smp {100*v[2]/ts:5.1f} %   threads {v[1]/max(v[0],1):5.1f}")
