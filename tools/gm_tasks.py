"""Reads the VF_GM_DBG dump of k_geom_march (worker start/end, per-task durations) and prints the
finish-time spread of the workers and the cost distribution of the tasks."""
import sys, numpy as np
a = np.fromfile(sys.argv[1], np.uint64)
nw, nt, nstrips = int(a[0]), int(a[1]), int(a[2])
segs1, rows, rows2 = int(a[3] >> np.uint64(32)), int((a[3] >> np.uint64(8)) & np.uint64(0xffffff)), int(a[3] & np.uint64(255))
w = a[4:4 + 2 * nw].reshape(nw, 2).astype(np.int64); t = a[4 + 2 * nw:4 + 2 * nw + nt].astype(np.int64)
t0 = w[:, 0].min(); st = (w[:, 0] - t0) / 1e3; en = (w[:, 1] - t0) / 1e3
print(f"workers {nw} tasks {nt} strips {nstrips} segs1 {segs1} rows {rows}/{rows2}")
print("worker start us: min %.1f med %.1f max %.1f" % (st.min(), np.median(st), st.max()))
print("worker end   us: min %.1f p10 %.1f med %.1f p90 %.1f max %.1f" % (en.min(), np.percentile(en, 10), np.median(en), np.percentile(en, 90), en.max()))
print("idle share before the last worker ends: %.1f %%" % (100 * (1 - ((en - st).sum() / (nw * (en.max() - st.min()))))))
tu = t / 1e3
print("task us: mean %.2f med %.2f p90 %.2f p99 %.2f max %.2f  sum/workers %.1f" % (tu.mean(), np.median(tu), np.percentile(tu, 90), np.percentile(tu, 99), tu.max(), tu.sum() / nw))
seg = np.arange(nt) // nstrips
print("mean task us by tenth of the task list:", np.round([tu[i * nt // 10:(i + 1) * nt // 10].mean() for i in range(10)], 1))
top = np.argsort(-tu)[:10]
print("longest tasks (task, seg, strip, us):", [(int(k), int(seg[k]), int(k % nstrips), round(float(tu[k]), 1)) for k in top])
cta_end = en.reshape(-1, 8).max(1)
print("CTA end us: min %.1f med %.1f max %.1f" % (cta_end.min(), np.median(cta_end), cta_end.max()))
acc = a[4 + 2 * nw + nt:4 + 6 * nw + nt].astype(np.int64).reshape(nw, 4)
if acc[:, 3].sum():
    n = acc[:, 3].sum()
    print("per task (us): setup before TMA wait %.3f, TMA wait %.3f, priming row %.3f; tasks %d" % (acc[:, 0].sum() / n / 1e3, acc[:, 1].sum() / n / 1e3, acc[:, 2].sum() / n / 1e3, n))
