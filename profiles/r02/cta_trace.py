#!/usr/bin/env python
"""profiles/r02/cta_trace.py <trace file> — reads the CTA table a -DZL_CTA_TRACE=1 build leaves (kind, cta, smid, clock64 in,
clock64 out; kind 0 gather_kernel of period 8, 1 plan_kernel and 2 walk_kernel of period 9) and prints, per SM and averaged:
how many gather CTAs are resident while planner CTAs are / are not on the same SM, and how long a gather CTA lives in
either case.  clock64 is per SM, so only intervals on the same SM are compared."""
import sys
import numpy as np

a = np.loadtxt(sys.argv[1], dtype=np.int64)
g = a[a[:, 0] == 0]; pl = a[a[:, 0] == 1]; wk = a[a[:, 0] == 2]
print(f"gather CTAs {len(g)}, plan CTAs {len(pl)}, walk CTAs {len(wk)}")
for name, k in (("plan", pl), ("walk", wk)):
    if len(k):
        d = (k[:, 4] - k[:, 3]) / 1965.0
        print(f"{name}: CTA life us mean {d.mean():.1f} p50 {np.median(d):.1f} max {d.max():.1f}; per SM {len(k) / 148:.1f}")
dg = (g[:, 4] - g[:, 3]) / 1965.0
print(f"gather: CTA life us mean {dg.mean():.1f} p50 {np.median(dg):.1f} p90 {np.percentile(dg, 90):.1f} max {dg.max():.1f}")
res_with, res_without, t_with, t_without = [], [], 0.0, 0.0
life_with, life_without = [], []
for sm in range(148):
    gs = g[g[:, 2] == sm]
    if not len(gs):
        continue
    ps = np.concatenate([pl[pl[:, 2] == sm], wk[wk[:, 2] == sm]])
    ps = ps[(ps[:, 4] - ps[:, 3]) > 2000]                       # CTAs that found an empty queue and left do not count
    lo, hi = gs[:, 3].min(), gs[:, 4].max()
    n = 2000
    grid = np.linspace(lo, hi, n)
    occ = ((gs[:, 3][None, :] <= grid[:, None]) & (gs[:, 4][None, :] > grid[:, None])).sum(1)
    pocc = ((ps[:, 3][None, :] <= grid[:, None]) & (ps[:, 4][None, :] > grid[:, None])).sum(1) if len(ps) else np.zeros(n, int)
    m = pocc > 0
    core = (grid > lo + 0.05 * (hi - lo)) & (grid < hi - 0.1 * (hi - lo))   # leave out ramp-up and tail
    if (m & core).any(): res_with.append(occ[m & core].mean())
    if (~m & core).any(): res_without.append(occ[~m & core].mean())
    t_with += (m & core).sum() / n * (hi - lo) / 1965.0; t_without += (~m & core).sum() / n * (hi - lo) / 1965.0
    if len(ps):
        ov = ((gs[:, 3][:, None] < ps[:, 4][None, :]) & (gs[:, 4][:, None] > ps[:, 3][None, :])).any(1)
    else:
        ov = np.zeros(len(gs), bool)
    life = (gs[:, 4] - gs[:, 3]) / 1965.0
    life_with += list(life[ov]); life_without += list(life[~ov])
print(f"resident gather CTAs per SM: with planner CTAs on the SM {np.mean(res_with) if res_with else float('nan'):.2f}, without {np.mean(res_without) if res_without else float('nan'):.2f}")
print(f"time per SM (us, core of the launch): with planner {t_with / 148:.1f}, without {t_without / 148:.1f}")
print(f"gather CTA life (us): overlapping a planner CTA on its SM {np.mean(life_with) if life_with else float('nan'):.1f} (n {len(life_with)}), not {np.mean(life_without) if life_without else float('nan'):.1f} (n {len(life_without)})")
span = [(g[g[:, 2] == sm][:, 4].max() - g[g[:, 2] == sm][:, 3].min()) / 1965.0 for sm in range(148) if (g[:, 2] == sm).any()]
print(f"gather span per SM us: mean {np.mean(span):.1f} min {np.min(span):.1f} max {np.max(span):.1f}")
