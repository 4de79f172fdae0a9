"""Per-CTA timeline of the tcgen05 plane-sweep kernel (RMVD_PSWEEP_STATS=1 RMVD_TC_CTA_DUMP=file): SM occupancy over time."""
import sys
import numpy as np
d = np.loadtxt(sys.argv[1], dtype=np.int64)
sm, t0, t1, cyc = d[:, 1], d[:, 2], d[:, 3], d[:, 4]
T0, T1 = t0.min(), t1.max()
print(f"ctas {len(d)}  kernel span {(T1 - T0) / 1e3:.1f} us  SMs used {len(np.unique(sm))}")
dur = (t1 - t0) / 1e3
print(f"CTA duration us: mean {dur.mean():.1f} min {dur.min():.1f} p50 {np.median(dur):.1f} p90 {np.percentile(dur, 90):.1f} max {dur.max():.1f}")
print(f"CTA cycles: mean {cyc.mean():.0f}  -> implied clock {cyc.mean() / dur.mean():.0f} MHz")
busy = np.zeros(sm.max() + 1)
last = np.zeros(sm.max() + 1)
cnt = np.zeros(sm.max() + 1, dtype=int)
for s in np.unique(sm):
    m = sm == s
    busy[s] = dur[m].sum()
    last[s] = (t1[m].max() - T0) / 1e3
    cnt[s] = m.sum()
span = (T1 - T0) / 1e3
used = np.unique(sm)
print(f"per SM: CTAs mean {cnt[used].mean():.2f} min {cnt[used].min()} max {cnt[used].max()};  busy (sum of CTA durations / 2 slots) mean {busy[used].mean() / 2:.1f} us "
      f"of span {span:.1f} us -> slot utilisation {busy[used].mean() / 2 / span:.3f}")
print(f"last CTA end per SM: min {last[used].min():.1f} p50 {np.median(last[used]):.1f} max {last[used].max():.1f} us")
first = np.array([(t0[sm == s].min() - T0) / 1e3 for s in used])
print(f"first CTA start per SM: min {first.min():.1f} p50 {np.median(first):.1f} max {first.max():.1f} us")
# concurrency over time
ev = np.concatenate([np.stack([t0, np.ones_like(t0)], 1), np.stack([t1, -np.ones_like(t1)], 1)])
ev = ev[np.argsort(ev[:, 0], kind="stable")]
conc = np.cumsum(ev[:, 1])
tt = (ev[:, 0] - T0) / 1e3
for q in np.linspace(0, span, 11)[1:]:
    i = np.searchsorted(tt, q) - 1
    print(f"  t={q:7.1f} us  resident CTAs {conc[max(i, 0)]}")
