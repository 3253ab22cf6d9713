"""Summarise a CTA residency trace (DLA_B200_TRACE=<file>, slices_path.cuh): for each kernel, CTA lifetimes, and per
SM the number of Voigt CTAs resident while a contraction CTA is resident there.

    python tools/trace_report.py gpurun_out/trace.bin
"""
import sys
import numpy as np

a = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(-1, 4)
kid, sm, t0, t1 = a[:, 0], a[:, 1].astype(int), a[:, 2].astype(np.int64), a[:, 3].astype(np.int64)
org = t0.min()
t0, t1 = (t0 - org) / 1e3, (t1 - org) / 1e3   # us
print(f"{len(a)} CTAs, span {t1.max():.0f} us, SMs {len(set(sm))}")
for k, name in ((1, "voigt_slices (K1s)"), (2, "contract_slices (K2s)")):
    m = kid == k
    if not m.any():
        continue
    d = t1[m] - t0[m]
    print(f"{name}: {m.sum()} CTAs, lifetime us min/median/max {d.min():.1f}/{np.median(d):.1f}/{d.max():.1f}, "
          f"first start {t0[m].min():.0f}, last end {t1[m].max():.0f}")
# K2s launches: group by start time gaps
m2 = kid == 2
if m2.any():
    order = np.argsort(t0[m2])
    s2, e2, sm2 = t0[m2][order], t1[m2][order], sm[m2][order]
    # launches = clusters of 148 CTAs
    bounds = [0] + [i for i in range(1, len(s2)) if s2[i] - s2[i - 1] > 200] + [len(s2)]
    print("K2s launches: start spread (first..last CTA start), duration of the longest CTA, Voigt CTAs resident on the same SM (time-average)")
    m1 = kid == 1
    s1, e1, sm1 = t0[m1], t1[m1], sm[m1]
    for b0, b1 in zip(bounds[:-1], bounds[1:]):
        ss, ee, sms = s2[b0:b1], e2[b0:b1], sm2[b0:b1]
        co = []
        for x, y, z in zip(ss[:20], ee[:20], sms[:20]):   # sample 20 SMs
            sel = sm1 == z
            ov = np.clip(np.minimum(e1[sel], y) - np.maximum(s1[sel], x), 0, None).sum()
            co.append(ov / max(y - x, 1e-9))
        # Voigt CTA lifetime inside vs outside this launch window
        inside = m1 & False
        w0, w1 = ss.min(), ee.max()
        mid = (s1 + e1) / 2
        ins = (mid > w0) & (mid < w1)
        life_in = np.median((e1 - s1)[ins]) if ins.any() else float("nan")
        print(f"  n={b1 - b0:4d} start {ss.min():8.0f}..{ss.max():8.0f} us  end {ee.max():8.0f}  dur {np.max(ee - ss):7.0f} us  "
              f"co-resident Voigt CTAs/SM {np.mean(co):.2f}  Voigt CTA lifetime inside {life_in:.0f} us")
    outside = np.ones(len(s1), bool)
    for b0, b1 in zip(bounds[:-1], bounds[1:]):
        w0, w1 = s2[b0:b1].min(), e2[b0:b1].max()
        mid = (s1 + e1) / 2
        outside &= ~((mid > w0) & (mid < w1))
    if outside.any():
        print(f"Voigt CTA lifetime with no contraction running: median {np.median((e1 - s1)[outside]):.0f} us ({outside.sum()} CTAs)")
