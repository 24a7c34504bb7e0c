"""Phase timeline of k_md_steps (library built with -DB200_PERSIST_TRACE, see md_steps_persist.cuh):
   HOOMD_B200_LIB=hoomd_b200/libhoomd_b200_trace.so python tools/trace_persist.py [workload]
marks per CTA and step: 0 step start, 1 thread 0's own walk done, 2 CTA's walk done, 3 integration + distance check done,
4 ticket atomic returned, 5 epoch seen, 6 acquire fence + control words read"""
import sys, os, json, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
import hoomd_b200
wl = sys.argv[1] if len(sys.argv) > 1 else "lj64k"
state, L = bench.make_state(wl)
s = bench.new_system(state, L, workload=wl)
s.run(400)
lib = hoomd_b200._lib.lib()
dims = (C.c_uint * 3)()
assert lib.b200_debug_persist_trace(None, dims) == 0   # reset the per-CTA step counters
nc, ns, nm = dims[0], dims[1], dims[2]
s.run(200)
buf = np.zeros((nc, ns, nm), np.uint64)
assert lib.b200_debug_persist_trace(buf.ctypes.data_as(C.POINTER(C.c_ulonglong)), dims) == 0
ncta = int((buf[:, 0, 0] != 0).sum())
t = buf[:ncta].astype(np.int64)
clk = 1.0 / 1.92   # ns per cycle (approx.)
d = lambda a, b: (t[:, :, b] - t[:, :, a]) * clk / 1000.0   # us
names = [("walk (thread 0)", 0, 1), ("walk (CTA)", 0, 2), ("integrate+check", 2, 3), ("ticket atomic", 3, 4),
         ("epoch wait", 4, 5), ("fence+ctl", 5, 6), ("whole step", 0, 6)]
out = {"workload": wl, "ctas": ncta, "steps": ns}
for n, a, b in names:
    x = d(a, b)[:, 2:]   # skip the first steps after a launch
    out[n] = {"mean_us": float(x.mean()), "p10": float(np.percentile(x, 10)), "p50": float(np.percentile(x, 50)),
              "p90": float(np.percentile(x, 90)), "max": float(x.max())}
# per step: spread of the CTA-walk finish times inside one SM clock domain is not comparable across SMs; report the
# distribution of per-CTA work (marks 0..3) instead
work = d(0, 3)[:, 2:]
out["work per CTA: mean over steps, then min/mean/max over CTAs"] = [float(work.mean(1).min()), float(work.mean()), float(work.mean(1).max())]
sm = t[:, 0, 7]
per_sm = np.bincount(sm.astype(np.int64))
out["CTAs per SM (min/max)"] = [int(per_sm[per_sm > 0].min()), int(per_sm.max())]
wk = work.mean(1)
out["work by CTAs-on-SM"] = {int(k): float(wk[per_sm[sm] == k].mean()) for k in sorted(set(per_sm[sm]))}
# what makes a CTA slow?  per-CTA: particles whose walk takes minimum images, longest row, mean row
pp = s.particles()
pos = pp["pos"][:, :3]
margin = 3.4
face = (np.abs(pos) > (L / 2 - margin)).any(axis=1)
nn = s.nlist(counts_only=True)[0].astype(np.int64)
ppb = 64
nct = min(ncta, (len(pos) + ppb - 1) // ppb)
fr = np.array([face[c * ppb:(c + 1) * ppb].mean() for c in range(nct)])
mx = np.array([nn[c * ppb:(c + 1) * ppb].max() for c in range(nct)])
mn = np.array([nn[c * ppb:(c + 1) * ppb].mean() for c in range(nct)])
w = wk[:nct]
cc = lambda a: float(np.corrcoef(a, w)[0, 1])
out["corr(work, face fraction / max row / mean row / blockIdx)"] = [cc(fr), cc(mx), cc(mn), cc(np.arange(nct))]
out["work: all-interior CTAs vs CTAs with face particles"] = [float(w[fr == 0].mean()) if (fr == 0).any() else None, float(w[fr > 0].mean())]
order = np.argsort(w)
out["slowest CTAs (blockIdx, work us, face fraction, max row, sm)"] = [[int(c), float(w[c]), float(fr[c]), int(mx[c]), int(sm[c])] for c in order[-8:]]
out["fastest CTAs"] = [[int(c), float(w[c]), float(fr[c]), int(mx[c]), int(sm[c])] for c in order[:8]]
walk = d(0, 2)[:, 2:]
out["per-step max over CTAs of walk / work (mean over steps)"] = [float(walk.max(0).mean()), float(work.max(0).mean())]
np.save(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "trace_work_per_cta.npy"), np.stack([w, fr, mx, mn, sm[:nct]]))
print(json.dumps(out, indent=1))
