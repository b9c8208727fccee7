"""Time the record-feature kernels under cramtools' quality models (BASELINE config 4) at 256 containers x 10 000 reads."""
import sys, os, time, ctypes as C
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cramtools_b200
from cramtools_b200 import synth, records as rec, device as dv
from cramtools_b200._lib import u64p
import bench

n_cont = int(sys.argv[1]) if len(sys.argv) > 1 else 256
ref, rd, extra = bench.unique_containers(32, 1000)
ln = bench.Lane(0, n_cont, ref, rd, extra, 32, torch.device('cuda', 0), 0, 1, None, False)
ctx = ln.ctx
for spec in [None, '*8', 'm5', 'N40-D8', 'R8-N40']:
    if spec is None:
        pol, npol = None, -1
    else:
        q = rec.QualityScorePreservation.lossyFromSpec(spec)
        pol, npol = q.policies, q.n
    F = rec.Features()
    for rep in range(2):
        ctx.profile_enable(True)
        t0 = time.perf_counter()
        ctx.check(ctx.lib.cramgpu_features_encode_dev(ctx.handle, C.byref(ln.R), pol, npol, C.byref(F), ln.soff.ctypes.data_as(u64p),
                                                      ln.slen.ctypes.data_as(u64p), ln._sp()))
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        sp = ctx.profile_read()
    agg = {}
    for k, v in sp:
        agg[k] = agg.get(k, 0) + v
    print(spec, 'wall %.1f ms' % dt, {k: round(v, 2) for k, v in agg.items() if v > 0.05})
