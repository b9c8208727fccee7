"""Time the coverage / mismatch tracks (selectors X and P) at scale: T tiles of bench.py's 32 unique containers laid
end to end on one synthetic sequence (T x 320 000 reads of 150 bp at 45x, coordinate-sorted), one call for all containers.
usage: time_tracks.py [tiles=8]    -> kernel times of cramgpu_tracks_update_dev and of the feature call that consumes its bytes"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cramtools_b200 import _lib, records as rec  # noqa: E402
from cramtools_b200._lib import u64p  # noqa: E402
import bench  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 8
ref1, rd1 = bench.unique_containers(32, 1000)
n1 = len(rd1['pos'])
ref = np.tile(ref1, T)
rd = {k: np.tile(rd1[k], T) for k in ('bases', 'quals', 'cigar', 'flag', 'mapq')}
rd['pos'] = (rd1['pos'][None, :].astype(np.int64) + (np.arange(T, dtype=np.int64) * ref1.size)[:, None]).reshape(-1).astype(np.int32)
nb, ncg = int(rd1['read_off'][-1]), int(rd1['cigar_off'][-1])
rd['read_off'] = np.concatenate([(rd1['read_off'][:-1][None, :] + (np.arange(T, dtype=np.int64) * nb)[:, None]).reshape(-1), [T * nb]]).astype(np.uint64)
rd['cigar_off'] = np.concatenate([(rd1['cigar_off'][:-1][None, :] + (np.arange(T, dtype=np.int64) * ncg)[:, None]).reshape(-1), [T * ncg]]).astype(np.uint64)
n = n1 * T
nc = n // bench.READS_PER_CONTAINER
cont_off = (np.arange(nc + 1, dtype=np.uint64) * np.uint64(bench.READS_PER_CONTAINER))
ctx = _lib.default_context()
dev = torch.device('cuda', 0)
t = {k: torch.from_numpy(np.ascontiguousarray(v).view(np.uint8).copy()).to(dev) for k, v in rd.items()}
t_ref = torch.from_numpy(ref.copy()).to(dev)
R = rec.Reads()
R.n_reads = n; R.n_containers = nc; R.cont_off = cont_off.ctypes.data_as(u64p)
roff = np.zeros(nc, np.uint64); rlen = np.full(nc, ref.size, np.uint64)
R.ref_off = roff.ctypes.data_as(u64p); R.ref_len = rlen.ctypes.data_as(u64p); R.ref = t_ref.data_ptr(); R.ref_total = ref.size
for k, v in t.items():
    setattr(R, k, v.data_ptr())
soff = np.zeros(nc * 12, np.uint64); slen = np.zeros(nc * 12, np.uint64)
print('%d containers, %d reads, %.1f Mbp' % (nc, n, ref.size / 1e6))
for spec in ('X20_8-P2_40', '*8'):
    q = rec.QualityScorePreservation.lossyFromSpec(spec)
    tracks = rec.ReferenceTracks(ref, 0, ctx=ctx)
    for rep in range(2):
        if rep:
            tracks.close(); tracks = rec.ReferenceTracks(ref, 0, ctx=ctx)
        torch.cuda.synchronize()
        ctx.profile_enable(True)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        st = np.zeros(nc, np.int32)
        ctx.check(ctx.lib.cramgpu_tracks_update_dev(ctx.handle, tracks.h, C.byref(R), q.policies, q.n, st.ctypes.data_as(C.POINTER(C.c_int32)), None))
        F = rec.Features()
        ctx.check(ctx.lib.cramgpu_features_encode_dev(ctx.handle, C.byref(R), q.policies, q.n, C.byref(F), soff.ctypes.data_as(u64p),
                                                      slen.ctypes.data_as(u64p), None))
        ctx.sync()
        sp = ctx.profile_read()
    agg = {}
    for k, v in sp:
        agg[k] = agg.get(k, 0) + v
    print(spec, 'window position', tracks.position, {k: round(v, 3) for k, v in agg.items() if v > 0.02},
          'tracks kernels %.2f ms' % sum(v for k, v in agg.items() if k.startswith('k_trk')))
    tracks.close()
