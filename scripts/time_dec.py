"""Order-1 decode of N quality-like blocks of 1.5 MB (the benchmark's dominant kernel), device resident: per-kernel CUDA-event
times (cramgpu_profile_*), cycles per decoder step, round trip checked.  usage: time_dec.py [N ...]"""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cramtools_b200
from cramtools_b200 import synth, device as dv
ctx = cramtools_b200.Context(0)
sizes = [int(a) for a in sys.argv[1:]] or [256, 740, 2048]
uniq = [synth.qs_stream(1500000, 100 + i) for i in range(8)]
for nb in sizes:
    blocks = [uniq[i % 8] for i in range(nb)]
    d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
    r = dv.rans_encode_dev(ctx, d_in, off, ln, np.ones(nb, np.uint8))
    torch.cuda.synchronize()
    comp, st = r.to_host()
    assert (st == 0).all()
    c_in, c_off, c_len = dv.pack_blocks(comp, 'cuda:0')
    pref = np.stack([np.frombuffer(c[:9], np.uint8) for c in comp])
    best = {}
    for rep in range(4):
        ctx.profile_enable(True)
        d = dv.rans_decode_dev(ctx, c_in, c_off, c_len, pref)
        torch.cuda.synchronize()
        for k, v in ctx.profile_read():
            best[k] = min(best.get(k, 1e9), v)
    ctx.profile_enable(False)
    back, st = d.to_host()
    ok = (st == 0).all() and all(back[i] == blocks[i].tobytes() for i in (0, 1, nb // 2, nb - 1))
    ms = best.get('k_dec1', 0.0)
    print('blocks %5d  ok %s  %s  k_dec1 %.2f ms = %.0f cycles/step @1.965 GHz, %.1f GB/s decoded' % (
        nb, ok, {k: round(v, 3) for k, v in best.items()}, ms, ms * 1e-3 * 1.965e9 / 375000, nb * 1.5e6 / (ms * 1e-3) / 1e9), flush=True)
    del d_in, c_in
