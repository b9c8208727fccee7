"""Order-1 decode of N blocks with all 256 byte values (the wide-alphabet path): per-kernel CUDA-event times.  usage: time_wide.py [N] [block bytes]"""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cramtools_b200
import bench
from cramtools_b200 import device as dv
ctx = cramtools_b200.Context(0)
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
size = int(sys.argv[2]) if len(sys.argv) > 2 else 524288


def itf8_wide(n, seed):
    r = np.random.default_rng(seed)
    v = np.concatenate([r.normal(400, 60, n // 6).astype(np.int64), -r.normal(400, 60, n // 6).astype(np.int64), r.integers(1, 1 << 27, n // 6)])
    r.shuffle(v)
    return bench.itf8_pack_np(v)[:n]


uniq = [np.ascontiguousarray(itf8_wide(size, 100 + i)) for i in range(8)]
blocks = [uniq[i % 8] for i in range(nb)]
d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
r = dv.rans_encode_dev(ctx, d_in, off, ln, np.ones(nb, np.uint8))
torch.cuda.synchronize()
comp, st = r.to_host()
c_in, c_off, c_len = dv.pack_blocks(comp, 'cuda:0')
pref = np.stack([np.frombuffer(c[:9], np.uint8) for c in comp])
best = {}
for rep in range(3):
    ctx.profile_enable(True)
    d = dv.rans_decode_dev(ctx, c_in, c_off, c_len, pref)
    torch.cuda.synchronize()
    for k, v in ctx.profile_read():
        best[k] = min(best.get(k, 1e9), v)
ctx.profile_enable(False)
back, st = d.to_host()
print('blocks', nb, 'bytes', size, 'ok', bool((st == 0).all() and back[0] == blocks[0].tobytes() and back[-1] == blocks[-1].tobytes()),
      {k: round(v, 3) for k, v in best.items()}, 'distinct', int(np.unique(uniq[0]).size))
