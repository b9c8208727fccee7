import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cramtools_b200
from cramtools_b200 import synth, device as dv
ctx = cramtools_b200.Context(0)
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 64
blocks = [synth.qs_stream(1500000, 100 + (i % 8)) for i in range(nb)]
d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
orders = np.ones(nb, np.uint8)
for rep in range(2):
    ctx.profile_enable(True)
    r = dv.rans_encode_dev(ctx, d_in, off, ln, orders)
    torch.cuda.synchronize()
    sp = ctx.profile_read()
print('DBG', os.environ.get('CRAMGPU_DBG', '0'), {k: round(v, 2) for k, v in sp if k in ('k_enc1', 'k_hist1')})
