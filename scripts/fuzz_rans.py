"""One-off fuzz of the rANS coder against the C oracle: skewed alphabets with clusters of rare symbols (long table searches in
the order-1 decoder), symbols of frequency 1 (two-byte renormalisations), alphabets around the shared-memory limits."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cramtools_b200
from cramtools_b200 import ExternalCompression as EC
from oracle import rans as orans

rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
ctx = cramtools_b200.default_context(0)
blocks = []
for t in range(240):
    n = int(rng.choice([5, 63, 64, 1000, 4097, 30000, 150000, 400000]))
    nsym = int(rng.choice([1, 2, 3, 8, 20, 41, 47, 48, 49, 64, 130, 256]))
    syms = np.sort(rng.choice(256, size=nsym, replace=False)).astype(np.uint8)
    kind = t % 4
    if kind == 0:      # one dominant symbol, the rest rare
        p = np.full(nsym, 0.02 / max(nsym - 1, 1)); p[rng.integers(nsym)] = 0.98
    elif kind == 1:    # geometric
        p = 0.6 ** np.arange(nsym)
    elif kind == 2:    # cluster of rare symbols at the low end, a few big ones
        p = np.full(nsym, 1e-4); p[-min(3, nsym):] = 1.0
    else:
        p = rng.random(nsym) ** 4 + 1e-5
    p = p / p.sum()
    if t % 3 == 0:     # first-order structure: the next symbol stays near the previous one
        x = np.empty(n, np.int64); x[0] = rng.integers(nsym)
        steps = rng.choice([-2, -1, 0, 0, 0, 0, 1, 2, 7], size=n)
        jump = rng.random(n) < 0.01
        r = rng.choice(nsym, size=n, p=p)
        for i in range(1, n):
            x[i] = r[i] if jump[i] else min(max(x[i - 1] + steps[i], 0), nsym - 1)
        b = syms[x]
    else:
        b = syms[rng.choice(nsym, size=n, p=p)]
    blocks.append(np.ascontiguousarray(b))
bad = 0
for order in (0, 1):
    got = EC.rans_batch(blocks, [order] * len(blocks), ctx)
    for i, (b, g) in enumerate(zip(blocks, got)):
        if g != orans.compress(b, order):
            bad += 1; print('ENCODE MISMATCH', order, i, b.size, np.unique(b).size)
    back = EC.unrans_batch(got, ctx)
    for i, (b, g) in enumerate(zip(blocks, back)):
        if g != b.tobytes():
            bad += 1; print('DECODE MISMATCH', order, i, b.size, np.unique(b).size)
print('fuzz done: %d blocks x 2 orders, %d mismatches' % (len(blocks), bad))
