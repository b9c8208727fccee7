"""BASELINE config[1]: rANS order-0 / order-1 block sweep, 64 KB - 8 MB, QS / BA / RN-like streams, device-resident,
byte-identical to the oracle (checked on the first block of every cell).  GB/s = raw bytes / wall time of the call between two device synchronisations
(all kernels of the call, best of 3), batch = as many blocks as fill ~1.5 GB (at most 2048); CPU = the C oracle on one host core."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cramtools_b200
from cramtools_b200 import synth, device as dv
from oracle import rans as orans

ctx = cramtools_b200.Context(0)
gens = {'QS': synth.qs_stream, 'BA': synth.ba_stream, 'RN': synth.rn_stream}
print('%-3s %-7s %-5s %6s %10s %10s %8s %9s %9s' % ('kind', 'size', 'order', 'blocks', 'enc GB/s', 'dec GB/s', 'ratio', 'cpu enc', 'cpu dec'))
for kind, gen in gens.items():
    for size in (64 << 10, 256 << 10, 1 << 20, 1500000, 4 << 20, 8 << 20):
        nb = int(max(8, min(2048, (1500 << 20) // size)))
        uniq = [gen(size, 100 + i) for i in range(8)]
        blocks = [uniq[i % 8] for i in range(nb)]
        d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
        for order in (0, 1):
            orders = np.full(nb, order, np.uint8)
            t_enc = 1e9
            for rep in range(3):
                torch.cuda.synchronize(); t0 = time.perf_counter()
                r = dv.rans_encode_dev(ctx, d_in, off, ln, orders)
                torch.cuda.synchronize(); t_enc = min(t_enc, time.perf_counter() - t0)
            comp, st = r.to_host()
            assert (st == 0).all()
            t0 = time.perf_counter(); want = orans.compress(uniq[0], order); t_ce = time.perf_counter() - t0
            assert comp[0] == want, (kind, size, order)
            c_in, c_off, c_len = dv.pack_blocks(comp, 'cuda:0')
            pref = np.stack([np.frombuffer(c[:9], np.uint8) for c in comp])
            t_dec = 1e9
            for rep in range(3):
                torch.cuda.synchronize(); t0 = time.perf_counter()
                d = dv.rans_decode_dev(ctx, c_in, c_off, c_len, pref)
                torch.cuda.synchronize(); t_dec = min(t_dec, time.perf_counter() - t0)
            back, st = d.to_host()
            assert (st == 0).all() and back[0] == blocks[0].tobytes() and back[-1] == blocks[-1].tobytes()
            t0 = time.perf_counter(); orans.uncompress(want); t_cd = time.perf_counter() - t0
            raw = float(nb) * size
            print('%-3s %-7d %-5d %6d %10.1f %10.1f %8.3f %9.3f %9.3f' % (kind, size, order, nb, raw / t_enc / 1e9,
                  raw / t_dec / 1e9, sum(len(c) for c in comp) / raw, size / t_ce / 1e9, size / t_cd / 1e9), flush=True)
        del d_in
