"""Host <-> device copy bandwidth of the box, the ceiling of bench.py's end-to-end leg: pinned H2D, D2H, and both at once,
on 1 GPU and on every GPU of the box at the same time (one thread per GPU).  usage: pcie_probe.py [--gpus N] [--mb 1024]
Prints one JSON line; profiles/ holds a run."""
import argparse, json, threading, time
import torch


def probe(dev, mb, out, barrier):
    torch.cuda.set_device(dev)
    n = mb << 20
    h_up, h_dn = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
    d_up, d_dn = torch.empty(n, dtype=torch.uint8, device=dev), torch.ones(n, dtype=torch.uint8, device=dev)
    s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(up, dn, reps=6):
        torch.cuda.synchronize(dev); barrier.wait(); t0 = time.perf_counter()
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s_up): d_up.copy_(h_up, non_blocking=True)
            if dn:
                with torch.cuda.stream(s_dn): h_dn.copy_(d_dn, non_blocking=True)
        torch.cuda.synchronize(dev); dt = time.perf_counter() - t0
        return reps * n / dt / 1e9
    run(True, True, 2)
    out[dev] = {'h2d_GBps': round(run(True, False), 1), 'd2h_GBps': round(run(False, True), 1), 'both_each_GBps': round(run(True, True), 1)}


if __name__ == '__main__':
    ap = argparse.ArgumentParser(); ap.add_argument('--gpus', type=int, default=1); ap.add_argument('--mb', type=int, default=1024)
    a = ap.parse_args()
    res = {}
    for ng in sorted({1, a.gpus}):
        out, bar = {}, threading.Barrier(ng)
        th = [threading.Thread(target=probe, args=(d, a.mb, out, bar)) for d in range(ng)]
        [t.start() for t in th]; [t.join() for t in th]
        res['%d_gpus_at_once' % ng] = out
    print(json.dumps({'probe': 'pinned host<->device copies, %d MB each, GB/s per GPU' % a.mb, 'results': res}))
