#!/usr/bin/env python
"""bench.py -- headline benchmark of the CRAM slice-codec hot path ("Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--containers C]

A "step" is one pass of the hot path over this GPU's containers (synthetic, seeded):
encode then decode of every container.  Containers are independent and are dealt round-robin to the
ranks (weak scaling: --containers is PER GPU); the only collective is an all-gather of per-container
compressed sizes from which every rank computes its containers' output offsets.

One JSON line is printed by rank 0; its keys are described in DESIGN.md.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

READS_PER_CONTAINER = 10000     # CRAMContainerStreamWriter.java:45, CramContext.java:33
READ_LEN = 150
UNIQUE_CONTAINERS = 32          # distinct synthetic containers; tiled to --containers (all copies live in HBM)


def measured_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), 'measured'
    return {'hbm_gbs': 6650.0, 'sm_max_mhz': 1965.0}, 'fallback'


# ------------------------------------------------------------------------------------------- workload


def slice_series(container_id):
    """The external data series of one lossless slice that htsjdk codes with rANS
    (CompressionHeaderFactory.build [CompressionHeaderFactory.java:84-113]): (name, order, bytes)."""
    from cramtools_b200 import synth
    seed = 1000 + container_id
    n = READS_PER_CONTAINER
    return [
        ('QS', 1, synth.qs_stream(n * READ_LEN, seed)),          # 1.5 MB, byte series
        ('BA', 1, synth.ba_stream(1500, seed + 1)),              # inserted bases
        ('BF', 1, synth.itf8_stream(n, seed + 2)),
        ('CF', 1, synth.itf8_stream(n, seed + 3)),
        ('RL', 1, synth.itf8_stream(2 * n, seed + 4)),
        ('NS', 1, synth.itf8_stream(n, seed + 5)),
        ('RG', 1, synth.itf8_stream(5 * n, seed + 6)),           # -1 -> 5-byte ITF8
        ('TS', 1, synth.itf8_stream(n, seed + 7)),
        ('AP', 0, synth.itf8_stream(n, seed + 8)),
    ]


def build_blocks(n_containers, rank):
    uniq = [slice_series(rank * UNIQUE_CONTAINERS + c) for c in range(min(UNIQUE_CONTAINERS, n_containers))]
    blocks, orders, names = [], [], []
    for c in range(n_containers):
        for name, order, data in uniq[c % len(uniq)]:
            blocks.append(data); orders.append(order); names.append(name)
    return blocks, np.array(orders, dtype=np.uint8), names


# ------------------------------------------------------------------------------------------- helpers


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region"""

    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                      '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(',')]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['unavailable']}
        sm = sorted(float(s[0]) for s in self.samples)
        reasons = []
        for i, nm in enumerate(['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']):
            if any(s[2 + i].lower().startswith('active') for s in self.samples):
                reasons.append(nm)
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': float(self.samples[0][1]), 'reasons': reasons,
                'samples': len(sm)}


def ptr_array(arrs):
    from cramtools_b200._lib import u8p
    pa = (u8p * len(arrs))()
    for i, a in enumerate(arrs):
        pa[i] = a.ctypes.data_as(u8p)
    return pa


# ------------------------------------------------------------------------------------------- reference arm


def run_reference(args):
    """CPU implementation of the same step on the host cores.  No JVM exists in this image, so the reference
    arm is the C restatement of htsjdk's rANS (oracle/, pinned to the jar by tests/golden) run on all host
    threads over a bounded sample of the workload."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from oracle import rans as orans
    lib = orans.lib()
    cores = os.cpu_count() or 1
    n_cont = min(args.containers, max(4, cores // 2))
    blocks, orders, _ = build_blocks(n_cont, 0)
    n = len(blocks)
    ins = [np.ascontiguousarray(b) for b in blocks]
    in_len = np.array([b.size for b in ins], dtype=np.uint32)
    caps = np.array([orans.bound(int(x)) for x in in_len], dtype=np.uint32)
    outs = [np.empty(int(c), dtype=np.uint8) for c in caps]
    out_len = np.zeros(n, dtype=np.uint32)
    rc = np.zeros(n, dtype=np.int32)
    back = [np.empty(max(int(x), 1), dtype=np.uint8) for x in in_len]
    back_len = np.zeros(n, dtype=np.uint32)
    u32 = ctypes.POINTER(ctypes.c_uint32)
    u8 = ctypes.POINTER(ctypes.c_uint8)
    pin, pout, pback = ptr_array(ins), ptr_array(outs), ptr_array(back)

    def step():
        r = lib.rans_oracle_many(1, cores, n, pin, in_len.ctypes.data_as(u32), orders.ctypes.data_as(u8), pout,
                                 caps.ctypes.data_as(u32), out_len.ctypes.data_as(u32),
                                 rc.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
        assert r == 0
        r = lib.rans_oracle_many(0, cores, n, pout, out_len.ctypes.data_as(u32), orders.ctypes.data_as(u8), pback,
                                 in_len.ctypes.data_as(u32), back_len.ctypes.data_as(u32),
                                 rc.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
        assert r == 0

    for _ in range(min(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    raw = float(in_len.sum())
    value = raw / dt / 1e9
    sample = '%d containers (%d rANS blocks, %.1f MB raw) per step, oracle C port on %d threads' % (
        n_cont, n, raw / 1e6, cores)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'u8/u32 integer (+f64 for order-1 normalisation)', 'data': 'synthetic',
        'config': workload_config(args, n_cont),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'note': 'no JVM in this image: reference arm = C restatement of htsjdk 2.4.1 rANS, byte-identical to the jar '
                'on tests/golden/rans_jvm.json',
    }
    print(json.dumps(line))


METRIC = 'rans_o0o1_encode_decode_GBps'
UNIT = 'GB/s of raw series bytes through encode+decode'


def workload_config(args, n_cont):
    return {'workload': 'rANS-coded external series of lossless CRAM slices (QS o1 1.5 MB + BA/BF/CF/RL/NS/RG/TS o1 + '
                        'AP o0 per 10k-read container), BASELINE configs[1]/[2] block mix',
            'containers_per_gpu': n_cont, 'reads_per_container': READS_PER_CONTAINER, 'read_len': READ_LEN,
            'unique_containers': min(UNIQUE_CONTAINERS, n_cont), 'l2': 'inputs larger than L2 (no flush needed)',
            'sharding': 'containers round-robin over ranks; all-gather of compressed sizes'}


# ------------------------------------------------------------------------------------------- our arm


def run_ours(args):
    import torch
    import torch.distributed as dist
    import cramtools_b200
    from cramtools_b200 import device as dv, sharding

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    dev = torch.device('cuda', local)
    ctx = cramtools_b200.Context(local)
    peaks, peak_kind = measured_peaks()

    n_cont = args.containers
    blocks, orders, names = build_blocks(n_cont, rank)
    n = len(blocks)
    per_cont = n // n_cont
    raw_bytes = float(sum(b.size for b in blocks))
    d_in, in_off, in_len = dv.pack_blocks(blocks, dev)

    # --- device-resident step -----------------------------------------------------------------
    enc_cap = int(sum(((9 + 1040 + int(l) + int(l) // 4 + 96 + 15) & ~15) for l in in_len)) + n_cont * 9 * 4096
    enc_out = torch.empty(enc_cap, dtype=torch.uint8, device=dev)
    res = dv.rans_encode_dev(ctx, d_in, in_off, in_len, orders, out=enc_out)
    torch.cuda.synchronize()
    blobs, st = res.to_host()
    assert (st == 0).all(), 'encode status'
    comp_bytes = float(sum(len(b) for b in blobs))
    c_off = res.off.cpu().numpy().view(np.uint64)[:-1].copy()
    c_len = res.len.cpu().numpy().view(np.uint32).copy()
    prefixes = np.zeros((n, 9), dtype=np.uint8)
    for i, b in enumerate(blobs):
        prefixes[i, :min(9, len(b))] = np.frombuffer(b[:9], dtype=np.uint8)
    dec_out = torch.empty(int(sum((int(l) + 15) & ~15 for l in in_len)) + 16, dtype=torch.uint8, device=dev)

    def step_dev():
        r = dv.rans_encode_dev(ctx, d_in, in_off, in_len, orders, out=enc_out)
        # per-container compressed size -> global output offsets (the path's only collective)
        sizes = r.len.view(n_cont, per_cont).sum(dim=1, dtype=torch.int64)
        offs = sharding.allgather_offsets(sizes, rank, world)
        r2 = dv.rans_decode_dev(ctx, enc_out, c_off, c_len, prefixes, out=dec_out)
        return r, r2, offs

    # correctness of what is being timed: decode(encode(x)) == x on device, sizes vs oracle on a sample
    r, r2, offs = step_dev()
    torch.cuda.synchronize()
    assert int(r.status.abs().sum()) == 0 and int(r2.status.abs().sum()) == 0
    raws, _ = r2.to_host()
    for i in range(0, n, max(1, n // 37)):
        assert raws[i] == blocks[i].tobytes(), 'round trip mismatch in block %d' % i
    del raws

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_dev()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ctx.launch_count()
    ctx.profile_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_dev()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    spans = ctx.profile_read()
    ctx.profile_enable(False)
    launches = (ctx.launch_count() - launches0) // args.steps
    clocks = sampler.summary()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())

    # per-kernel average duration inside the timed region
    kt = {}
    for name, m in spans:
        kt.setdefault(name, []).append(m)
    kavg = {k: float(np.mean(v)) for k, v in kt.items()}
    top = max(kavg, key=kavg.get)

    # --- end-to-end step through the host-buffer C ABI ------------------------------------------
    ins = [np.ascontiguousarray(b) for b in blocks]
    from cramtools_b200 import rans_bound
    caps = np.array([rans_bound(b.size) for b in ins], dtype=np.uint32)
    outs = [np.empty(int(c), dtype=np.uint8) for c in caps]
    out_len = np.zeros(n, dtype=np.uint32)
    status = np.zeros(n, dtype=np.int32)
    back = [np.empty(max(b.size, 1), dtype=np.uint8) for b in ins]
    back_len = np.zeros(n, dtype=np.uint32)
    in_len32 = np.array([b.size for b in ins], dtype=np.uint32)
    from cramtools_b200._lib import u32p, u8p, i32p
    pin, pout, pback = ptr_array(ins), ptr_array(outs), ptr_array(back)

    def step_e2e():
        rc = ctx.lib.cramgpu_rans_encode_batch(ctx.handle, n, pin, in_len32.ctypes.data_as(u32p), orders.ctypes.data_as(u8p),
                                               pout, caps.ctypes.data_as(u32p), out_len.ctypes.data_as(u32p),
                                               status.ctypes.data_as(i32p))
        ctx.check(rc)
        rc = ctx.lib.cramgpu_rans_decode_batch(ctx.handle, n, pout, out_len.ctypes.data_as(u32p), pback,
                                               in_len32.ctypes.data_as(u32p), back_len.ctypes.data_as(u32p),
                                               status.ctypes.data_as(i32p))
        ctx.check(rc)

    step_e2e()
    assert bytes(back[0][:back_len[0]]) == blocks[0].tobytes()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 3))
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    e2e_ms = (time.perf_counter() - t0) / e2e_steps * 1e3
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())

    if rank == 0:
        value = raw_bytes * world / (ms * 1e-3) / 1e9
        # roofline of the dominant kernel: algorithmic bytes per launch / its average duration
        algo = {'k_enc1': None}
        raw1 = float(sum(b.size for b, o in zip(blocks, orders) if o == 1 and b.size >= 4))
        raw0 = raw_bytes - raw1
        comp1 = float(sum(len(b) for b, o in zip(blobs, orders) if o == 1))
        comp0 = comp_bytes - comp1
        algo_bytes = {'k_enc1': raw1 + comp1, 'k_enc0': raw0 + comp0, 'k_dec1': raw1 + comp1, 'k_dec0': raw0 + comp0,
                      'k_hist0': raw_bytes, 'k_hist1': raw1, 'k_pack': 2 * comp_bytes}
        ab = algo_bytes.get(top)
        achieved = ab / (kavg[top] * 1e-3) / 1e9 if ab else None
        peak = float(peaks['hbm_gbs'])
        lanes1 = 4 * sum(1 for o in orders if o == 1)
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'u8/u32 integer (+f64 for order-1 normalisation)', 'data': 'synthetic',
            'config': workload_config(args, n_cont),
            'compression_ratio': comp_bytes / raw_bytes,
            'kernels_ms': kavg, 'gpu_launches': int(launches),
            'roofline': {'bound': 'hbm', 'kernel': top, 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                         'frac': achieved / peak if achieved else None, 'traffic': None, 'peak_kind': peak_kind,
                         'lanes_in_flight': lanes1,
                         'note': 'rANS is a per-state recurrence: attainable rate = lanes x clock / cycles-per-symbol, see DESIGN.md'},
            'e2e': {'value': raw_bytes * world / (e2e_ms * 1e-3) / 1e9, 'unit': UNIT,
                    'h2d_bytes_per_step': int(raw_bytes + comp_bytes), 'd2h_bytes_per_step': int(comp_bytes + raw_bytes),
                    'ms_per_step': e2e_ms},
            'clocks': clocks,
        }
        if world == 1 and not args.no_cpu:
            line['cpu_baseline'] = cpu_baseline(args)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline(args):
    """the oracle (C port of htsjdk's rANS) timed on this box's host cores on a bounded sample"""
    from oracle import rans as orans
    lib = orans.lib()
    cores = os.cpu_count() or 1
    n_cont = max(2, min(args.containers, cores // 2))
    blocks, orders, _ = build_blocks(n_cont, 0)
    n = len(blocks)
    ins = [np.ascontiguousarray(b) for b in blocks]
    in_len = np.array([b.size for b in ins], dtype=np.uint32)
    caps = np.array([orans.bound(int(x)) for x in in_len], dtype=np.uint32)
    outs = [np.empty(int(c), dtype=np.uint8) for c in caps]
    out_len = np.zeros(n, dtype=np.uint32)
    rc = np.zeros(n, dtype=np.int32)
    back = [np.empty(max(int(x), 1), dtype=np.uint8) for x in in_len]
    back_len = np.zeros(n, dtype=np.uint32)
    u32 = ctypes.POINTER(ctypes.c_uint32)
    u8 = ctypes.POINTER(ctypes.c_uint8)
    pin, pout, pback = ptr_array(ins), ptr_array(outs), ptr_array(back)
    reps = 0
    t0 = time.perf_counter()
    while True:
        lib.rans_oracle_many(1, cores, n, pin, in_len.ctypes.data_as(u32), orders.ctypes.data_as(u8), pout,
                             caps.ctypes.data_as(u32), out_len.ctypes.data_as(u32), rc.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
        lib.rans_oracle_many(0, cores, n, pout, out_len.ctypes.data_as(u32), orders.ctypes.data_as(u8), pback,
                             in_len.ctypes.data_as(u32), back_len.ctypes.data_as(u32), rc.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
        reps += 1
        if time.perf_counter() - t0 > 10.0 or reps >= 20:
            break
    dt = (time.perf_counter() - t0) / reps
    raw = float(in_len.sum())
    return {'value': raw / dt / 1e9, 'unit': UNIT, 'cores': cores, 'kind': 'port',
            'sample': '%d containers (%d blocks, %.1f MB raw) x %d repetitions, oracle C port, %d threads'
                      % (n_cont, n, raw / 1e6, reps, cores)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--containers', type=int, default=1024, help='containers PER GPU (10 000 reads each)')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
