#!/usr/bin/env python
"""bench.py -- headline benchmark of the CRAM slice-codec hot path ("Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--containers C]

Workload (BASELINE.json configs[2], "chr20-scale"): C containers PER GPU of 10 000 simulated 150 bp reads each at
~45x over a random reference, lossless quality scores.  A "step" is one pass of the whole hot path over this GPU's
containers:
    encode  feature extraction + substitution matrix + series packing  ->  rANS of the slice's rANS-coded series
            (QS, BA from the device; BF, CF, RL, RG, NS, TS order 1 and AP order 0 as the Java host hands them over)
    decode  rANS decode of the same blocks  ->  base and quality restoration of every record
Containers are independent and dealt round-robin to the ranks (weak scaling); the only collective is an all-gather
of per-container compressed sizes, from which every rank derives its containers' output offsets.

One JSON line is printed by rank 0; the keys are explained in DESIGN.md.
"""
import argparse
import ctypes as C
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
COVERAGE = 45
UNIQUE = 32                     # distinct synthetic containers; tiled to --containers (every copy lives in HBM)
METRIC = 'cram_encode_decode_reads_per_s'
UNIT = 'reads/s'
EXTRA = [('BF', 1), ('CF', 1), ('RL', 1), ('RG', 1), ('NS', 1), ('TS', 1), ('AP', 0)]   # host-produced rANS series (R0)


def measured_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), 'measured'
    return {'hbm_gbs': 6650.0, 'sm_max_mhz': 1965.0}, 'fallback'


def ncu_traffic(kernel, n_cont):
    """DRAM bytes of one launch of `kernel` from the committed ncu capture (profiles/r1_traffic.json), if it was taken at this size"""
    try:
        with open(os.path.join(ROOT, 'profiles', 'r1_traffic.json')) as f:
            t = json.load(f).get(kernel)
        if t and int(t['containers_per_gpu']) == int(n_cont):
            return int(t['dram_bytes_read']) + int(t['dram_bytes_write'])
    except (OSError, ValueError, KeyError):
        pass
    return None


# ------------------------------------------------------------------------------------------- workload


def itf8_pack_np(v):
    v = np.asarray(v, np.int64) & 0xFFFFFFFF
    out = []
    n1 = v < (1 << 7); n2 = (~n1) & (v < (1 << 14)); n3 = (~n1) & (~n2) & (v < (1 << 21)); n4 = (~n1) & (~n2) & (~n3) & (v < (1 << 28))
    n5 = ~(n1 | n2 | n3 | n4)
    ln = n1 * 1 + n2 * 2 + n3 * 3 + n4 * 4 + n5 * 5
    off = np.concatenate([[0], np.cumsum(ln)])
    buf = np.zeros(int(off[-1]), np.uint8)
    o = off[:-1]
    buf[o[n1]] = v[n1]
    buf[o[n2]] = (v[n2] >> 8) | 0x80; buf[o[n2] + 1] = v[n2] & 0xFF
    buf[o[n3]] = (v[n3] >> 16) | 0xC0; buf[o[n3] + 1] = (v[n3] >> 8) & 0xFF; buf[o[n3] + 2] = v[n3] & 0xFF
    buf[o[n4]] = (v[n4] >> 24) | 0xE0; buf[o[n4] + 1] = (v[n4] >> 16) & 0xFF; buf[o[n4] + 2] = (v[n4] >> 8) & 0xFF; buf[o[n4] + 3] = v[n4] & 0xFF
    buf[o[n5]] = ((v[n5] >> 28) | 0xF0) & 0xFF; buf[o[n5] + 1] = (v[n5] >> 20) & 0xFF; buf[o[n5] + 2] = (v[n5] >> 12) & 0xFF
    buf[o[n5] + 3] = (v[n5] >> 4) & 0xFF; buf[o[n5] + 4] = v[n5] & 0xFF
    return buf


def unique_containers(n_unique, seed):
    """n_unique coordinate-sorted containers over one synthetic contig; returns (ref, rd, extra) where extra[c] is the list of
    host-produced series of container c in EXTRA order (what htsjdk's Writer would hand to the rANS block codec)."""
    from cramtools_b200 import synth
    n = n_unique * READS_PER_CONTAINER
    ref_len = n * READ_LEN // COVERAGE + 4096
    ref = synth.reference(ref_len, seed)
    rd = synth.reads(ref, n, seed + 1, read_len=READ_LEN, with_names=False)
    extra = []
    for c in range(n_unique):
        a, b = c * READS_PER_CONTAINER, (c + 1) * READS_PER_CONTAINER
        flag, pos = rd['flag'][a:b].astype(np.int64), rd['pos'][a:b].astype(np.int64)
        bf = np.where(flag & 16, 0x10, 0)                                   # CRAM BF: negative strand bit
        cf = np.full(b - a, 3, np.uint8)                                    # force-preserve QS | detached
        rl = np.full(b - a, READ_LEN)
        rg = np.full(b - a, -1)                                             # no read group -> 5-byte ITF8
        ns = np.full(b - a, -1)
        ts = np.zeros(b - a, np.int64)
        ap = np.diff(np.concatenate([[pos[0]], pos]))                       # AP delta vs previous record
        extra.append([itf8_pack_np(bf), cf, itf8_pack_np(rl), itf8_pack_np(rg), itf8_pack_np(ns), itf8_pack_np(ts), itf8_pack_np(ap)])
    return ref, rd, extra


def tile_reads(rd, n_unique, n_cont):
    """host arrays for n_cont containers by tiling the unique ones"""
    reps = (n_cont + n_unique - 1) // n_unique
    nr = n_cont * READS_PER_CONTAINER
    nb = int(rd['read_off'][-1])
    ncg = int(rd['cigar_off'][-1])
    out = {}
    for k in ('bases', 'quals', 'cigar', 'pos', 'flag', 'mapq'):
        out[k] = np.tile(rd[k], reps)
    ro = (rd['read_off'][:-1][None, :] + (np.arange(reps, dtype=np.int64) * nb)[:, None]).reshape(-1)
    co = (rd['cigar_off'][:-1][None, :] + (np.arange(reps, dtype=np.int64) * ncg)[:, None]).reshape(-1)
    out['read_off'] = np.concatenate([ro[:nr], [ro[nr] if nr < ro.size else reps * nb]])
    out['cigar_off'] = np.concatenate([co[:nr], [co[nr] if nr < co.size else reps * ncg]])
    for k in ('pos', 'flag', 'mapq'):
        out[k] = out[k][:nr]
    out['bases'] = out['bases'][:int(out['read_off'][-1])]
    out['quals'] = out['quals'][:int(out['read_off'][-1])]
    out['cigar'] = out['cigar'][:int(out['cigar_off'][-1])]
    return out


# ------------------------------------------------------------------------------------------- helpers


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region"""

    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

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
        reasons = [nm for i, nm in enumerate(['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'])
                   if any(s[2 + i].lower().startswith('active') for s in self.samples)]
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': float(self.samples[0][1]), 'reasons': reasons, 'samples': len(sm)}


def workload_config(args, n_cont, note=''):
    return {'workload': 'chr20-scale slices (BASELINE configs[2]): %d containers/GPU x 10000 simulated 150 bp reads at %dx, '
                        'lossless QS; encode = features + series packing + rANS (QS, BA, BF, CF, RL, RG, NS, TS o1; AP o0), '
                        'decode = rANS + base/quality restoration' % (n_cont, COVERAGE),
            'containers_per_gpu': n_cont, 'reads_per_container': READS_PER_CONTAINER, 'read_len': READ_LEN,
            'unique_containers': min(UNIQUE, n_cont), 'quality_model': 'lossless',
            'l2': 'inputs larger than L2 (no flush needed)' if n_cont >= 64 else 'inputs may fit L2',
            'sharding': 'containers round-robin over ranks; all-gather of compressed sizes', 'note': note}


# ------------------------------------------------------------------------------------------- CPU arm


def cpu_step_setup(n_cont, seed):
    from oracle import rans as orans
    lib = orans.lib()
    nu = min(UNIQUE, n_cont)
    ref, rd, extra = unique_containers(nu, seed)
    trd = tile_reads(rd, nu, n_cont)
    i64 = C.POINTER(C.c_int64)
    keep = dict(cont=np.arange(n_cont + 1, dtype=np.int64) * READS_PER_CONTAINER, bases=trd['bases'], quals=trd['quals'],
                ro=trd['read_off'].astype(np.int64), cg=trd['cigar'], co=trd['cigar_off'].astype(np.int64), pos=trd['pos'].astype(np.int32),
                flag=trd['flag'].astype(np.uint16), mapq=trd['mapq'].astype(np.uint8), ref=ref)
    ex = [np.ascontiguousarray(e, np.uint8) for c in range(n_cont) for e in extra[c % nu]]
    ex_ptr = np.array([e.ctypes.data for e in ex], np.uint64)
    ex_len = np.array([e.size for e in ex], np.uint32)
    ex_ord = np.array([o for _, o in EXTRA], np.uint8)
    keep.update(ex=ex, ex_ptr=ex_ptr, ex_len=ex_len, ex_ord=ex_ord)
    lib.oracle_cram_step.restype = C.c_int
    comp = C.c_int64(0); chk = C.c_uint64(0)

    def step(threads):
        rc = lib.oracle_cram_step(
            C.c_int(threads), C.c_int(n_cont), keep['cont'].ctypes.data_as(i64), C.c_void_p(keep['bases'].ctypes.data),
            C.c_void_p(keep['quals'].ctypes.data), keep['ro'].ctypes.data_as(i64), C.c_void_p(keep['cg'].ctypes.data),
            keep['co'].ctypes.data_as(i64), C.c_void_p(keep['pos'].ctypes.data), C.c_void_p(keep['flag'].ctypes.data),
            C.c_void_p(keep['mapq'].ctypes.data), C.c_void_p(keep['ref'].ctypes.data), C.c_int64(keep['ref'].size), C.c_int(len(EXTRA)),
            C.c_void_p(keep['ex_ptr'].ctypes.data), C.c_void_p(keep['ex_len'].ctypes.data), C.c_void_p(keep['ex_ord'].ctypes.data),
            C.c_char_p(b'lossless'), C.byref(comp), C.byref(chk))
        assert rc == 0, 'CPU oracle round trip failed (%d)' % rc
        return comp.value
    return step, keep


def cpu_baseline(args, seconds=12.0):
    """the oracle (C restatement of the reference's Java, pinned to the jar by tests/golden) on all host cores, bounded sample"""
    cores = os.cpu_count() or 1
    n_cont = max(2, min(args.containers, cores))
    step, _ = cpu_step_setup(n_cont, 1000)
    step(cores)
    reps, t0 = 0, time.perf_counter()
    while True:
        step(cores); reps += 1
        if time.perf_counter() - t0 > seconds or reps >= 50:
            break
    dt = (time.perf_counter() - t0) / reps
    return {'value': n_cont * READS_PER_CONTAINER / dt, 'unit': UNIT, 'cores': cores, 'kind': 'port',
            'sample': '%d containers (%d reads) x %d repetitions, one container per thread, %d threads; C restatement of the Java '
                      'reference (no JVM in the image)' % (n_cont, n_cont * READS_PER_CONTAINER, reps, cores)}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_cont = max(2, min(args.containers, 2 * cores))
    step, _ = cpu_step_setup(n_cont, 1000)
    for _ in range(min(args.warmup, 1)):
        step(cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(cores)
    dt = (time.perf_counter() - t0) / args.steps
    value = n_cont * READS_PER_CONTAINER / dt
    sample = '%d containers (%d reads) per step, one container per thread, %d threads' % (n_cont, n_cont * READS_PER_CONTAINER, cores)
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'u8/i32 integer (+f64 in the order-1 frequency normalisation)', 'data': 'synthetic',
            'config': workload_config(args, n_cont, 'bounded sample of the same workload'),
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'note': 'no JVM in this image: the reference arm is the C restatement of cramtools/htsjdk (oracle/), byte-identical to '
                    'the jar on tests/golden/*.json, run the way paralell/BamToCram + CramToBam do: one container per host thread'}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------- our arm


class Lane(object):
    """One CUDA stream + one cramgpu context + a share of the rank's containers, resident in HBM.  The rank's step is
    the step of all its lanes, each driven by its own host thread, so that one lane's latency-bound coder kernels
    (a few warps per SM) overlap the other lanes' record kernels and copies."""

    def __init__(self, lane_id, n_cont, ref, rd, extra, nu, dev, rank, world, group, use_unpack=True):
        import torch
        import cramtools_b200
        from cramtools_b200 import device as dv, records as rec
        from cramtools_b200._lib import u64p
        self.torch, self.rec, self.dv = torch, rec, dv
        self.n_cont, self.rank, self.world, self.group, self.dev = n_cont, rank, world, group, dev
        self.use_unpack = use_unpack
        self.stream = torch.cuda.Stream(device=dev)
        self.ctx = cramtools_b200.Context(dev.index)
        vp = C.c_void_p
        with torch.cuda.stream(self.stream):
            reps = (n_cont + nu - 1) // nu
            first = (lane_id * 7) % nu                    # lanes start at different unique containers
            n_reads = self.n_reads = n_cont * READS_PER_CONTAINER
            ncg_u = int(rd['cigar_off'][-1])

            def up(a, dt):
                return torch.from_numpy(np.ascontiguousarray(a, dt).view(np.uint8)).to(dev)

            def tile_u8(a, dt, n_elems):
                return up(a, dt).repeat(reps)[:n_elems * np.dtype(dt).itemsize].contiguous()
            self.t_bases = tile_u8(rd['bases'], np.uint8, n_reads * READ_LEN)
            self.t_quals = tile_u8(rd['quals'], np.uint8, n_reads * READ_LEN)
            self.t_pos = tile_u8(rd['pos'], np.int32, n_reads); self.t_flag = tile_u8(rd['flag'], np.uint16, n_reads)
            self.t_mapq = tile_u8(rd['mapq'], np.uint8, n_reads)
            cig_all = np.tile(rd['cigar'], reps)
            co_all = (rd['cigar_off'][:-1][None, :] + (np.arange(reps, dtype=np.int64) * ncg_u)[:, None]).reshape(-1)
            co_all = np.concatenate([co_all[:n_reads], [co_all[n_reads] if n_reads < co_all.size else reps * ncg_u]]).astype(np.uint64)
            self.t_cigar = up(cig_all[:int(co_all[-1])], np.uint32)
            self.t_coff = up(co_all, np.uint64)
            self.t_roff = up(np.arange(n_reads + 1, dtype=np.uint64) * READ_LEN, np.uint64)
            self.t_ref = up(ref, np.uint8)
            self.cont_off = (np.arange(n_cont + 1, dtype=np.uint64) * READS_PER_CONTAINER)
            self.ref_off = np.zeros(n_cont, np.uint64); self.ref_len = np.full(n_cont, ref.size, np.uint64)
            R = self.R = rec.Reads()
            R.n_reads = n_reads; R.n_containers = n_cont
            R.cont_off = self.cont_off.ctypes.data_as(u64p); R.ref_off = self.ref_off.ctypes.data_as(u64p)
            R.ref_len = self.ref_len.ctypes.data_as(u64p)
            R.ref = vp(self.t_ref.data_ptr()); R.ref_total = ref.size
            R.bases = vp(self.t_bases.data_ptr()); R.quals = vp(self.t_quals.data_ptr()); R.read_off = vp(self.t_roff.data_ptr())
            R.cigar = vp(self.t_cigar.data_ptr()); R.cigar_off = vp(self.t_coff.data_ptr()); R.pos = vp(self.t_pos.data_ptr())
            R.flag = vp(self.t_flag.data_ptr()); R.mapq = vp(self.t_mapq.data_ptr()); R.rflags = None
            # host-produced series, resident: one buffer, 16-byte aligned blocks
            ex_blocks = [np.ascontiguousarray(e, np.uint8) for c in range(n_cont) for e in extra[c % nu]]
            self.t_extra, ex_off, ex_len = dv.pack_blocks(ex_blocks, dev)
            ex_addr = ex_off + np.uint64(self.t_extra.data_ptr())
            NX = len(EXTRA)
            NB = self.NB = 2 + NX                               # rANS blocks per container: QS, BA, then EXTRA
            nblk = self.nblk = n_cont * NB
            self.order = np.tile(np.array([1, 1] + [o for _, o in EXTRA], np.uint8), n_cont)
            self.KQS, self.KBA = rec.SERIES.index('QS'), rec.SERIES.index('BA')
            self.soff = np.zeros((n_cont, 12), np.uint64); self.slen = np.zeros((n_cont, 12), np.uint64)
            self.F = rec.Features()
            self.in_addr = np.zeros(nblk, np.uint64); self.in_len = np.zeros(nblk, np.uint32)
            self.in_addr.reshape(n_cont, NB)[:, 2:] = ex_addr.reshape(n_cont, NX)
            self.in_len.reshape(n_cont, NB)[:, 2:] = ex_len.reshape(n_cont, NX)
            self.enc_cap = int(n_reads * READ_LEN * 0.6) + nblk * 4096 + int(ex_len.sum())
            self.enc_out = torch.empty(self.enc_cap, dtype=torch.uint8, device=dev)
            self.e_off = torch.empty(nblk + 1, dtype=torch.int64, device=dev); self.e_len = torch.empty(nblk, dtype=torch.int32, device=dev)
            self.e_st = torch.empty(nblk, dtype=torch.int32, device=dev)
            # region per container = [min start, max end] of its records (CramNormalizer.java:212-225)
            p = rd['pos'].astype(np.int64).reshape(nu, READS_PER_CONTAINER)
            ws = (p.min(axis=1) - 1)
            we = np.minimum(p.max(axis=1) - 1 + READ_LEN + 8, ref.size)
            self.win_off = np.tile(ws, reps)[:n_cont].astype(np.uint64); self.win_len = np.tile(we - ws, reps)[:n_cont].astype(np.uint64)
            self.win_s0 = np.tile(ws, reps)[:n_cont].astype(np.int64)
            self.o_bases = torch.empty(n_reads * READ_LEN, dtype=torch.uint8, device=dev); self.o_quals = torch.empty_like(self.o_bases)
            self.o_star = torch.empty(n_reads, dtype=torch.uint8, device=dev); self.o_stat = torch.empty(n_reads, dtype=torch.int32, device=dev)
            self.t_rfl = torch.ones(n_reads, dtype=torch.uint8, device=dev)          # lossless, mapped: force-preserve QS
        self.prepared = False

    def _sp(self):
        return C.c_void_p(self.stream.cuda_stream)

    def encode(self):
        from cramtools_b200 import sharding
        from cramtools_b200._lib import u8p, u32p, u64p
        ctx, F, vp = self.ctx, self.F, C.c_void_p
        ctx.check(ctx.lib.cramgpu_features_encode_dev(ctx.handle, C.byref(self.R), None, -1, C.byref(F), self.soff.ctypes.data_as(u64p),
                                                      self.slen.ctypes.data_as(u64p), self._sp()))
        base = np.uint64(F.streams)
        v = self.in_addr.reshape(self.n_cont, self.NB); w = self.in_len.reshape(self.n_cont, self.NB)
        v[:, 0] = base + self.soff[:, self.KQS]; w[:, 0] = self.slen[:, self.KQS]
        v[:, 1] = base + self.soff[:, self.KBA]; w[:, 1] = self.slen[:, self.KBA]
        ctx.check(ctx.lib.cramgpu_rans_encode_dev(ctx.handle, self.nblk, None, self.in_addr.ctypes.data_as(u64p),
                                                  self.in_len.ctypes.data_as(u32p), self.order.ctypes.data_as(u8p),
                                                  vp(self.enc_out.data_ptr()), C.c_uint64(self.enc_cap), vp(self.e_off.data_ptr()),
                                                  vp(self.e_len.data_ptr()), vp(self.e_st.data_ptr()), self._sp()))
        with self.torch.cuda.stream(self.stream):
            sizes = self.e_len.view(self.n_cont, self.NB).sum(dim=1, dtype=self.torch.int64)
            return sharding.allgather_offsets(sizes, self.rank, self.world, group=self.group)   # the path's only collective

    def prepare(self):
        """one verified pass to learn what the Java host would know on the decode side (block sizes, prefixes, offsets)"""
        torch, rec, dv = self.torch, self.rec, self.dv
        from cramtools_b200._lib import u64p
        vp = C.c_void_p
        self.encode()
        self.stream.synchronize()
        assert int(self.e_st.abs().sum()) == 0, 'rANS encode status'
        nblk, n_cont, NB, n_reads = self.nblk, self.n_cont, self.NB, self.n_reads
        with torch.cuda.stream(self.stream):
            self.c_off = self.e_off.cpu().numpy().view(np.uint64)[:-1].copy(); self.c_len = self.e_len.cpu().numpy().view(np.uint32).copy()
            self.comp_bytes = float(self.c_len.sum())
            enc_host = self.enc_out[:int(self.c_off[-1]) + int(self.c_len[-1])].cpu().numpy()
            self.prefixes = np.zeros((nblk, 9), np.uint8)
            idx = self.c_off[:, None].astype(np.int64) + np.arange(9)[None, :]
            ok = self.c_len >= 9
            self.prefixes[ok] = enc_host[idx[ok]]
            del enc_host
            raw = self.in_len.astype(np.int64)
            d_off = np.concatenate([[0], np.cumsum((raw + 15) & ~15)]).astype(np.uint64)
            self.dec_out = torch.empty(int(d_off[-1]) + 64, dtype=torch.uint8, device=self.dev)
            self.dd_off = torch.empty(nblk + 1, dtype=torch.int64, device=self.dev); self.dd_len = torch.empty(nblk, dtype=torch.int32, device=self.dev)
            self.dd_st = torch.empty(nblk, dtype=torch.int32, device=self.dev)
            # per-read offset of its score array inside the decoded QS block (what CramRecordReader knows while parsing)
            rq = np.empty(n_reads, np.uint64)
            self.ctx.check(self.ctx.lib.cramgpu_copy_to_host(self.ctx.handle, vp(rq.ctypes.data), vp(self.F.read_qs_off),
                                                             C.c_uint64(rq.nbytes), self._sp()))
            cidx = np.repeat(np.arange(n_cont), READS_PER_CONTAINER)
            rq_new = rq - self.soff[cidx, self.KQS] + d_off[cidx * NB + 0]
            self.t_rq = torch.from_numpy(rq_new.view(np.uint8)).to(self.dev)
            self.t_rba = torch.full((n_reads,), -1, dtype=torch.int64, device=self.dev)
            self.raw_bytes = float(raw.sum())
            self.raw_o0 = float(self.in_len.reshape(n_cont, NB)[:, -1].sum())
            self.qs_bytes = float(self.slen[:, self.KQS].sum())
        N = self.N = rec.NormIn()
        N.n_reads = n_reads; N.n_containers = n_cont
        N.cont_off = self.cont_off.ctypes.data_as(u64p); N.win_off = self.win_off.ctypes.data_as(u64p); N.win_len = self.win_len.ctypes.data_as(u64p)
        N.win_start0 = self.win_s0.ctypes.data_as(C.POINTER(C.c_int64)); N.ref = vp(self.t_ref.data_ptr()); N.ref_total = self.t_ref.numel()
        N.pos = vp(self.t_pos.data_ptr()); N.read_off = vp(self.t_roff.data_ptr()); N.rflags = vp(self.t_rfl.data_ptr())
        N.qs = vp(self.dec_out.data_ptr()); N.read_qs_off = vp(self.t_rq.data_ptr()); N.ba = vp(self.dec_out.data_ptr())
        N.read_ba_off = vp(self.t_rba.data_ptr())
        if self.use_unpack:
            # series table of the decode side: absolute device addresses (QS / BA inside dec_out, the others inside F.streams)
            sa = (np.uint64(self.F.streams) + self.soff).astype(np.uint64)
            sa[:, self.KQS] = np.uint64(self.dec_out.data_ptr()) + d_off[np.arange(n_cont) * NB + 0]
            sa[:, self.KBA] = np.uint64(self.dec_out.data_ptr()) + d_off[np.arange(n_cont) * NB + 1]
            self.s_addr = np.ascontiguousarray(sa); self.s_len = np.ascontiguousarray(self.slen.astype(np.uint64))
            SR = self.SR = rec.Series()
            SR.n_reads = n_reads; SR.n_containers = n_cont
            SR.cont_off = self.cont_off.ctypes.data_as(u64p); SR.streams = None; SR.streams_len = 0
            SR.stream_off = self.s_addr.ctypes.data_as(u64p); SR.stream_len = self.s_len.ctypes.data_as(u64p)
            SR.read_off = vp(self.t_roff.data_ptr()); SR.rflags = vp(self.t_rfl.data_ptr())
            self.U = rec.Features(); self.d_mapq = C.c_void_p(); self.u_status = np.zeros(n_cont, np.int32)
        # correctness of what is timed: the round trip restores every base and quality of every read
        self.decode()
        self.stream.synchronize()
        assert int(self.dd_st.abs().sum()) == 0 and int(self.o_stat.abs().sum()) == 0 and int(self.o_star.sum()) == 0
        assert torch.equal(self.o_bases, self.t_bases) and torch.equal(self.o_quals, self.t_quals), 'round trip mismatch'
        if self.use_unpack:
            got = np.empty(n_reads, np.uint8)
            self.ctx.check(self.ctx.lib.cramgpu_copy_to_host(self.ctx.handle, vp(got.ctypes.data), self.d_mapq, C.c_uint64(n_reads), self._sp()))
            assert np.array_equal(got, self.t_mapq.cpu().numpy()) and int(np.abs(self.u_status).sum()) == 0, 'series -> features mismatch'
        self.prepared = True

    def decode(self):
        from cramtools_b200._lib import u8p, u32p, u64p, i32p
        ctx, F, N, vp = self.ctx, self.F, self.N, C.c_void_p
        ctx.check(ctx.lib.cramgpu_rans_decode_dev(ctx.handle, self.nblk, vp(self.enc_out.data_ptr()), self.c_off.ctypes.data_as(u64p),
                                                  self.c_len.ctypes.data_as(u32p), self.prefixes.ctypes.data_as(u8p),
                                                  vp(self.dec_out.data_ptr()), C.c_uint64(self.dec_out.numel()), vp(self.dd_off.data_ptr()),
                                                  vp(self.dd_len.data_ptr()), vp(self.dd_st.data_ptr()), self._sp()))
        if self.use_unpack:
            # row P on the device: the read features come from the series themselves (QS / BA as just decoded, the series the
            # Java host inflates -- FN FC FP BS DL SC RS HC PD MQ -- as the encoder left them), not from the encoder's arrays
            U = self.U
            sa = self.s_addr; keepq, keepb = sa[:, self.KQS].copy(), sa[:, self.KBA].copy()
            sa[:] = np.uint64(F.streams) + self.soff; sa[:, self.KQS] = keepq; sa[:, self.KBA] = keepb
            ctx.check(ctx.lib.cramgpu_features_unpack_dev(ctx.handle, C.byref(self.SR), C.byref(U), C.byref(self.d_mapq),
                                                          self.u_status.ctypes.data_as(i32p), self._sp()))
            N.feat_off = U.feat_off; N.f_op = U.f_op; N.f_pos = U.f_pos; N.f_val = U.f_val; N.f_aux = U.f_aux; N.payload = U.payload; N.payload_len = U.payload_len
            N.qs = None; N.read_qs_off = U.read_qs_off; N.ba = None; N.read_ba_off = U.read_ba_off      # absolute addresses
            N.matrix = F.matrix
        else:
            N.matrix = F.matrix; N.feat_off = F.feat_off; N.f_op = F.f_op; N.f_pos = F.f_pos; N.f_val = F.f_val; N.f_aux = F.f_aux
            N.payload = F.payload; N.payload_len = F.payload_len
        ctx.check(ctx.lib.cramgpu_normalize_decode_dev(ctx.handle, C.byref(N), vp(self.o_bases.data_ptr()), vp(self.o_quals.data_ptr()),
                                                       vp(self.o_star.data_ptr()), vp(self.o_stat.data_ptr()), self._sp()))

    def steps(self, k):
        for _ in range(k):
            self.encode()
            self.decode()


def run_ours(args):
    import torch
    import torch.distributed as dist
    import cramtools_b200  # noqa: F401  (fails loudly here when libcramgpu.so is missing)

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    dev = torch.device('cuda', local)
    peaks, peak_kind = measured_peaks()

    n_cont = args.containers
    nu = min(UNIQUE, n_cont)
    ref, rd, extra = unique_containers(nu, 1000 + 17 * rank)
    # Software pipeline over consecutive steps: `depth` contexts, each with its own stream and the whole batch; step i runs
    # on context i % depth.  The order-1 decoder is a recurrence per rANS state -- a few warps per SM for tens of
    # milliseconds, whatever the batch size -- so the encode of step i+1 (record kernels, histograms, encoder) runs beside
    # the decode of step i instead of after it, the way consecutive batches of a whole-genome stream would be fed.
    # Two host locks keep at most one context in each phase.
    n_lanes = depth = max(1, args.pipeline)
    groups = [dist.new_group() if world > 1 else None for _ in range(depth)]
    lanes = [Lane(i, n_cont, ref, rd, extra, nu, dev, rank, world, groups[i], not args.no_unpack) for i in range(depth)]
    for ln in lanes:
        ln.prepare()
    n_reads = lanes[0].n_reads

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    enc_lock, dec_lock = threading.Lock(), threading.Lock()

    def run_steps(k):
        """k steps in all; with depth > 1 each context is driven by its own host thread (the C calls release the GIL)"""
        if depth == 1:
            lanes[0].steps(k)
            return
        errs = []

        def work(ln, i):
            try:
                torch.cuda.set_device(local)
                for _ in range(i, k, depth):
                    with enc_lock:
                        ln.encode()
                    with dec_lock:
                        ln.decode()
            except Exception as e:      # noqa: BLE001
                errs.append(e)
        th = [threading.Thread(target=work, args=(ln, i)) for i, ln in enumerate(lanes)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]

    run_steps(args.warmup)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = sum(ln.ctx.launch_count() for ln in lanes)
    for ln in lanes:
        ln.ctx.profile_enable(True)
    # device time of the timed region: first start .. last end over the lanes' streams
    e0 = [torch.cuda.Event(enable_timing=True) for _ in lanes]; e1 = [torch.cuda.Event(enable_timing=True) for _ in lanes]
    for ln, ev in zip(lanes, e0):
        ev.record(ln.stream)
    run_steps(args.steps)
    for ln, ev in zip(lanes, e1):
        ev.record(ln.stream)
    barrier()
    ms = max(e0[0].elapsed_time(b) for b in e1) / args.steps      # all e0 are recorded back to back right after a device-wide sync
    spans = []
    for ln in lanes:
        spans += ln.ctx.profile_read()
        ln.ctx.profile_enable(False)
    launches = (sum(ln.ctx.launch_count() for ln in lanes) - launches0) // args.steps
    clocks = sampler.summary()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    kt = {}
    for name, m in spans:
        kt.setdefault(name, []).append(m)
    kavg = {k: float(np.sum(v)) / args.steps for k, v in kt.items()}       # summed over lanes and launches, per step
    kcount = {k: len(v) / args.steps for k, v in kt.items()}
    top = max(kavg, key=kavg.get)

    # ---- end-to-end through the host-buffer C ABI (bounded to keep pinned host memory modest at 8 ranks)
    # every rank page-locks its own buffers (~21 MB per container with four contexts): keep all ranks of the node together
    # under 40 % of the free host memory
    n_e2e = min(n_cont, args.e2e_containers)
    try:
        import psutil
        per_rank = 0.4 * psutil.virtual_memory().available / max(world, 1)
        n_e2e = int(max(128, min(n_e2e, per_rank // (21 << 20))))
    except Exception:      # noqa: BLE001
        n_e2e = min(n_e2e, max(128, 2 * args.e2e_containers // max(world, 2))) if world > 1 else n_e2e
    if world > 1:
        t_n = torch.tensor([n_e2e], dtype=torch.int64, device=dev)
        dist.all_reduce(t_n, op=dist.ReduceOp.MIN)
        n_e2e = int(t_n.item())
    st0 = lanes[0]
    stats = dict(raw_bytes=st0.raw_bytes, comp_bytes=st0.comp_bytes, raw_o0=st0.raw_o0, qs_bytes=st0.qs_bytes, nblk=st0.nblk)
    comp_hint = st0.c_len[:n_e2e * st0.NB].copy()
    # the device-resident leg is over: its buffers and contexts go before the end-to-end leg allocates its own
    for ln in lanes:
        ln.ctx.close()
    del lanes, st0, ln
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    e2e = run_e2e(args, ref, rd, extra, nu, n_e2e, world, barrier, args.e2e_pipeline, comp_hint, local)

    if rank == 0:
        value = n_reads * world / (ms * 1e-3)
        raw_bytes = stats['raw_bytes']; comp_bytes = stats['comp_bytes']
        o1_raw = raw_bytes - stats['raw_o0']
        comp1 = comp_bytes * (o1_raw / raw_bytes)
        algo = {'k_enc1': o1_raw + comp1, 'k_dec1': o1_raw + comp1, 'k_hist1': o1_raw, 'k_hist0': raw_bytes,
                'k_rec_simple_count': n_reads * (2 * READ_LEN + 40.0), 'k_rec_simple_write': n_reads * (2 * READ_LEN + 60.0),
                'k_norm_simple': n_reads * (4 * READ_LEN + 40.0), 'k_pack': 2 * comp_bytes}
        ab = algo.get(top)
        # kavg sums the lanes' launches, which overlap in time; the per-launch figure is what the roofline wants
        per_launch_ms = kavg[top] / max(kcount[top], 1)
        achieved = (ab / max(kcount[top], 1)) / (per_launch_ms * 1e-3) / 1e9 if ab else None
        peak = float(peaks['hbm_gbs'])
        rans_ms_enc = sum(v for k, v in kavg.items() if k in ('k_hist0', 'k_prep_o1', 'k_hist1', 'k_tables0', 'k_tables1', 'k_enc0', 'k_enc1', 'k_scan_out', 'k_pack'))
        rans_ms_dec = sum(v for k, v in kavg.items() if k in ('k_dec_tables', 'k_dec0', 'k_dec1', 'k_dec1_big'))
        cfg = workload_config(args, n_cont)
        cfg['decode_features'] = 'taken from the encoder' if args.no_unpack else 'unpacked from the decoded series on the device (k_unpack, row P)'
        cfg['pipeline'] = ('%d contexts/streams per GPU, step i on context i %% %d: encode of step i+1 overlaps decode of step i' % (depth, depth)) if depth > 1 else 'none: one stream, phases back to back'
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'u8/i32 integer (+f64 in the order-1 frequency normalisation)', 'data': 'synthetic', 'config': cfg,
            'rans': {'raw_GB_per_step_per_gpu': raw_bytes / 1e9, 'compressed_ratio': comp_bytes / raw_bytes,
                     'encode_GBps_per_gpu_kernel_time': raw_bytes / (rans_ms_enc * 1e-3) / 1e9,
                     'decode_GBps_per_gpu_kernel_time': raw_bytes / (rans_ms_dec * 1e-3) / 1e9,
                     'qs_block_bytes': stats['qs_bytes'] / n_cont, 'state_lanes_in_flight': int(4 * stats['nblk'])},
            'kernels_ms': {k: round(v, 4) for k, v in kavg.items()}, 'kernel_launches_per_step': {k: v for k, v in kcount.items()},
            'gpu_launches': int(launches),
            'roofline': {'bound': 'hbm', 'kernel': top, 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                         'frac': achieved / peak if achieved else None, 'traffic': ncu_traffic(top, n_cont), 'peak_kind': peak_kind,
                         'algorithmic_bytes_per_launch': (ab / max(kcount[top], 1)) if ab else None,
                         'note': 'per launch; traffic = dram bytes read + written by one launch in the ncu --set full capture of '
                                 'profiles/ (null at other sizes); rANS is a recurrence per state: attainable rate = resident states x '
                                 'clock / cycles per step (DESIGN.md, "Why the coder cannot reach the HBM roofline")'},
            'e2e': e2e, 'clocks': clocks,
        }
        if world == 1 and not args.no_cpu:
            line['cpu_baseline'] = cpu_baseline(args)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


class E2EContext(object):
    """One context of the end-to-end leg: its own cramgpu context, output buffers in pinned host memory and the device
    buffer the decoded QS / BA series stay in.  The input reads (pinned host memory) are shared, read-only."""

    def __init__(self, ctx, inp, pinned_empty):
        import torch
        from cramtools_b200 import records as rec, rans_bound
        from cramtools_b200._lib import u64p
        self.ctx, self.inp, self.rec, self.torch = ctx, inp, rec, torch
        vp = C.c_void_p
        n_cont, n_reads, NB = inp['n_cont'], inp['n_reads'], inp['NB']
        nblk = self.nblk = n_cont * NB
        self.F, self.D = rec.Features(), rec.Features()
        self.in_ptr = np.zeros(nblk, np.uint64); self.in_len = np.zeros(nblk, np.uint32)
        self.in_ptr.reshape(n_cont, NB)[:, 2:] = inp['ex_ptr']; self.in_len.reshape(n_cont, NB)[:, 2:] = inp['ex_len']
        # compressed blocks, the product of the encode side: one pinned buffer, a slot per block.  (The reference allocates
        # rans_bound(n) per block and trims; here the slots are the sizes the device-resident leg saw plus a margin, so that
        # the page-locked memory stays in proportion to the data.)
        self.caps = ((inp['comp_hint'].astype(np.int64) * 9 // 8 + 4096 + 15) & ~15).astype(np.uint32)
        self.cap_off = np.concatenate([[0], np.cumsum(self.caps.astype(np.int64))])
        self.comp = pinned_empty(self.cap_off[-1], np.uint8)
        self.out_ptr = (np.uint64(self.comp.ctypes.data) + self.cap_off[:-1].astype(np.uint64))
        self.out_len = np.zeros(nblk, np.uint32); self.status = np.zeros(nblk, np.int32)
        # decoded series: QS and BA stay in device memory, the host-parsed series come back to the host
        raw_qs = READS_PER_CONTAINER * READ_LEN + 64
        self.dcap = np.tile(np.array([raw_qs, raw_qs] + [60000 + 64] * (NB - 2), np.uint32), n_cont)
        dslot = np.concatenate([[0], np.cumsum(self.dcap.astype(np.int64))])
        self.d_back = torch.empty(int(dslot[-1]), dtype=torch.uint8, device='cuda')
        self.back = pinned_empty(int((60000 + 64) * (NB - 2) * n_cont), np.uint8)
        hslot = (np.arange(n_cont * (NB - 2), dtype=np.int64) * (60000 + 64)).reshape(n_cont, NB - 2)
        bp = np.zeros((n_cont, NB), np.uint64)
        bp[:, :2] = (np.uint64(self.d_back.data_ptr()) + dslot[:-1].astype(np.uint64)).reshape(n_cont, NB)[:, :2]
        bp[:, 2:] = np.uint64(self.back.ctypes.data) + hslot.astype(np.uint64)
        self.back_ptr = np.ascontiguousarray(bp.reshape(-1))
        self.back_len = np.zeros(nblk, np.uint32)
        self.qs_slot = dslot[np.arange(n_cont) * NB].astype(np.int64)             # where container c's QS lands in d_back
        self.rfl = np.ones(n_reads, np.uint8)
        self.rba = np.full(n_reads, 0xFFFFFFFFFFFFFFFF, np.uint64)
        self.rq_new = np.zeros(n_reads, np.uint64)
        self.o_b = pinned_empty(n_reads * READ_LEN, np.uint8); self.o_q = pinned_empty(n_reads * READ_LEN, np.uint8)
        self.o_s = np.empty(n_reads, np.uint8); self.o_t = np.empty(n_reads, np.int32)
        N = self.N = rec.NormIn()
        N.n_reads = n_reads; N.n_containers = n_cont
        N.cont_off = inp['cont_off'].ctypes.data_as(u64p); N.win_off = inp['win_off'].ctypes.data_as(u64p); N.win_len = inp['win_len'].ctypes.data_as(u64p)
        N.win_start0 = inp['win_s0'].ctypes.data_as(C.POINTER(C.c_int64)); N.ref = vp(inp['ref'].ctypes.data); N.ref_total = inp['ref'].size
        N.pos = vp(inp['a']['pos'].ctypes.data); N.read_off = vp(inp['a']['read_off'].ctypes.data); N.rflags = vp(self.rfl.ctypes.data)
        N.ba = vp(self.d_back.data_ptr()); N.read_ba_off = vp(self.rba.ctypes.data); N.qs = vp(self.d_back.data_ptr())
        self.ms = {'features_encode_resident': 0.0, 'rans_encode_batch': 0.0, 'rans_decode_batch': 0.0, 'normalize_decode': 0.0, 'host_glue': 0.0}
        self.h2d = self.d2h = 0
        self.stages = (self.stage_features, self.stage_rans_encode, self.stage_rans_decode, self.stage_normalize)

    def step(self, locks=None):
        for k, stage in enumerate(self.stages):
            if locks is None:
                stage()
            else:
                with locks[k]:
                    stage()

    def stage_features(self):
        """reads (host) -> the small series and the feature arrays (host); QS / BA stay on the device"""
        ctx, inp, F, D = self.ctx, self.inp, self.F, self.D
        t0 = time.perf_counter()
        ctx.check(ctx.lib.cramgpu_features_encode_resident(ctx.handle, C.byref(inp['R']), None, -1, inp['host_series'], C.byref(F), C.byref(D)))
        self.ms['features_encode_resident'] += (time.perf_counter() - t0) * 1e3

    def stage_rans_encode(self):
        """series -> compressed blocks (host)"""
        from cramtools_b200._lib import u8p, u32p, u64p, i32p, u8pp
        ctx, inp, rec, F, D = self.ctx, self.inp, self.rec, self.F, self.D
        n_cont, NB = inp['n_cont'], inp['NB']
        t1 = time.perf_counter()
        self.soff = np.ctypeslib.as_array(C.cast(F.stream_off, u64p), (n_cont, 12)); self.slen = np.ctypeslib.as_array(C.cast(F.stream_len, u64p), (n_cont, 12))
        v = self.in_ptr.reshape(n_cont, NB); w = self.in_len.reshape(n_cont, NB)
        KQS, KBA = inp['KQS'], inp['KBA']
        v[:, 0] = np.uint64(D.streams) + self.soff[:, KQS]; w[:, 0] = self.slen[:, KQS]      # QS and BA: coded where they are, on the device
        v[:, 1] = np.uint64(D.streams) + self.soff[:, KBA]; w[:, 1] = self.slen[:, KBA]
        t2 = time.perf_counter()
        ctx.check(ctx.lib.cramgpu_rans_encode_batch(ctx.handle, self.nblk, C.cast(self.in_ptr.ctypes.data, u8pp), self.in_len.ctypes.data_as(u32p),
                                                    inp['order'].ctypes.data_as(u8p), C.cast(self.out_ptr.ctypes.data, u8pp), self.caps.ctypes.data_as(u32p),
                                                    self.out_len.ctypes.data_as(u32p), self.status.ctypes.data_as(i32p)))
        t3 = time.perf_counter()
        self.ms['rans_encode_batch'] += (t3 - t2) * 1e3; self.ms['host_glue'] += (t2 - t1) * 1e3

    def stage_rans_decode(self):
        """compressed blocks (host) -> series: QS / BA into device memory, the host-parsed ones into host memory"""
        from cramtools_b200._lib import u32p, i32p, u8pp
        ctx = self.ctx
        t3 = time.perf_counter()
        ctx.check(ctx.lib.cramgpu_rans_decode_batch(ctx.handle, self.nblk, C.cast(self.out_ptr.ctypes.data, u8pp), self.out_len.ctypes.data_as(u32p),
                                                    C.cast(self.back_ptr.ctypes.data, u8pp), self.dcap.ctypes.data_as(u32p),
                                                    self.back_len.ctypes.data_as(u32p), self.status.ctypes.data_as(i32p)))
        self.ms['rans_decode_batch'] += (time.perf_counter() - t3) * 1e3

    def stage_normalize(self):
        """records as CramRecordReader would hand them over + the decoded QS / BA on the device -> bases and qualities (host)"""
        from cramtools_b200._lib import u8p, u64p, i32p
        ctx, inp, F, N = self.ctx, self.inp, self.F, self.N
        n_reads = inp['n_reads']
        vp = C.c_void_p
        t4 = time.perf_counter()
        # score arrays sit inside the decoded QS series: per read, its offset there (the encoder's offset, moved to the QS slot)
        rq = np.ctypeslib.as_array(C.cast(F.read_qs_off, u64p), (n_reads,))
        delta = (self.qs_slot - self.soff[:, inp['KQS']].astype(np.int64)).astype(np.int64)
        np.add(rq.view(np.int64).reshape(inp['n_cont'], READS_PER_CONTAINER), delta[:, None],
               out=self.rq_new.view(np.int64).reshape(inp['n_cont'], READS_PER_CONTAINER))
        N.read_qs_off = vp(self.rq_new.ctypes.data)
        N.matrix = F.matrix; N.feat_off = F.feat_off; N.f_op = F.f_op; N.f_pos = F.f_pos; N.f_val = F.f_val; N.f_aux = F.f_aux
        N.payload = F.payload; N.payload_len = F.payload_len
        t5 = time.perf_counter()
        ctx.check(ctx.lib.cramgpu_normalize_decode(ctx.handle, C.byref(N), self.o_b.ctypes.data_as(u8p), self.o_q.ctypes.data_as(u8p),
                                                   self.o_s.ctypes.data_as(u8p), self.o_t.ctypes.data_as(i32p)))
        t6 = time.perf_counter()
        self.ms['normalize_decode'] += (t6 - t5) * 1e3; self.ms['host_glue'] += (t5 - t4) * 1e3
        # bytes that crossed the bus in this step
        a, nf, nb = inp['a'], int(F.n_feat), n_reads * READ_LEN
        KQS, KBA = inp['KQS'], inp['KBA']
        small = int(self.slen.sum() - self.slen[:, KQS].sum() - self.slen[:, KBA].sum())
        ex_bytes = int(inp['ex_len'].sum())
        comp_bytes = int(self.out_len.sum())
        self.h2d = int(2 * nb + inp['ref'].size + a['read_off'].nbytes + a['cigar_off'].nbytes + a['cigar'].nbytes + 7 * n_reads      # reads
                       + ex_bytes                                                                                         # host-produced series
                       + comp_bytes                                                                                       # compressed blocks, decode side
                       + inp['ref'].size + 37 * n_reads + 17 * nf + int(F.payload_len))                                   # records + features, decode side
        self.d2h = int(17 * nf + 25 * n_reads + int(F.payload_len) + small                                                # features, tables, small series
                       + comp_bytes + ex_bytes + 2 * nb + 5 * n_reads)


def run_e2e(args, ref, rd, extra, nu, n_cont, world, barrier, depth, comp_hint, device):
    """the same step through the host-memory entry points: inputs start in (pinned) host memory, the compressed blocks and
    the restored bases / qualities end in host memory, every copy inside the timed region.  Intermediate series the host
    does not parse (QS, BA) stay on the device between the calls.  `depth` contexts take consecutive steps in turn, each call
    guarded by its own lock, so the host-to-device traffic of one step's encode overlaps the device-to-host traffic and the
    coder kernels of the steps before it."""
    import cramtools_b200
    from cramtools_b200 import records as rec
    from cramtools_b200._lib import u64p
    import torch
    vp = C.c_void_p
    keep_pinned = []

    def pinned(x):
        # page-locked host arrays, as the contract asks ("inputs from pinned host memory"); the library sends them to the DMA engine as they are
        t = torch.empty(x.nbytes, dtype=torch.uint8, pin_memory=True)
        keep_pinned.append(t)
        out = t.numpy().view(x.dtype).reshape(x.shape)
        out[...] = x
        return out

    def pinned_empty(n, dt):
        t = torch.empty(int(n) * np.dtype(dt).itemsize, dtype=torch.uint8, pin_memory=True)
        keep_pinned.append(t)
        return t.numpy().view(dt)
    trd = tile_reads(rd, nu, n_cont)
    n_reads = n_cont * READS_PER_CONTAINER
    ref = pinned(ref)
    cont_off = (np.arange(n_cont + 1, dtype=np.uint64) * READS_PER_CONTAINER)
    ref_off = np.zeros(n_cont, np.uint64); ref_len = np.full(n_cont, ref.size, np.uint64)
    a = dict(bases=pinned(trd['bases']), quals=pinned(trd['quals']), read_off=pinned(trd['read_off'].astype(np.uint64)),
             cigar=pinned(trd['cigar'].astype(np.uint32)), cigar_off=pinned(trd['cigar_off'].astype(np.uint64)),
             pos=pinned(trd['pos'].astype(np.int32)), flag=pinned(trd['flag'].astype(np.uint16)), mapq=pinned(trd['mapq'].astype(np.uint8)))
    R = rec.Reads()
    R.n_reads = n_reads; R.n_containers = n_cont
    R.cont_off = cont_off.ctypes.data_as(u64p); R.ref_off = ref_off.ctypes.data_as(u64p); R.ref_len = ref_len.ctypes.data_as(u64p)
    R.ref = vp(ref.ctypes.data); R.ref_total = ref.size
    for k, v in a.items():
        setattr(R, k, vp(v.ctypes.data))
    R.rflags = None
    NX = len(EXTRA); NB = 2 + NX
    ex = [np.ascontiguousarray(e, np.uint8) for c in range(n_cont) for e in extra[c % nu]]
    p = rd['pos'].astype(np.int64).reshape(nu, READS_PER_CONTAINER)
    reps = (n_cont + nu - 1) // nu
    ws = (p.min(axis=1) - 1); we = np.minimum(p.max(axis=1) - 1 + READ_LEN + 8, ref.size)
    KQS, KBA = rec.SERIES.index('QS'), rec.SERIES.index('BA')
    inp = dict(n_cont=n_cont, n_reads=n_reads, NB=NB, R=R, a=a, ref=ref, cont_off=cont_off, keep=[ref_off, ref_len, ex],
               ex_ptr=np.array([e.ctypes.data for e in ex], np.uint64).reshape(n_cont, NX),
               ex_len=np.array([e.size for e in ex], np.uint32).reshape(n_cont, NX),
               order=np.tile(np.array([1, 1] + [o for _, o in EXTRA], np.uint8), n_cont), KQS=KQS, KBA=KBA,
               host_series=sum(1 << k for k in range(len(rec.SERIES)) if k not in (KQS, KBA)), comp_hint=comp_hint,
               win_off=np.tile(ws, reps)[:n_cont].astype(np.uint64), win_len=np.tile(we - ws, reps)[:n_cont].astype(np.uint64),
               win_s0=np.tile(ws, reps)[:n_cont].astype(np.int64))
    depth = max(1, depth)
    ctxs = [cramtools_b200.Context(device) for _ in range(depth)]
    lanes = [E2EContext(c, inp, pinned_empty) for c in ctxs]
    for ln in lanes:
        ln.step()
        assert np.array_equal(ln.o_b, a['bases']) and np.array_equal(ln.o_q, a['quals']), 'e2e round trip mismatch'
        ln.o_b[:] = 0; ln.o_q[:] = 0
    barrier()
    k = max(3 * depth, args.steps)        # long enough for the pipeline to run full most of the time (its fill and drain are inside the timed region)
    for ln in lanes:
        for key in ln.ms:
            ln.ms[key] = 0.0
    # one lock per call: at most one context inside each of the four calls, different calls of different steps overlap
    locks = [threading.Lock() for _ in range(4)]
    errs = []
    dev_index = torch.cuda.current_device()

    def runner(ln, i):
        try:
            torch.cuda.set_device(dev_index)
            for _ in range(i, k, depth):
                ln.step(None if os.environ.get('CRAMGPU_E2E_NOLOCK') else locks)
        except Exception as e:      # noqa: BLE001
            errs.append(e)
    t0 = time.perf_counter()
    if depth == 1:
        for _ in range(k):
            lanes[0].step()
    else:
        th = [threading.Thread(target=runner, args=(ln, i)) for i, ln in enumerate(lanes)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]
    barrier()
    ms = (time.perf_counter() - t0) / k * 1e3
    t = torch.tensor([ms], dtype=torch.float64, device='cuda')
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    calls = {kk: round(sum(ln.ms[kk] for ln in lanes) / k, 2) for kk in lanes[0].ms}
    return {'value': n_reads * world / (ms * 1e-3), 'unit': UNIT, 'h2d_bytes_per_step': lanes[0].h2d, 'd2h_bytes_per_step': lanes[0].d2h,
            'ms_per_step': ms, 'containers_per_gpu': n_cont, 'steps': k, 'contexts': depth, 'calls_ms': calls,
            'path': 'cramgpu_features_encode_resident -> cramgpu_rans_encode_batch -> cramgpu_rans_decode_batch -> cramgpu_normalize_decode: '
                    'reads in and compressed blocks + restored bases/qualities out through host buffers, every copy inside the timed '
                    'region; the QS and BA series stay on the device between the calls'}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--containers', type=int, default=2048, help='containers PER GPU (10 000 reads each)')
    ap.add_argument('--e2e-containers', type=int, default=1024, help='containers per GPU of the end-to-end (host buffer) leg')
    ap.add_argument('--pipeline', type=int, default=2, help='contexts per GPU taking consecutive steps in turn (1 = no overlap of encode and decode)')
    ap.add_argument('--e2e-pipeline', type=int, default=4, help='contexts of the end-to-end leg taking consecutive steps in turn')
    ap.add_argument('--no-unpack', action='store_true', help='decode side: take the read features from the encoder instead of unpacking the series (row P)')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
