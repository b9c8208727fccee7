#!/usr/bin/env python
"""bench.py -- headline benchmark of the CRAM slice-codec hot path ("Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--containers C]

Workload (BASELINE.json configs[2], "chr20-scale"): C containers PER GPU of 10 000 simulated 150 bp reads each at
~45x over a random reference, lossless quality scores.  A "step" is one pass of the whole hot path over this GPU's
containers:
    encode  feature extraction + substitution matrix + series packing (read features and the record fields in front of
            them: flags, lengths, positions, read groups, names, mate fields)  ->  rANS of the slice's rANS-coded series
            (QS, BA, BF, CF, RL, RG, NS, TS order 1 and AP order 0)
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
# the record-field series the reference codes with rANS (CompressionHeaderFactory.build [:84-113]); the others
# (RN MF NP NF TL, RI of single-reference slices is empty) go to gzip on the Java host
EXTRA = [('BF', 1), ('CF', 1), ('RL', 1), ('RG', 1), ('NS', 1), ('TS', 1), ('AP', 0)]


def measured_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), 'measured'
    return {'hbm_gbs': 6650.0, 'sm_max_mhz': 1965.0}, 'fallback'


def ncu_traffic(kernel, n_cont):
    """DRAM bytes of one launch of `kernel` from the committed ncu capture (profiles/r2_traffic.json), if it was taken at this size"""
    try:
        with open(os.path.join(ROOT, 'profiles', 'r2_traffic.json')) as f:
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
    """n_unique coordinate-sorted containers over one synthetic contig: (ref, rd).  rd also carries what the record fields
    in front of the read features are made of (Writer.write [Writer.java:139-172]): read names (`names` u8 concatenated,
    `name_off`) and a read-group index per read.  The reads are single-end, so every record is detached
    (CramSerilization.setMateInfo, :278-285): MF 0, NS -1, NP 0, TS 0."""
    from cramtools_b200 import synth
    n = n_unique * READS_PER_CONTAINER
    ref_len = n * READ_LEN // COVERAGE + 4096
    ref = synth.reference(ref_len, seed)
    rd = synth.reads(ref, n, seed + 1, read_len=READ_LEN, with_names=False)
    rng = np.random.default_rng(seed + 2)
    # SIM:1:FC01:<lane>:<tile>:<x>:<y>, built vectorised (synth.read_names makes the same kind of name one by one)
    lane = rng.integers(1, 9, size=n); tile = rng.integers(1101, 2229, size=n); x = rng.integers(1000, 30000, size=n); y = rng.integers(1000, 200000, size=n)
    names = np.char.add(np.char.add(np.char.add(np.char.add('SIM:1:FC01:', lane.astype('U1')), np.char.add(':', tile.astype('U4'))),
                                    np.char.add(':', x.astype('U5'))), np.char.add(':', y.astype('U6')))
    nb = np.char.encode(names, 'ascii')
    lens = np.char.str_len(names).astype(np.int64)
    rd['names'] = np.frombuffer(b''.join(nb.tolist()), np.uint8).copy()
    rd['name_off'] = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    rd['read_group'] = rng.integers(0, 4, size=n).astype(np.int32)
    return ref, rd


def tile_reads(rd, n_unique, n_cont):
    """host arrays for n_cont containers by tiling the unique ones"""
    reps = (n_cont + n_unique - 1) // n_unique
    nr = n_cont * READS_PER_CONTAINER
    nb = int(rd['read_off'][-1])
    ncg = int(rd['cigar_off'][-1])
    nnm = int(rd['name_off'][-1])
    out = {}
    for k in ('bases', 'quals', 'cigar', 'pos', 'flag', 'mapq', 'names', 'read_group'):
        out[k] = np.tile(rd[k], reps)

    def tile_off(off, total):
        o = (off[:-1][None, :] + (np.arange(reps, dtype=np.int64) * total)[:, None]).reshape(-1)
        return np.concatenate([o[:nr], [o[nr] if nr < o.size else reps * total]])
    out['read_off'] = tile_off(rd['read_off'], nb)
    out['cigar_off'] = tile_off(rd['cigar_off'], ncg)
    out['name_off'] = tile_off(rd['name_off'], nnm)
    for k in ('pos', 'flag', 'mapq', 'read_group'):
        out[k] = out[k][:nr]
    out['bases'] = out['bases'][:int(out['read_off'][-1])]
    out['quals'] = out['quals'][:int(out['read_off'][-1])]
    out['cigar'] = out['cigar'][:int(out['cigar_off'][-1])]
    out['names'] = out['names'][:int(out['name_off'][-1])]
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
                        'lossless QS; encode = features + packing of all series (read features and record fields) + rANS (QS, BA, BF, CF, RL, RG, NS, TS o1; AP o0), '
                        'decode = rANS + base/quality restoration' % (n_cont, COVERAGE),
            'containers_per_gpu': n_cont, 'reads_per_container': READS_PER_CONTAINER, 'read_len': READ_LEN,
            'unique_containers': min(UNIQUE, n_cont), 'quality_model': 'lossless',
            'reference': 'a separate copy of the synthetic contig per tile of %d containers (no cache reuse between tiles); resident on the device' % min(UNIQUE, n_cont),
            'l2': 'inputs larger than L2 (no flush needed)' if n_cont >= 64 else 'inputs may fit L2',
            'sharding': 'containers round-robin over ranks; all-gather of compressed sizes', 'note': note}


# ------------------------------------------------------------------------------------------- CPU arm


def cpu_step_setup(n_cont, seed, spec=b'lossless'):
    from oracle import rans as orans
    lib = orans.lib()
    nu = min(UNIQUE, n_cont)
    ref, rd = unique_containers(nu, seed)
    trd = tile_reads(rd, nu, n_cont)
    i64 = C.POINTER(C.c_int64)
    keep = dict(cont=np.arange(n_cont + 1, dtype=np.int64) * READS_PER_CONTAINER, bases=trd['bases'], quals=trd['quals'],
                ro=trd['read_off'].astype(np.int64), cg=trd['cigar'], co=trd['cigar_off'].astype(np.int64), pos=trd['pos'].astype(np.int32),
                flag=trd['flag'].astype(np.uint16), mapq=trd['mapq'].astype(np.uint8), ref=ref,
                rg=np.ascontiguousarray(trd['read_group'], np.int32), names=np.ascontiguousarray(trd['names'], np.uint8),
                no=trd['name_off'].astype(np.int64))
    lib.oracle_cram_step.restype = C.c_int
    comp = C.c_int64(0); chk = C.c_uint64(0)

    def step(threads):
        rc = lib.oracle_cram_step(
            C.c_int(threads), C.c_int(n_cont), keep['cont'].ctypes.data_as(i64), C.c_void_p(keep['bases'].ctypes.data),
            C.c_void_p(keep['quals'].ctypes.data), keep['ro'].ctypes.data_as(i64), C.c_void_p(keep['cg'].ctypes.data),
            keep['co'].ctypes.data_as(i64), C.c_void_p(keep['pos'].ctypes.data), C.c_void_p(keep['flag'].ctypes.data),
            C.c_void_p(keep['mapq'].ctypes.data), C.c_void_p(keep['ref'].ctypes.data), C.c_int64(keep['ref'].size),
            C.c_void_p(keep['rg'].ctypes.data), C.c_void_p(keep['names'].ctypes.data), keep['no'].ctypes.data_as(i64),
            C.c_char_p(spec), C.byref(comp), C.byref(chk))
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

    def __init__(self, lane_id, n_cont, ref, rd, nu, dev, rank, world, group, use_unpack=True, spec=None):
        import torch
        import cramtools_b200
        from cramtools_b200 import device as dv, records as rec
        from cramtools_b200._lib import u64p, u8p
        self.torch, self.rec, self.dv = torch, rec, dv
        self.n_cont, self.rank, self.world, self.group, self.dev = n_cont, rank, world, group, dev
        self.use_unpack = use_unpack
        self.spec = spec
        self.pol, self.npol = rec._policy_args(spec)
        self.stream = torch.cuda.Stream(device=dev)
        self.wait_event = None; self.d1_event = None; self.before_rans_encode = None      # set per step by run_ours' schedule
        self.side = torch.cuda.Stream(device=dev)          # the slice digests run beside the record-field and histogram kernels
        self.ev_fork, self.ev_join = torch.cuda.Event(), torch.cuda.Event()
        self.ctx = cramtools_b200.Context(dev.index)
        self.with_digests = False          # slice header fields + BD/SD digests: a leg of their own, see run_ours
        vp = C.c_void_p
        with torch.cuda.stream(self.stream):
            reps = (n_cont + nu - 1) // nu
            n_reads = self.n_reads = n_cont * READS_PER_CONTAINER
            ncg_u = int(rd['cigar_off'][-1]); nnm_u = int(rd['name_off'][-1])

            def up(a, dt):
                return torch.from_numpy(np.ascontiguousarray(a, dt).view(np.uint8)).to(dev)

            def tile_u8(a, dt, n_elems):
                return up(a, dt).repeat(reps)[:n_elems * np.dtype(dt).itemsize].contiguous()

            def tile_off(off, total):
                o = (off[:-1][None, :] + (np.arange(reps, dtype=np.int64) * total)[:, None]).reshape(-1)
                return np.concatenate([o[:n_reads], [o[n_reads] if n_reads < o.size else reps * total]]).astype(np.uint64)
            self.t_bases = tile_u8(rd['bases'], np.uint8, n_reads * READ_LEN)
            self.t_quals = tile_u8(rd['quals'], np.uint8, n_reads * READ_LEN)
            self.t_pos = tile_u8(rd['pos'], np.int32, n_reads); self.t_flag = tile_u8(rd['flag'], np.uint16, n_reads)
            self.t_mapq = tile_u8(rd['mapq'], np.uint8, n_reads)
            self.t_rg = tile_u8(rd['read_group'], np.int32, n_reads)
            co_all = tile_off(rd['cigar_off'], ncg_u)
            self.t_cigar = up(np.tile(rd['cigar'], reps)[:int(co_all[-1])], np.uint32)
            self.t_coff = up(co_all, np.uint64)
            no_all = tile_off(rd['name_off'], nnm_u)
            self.t_names = tile_u8(rd['names'], np.uint8, int(no_all[-1]))
            self.t_noff = up(no_all, np.uint64)
            self.t_roff = up(np.arange(n_reads + 1, dtype=np.uint64) * READ_LEN, np.uint64)
            # every tile of the distinct containers reads its OWN copy of the reference (separate memory: no L2 reuse between tiles)
            self.t_ref = up(ref, np.uint8).repeat(reps).contiguous()
            self.cont_off = (np.arange(n_cont + 1, dtype=np.uint64) * READS_PER_CONTAINER)
            tile_of = (np.arange(n_cont) // nu).astype(np.uint64)
            self.ref_off = tile_of * np.uint64(ref.size); self.ref_len = np.full(n_cont, ref.size, np.uint64)
            R = self.R = rec.Reads()
            R.n_reads = n_reads; R.n_containers = n_cont
            R.cont_off = self.cont_off.ctypes.data_as(u64p); R.ref_off = self.ref_off.ctypes.data_as(u64p)
            R.ref_len = self.ref_len.ctypes.data_as(u64p)
            R.ref = vp(self.t_ref.data_ptr()); R.ref_total = self.t_ref.numel()
            R.bases = vp(self.t_bases.data_ptr()); R.quals = vp(self.t_quals.data_ptr()); R.read_off = vp(self.t_roff.data_ptr())
            R.cigar = vp(self.t_cigar.data_ptr()); R.cigar_off = vp(self.t_coff.data_ptr()); R.pos = vp(self.t_pos.data_ptr())
            R.flag = vp(self.t_flag.data_ptr()); R.mapq = vp(self.t_mapq.data_ptr()); R.rflags = None
            # region per container = [min start, max end] of its records (CramNormalizer.java:212-225); Slice.alignmentStart = min start
            p = rd['pos'].astype(np.int64).reshape(nu, READS_PER_CONTAINER)
            ws = (p.min(axis=1) - 1)
            we = np.minimum(p.max(axis=1) - 1 + READ_LEN + 8, ref.size)
            self.win_off = np.tile(ws, reps)[:n_cont].astype(np.uint64) + self.ref_off; self.win_len = np.tile(we - ws, reps)[:n_cont].astype(np.uint64)
            self.win_s0 = np.tile(ws, reps)[:n_cont].astype(np.int64)
            self.slice_start = np.tile(p.min(axis=1), reps)[:n_cont].astype(np.int32)
            # the record fields in front of the read features, packed on the device (cramgpu_host_series_pack_dev)
            H = self.H = rec.HostFields()
            H.n_reads = n_reads; H.n_containers = n_cont
            H.cont_off = self.cont_off.ctypes.data_as(u64p); H.slice_start = self.slice_start.ctypes.data_as(C.POINTER(C.c_int32))
            H.multi_ref = None; H.ap_delta = 1; H.names_included = 0
            H.flag = vp(self.t_flag.data_ptr()); H.cf = None; H.mate_cf = None; H.force_qs = None; H.rflags = None
            H.ref_id = None; H.pos = vp(self.t_pos.data_ptr()); H.read_off = vp(self.t_roff.data_ptr()); H.read_group = vp(self.t_rg.data_ptr())
            H.names = vp(self.t_names.data_ptr()); H.name_off = vp(self.t_noff.data_ptr())
            H.mate_ref_id = None; H.mate_pos = None; H.tlen = None; H.nf = None; H.tl = None
            self.HS = rec.HostSeries()
            # slice header fields + BD / SD digests (ContainerFactory.buildSlice :138-176), also inside the step
            SI = self.SI = rec.SliceIn()
            SI.n_reads = n_reads; SI.n_containers = n_cont; SI.cont_off = self.cont_off.ctypes.data_as(u64p)
            SI.bases = vp(self.t_bases.data_ptr()); SI.read_off = vp(self.t_roff.data_ptr()); SI.rflags = None; SI.ref_id = None
            SI.pos = vp(self.t_pos.data_ptr()); SI.flag = vp(self.t_flag.data_ptr())
            SI.scores = vp(self.t_quals.data_ptr()) if self.npol < 0 else None      # lossless: record.qualityScores = the qualities
            SI.ref = vp(self.t_ref.data_ptr()); SI.ref_total = self.t_ref.numel()   # Slice.refMD5
            SI.ref_off = self.ref_off.ctypes.data_as(u64p); SI.ref_len = self.ref_len.ctypes.data_as(u64p)
            self.ref_unique = ref
            self.slice_hdr = torch.zeros(n_cont * C.sizeof(rec.SliceHeader), dtype=torch.uint8, device=dev)
            self.uniq_reads = (rd, nu)
            NHS = len(rec.HOST_SERIES)
            self.hoff = np.zeros((n_cont, NHS), np.uint64); self.hlen = np.zeros((n_cont, NHS), np.uint64)
            self.KX = [rec.HOST_SERIES.index(nm) for nm, _ in EXTRA]
            NX = len(EXTRA)
            NB = self.NB = 2 + NX                               # rANS blocks per container: QS, BA, then EXTRA
            nblk = self.nblk = n_cont * NB
            self.order = np.tile(np.array([1, 1] + [o for _, o in EXTRA], np.uint8), n_cont)
            self.KQS, self.KBA = rec.SERIES.index('QS'), rec.SERIES.index('BA')
            self.soff = np.zeros((n_cont, 12), np.uint64); self.slen = np.zeros((n_cont, 12), np.uint64)
            self.F = rec.Features()
            self.in_addr = np.zeros(nblk, np.uint64); self.in_len = np.zeros(nblk, np.uint32)
            self.enc_cap = int(n_reads * READ_LEN * 0.6) + nblk * 4096 + n_reads * 16
            self.enc_out = torch.empty(self.enc_cap, dtype=torch.uint8, device=dev)
            self.e_off = torch.empty(nblk + 1, dtype=torch.int64, device=dev); self.e_len = torch.empty(nblk, dtype=torch.int32, device=dev)
            self.e_st = torch.empty(nblk, dtype=torch.int32, device=dev)
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
        if self.wait_event is not None:
            self.stream.wait_event(self.wait_event)              # the record kernels start when that step's order-1 decode has ended
        ctx.check(ctx.lib.cramgpu_features_encode_dev(ctx.handle, C.byref(self.R), self.pol, self.npol, C.byref(F), self.soff.ctypes.data_as(u64p),
                                                      self.slen.ctypes.data_as(u64p), self._sp()))
        self.H.force_qs = F.force_qs                     # CF bit 0 comes from the quality model's decision
        if self.with_digests:
            SI = self.SI
            SI.f_op = F.f_op; SI.f_pos = F.f_pos; SI.f_val = F.f_val; SI.f_aux = F.f_aux; SI.feat_off = F.feat_off
            SI.force_qs = F.force_qs; SI.read_qs_off = F.read_qs_off; SI.streams = F.streams
            # on a side stream: CRC32 is bound by shared-memory lookups, what follows on the main stream by HBM and by latency
            self.ev_fork.record(self.stream)
            self.side.wait_event(self.ev_fork)
            ctx.check(ctx.lib.cramgpu_slice_headers_dev(ctx.handle, C.byref(SI), vp(self.slice_hdr.data_ptr()), C.c_void_p(self.side.cuda_stream)))
            self.ev_join.record(self.side)
        ctx.check(ctx.lib.cramgpu_host_series_pack_dev(ctx.handle, C.byref(self.H), C.byref(self.HS), self.hoff.ctypes.data_as(u64p),
                                                       self.hlen.ctypes.data_as(u64p), self._sp()))
        base = np.uint64(F.streams); hbase = np.uint64(self.HS.streams)
        v = self.in_addr.reshape(self.n_cont, self.NB); w = self.in_len.reshape(self.n_cont, self.NB)
        v[:, 0] = base + self.soff[:, self.KQS]; w[:, 0] = self.slen[:, self.KQS]
        v[:, 1] = base + self.soff[:, self.KBA]; w[:, 1] = self.slen[:, self.KBA]
        v[:, 2:] = hbase + self.hoff[:, self.KX]; w[:, 2:] = self.hlen[:, self.KX]
        if self.before_rans_encode is not None:
            self.before_rans_encode()
        ctx.check(ctx.lib.cramgpu_rans_encode_dev(ctx.handle, self.nblk, None, self.in_addr.ctypes.data_as(u64p),
                                                  self.in_len.ctypes.data_as(u32p), self.order.ctypes.data_as(u8p),
                                                  vp(self.enc_out.data_ptr()), C.c_uint64(self.enc_cap), vp(self.e_off.data_ptr()),
                                                  vp(self.e_len.data_ptr()), vp(self.e_st.data_ptr()), self._sp()))
        if self.with_digests:
            self.stream.wait_event(self.ev_join)         # the step's results include the slice headers
        with self.torch.cuda.stream(self.stream):
            sizes = self.e_len.view(self.n_cont, self.NB).sum(dim=1, dtype=self.torch.int64)
            return sharding.allgather_offsets(sizes, self.rank, self.world, group=self.group)   # the path's only collective

    def prepare(self):
        """one verified pass to learn what the Java host would know on the decode side (block sizes, prefixes, offsets)"""
        torch, rec, dv = self.torch, self.rec, self.dv
        from cramtools_b200._lib import u64p
        vp = C.c_void_p
        self.with_digests = True                         # verified here once; timed only in the digest leg
        self.encode()
        self.with_digests = False
        self.stream.synchronize()
        assert int(self.e_st.abs().sum()) == 0, 'rANS encode status'
        nblk, n_cont, NB, n_reads = self.nblk, self.n_cont, self.NB, self.n_reads
        lossless = self.npol < 0
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
            rq_new = np.where(rq == np.uint64(0xFFFFFFFFFFFFFFFF), rq, rq - self.soff[cidx, self.KQS] + d_off[cidx * NB + 0])
            self.t_rq = torch.from_numpy(rq_new.view(np.uint8)).to(self.dev)
            self.t_rba = torch.full((n_reads,), -1, dtype=torch.int64, device=self.dev)
            if not lossless:
                # CF bit 0 of every record, as the host reads it from the CF series
                fq = np.empty(n_reads, np.uint8)
                self.ctx.check(self.ctx.lib.cramgpu_copy_to_host(self.ctx.handle, vp(fq.ctypes.data), vp(self.F.force_qs),
                                                                 C.c_uint64(n_reads), self._sp()))
                self.t_rfl = torch.from_numpy((fq & 1).astype(np.uint8)).to(self.dev)
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
        # correctness of what is timed: the round trip restores every base and (lossless) every quality of every read
        self.decode()
        self.stream.synchronize()
        assert int(self.dd_st.abs().sum()) == 0 and int(self.o_stat.abs().sum()) == 0
        assert torch.equal(self.o_bases, self.t_bases), 'round trip mismatch (bases)'
        if lossless:
            assert int(self.o_star.sum()) == 0 and torch.equal(self.o_quals, self.t_quals), 'round trip mismatch (qualities)'
        # the decoded record-field series are the packed ones
        for k in range(2, NB):
            a = int(self.in_addr[k]) - int(self.HS.streams); ln = int(self.in_len[k])
            want = self.dv.dev_to_numpy(self.ctx, int(self.HS.streams) + a, ln, np.uint8)
            got = self.dec_out[int(d_off[k]):int(d_off[k]) + ln].cpu().numpy()
            assert np.array_equal(got, want), 'record-field series round trip mismatch'
        # slice headers: record / base counts of every slice, and the digests of the first distinct containers against zlib
        hdr = np.frombuffer(self.slice_hdr.cpu().numpy().tobytes(), dtype=np.dtype([('seq', '<i4'), ('start', '<i4'), ('span', '<i4'), ('nrec', '<u4'),
                                                                                   ('bases', '<u8'), ('bd', '<u4'), ('sd', '<u4'), ('md5', 'V16')]))
        assert (hdr['nrec'] == READS_PER_CONTAINER).all() and (hdr['bases'] == READS_PER_CONTAINER * READ_LEN).all() and (hdr['seq'] == 0).all()
        assert (hdr['start'] == self.slice_start).all(), 'slice alignment start'
        import zlib
        urd, nu_ = self.uniq_reads
        for c in range(min(2, n_cont)):
            u = c % nu_
            lo = u * READS_PER_CONTAINER * READ_LEN
            b = urd['bases'][lo:lo + READS_PER_CONTAINER * READ_LEN].reshape(READS_PER_CONTAINER, READ_LEN)
            assert int(hdr['bd'][c]) == sum(zlib.crc32(x.tobytes()) for x in b) & 0xFFFFFFFF, 'BD digest'
            if lossless:
                q = urd['quals'][lo:lo + READS_PER_CONTAINER * READ_LEN].reshape(READS_PER_CONTAINER, READ_LEN)
                assert int(hdr['sd'][c]) == sum(zlib.crc32(x.tobytes()) for x in q) & 0xFFFFFFFF, 'SD digest'
            import hashlib
            s0, sp = int(hdr['start'][c]), int(hdr['span'][c])
            refb = self.ref_unique.tobytes()
            assert bytes(hdr['md5'][c]) == hashlib.md5(refb[s0 - 1:s0 - 1 + min(sp, len(refb) - s0 + 1)]).digest(), 'slice reference MD5'
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
        if self.d1_event is not None:
            self.d1_event.record(self.stream)
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

    def timed(self, steps, warmup):
        """`steps` unpipelined steps on this lane's stream, device-timed: (ms per step, isolated kernel times of one step)"""
        torch = self.torch
        self.steps(warmup)
        self.stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        self.steps(steps)
        e1.record(self.stream)
        self.stream.synchronize()
        ms = e0.elapsed_time(e1) / steps
        self.ctx.profile_enable(True)
        self.steps(1)
        self.stream.synchronize()
        kt = {}
        for name, m in self.ctx.profile_read():
            kt[name] = kt.get(name, 0.0) + m
        self.ctx.profile_enable(False)
        return ms, kt

    def close(self):
        self.ctx.close()


def _itf8_bytes(v):
    v &= 0xFFFFFFFF
    if v < 1 << 7: return bytes([v])
    if v < 1 << 14: return bytes([(v >> 8) | 0x80, v & 0xFF])
    if v < 1 << 21: return bytes([(v >> 16) | 0xC0, (v >> 8) & 0xFF, v & 0xFF])
    if v < 1 << 28: return bytes([(v >> 24) | 0xE0, (v >> 16) & 0xFF, (v >> 8) & 0xFF, v & 0xFF])
    return bytes([((v >> 28) | 0xF0) & 0xFF, (v >> 20) & 0xFF, (v >> 12) & 0xFF, (v >> 4) & 0xFF, v & 0xFF])


def block_framing_leg(ln):
    """SURVEY 8f-5, block part: the compressed blocks of one encode pass, wrapped as CRAM v3 blocks (header, data, CRC32) on the
    device -- Block.write [Block.java:272-294].  Verified against zlib on a few blocks; CUDA-event timed."""
    import zlib
    torch, rec, ctx = ln.torch, ln.rec, ln.ctx
    vp = C.c_void_p
    ln.encode()
    ln.stream.synchronize()
    nblk = ln.nblk
    cids = np.tile(np.array([rec.SERIES_CONTENT_ID['QS'], rec.SERIES_CONTENT_ID['BA']] + [rec.HOST_SERIES_CONTENT_ID[nm] for nm, _ in EXTRA], np.int32), ln.n_cont)
    method = np.full(nblk, 4, np.uint8); ctype = np.full(nblk, 4, np.uint8)          # RANS, EXTERNAL
    with torch.cuda.stream(ln.stream):
        t_raw = torch.from_numpy(ln.in_len.view(np.uint8).copy()).to(ln.dev)
    out, olen, ooff = vp(), C.c_uint64(), vp()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(3):
        ln.stream.synchronize()
        e0.record(ln.stream)
        ctx.check(ctx.lib.cramgpu_blocks_frame_dev(ctx.handle, nblk, vp(ln.enc_out.data_ptr()), vp(ln.e_off.data_ptr()), vp(ln.e_len.data_ptr()),
                                                   vp(t_raw.data_ptr()), cids.ctypes.data_as(C.POINTER(C.c_int32)),
                                                   method.ctypes.data_as(C.POINTER(C.c_uint8)), ctype.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                   C.byref(out), C.byref(olen), C.byref(ooff), ln._sp()))
        e1.record(ln.stream)
        ln.stream.synchronize()
        best = min(best, e0.elapsed_time(e1))
    off = ln.dv.dev_to_numpy(ctx, ooff.value, nblk + 1, np.uint64)
    c_off = ln.e_off.cpu().numpy().view(np.uint64); c_len = ln.e_len.cpu().numpy().view(np.uint32)
    for i in (0, 1, 2, nblk - 1):
        comp = ln.enc_out[int(c_off[i]):int(c_off[i]) + int(c_len[i])].cpu().numpy().tobytes()
        want = bytes([4, 4]) + _itf8_bytes(int(cids[i])) + _itf8_bytes(len(comp)) + _itf8_bytes(int(ln.in_len[i])) + comp
        want += int(zlib.crc32(want) & 0xFFFFFFFF).to_bytes(4, 'little')
        got = ln.dv.dev_to_numpy(ctx, out.value + int(off[i]), len(want), np.uint8).tobytes()
        assert got == want, 'framed block %d' % i
    return {'row': 'Block.write :272-294 for every rANS block of the step (method, content type, ITF8 id / sizes, data, CRC32)',
            'blocks': int(nblk), 'bytes_out': int(olen.value), 'ms': round(best, 4), 'GBps': olen.value / (best * 1e-3) / 1e9,
            'verified': 'blocks 0, 1, 2 and the last against zlib.crc32', 'in_default_timed_step': False}


def container_assembly_leg(ln):
    """SURVEY 8f-5, container part: whole CRAM v3 containers (container header, compression header block, slice header block with
    the digest tags, empty core block, the rANS blocks of the step) laid out on the device -- ContainerIO.writeContainer
    [ContainerIO.java:152-209].  Verified structurally on the first and the last container (header CRC32, size field, block
    count, every block's CRC32, the walk ends at the container's end); CUDA-event timed."""
    import zlib
    torch, rec, ctx = ln.torch, ln.rec, ln.ctx
    vp = C.c_void_p
    ln.with_digests = True
    ln.encode()                                        # fills ln.slice_hdr (device) beside the coded blocks
    ln.with_digests = False
    ln.stream.synchronize()
    n_cont, NB = ln.n_cont, ln.NB
    ids = [rec.SERIES_CONTENT_ID['QS'], rec.SERIES_CONTENT_ID['BA']] + [rec.HOST_SERIES_CONTENT_ID[nm] for nm, _ in EXTRA]
    cids = np.tile(np.array([0] + ids, np.int32), n_cont)                    # per container: the (empty) core block, then the series
    method = np.tile(np.array([0] + [4] * NB, np.uint8), n_cont); ctype = np.tile(np.array([5] + [4] * NB, np.uint8), n_cont)
    with torch.cuda.stream(ln.stream):
        z64 = torch.zeros(n_cont, 1, dtype=torch.int64, device=ln.dev); z32 = torch.zeros(n_cont, 1, dtype=torch.int32, device=ln.dev)
        t_off = torch.cat([z64, ln.e_off[:-1].view(n_cont, NB)], dim=1).contiguous()
        t_len = torch.cat([z32, ln.e_len.view(n_cont, NB)], dim=1).contiguous()
        t_raw = torch.cat([z32, torch.from_numpy(ln.in_len.view(np.int32).reshape(n_cont, NB).copy()).to(ln.dev)], dim=1).contiguous()
    comp_header = bytes(range(170))                    # stands for CompressionHeader.toByteArray(): opaque to the device
    grc = (np.arange(n_cont, dtype=np.uint64) * np.uint64(10000))
    bf = (np.arange(n_cont + 1, dtype=np.uint32) * np.uint32(NB + 1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(3):
        ln.stream.synchronize()
        e0.record(ln.stream)
        out, olen, ooff = rec.containers_assemble(ln.slice_hdr.data_ptr(), grc, [comp_header] * n_cont, bf, ln.enc_out.data_ptr(), t_off.data_ptr(),
                                                  t_len.data_ptr(), t_raw.data_ptr(), cids, method, ctype, tags=True, ctx=ctx, stream=ln._sp())
        e1.record(ln.stream)
        ln.stream.synchronize()
        best = min(best, e0.elapsed_time(e1))
    off = ln.dv.dev_to_numpy(ctx, ooff, n_cont + 1, np.uint64)

    def itf8(b, o):
        v = b[o]
        if v < 0x80: return v, o + 1
        if v < 0xC0: return ((v & 0x7F) << 8) | b[o + 1], o + 2
        if v < 0xE0: return ((v & 0x3F) << 16) | (b[o + 1] << 8) | b[o + 2], o + 3
        if v < 0xF0: return ((v & 0x1F) << 24) | (b[o + 1] << 16) | (b[o + 2] << 8) | b[o + 3], o + 4
        return ((v & 0x0F) << 28) | (b[o + 1] << 20) | (b[o + 2] << 12) | (b[o + 3] << 4) | (b[o + 4] & 0x0F), o + 5

    def ltf8(b, o):
        n = 0
        while n < 8 and b[o] & (0x80 >> n): n += 1
        v = b[o] & (0xFF >> (n + 1)) if n < 8 else 0
        for k in range(n): v = (v << 8) | b[o + 1 + k]
        return v, o + 1 + n

    for c in (0, n_cont - 1):
        b = ln.dv.dev_to_numpy(ctx, out + int(off[c]), int(off[c + 1] - off[c]), np.uint8).tobytes()
        size = int.from_bytes(b[:4], 'little'); o = 4
        for _ in range(4): _, o = itf8(b, o)
        g, o = ltf8(b, o); _, o = ltf8(b, o)
        nb, o = itf8(b, o); nl, o = itf8(b, o); lm, o = itf8(b, o)
        assert g == int(grc[c]) and nb == NB + 3 and nl == 1, 'container header fields'
        assert int.from_bytes(b[o:o + 4], 'little') == zlib.crc32(b[:o]) & 0xFFFFFFFF, 'container header CRC32'
        o += 4
        assert size == len(b) - o, 'container byte size'
        first = o
        for k in range(nb):
            s0 = o; o += 2
            _, o = itf8(b, o); cl, o = itf8(b, o); _, o = itf8(b, o)
            o += cl
            assert int.from_bytes(b[o:o + 4], 'little') == zlib.crc32(b[s0:o]) & 0xFFFFFFFF, 'block %d CRC32' % k
            o += 4
            if k == 0: assert o - first == lm, 'landmark'
        assert o == len(b), 'the blocks fill the container'
    return {'row': 'ContainerIO.writeContainer :152-209 for every container of the step (header + CRC32, compression header block, slice header block with digest tags, core, rANS blocks)',
            'containers': int(n_cont), 'blocks': int(n_cont * (NB + 3)), 'bytes_out': int(olen), 'ms': round(best, 4), 'GBps': olen / (best * 1e-3) / 1e9,
            'verified': 'first and last container: header fields and CRC32, byte size, landmark, the CRC32 of every block, exact length', 'in_default_timed_step': False}


def bam_records_leg(ln):
    """the decoded records of the step as BAM record bytes, on the device -- CIGAR rebuilt from the unpacked read
    features (Cram2SamRecordFactory.getCigar2), the decoder's pairing restoration (CramNormalizer.normalize), the record layout of
    BAMRecordView -- and back again as Bam_OBA_Supplier's stripes are read (cramgpu_bam_unpack_dev).  Verified: the rebuilt CIGARs
    are the input CIGARs where those hold only M / I / D / S (= / X come back as M by design), and unpacking the packed records
    returns the very bases, qualities, positions and flags that went in; CUDA-event timed."""
    torch, rec, ctx = ln.torch, ln.rec, ln.ctx
    vp = C.c_void_p
    ln.encode(); ln.decode()
    ln.stream.synchronize()
    n = ln.n_reads
    U = ln.U
    dev = ln.dev
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    z32 = torch.zeros(n, dtype=torch.int32, device=dev); m1 = torch.full((n,), -1, dtype=torch.int32, device=dev)
    cf = torch.full((n,), 3, dtype=torch.uint8, device=dev); mf = torch.zeros(n, dtype=torch.uint8, device=dev)       # single-end reads: detached
    nf = torch.full((n,), -1, dtype=torch.int32, device=dev); st_c = torch.zeros(ln.n_cont, dtype=torch.int32, device=dev)
    st_r = torch.zeros(n, dtype=torch.int32, device=dev); st_u = torch.zeros(n, dtype=torch.int32, device=dev)
    pos = ln.t_pos.clone()
    cont_off = (np.arange(ln.n_cont + 1, dtype=np.uint64) * np.uint64(READS_PER_CONTAINER))
    best = None
    kern = {}
    for rep in range(3):
        ln.stream.synchronize()
        if rep == 2:
            ctx.profile_enable(True)                    # a third pass for the kernels' own durations
        ev[0].record(ln.stream)
        d_cig, d_coff, cig_len = vp(), vp(), C.c_uint64()
        ctx.check(ctx.lib.cramgpu_cigar_rebuild_dev(ctx.handle, n, vp(ln.t_flag.data_ptr()), vp(ln.t_roff.data_ptr()), U.feat_off, U.f_op, U.f_pos,
                                                    U.f_val, U.f_aux, C.byref(d_cig), C.byref(d_coff), C.byref(cig_len), ln._sp()))
        ev[1].record(ln.stream)
        M = rec.MateRestore()
        M.n_reads = n; M.n_containers = ln.n_cont; M.cont_off = cont_off.ctypes.data_as(C.POINTER(C.c_uint64))
        M.flag = ln.t_flag.data_ptr(); M.cf = cf.data_ptr(); M.mate_flags = mf.data_ptr(); M.nf = nf.data_ptr(); M.ref_id = z32.data_ptr()
        M.pos = pos.data_ptr(); M.mate_ref_id = m1.data_ptr(); M.mate_pos = z32.data_ptr(); M.tlen = z32.data_ptr()
        M.read_off = ln.t_roff.data_ptr(); M.feat_off = U.feat_off; M.f_op = U.f_op; M.f_val = U.f_val; M.f_aux = U.f_aux
        ctx.check(ctx.lib.cramgpu_mate_restore_dev(ctx.handle, C.byref(M), vp(st_c.data_ptr()), ln._sp()))
        ev[2].record(ln.stream)
        B = rec.BamRecords()
        B.n_reads = n; B.ref_id = z32.data_ptr(); B.pos = pos.data_ptr(); B.mate_ref_id = m1.data_ptr(); B.mate_pos = z32.data_ptr(); B.tlen = z32.data_ptr()
        B.flag = ln.t_flag.data_ptr(); B.bin = None; B.mapq = ln.d_mapq.value if isinstance(ln.d_mapq, C.c_void_p) else ln.d_mapq; B.rflags = None
        B.names = ln.t_names.data_ptr(); B.name_off = ln.t_noff.data_ptr(); B.cigar = d_cig.value; B.cigar_off = d_coff.value
        B.bases = ln.o_bases.data_ptr(); B.quals = ln.o_quals.data_ptr(); B.read_off = ln.t_roff.data_ptr(); B.tags = None; B.tag_off = None
        B.names_len = ln.t_names.numel(); B.cigar_len = cig_len.value; B.bases_len = ln.o_bases.numel(); B.tags_len = 0
        d_bam, bam_len, d_roff = vp(), C.c_uint64(), vp()
        ctx.check(ctx.lib.cramgpu_bam_pack_dev(ctx.handle, C.byref(B), C.byref(d_bam), C.byref(bam_len), C.byref(d_roff), vp(st_r.data_ptr()), ln._sp()))
        ev[3].record(ln.stream)
        Ub = rec.BamRecords()
        ctx.check(ctx.lib.cramgpu_bam_unpack_dev(ctx.handle, d_bam, d_roff, n, C.byref(Ub), vp(st_u.data_ptr()), ln._sp()))
        ev[4].record(ln.stream)
        ln.stream.synchronize()
        if rep == 2:
            for name, m in ctx.profile_read():
                kern[name] = round(kern.get(name, 0.0) + m, 4)
            ctx.profile_enable(False)
            break
        t = [ev[i].elapsed_time(ev[i + 1]) for i in range(4)]
        best = t if best is None else [min(a, b) for a, b in zip(best, t)]
    assert int(st_c.abs().sum().item()) == 0 and int(st_r.abs().sum().item()) == 0 and int(st_u.abs().sum().item()) == 0, 'BAM leg: status'
    # round trip: what was packed comes back
    nb = ln.o_bases.numel()
    got_b = ln.dv.dev_to_numpy(ctx, Ub.bases, nb, np.uint8); got_q = ln.dv.dev_to_numpy(ctx, Ub.quals, nb, np.uint8)
    assert np.array_equal(got_b, ln.o_bases.cpu().numpy()) and np.array_equal(got_q, ln.o_quals.cpu().numpy()), 'BAM leg: bases / qualities'
    assert np.array_equal(ln.dv.dev_to_numpy(ctx, Ub.pos, n, np.int32), ln.t_pos.cpu().numpy().view(np.int32)), 'BAM leg: positions'
    assert np.array_equal(ln.dv.dev_to_numpy(ctx, Ub.flag, n, np.uint16), ln.t_flag.cpu().numpy().view(np.uint16)), 'BAM leg: flags'
    # the rebuilt CIGARs against the input ones (the simulated reads use M I D S only)
    cg = ln.dv.dev_to_numpy(ctx, d_cig.value, int(cig_len.value), np.uint32)
    want = ln.t_cigar.cpu().numpy().view(np.uint32)
    assert cg.size == want.size and np.array_equal(cg, want), 'BAM leg: rebuilt CIGARs'
    total = int(bam_len.value)
    return {'row': 'Cram2SamRecordFactory.getCigar2 :119-214 + CramNormalizer.normalize pairing :67-102 + BAMRecordView :70-241 (pack) + Bam_OBA_Supplier :45-82 (unpack) for every record of the step',
            'reads': int(n), 'bam_bytes': total, 'ms': {'cigar_rebuild': round(best[0], 4), 'mate_restore': round(best[1], 4), 'bam_pack': round(best[2], 4), 'bam_unpack': round(best[3], 4)},
            'kernels_ms': kern, 'bam_pack_GBps': total / (best[2] * 1e-3) / 1e9, 'bam_unpack_GBps': total / (best[3] * 1e-3) / 1e9,
            'verified': 'rebuilt CIGARs equal the input CIGARs; packed then unpacked records return the restored bases, qualities, positions, flags', 'in_default_timed_step': False}


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
    ref, rd = unique_containers(nu, 1000 + 17 * rank)
    # Software pipeline over consecutive steps: `depth` contexts, each with its own stream and the whole batch; step i runs
    # on context i % depth.  The order-1 decoder is a recurrence per rANS state -- a few warps per SM for tens of
    # milliseconds, whatever the batch size -- so the encode of step i+1 (record kernels, histograms, encoder) runs beside
    # the decode of step i instead of after it, the way consecutive batches of a whole-genome stream would be fed.
    # Two host locks keep at most one context in each phase.
    n_lanes = depth = max(1, args.pipeline)
    groups = [dist.new_group() if world > 1 else None for _ in range(depth)]
    lanes = [Lane(i, n_cont, ref, rd, nu, dev, rank, world, groups[i], not args.no_unpack) for i in range(depth)]
    for ln in lanes:
        ln.prepare()
    # Two schedules that were measured and are NOT the default ("pipeline arithmetic"): CRAMGPU_SERIAL_CODERS=1
    # makes the encoder call of step j + 1 wait for the order-1 decode of step j (k_dec1 then runs at its isolated 39.8 ms, 0.0147
    # of the roofline, but hist / tables / pack and the call's host synchronisation no longer hide beside it: 78.9 ms per step
    # against 72.5); CRAMGPU_STAGGER=1 (three contexts) makes the record kernels of step j + 2 wait for that decode too (73.0 ms).
    # CRAMGPU_TRACE_STEPS=1 prints when each phase of each step ended on the device.
    stagger = depth >= 3 and bool(os.environ.get('CRAMGPU_STAGGER'))
    serial_coders = depth >= 2 and bool(os.environ.get('CRAMGPU_SERIAL_CODERS'))
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
        trace = bool(os.environ.get('CRAMGPU_TRACE_STEPS'))
        d1_done = [torch.cuda.Event(enable_timing=trace) for _ in range(k)]      # order-1 decode of step j has ended (device)
        d1_sent = [threading.Event() for _ in range(k)]              # ... and its event has been recorded (host)
        enc_sent = [threading.Event() for _ in range(k)]             # the encode calls of step j have returned (host)
        tr = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(k)] if trace else None

        def work(ln, i):
            try:
                torch.cuda.set_device(local)
                for j in range(i, k, depth):
                    ln.wait_event = None
                    if stagger and j >= 2:
                        d1_sent[j - 2].wait()
                        ln.wait_event = d1_done[j - 2]
                    ln.d1_event = d1_done[j]
                    ln.before_rans_encode = None
                    if serial_coders and j >= 1:
                        def gate(ln=ln, j=j):
                            d1_sent[j - 1].wait()
                            ln.stream.wait_event(d1_done[j - 1])
                            if trace:
                                tr[j][1].record(ln.stream)
                        ln.before_rans_encode = gate
                    if (serial_coders or stagger) and j >= 1:
                        enc_sent[j - 1].wait()                       # these schedules need the steps to encode in order
                    with enc_lock:
                        if trace:
                            tr[j][0].record(ln.stream)
                        ln.encode()
                        if trace:
                            tr[j][2].record(ln.stream)
                    enc_sent[j].set()
                    with dec_lock:
                        ln.decode()
                        if trace:
                            tr[j][3].record(ln.stream)
                    d1_sent[j].set()
            except Exception as e:      # noqa: BLE001
                errs.append(e)
                for ev in d1_sent + enc_sent:
                    ev.set()
        th = [threading.Thread(target=work, args=(ln, i)) for i, ln in enumerate(lanes)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]
        if trace and k >= 4:
            torch.cuda.synchronize()
            z = tr[0][0]
            for j in range(k):
                g = z.elapsed_time(tr[j][1]) if j >= 1 and serial_coders else float('nan')
                sys.stderr.write('step %d lane %d: pre %.1f gate %.1f enc_end %.1f d1_end %.1f dec_end %.1f\n' % (
                    j, j % depth, z.elapsed_time(tr[j][0]), g, z.elapsed_time(tr[j][2]), z.elapsed_time(d1_done[j]), z.elapsed_time(tr[j][3])))

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

    def max_over_ranks(x):
        t_ = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t_, op=dist.ReduceOp.MAX)
        return float(t_.item())

    # ---- what the pipelined spans cannot say: one context alone, phases back to back -> true kernel durations
    for ln in lanes[1:]:
        ln.close()
    ms_unpip, kiso = lanes[0].timed(2, 1)
    ms_unpip = max_over_ranks(ms_unpip)
    # ---- SURVEY 8f-3, the next row: slice header fields + BD / SD content digests as part of the step (side stream)
    lanes[0].with_digests = True
    ms_dig, kdig = lanes[0].timed(2, 1)
    lanes[0].with_digests = False
    ms_dig = max_over_ranks(ms_dig)
    framing_leg = block_framing_leg(lanes[0])
    container_leg = container_assembly_leg(lanes[0])
    bam_leg = bam_records_leg(lanes[0]) if lanes[0].use_unpack else None
    digest_leg = {'row': 'ContainerFactory.buildSlice :138-176 (nofRecords, bases, sequenceId, alignment start/span, BD, SD) + Slice.setRefMD5',
                  'ms_per_step_unpipelined_with_digests': ms_dig, 'ms_per_step_unpipelined_without': ms_unpip,
                  'k_slice_digest_ms': round(kdig.get('k_slice_digest', 0.0), 4), 'k_slice_finish_ms': round(kdig.get('k_slice_finish', 0.0), 4), 'k_slice_md5_ms': round(kdig.get('k_slice_md5', 0.0), 4),
                  'bytes_read_per_step': 2.0 * n_reads * READ_LEN,
                  'GBps': 2.0 * n_reads * READ_LEN / max(kdig.get('k_slice_digest', 1e9), 1e-9) / 1e6,
                  'verified': 'record / base counts and alignment start of every slice; BD and SD of the first two slices against zlib.crc32, their reference MD5 against hashlib',
                  'in_default_timed_step': False}

    # ---- the other configurations north_star names, each a short unpipelined run of the same step
    def leg(nc, spec, steps=2, warmup=1):
        ln_ = Lane(0, nc, ref, rd, nu, dev, rank, world, groups[0], not args.no_unpack, spec=spec)
        ln_.prepare()
        m_, k_ = ln_.timed(steps, warmup)
        ln_.close()
        del ln_
        torch.cuda.empty_cache()
        m_ = max_over_ranks(m_)
        return {'containers_per_gpu': nc, 'ms_per_step': m_, 'reads_per_s': nc * READS_PER_CONTAINER * world / (m_ * 1e-3),
                'kernels_ms': {k: round(v, 3) for k, v in k_.items() if v >= 0.05}}
    extra_legs = {}
    if not args.no_configs:
        # configs[3]: the lossy quality models on the same workload
        extra_legs['config4_quality_models'] = {sp: leg(n_cont, sp) for sp in ('*8', 'm5', 'N40-D8')}
        extra_legs['config4_quality_models']['lossless'] = {'containers_per_gpu': n_cont, 'ms_per_step': ms_unpip,
                                                            'reads_per_s': n_reads * world / (ms_unpip * 1e-3)}
        # configs[0]: 200 000 reads = 20 containers in all: one coder block's latency whatever the GPU count
        nc1 = max(1, -(-20 // world))
        extra_legs['config1_20_containers'] = leg(nc1, None, steps=3, warmup=2)
        # configs[2] at its stated FIXED size: 2 048 containers in all, 2048 / N per GPU (strong scaling); at N = 1 that is the main run
        if world > 1 and n_cont == 2048:
            st_ = leg(max(1, 2048 // world), None, steps=3, warmup=2)
            st_['speedup_vs_one_gpu_unpipelined'] = ms_unpip / st_['ms_per_step']
            st_['limiter'] = 'single-block coder latency: a call takes one quality block\'s chain (375 000 dependent steps per rANS state) however few blocks a GPU holds'
            extra_legs['config3_fixed_total_2048_containers'] = st_
        else:
            extra_legs['config3_fixed_total_2048_containers'] = {'containers_per_gpu': n_cont, 'ms_per_step': ms_unpip,
                                                                 'reads_per_s': n_reads * world / (ms_unpip * 1e-3),
                                                                 'note': 'N = 1: the fixed-size configuration is the main run (unpipelined figure)'}

    # ---- end-to-end through the host-buffer C ABI (bounded to keep pinned host memory modest at 8 ranks)
    # every rank page-locks its own buffers (~21 MB per container with four contexts): keep all ranks of the node together
    # under 40 % of the free host memory
    n_e2e = min(n_cont, args.e2e_containers)
    try:
        import psutil
        per_rank = 0.4 * psutil.virtual_memory().available / max(world, 1)
        n_e2e = min(n_cont, int(max(128, min(n_e2e, per_rank // (21 << 20)))))
    except Exception:      # noqa: BLE001
        n_e2e = min(n_e2e, max(128, 2 * args.e2e_containers // max(world, 2))) if world > 1 else n_e2e
    if world > 1:
        t_n = torch.tensor([n_e2e], dtype=torch.int64, device=dev)
        dist.all_reduce(t_n, op=dist.ReduceOp.MIN)
        n_e2e = int(t_n.item())
    st0 = lanes[0]
    stats = dict(raw_bytes=st0.raw_bytes, comp_bytes=st0.comp_bytes, raw_o0=st0.raw_o0, qs_bytes=st0.qs_bytes, nblk=st0.nblk)
    stats['comp_qs'] = float(st0.c_len.reshape(st0.n_cont, st0.NB)[:, 0].sum())
    comp_hint = st0.c_len[:n_e2e * st0.NB].copy()
    # the device-resident leg is over: its buffers and contexts go before the end-to-end leg allocates its own
    lanes[0].close()
    del lanes, st0, ln
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    # contexts of the end-to-end leg: 8 on one GPU (five calls per step, each holding the host for 100+ ms: more contexts than
    # calls keep every stage busy); fewer per GPU when several ranks share the host's cores and page-locked memory
    e2e_depth = args.e2e_pipeline if args.e2e_pipeline > 0 else (8 if world == 1 else 6 if world == 2 else 4)
    e2e = None if args.no_e2e else run_e2e(args, ref, rd, nu, n_e2e, world, barrier, e2e_depth, comp_hint, local)

    if rank == 0:
        value = n_reads * world / (ms * 1e-3)
        raw_bytes = stats['raw_bytes']; comp_bytes = stats['comp_bytes']
        o1_raw = raw_bytes - stats['raw_o0']
        comp1 = comp_bytes * (o1_raw / raw_bytes)
        # k_dec1 decodes the long blocks (the quality series) when the short ones ride beside it in k_dec1s
        dec1_bytes = (stats['qs_bytes'] + stats['comp_qs']) if 'k_dec1s' in kavg else (o1_raw + comp1)
        algo = {'k_enc1': o1_raw + comp1, 'k_dec1': dec1_bytes, 'k_hist1': o1_raw, 'k_hist0': raw_bytes,
                'k_rec_simple_count': n_reads * (2 * READ_LEN + 40.0), 'k_rec_simple_write': n_reads * (2 * READ_LEN + 60.0),
                'k_norm_simple': n_reads * (4 * READ_LEN + 40.0), 'k_pack': 2 * comp_bytes}
        ab = algo.get(top)
        # kavg sums the lanes' launches, which overlap in time; the per-launch figure is what the roofline wants
        per_launch_ms = kavg[top] / max(kcount[top], 1)
        achieved = (ab / max(kcount[top], 1)) / (per_launch_ms * 1e-3) / 1e9 if ab else None
        peak = float(peaks['hbm_gbs'])
        # rANS GB/s from the ISOLATED kernel durations (one context, phases back to back); k_dec1s / k_dec0 / k_dec1_big run
        # beside k_dec1, their spans lie inside its
        rans_ms_enc = sum(v for k, v in kiso.items() if k in ('k_hist0', 'k_prep_o1', 'k_hist1', 'k_tables0', 'k_tables1', 'k_enc0', 'k_enc1', 'k_scan_out', 'k_pack'))
        rans_ms_dec = sum(v for k, v in kiso.items() if k in ('k_dec_tables', 'k_dec_plan', 'k_dec1'))
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
            'kernels_ms': {k: round(v, 4) for k, v in kiso.items()},
            'kernels_ms_note': 'durations of one unpipelined step on one context (CUDA events around every launch); kernels that '
                               'run beside k_dec1 on another stream (k_dec1s, k_dec1_big, k_dec0) include their wait for SM resources',
            'ms_per_step_unpipelined': ms_unpip,
            'pipelined_spans_ms': {k: round(v, 4) for k, v in kavg.items()}, 'kernel_launches_per_step': {k: v for k, v in kcount.items()},
            'gpu_launches': int(launches),
            'roofline': {'bound': 'hbm', 'kernel': top, 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                         'frac': achieved / peak if achieved else None, 'traffic': ncu_traffic(top, n_cont), 'peak_kind': peak_kind,
                         'algorithmic_bytes_per_launch': (ab / max(kcount[top], 1)) if ab else None,
                         'note': 'per launch; traffic = dram bytes read + written by one launch in the ncu --set full capture of '
                                 'profiles/ (null at other sizes); rANS is a recurrence per state: attainable rate = resident states x '
                                 'clock / cycles per step (DESIGN.md, "Why the coder cannot reach the HBM roofline")'},
            'e2e': e2e, 'clocks': clocks,
        }
        line['slice_digests'] = digest_leg
        line['block_framing'] = framing_leg
        line['container_assembly'] = container_leg
        if bam_leg: line['bam_records'] = bam_leg
        line.update(extra_legs)
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
        self.HS, self.HD = rec.HostSeries(), rec.HostSeries()
        self.H = rec.HostFields()
        C.memmove(C.byref(self.H), C.byref(inp['H']), C.sizeof(rec.HostFields))       # force_qs is set per step
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
        self.ba_slot = dslot[np.arange(n_cont) * NB + 1].astype(np.int64)
        self.rfl = np.ones(n_reads, np.uint8)
        # series table of the decode side (row P on the device): absolute addresses -- QS / BA where the block codec left them
        # in device memory, the series the Java host inflates (FN FC FP BS DL SC RS HC PD MQ) in host memory
        self.s_addr = np.zeros((n_cont, 12), np.uint64); self.s_len = np.zeros((n_cont, 12), np.uint64)
        SR = self.SR = rec.Series()
        SR.n_reads = n_reads; SR.n_containers = n_cont; SR.cont_off = inp['cont_off'].ctypes.data_as(u64p); SR.streams = None; SR.streams_len = 0
        SR.stream_off = self.s_addr.ctypes.data_as(u64p); SR.stream_len = self.s_len.ctypes.data_as(u64p)
        SR.read_off = vp(inp['a']['read_off'].ctypes.data); SR.rflags = vp(self.rfl.ctypes.data)
        self.U = rec.Features(); self.u_status = np.zeros(n_cont, np.int32); self.u_mapq = np.zeros(n_reads, np.uint8)
        self.o_b = pinned_empty(n_reads * READ_LEN, np.uint8); self.o_q = pinned_empty(n_reads * READ_LEN, np.uint8)
        self.o_s = np.empty(n_reads, np.uint8); self.o_t = np.empty(n_reads, np.int32)
        N = self.N = rec.NormIn()
        N.n_reads = n_reads; N.n_containers = n_cont
        N.cont_off = inp['cont_off'].ctypes.data_as(u64p); N.win_off = inp['win_off'].ctypes.data_as(u64p); N.win_len = inp['win_len'].ctypes.data_as(u64p)
        N.win_start0 = inp['win_s0'].ctypes.data_as(C.POINTER(C.c_int64)); N.ref = vp(inp['d_ref'].data_ptr()); N.ref_total = inp['d_ref'].numel()
        N.pos = vp(inp['a']['pos'].ctypes.data); N.read_off = vp(inp['a']['read_off'].ctypes.data); N.rflags = vp(self.rfl.ctypes.data)
        N.ba = None; N.qs = None                        # absolute device addresses in the unpacked read_qs_off / read_ba_off
        self.ms = {'features_encode_resident': 0.0, 'host_series_pack_resident': 0.0, 'rans_encode_batch': 0.0, 'rans_decode_batch': 0.0,
                   'features_unpack_resident': 0.0, 'normalize_decode': 0.0, 'host_glue': 0.0}
        self.h2d = self.d2h = 0
        self.stages = (self.stage_features, self.stage_host_series, self.stage_rans_encode, self.stage_rans_decode, self.stage_unpack,
                       self.stage_normalize)

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
        # the feature arrays stay on the device: the host hands the series to the block layer and never looks at them
        ctx.check(ctx.lib.cramgpu_features_encode_resident(ctx.handle, C.byref(inp['R']), None, -1, inp['host_series'] | self.rec.FEATURES_ON_DEVICE_ONLY,
                                                           C.byref(F), C.byref(D)))
        t1 = time.perf_counter()
        self.ms['features_encode_resident'] += (t1 - t0) * 1e3

    def stage_host_series(self):
        """the record fields in front of the features: the rANS-coded series stay on the device, the gzip ones come back
        (RN is not sent at all: packing names is a plain copy the Java host does itself)"""
        ctx, inp, D = self.ctx, self.inp, self.D
        t1 = time.perf_counter()
        self.H.force_qs = D.force_qs
        ctx.check(ctx.lib.cramgpu_host_series_pack_resident(ctx.handle, C.byref(self.H), inp['hs_mask'], C.byref(self.HS), C.byref(self.HD)))
        t2 = time.perf_counter()
        self.ms['host_series_pack_resident'] += (t2 - t1) * 1e3

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
        NHS = len(rec.HOST_SERIES)
        hoff = np.ctypeslib.as_array(C.cast(self.HS.stream_off, u64p), (n_cont, NHS)); hlen = np.ctypeslib.as_array(C.cast(self.HS.stream_len, u64p), (n_cont, NHS))
        v[:, 2:] = np.uint64(self.HD.streams) + hoff[:, inp['KX']]; w[:, 2:] = hlen[:, inp['KX']]
        self.hs_down = int(sum(int(hlen[:, k].sum()) for k in range(NHS) if inp['hs_mask'] & (1 << k)))
        self.ex_bytes = int(w[:, 2:].sum())
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

    def stage_unpack(self):
        """row P: the series as the decode side holds them (small ones in host memory, QS / BA on the device) -> read features,
        left on the device"""
        from cramtools_b200._lib import u8p, i32p
        ctx, inp, F = self.ctx, self.inp, self.F
        t4 = time.perf_counter()
        KQS, KBA = inp['KQS'], inp['KBA']
        sa = self.s_addr
        sa[:] = np.uint64(F.streams) + self.soff                       # F.streams: the host copy of the small series (encode side)
        sa[:, KQS] = np.uint64(self.d_back.data_ptr()) + self.qs_slot.astype(np.uint64)
        sa[:, KBA] = np.uint64(self.d_back.data_ptr()) + self.ba_slot.astype(np.uint64)
        self.s_len[:] = self.slen
        t5 = time.perf_counter()
        ctx.check(ctx.lib.cramgpu_features_unpack_resident(ctx.handle, C.byref(self.SR), C.byref(self.U), self.u_mapq.ctypes.data_as(u8p),
                                                           self.u_status.ctypes.data_as(i32p)))
        self.ms['features_unpack_resident'] += (time.perf_counter() - t5) * 1e3; self.ms['host_glue'] += (t5 - t4) * 1e3

    def stage_normalize(self):
        """records as CramRecordReader would hand them over (positions, lengths; the features and the decoded QS / BA are on the
        device) -> bases and qualities (host)"""
        from cramtools_b200._lib import u8p, u64p, i32p
        ctx, inp, F, N, U = self.ctx, self.inp, self.F, self.N, self.U
        n_reads = inp['n_reads']
        t5 = time.perf_counter()
        N.read_qs_off = U.read_qs_off; N.read_ba_off = U.read_ba_off
        N.matrix = F.matrix; N.feat_off = U.feat_off; N.f_op = U.f_op; N.f_pos = U.f_pos; N.f_val = U.f_val; N.f_aux = U.f_aux
        N.payload = U.payload; N.payload_len = U.payload_len
        ctx.check(ctx.lib.cramgpu_normalize_decode(ctx.handle, C.byref(N), self.o_b.ctypes.data_as(u8p), self.o_q.ctypes.data_as(u8p),
                                                   self.o_s.ctypes.data_as(u8p), self.o_t.ctypes.data_as(i32p)))
        t6 = time.perf_counter()
        self.ms['normalize_decode'] += (t6 - t5) * 1e3
        # bytes that crossed the bus in this step
        a, nb = inp['a'], n_reads * READ_LEN
        KQS, KBA = inp['KQS'], inp['KBA']
        small = int(self.slen.sum() - self.slen[:, KQS].sum() - self.slen[:, KBA].sum())
        ex_bytes = self.ex_bytes
        comp_bytes = int(self.out_len.sum())
        self.h2d = int(2 * nb + a['read_off'].nbytes + a['cigar_off'].nbytes + a['cigar'].nbytes + 7 * n_reads      # reads
                       + (2 + 4 + 8 + 4) * n_reads                                                                        # record fields: flag, pos, read_off, read group
                       + comp_bytes                                                                                       # compressed blocks, decode side
                       + small                                                                                            # the series the host inflated, for row P
                       + (4 + 8 + 1) * n_reads + 8 * (n_reads + 1) + n_reads)                          # decode side: pos, read_off, rflags (twice: unpack, normaliser)
        self.d2h = int(n_reads + 8 * inp['n_cont'] + small                                                                # force_qs, matrix, small series
                       + self.hs_down                                                                                     # record-field series the host gzips
                       + comp_bytes + ex_bytes + 2 * nb + 5 * n_reads + n_reads)                                          # blocks; bases, qualities, star, status; MAPQ


def run_e2e(args, ref, rd, nu, n_cont, world, barrier, depth, comp_hint, device):
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
    reps = (n_cont + nu - 1) // nu
    ref_len1 = int(ref.size)
    # the reference is static: uploaded once, a copy per tile of containers, and handed to the calls as a device pointer
    d_ref = torch.from_numpy(np.ascontiguousarray(ref)).to('cuda').repeat(reps).contiguous()
    keep_pinned.append(d_ref)
    cont_off = (np.arange(n_cont + 1, dtype=np.uint64) * READS_PER_CONTAINER)
    ref_off = (np.arange(n_cont) // nu).astype(np.uint64) * np.uint64(ref_len1); ref_len = np.full(n_cont, ref_len1, np.uint64)
    a = dict(bases=pinned(trd['bases']), quals=pinned(trd['quals']), read_off=pinned(trd['read_off'].astype(np.uint64)),
             cigar=pinned(trd['cigar'].astype(np.uint32)), cigar_off=pinned(trd['cigar_off'].astype(np.uint64)),
             pos=pinned(trd['pos'].astype(np.int32)), flag=pinned(trd['flag'].astype(np.uint16)), mapq=pinned(trd['mapq'].astype(np.uint8)))
    R = rec.Reads()
    R.n_reads = n_reads; R.n_containers = n_cont
    R.cont_off = cont_off.ctypes.data_as(u64p); R.ref_off = ref_off.ctypes.data_as(u64p); R.ref_len = ref_len.ctypes.data_as(u64p)
    R.ref = vp(d_ref.data_ptr()); R.ref_total = d_ref.numel()
    for k, v in a.items():
        setattr(R, k, vp(v.ctypes.data))
    R.rflags = None
    NX = len(EXTRA); NB = 2 + NX
    rg = pinned(trd['read_group'].astype(np.int32))
    p = rd['pos'].astype(np.int64).reshape(nu, READS_PER_CONTAINER)
    ws = (p.min(axis=1) - 1); we = np.minimum(p.max(axis=1) - 1 + READ_LEN + 8, ref_len1)
    KQS, KBA = rec.SERIES.index('QS'), rec.SERIES.index('BA')
    slice_start = np.tile(p.min(axis=1), reps)[:n_cont].astype(np.int32)
    H = rec.HostFields()
    H.n_reads = n_reads; H.n_containers = n_cont
    H.cont_off = cont_off.ctypes.data_as(u64p); H.slice_start = slice_start.ctypes.data_as(C.POINTER(C.c_int32)); H.multi_ref = None
    H.ap_delta = 1; H.names_included = 0
    H.flag = vp(a['flag'].ctypes.data); H.pos = vp(a['pos'].ctypes.data); H.read_off = vp(a['read_off'].ctypes.data); H.read_group = vp(rg.ctypes.data)
    KX = [rec.HOST_SERIES.index(nm) for nm, _ in EXTRA]
    hs_mask = sum(1 << k for k in range(len(rec.HOST_SERIES)) if k not in KX and rec.HOST_SERIES[k] != 'RN')
    inp = dict(n_cont=n_cont, n_reads=n_reads, NB=NB, R=R, a=a, d_ref=d_ref, cont_off=cont_off, keep=[ref_off, ref_len, rg, slice_start],
               H=H, KX=KX, hs_mask=hs_mask,
               order=np.tile(np.array([1, 1] + [o for _, o in EXTRA], np.uint8), n_cont), KQS=KQS, KBA=KBA,
               host_series=sum(1 << k for k in range(len(rec.SERIES)) if k not in (KQS, KBA)), comp_hint=comp_hint,
               win_off=np.tile(ws, reps)[:n_cont].astype(np.uint64) + ref_off, win_len=np.tile(we - ws, reps)[:n_cont].astype(np.uint64),
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
    # one lock per call: at most one context inside each of the six calls, different calls of different steps overlap
    permits = int(os.environ.get('CRAMGPU_E2E_PERMITS', '2'))      # contexts admitted to one call at a time
    locks = [threading.Semaphore(permits) for _ in range(6)]
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
            'path': 'cramgpu_features_encode_resident (feature arrays stay on the device) -> cramgpu_host_series_pack_resident -> '
                    'cramgpu_rans_encode_batch -> cramgpu_rans_decode_batch -> cramgpu_features_unpack_resident -> cramgpu_normalize_decode: '
                    'reads in and compressed blocks + the series the host gzips out; compressed blocks + those series in and restored '
                    'bases/qualities out; every copy inside the timed region; QS / BA and the read features never leave the device'}


# ------------------------------------------------------------------------------------------- configs[1]: block sweep


def run_sweep(args):
    """BASELINE configs[1]: rANS order-0 / order-1 over blocks of 64 KB - 8 MB of QS-, BA-, RN- and ITF8-like bytes, device
    resident, byte-identical to the oracle (first block of every cell), CUDA-event timed (best of 3 calls).  Blocks per cell:
    as many as fill ~1.5 GB, at most 2048.  The ITF8-like stream has > 100 distinct bytes: it takes the large-alphabet
    decoder.  Prints one JSON line; profiles/r2_rans_block_sweep.txt holds a run."""
    import torch
    import cramtools_b200
    from cramtools_b200 import synth, device as dv
    from oracle import rans as orans
    ctx = cramtools_b200.Context(0)
    rng = np.random.default_rng(5)

    def itf8_wide(n, seed):
        # template-size-like values: ITF8 of ints around +-400 and mate positions -> all 256 byte values occur
        r = np.random.default_rng(seed)
        v = np.concatenate([r.normal(400, 60, n // 6).astype(np.int64), -r.normal(400, 60, n // 6).astype(np.int64), r.integers(1, 1 << 27, n // 6)])
        r.shuffle(v)
        return itf8_pack_np(v)[:n]
    gens = {'QS': synth.qs_stream, 'BA': synth.ba_stream, 'RN': synth.rn_stream, 'ITF8': itf8_wide}
    sizes = [64 << 10, 128 << 10, 256 << 10, 512 << 10, 1 << 20, 2 << 20, 4 << 20, 8 << 20]
    rows = []
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    # a stream of our own: the legacy default stream (handle 0) would mean "the context's stream" to the library and the events
    # would then bracket nothing
    side = torch.cuda.Stream(device='cuda:0')

    def timed(fn):
        best = 1e9
        for _ in range(3):
            torch.cuda.synchronize()
            with torch.cuda.stream(side):
                ev[0].record(side)
                r_ = fn()
                ev[1].record(side)
            torch.cuda.synchronize()
            best = min(best, ev[0].elapsed_time(ev[1]))
        return best, r_
    for kind, gen in gens.items():
        if args.sweep_kinds and kind not in args.sweep_kinds.split(','):
            continue
        for size in sizes:
            nb = int(max(8, min(2048, (1500 << 20) // size)))
            uniq = [np.ascontiguousarray(gen(size, 100 + i)) for i in range(8)]
            blocks = [uniq[i % 8] for i in range(nb)]
            d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
            for order in (0, 1):
                orders = np.full(nb, order, np.uint8)
                t_enc, r = timed(lambda: dv.rans_encode_dev(ctx, d_in, off, ln, orders))
                comp, st = r.to_host()
                assert (st == 0).all()
                t0 = time.perf_counter(); want = orans.compress(uniq[0], order); t_ce = time.perf_counter() - t0
                assert comp[0] == want, (kind, size, order)
                c_in, c_off, c_len = dv.pack_blocks(comp, 'cuda:0')
                pref = np.stack([np.frombuffer(c[:9], np.uint8) for c in comp])
                t_dec, d = timed(lambda: dv.rans_decode_dev(ctx, c_in, c_off, c_len, pref))
                back, st = d.to_host()
                assert (st == 0).all() and back[0] == blocks[0].tobytes() and back[-1] == blocks[-1].tobytes()
                t0 = time.perf_counter(); orans.uncompress(want); t_cd = time.perf_counter() - t0
                raw = float(nb) * size
                rows.append({'kind': kind, 'block_bytes': size, 'order': order, 'blocks': nb, 'state_lanes': 4 * nb,
                             'distinct_bytes': int(np.unique(uniq[0]).size),
                             'encode_GBps': raw / (t_enc * 1e-3) / 1e9, 'decode_GBps': raw / (t_dec * 1e-3) / 1e9,
                             'ratio': sum(len(c_) for c_ in comp) / raw, 'cpu_encode_GBps_1core': size / t_ce / 1e9,
                             'cpu_decode_GBps_1core': size / t_cd / 1e9})
                del c_in
            del d_in
            torch.cuda.empty_cache()
    print(json.dumps({'sweep': 'rANS block codec, BASELINE configs[1]', 'timing': 'CUDA events around the call, best of 3; GB/s = raw bytes / time',
                      'rows': rows}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--containers', type=int, default=2048, help='containers PER GPU (10 000 reads each)')
    ap.add_argument('--e2e-containers', type=int, default=1024, help='containers per GPU of the end-to-end (host buffer) leg')
    ap.add_argument('--pipeline', type=int, default=2, help='contexts per GPU taking consecutive steps in turn (1 = no overlap of encode and decode)')
    ap.add_argument('--e2e-pipeline', type=int, default=0, help='contexts of the end-to-end leg taking consecutive steps in turn (0 = 8 on one GPU, 6 on two, 4 beyond)')
    ap.add_argument('--no-unpack', action='store_true', help='decode side: take the read features from the encoder instead of unpacking the series (row P)')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--no-e2e', action='store_true', help='skip the end-to-end leg (profiling runs under ncu only: the line then has no e2e number)')
    ap.add_argument('--no-configs', action='store_true', help='skip the short runs of the other north_star configurations (quality models, 20 containers, fixed total)')
    ap.add_argument('--sweep-kinds', default='', help='--sweep: only these kinds (comma separated: QS,BA,RN,ITF8)')
    ap.add_argument('--sweep', action='store_true', help='BASELINE configs[1]: the rANS block sweep instead of the step benchmark')
    args = ap.parse_args()
    if args.sweep:
        run_sweep(args)
        return
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
