"""GPU parity tests for the rANS block codec, through the C ABI (libcramgpu.so)."""
import numpy as np
import pytest

from cramtools_b200 import synth, _lib

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def EC():
    from cramtools_b200 import ExternalCompression
    return ExternalCompression


def test_golden_vectors_encode_byte_identical(EC, rans_golden, gpu_ctx):
    vec, _ = rans_golden
    got = EC.rans_batch([v[2] for v in vec], [v[1] for v in vec], gpu_ctx)
    for (name, order, data, blob), g in zip(vec, got):
        assert g == blob, (name, order)


def test_golden_vectors_decode(EC, rans_golden, gpu_ctx):
    vec, _ = rans_golden
    got = EC.unrans_batch([v[3] for v in vec], gpu_ctx)
    for (name, order, data, blob), g in zip(vec, got):
        assert g == data, (name, order)


def test_single_block_calls_like_the_reference(EC, rans_golden, gpu_ctx):
    vec, _ = rans_golden
    for name, order, data, blob in vec[:12]:
        assert EC.rans(data, order, gpu_ctx) == blob
        assert EC.unrans(blob, gpu_ctx) == data


def test_error_behaviour_matches_reference(EC, rans_golden, gpu_ctx):
    _, err = rans_golden
    for name, blob, throws in err:
        with pytest.raises(_lib.CramGpuError) as e:
            EC.unrans(blob, gpu_ctx)
        msg = throws.split(': ', 1)[1]
        assert msg in str(e.value), (name, str(e.value))
    with pytest.raises(_lib.CramGpuError):
        EC.rans(b'abcdef', 2, gpu_ctx)
    # a caller's output buffer that cannot hold the block: the wording of RANS.allocateIfNeeded [RANS.java:83]
    import ctypes
    data = np.frombuffer(bytes(range(256)) * 8, np.uint8)
    out = np.zeros(64, np.uint8)
    pin = (_lib.u8p * 1)(data.ctypes.data_as(_lib.u8p)); pout = (_lib.u8p * 1)(out.ctypes.data_as(_lib.u8p))
    ln = (ctypes.c_uint32 * 1)(data.size); cap = (ctypes.c_uint32 * 1)(out.size); got = (ctypes.c_uint32 * 1)(7)
    order = (ctypes.c_uint8 * 1)(0); st = (ctypes.c_int32 * 1)(0)
    rc = gpu_ctx.lib.cramgpu_rans_encode_batch(gpu_ctx.handle, 1, pin, ln, order, pout, cap, got, st)
    assert rc == _lib.ERR_OUT_TOO_SMALL and st[0] == rc and got[0] == 0 and not out.any()
    assert 'Insufficient buffer size.' in gpu_ctx.last_error()


def _oracle():
    from oracle import rans as orans
    return orans


@pytest.mark.parametrize('order', [0, 1])
def test_parity_with_oracle_size_sweep(EC, gpu_ctx, order):
    orans = _oracle()
    blocks = []
    for kind in sorted(synth.STREAM_KINDS):
        for n in (4, 5, 63, 64, 65, 4095, 4096, 4097, 65536, 65537, 131072 + 1, 262144 + 2, 1 << 20):
            blocks.append(synth.STREAM_KINDS[kind](n, 7 * n + order))
    got = EC.rans_batch(blocks, [order] * len(blocks), gpu_ctx)
    for b, g in zip(blocks, got):
        assert g == orans.compress(b, order), (len(b), order)
    back = EC.unrans_batch(got, gpu_ctx)
    for b, r in zip(blocks, back):
        assert r == b.tobytes()


def test_parity_adversarial_alphabets(EC, gpu_ctx):
    orans = _oracle()
    rng = np.random.default_rng(5)
    blocks = [
        rng.integers(0, 256, size=300000, dtype=np.uint8),                    # all 256 symbols, 256 contexts
        np.repeat(np.arange(256, dtype=np.uint8), 40),                        # long runs, every symbol
        np.tile(np.arange(256, dtype=np.uint8), 50),                          # every (i, i+1) pair
        np.zeros(100000, dtype=np.uint8),                                     # single symbol 0
        np.full(77777, 255, dtype=np.uint8),                                  # single symbol 255
        (rng.random(500000) < 0.001).astype(np.uint8) * 3,                    # very skewed
        np.minimum(rng.geometric(0.02, size=200000), 255).astype(np.uint8),   # ~200 symbols, skewed
        rng.integers(0, 130, size=100000, dtype=np.uint8),                    # alphabet just above the smem limit
    ]
    for order in (0, 1):
        got = EC.rans_batch(blocks, [order] * len(blocks), gpu_ctx)
        for i, (b, g) in enumerate(zip(blocks, got)):
            assert g == orans.compress(b, order), (i, order)
        back = EC.unrans_batch(got, gpu_ctx)
        for b, r in zip(blocks, back):
            assert r == b.tobytes()


def test_mixed_batch_with_empty_and_tiny_blocks(EC, gpu_ctx):
    orans = _oracle()
    blocks = [b'', b'A', b'AC', b'ACG', b'ACGT', b'', synth.qs_stream(150000, 1).tobytes(), b'\x00' * 7,
              synth.rn_stream(30000, 2).tobytes(), b'', synth.itf8_stream(10000, 3).tobytes()]
    orders = [1, 1, 0, 1, 1, 0, 1, 0, 1, 0, 0]
    got = EC.rans_batch(blocks, orders, gpu_ctx)
    for b, o, g in zip(blocks, orders, got):
        assert g == orans.compress(b, o)
    assert EC.unrans_batch(got, gpu_ctx) == [bytes(b) for b in blocks]


def test_gpu_decodes_oracle_blocks_and_oracle_decodes_gpu_blocks(EC, gpu_ctx):
    orans = _oracle()
    data = [synth.STREAM_KINDS[k](200000 + i, 50 + i) for i, k in enumerate(sorted(synth.STREAM_KINDS))]
    for order in (0, 1):
        blobs = [orans.compress(d, order) for d in data]
        assert EC.unrans_batch(blobs, gpu_ctx) == [d.tobytes() for d in data]
        for d, g in zip(data, EC.rans_batch(data, [order] * len(data), gpu_ctx)):
            assert orans.uncompress(g) == d.tobytes()


def test_corrupt_blocks_are_reported_not_crashing(EC, gpu_ctx):
    orans = _oracle()
    good = bytearray(orans.compress(synth.qs_stream(20000, 9), 1))
    # keep the prefix consistent but scribble over the frequency table
    bad = bytearray(good)
    for i in range(9, 60):
        bad[i] = 0xFF
    with pytest.raises(_lib.CramGpuError) as e:
        EC.unrans(bytes(bad), gpu_ctx)
    assert e.value.status == _lib.ERR_CORRUPT
    # the context still works afterwards
    assert EC.unrans(bytes(good), gpu_ctx) == synth.qs_stream(20000, 9).tobytes()


def test_large_many_block_round_trip_checksum(EC, gpu_ctx):
    """Full-size property check: many slice-sized QS blocks, decode(encode(x)) == x and block sizes
    equal the oracle's on a sample."""
    orans = _oracle()
    q = synth.qualities(10000 * 12, 150, 77).reshape(12, -1)     # 12 slices' QS blocks of 1.5 MB
    blocks = [np.ascontiguousarray(q[i]) for i in range(12)]
    got = EC.rans_batch(blocks, [1] * 12, gpu_ctx)
    assert got[0] == orans.compress(blocks[0], 1)
    assert got[11] == orans.compress(blocks[11], 1)
    back = EC.unrans_batch(got, gpu_ctx)
    for b, r in zip(blocks, back):
        assert r == b.tobytes()


def test_device_resident_entry_points(gpu_ctx):
    import torch
    from cramtools_b200 import device as dv
    orans = _oracle()
    blocks = [synth.qs_stream(100000 + 17 * i, i) for i in range(20)] + [synth.ba_stream(5000, 3), np.zeros(0, np.uint8)]
    orders = [i % 2 for i in range(len(blocks))]
    d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
    res = dv.rans_encode_dev(gpu_ctx, d_in, off, ln, orders)
    torch.cuda.synchronize()
    blobs, st = res.to_host()
    assert (st == 0).all()
    for b, o, g in zip(blocks, orders, blobs):
        assert g == orans.compress(b, o)
    # decode straight from the encoder's device output
    off2 = res.off.cpu().numpy().view(np.uint64)[:-1]
    ln2 = res.len.cpu().numpy().view(np.uint32)
    pre = np.zeros((len(blocks), 9), dtype=np.uint8)
    for i, g in enumerate(blobs):
        pre[i, :min(9, len(g))] = np.frombuffer(g[:9], dtype=np.uint8)
    res2 = dv.rans_decode_dev(gpu_ctx, res.out, off2, ln2, pre)
    torch.cuda.synchronize()
    raws, st2 = res2.to_host()
    assert (st2 == 0).all()
    assert raws == [b.tobytes() for b in blocks]


def test_fuzz_skewed_alphabets_against_oracle(gpu_ctx):
    """skewed alphabets with clusters of rare symbols (long table searches of the order-1 decoder), symbols of frequency 1
    (two-byte renormalisations), alphabets around the shared-memory table limits: encode byte-identical to the oracle, decode
    round trip (scripts/fuzz_rans.py runs more seeds)"""
    import cramtools_b200
    from cramtools_b200 import ExternalCompression as EC
    from oracle import rans as orans
    rng = np.random.default_rng(7)
    blocks = []
    for t in range(96):
        n = int(rng.choice([5, 63, 64, 1000, 4097, 30000, 150000]))
        nsym = int(rng.choice([1, 2, 3, 8, 20, 41, 47, 48, 49, 64, 130, 256]))
        syms = np.sort(rng.choice(256, size=nsym, replace=False)).astype(np.uint8)
        kind = t % 4
        if kind == 0:
            p = np.full(nsym, 0.02 / max(nsym - 1, 1)); p[rng.integers(nsym)] = 0.98
        elif kind == 1:
            p = 0.6 ** np.arange(nsym)
        elif kind == 2:
            p = np.full(nsym, 1e-4); p[-min(3, nsym):] = 1.0
        else:
            p = rng.random(nsym) ** 4 + 1e-5
        p = p / p.sum()
        x = rng.choice(nsym, size=n, p=p)
        if t % 3 == 0:                    # first-order structure: runs
            x = np.repeat(x[:max(1, n // 7)], 7)[:n] if n >= 7 else x
            if x.size < n:
                x = np.concatenate([x, np.zeros(n - x.size, x.dtype)])
        blocks.append(np.ascontiguousarray(syms[x]))
    for order in (0, 1):
        got = EC.rans_batch(blocks, [order] * len(blocks), gpu_ctx)
        for i, (b, g) in enumerate(zip(blocks, got)):
            assert g == orans.compress(b, order), (order, i, b.size, np.unique(b).size)
        back = EC.unrans_batch(got, gpu_ctx)
        for i, (b, g) in enumerate(zip(blocks, back)):
            assert g == b.tobytes(), (order, i)


def test_eight_megabyte_blocks(gpu_ctx):
    """the top of BASELINE config[1]'s sweep: 8 MB blocks of the three stream kinds, both orders, byte-identical to the oracle"""
    from cramtools_b200 import ExternalCompression as EC, synth
    from oracle import rans as orans
    blocks = [synth.qs_stream(8 << 20, 301), synth.ba_stream(8 << 20, 302), synth.rn_stream((8 << 20) - 3, 303)]
    for order in (0, 1):
        got = EC.rans_batch(blocks, [order] * 3, gpu_ctx)
        for b, g in zip(blocks, got):
            assert g == orans.compress(b, order)
        assert EC.unrans_batch(got, gpu_ctx) == [b.tobytes() for b in blocks]


def test_decode_dev_rejects_misaligned_blocks(EC, gpu_ctx):
    """the decoders fetch a block as aligned words: an offset that is not a multiple of 16 must be refused with
    CRAMGPU_ERR_ARG (the header's contract) instead of faulting the device"""
    import torch
    from cramtools_b200 import device as dv, _lib
    blocks = [synth.qs_stream(5000, 7)]
    comp = EC.rans_batch(blocks, [1], gpu_ctx)
    pad = b'\0' * 4 + comp[0]                               # the block starts 4 bytes into the buffer
    d_in, off, ln = dv.pack_blocks([pad], 'cuda:0')
    pref = np.frombuffer(comp[0][:9], np.uint8).reshape(1, 9)
    with pytest.raises(_lib.CramGpuError) as e:
        dv.rans_decode_dev(gpu_ctx, d_in, np.array([4], np.uint64), np.array([len(comp[0])], np.uint32), pref)
    assert e.value.status == _lib.ERR_ARG
    # and the context stays usable
    back = EC.unrans_batch(comp, gpu_ctx)
    assert back[0] == blocks[0].tobytes()


def test_wide_alphabets_pool_and_overflow(EC, gpu_ctx):
    """order-1 blocks with more than 64 distinct bytes: decoded from pooled table images (k_dec1w), and -- when a call has more
    of them than the pool holds -- the rest through the slab path (k_dec1_big).  Same bytes either way; corrupt wide blocks
    are reported."""
    import cramtools_b200
    orans = _oracle()
    rng = np.random.default_rng(11)
    blocks = []
    for i in range(40):
        kind = i % 4
        n = int(rng.integers(2000, 90000))
        if kind == 0: b = rng.integers(0, 65 + i, size=n, dtype=np.uint8)                       # 65 .. 104 distinct
        elif kind == 1: b = np.minimum(rng.geometric(0.03, size=n), 255).astype(np.uint8)       # skewed, ~150 distinct
        elif kind == 2: b = rng.integers(0, 256, size=n, dtype=np.uint8)                        # all 256
        else: b = synth.qs_stream(n, i)                                                          # a narrow one in between
        blocks.append(b)
    comp = [orans.compress(b, 1) for b in blocks]
    ctx = cramtools_b200.Context(0)
    try:
        for pool in (2048, 3, 0):
            ctx.set_option('dec_wide_pool_blocks', pool)
            back = EC.unrans_batch(comp, ctx)
            for b, r in zip(blocks, back):
                assert r == b.tobytes(), (pool, len(b))
        ctx.set_option('dec_wide_pool_blocks', 2048)
        # a wide block cut short inside its payload, and one whose table is cut: reported, neighbours unharmed
        bad = list(comp)
        c2 = bytearray(comp[2]); del c2[-len(c2) // 3:]; c2[1:5] = (len(c2) - 9).to_bytes(4, 'little'); bad[2] = bytes(c2)
        ins = [np.frombuffer(x, np.uint8) for x in bad]
        import ctypes
        n = len(ins)
        in_len = np.array([a.size for a in ins], np.uint32); raw = np.array([len(b) for b in blocks], np.uint32)
        outs = [np.zeros(int(r) + 16, np.uint8) for r in raw]
        pin = (_lib.u8p * n)(*[a.ctypes.data_as(_lib.u8p) for a in ins]); pout = (_lib.u8p * n)(*[a.ctypes.data_as(_lib.u8p) for a in outs])
        out_len = np.zeros(n, np.uint32); st = np.zeros(n, np.int32)
        rc = ctx.lib.cramgpu_rans_decode_batch(ctx.handle, n, pin, in_len.ctypes.data_as(_lib.u32p), pout, raw.ctypes.data_as(_lib.u32p),
                                               out_len.ctypes.data_as(_lib.u32p), st.ctypes.data_as(_lib.i32p))
        assert rc == _lib.ERR_CORRUPT and st[2] == _lib.ERR_CORRUPT and (np.delete(st, 2) == 0).all()
        for i in range(n):
            if i != 2:
                assert outs[i][:raw[i]].tobytes() == blocks[i].tobytes()
        with pytest.raises(_lib.CramGpuError):
            ctx.set_option('no_such_option', 1)
    finally:
        ctx.close()
