"""CPU: the C oracle (oracle/rans_oracle.c) against the golden vectors that came out of the reference jar."""
import numpy as np
import pytest

from oracle import rans as orans
from cramtools_b200 import synth


def test_oracle_matches_reference_jar_bytes(rans_golden):
    vec, _ = rans_golden
    assert len(vec) >= 50
    for name, order, data, blob in vec:
        assert orans.compress(data, order) == blob, (name, order)


def test_oracle_decodes_reference_jar_blocks(rans_golden):
    vec, _ = rans_golden
    for name, order, data, blob in vec:
        assert orans.uncompress(blob) == data, (name, order)


def test_oracle_error_behaviour(rans_golden):
    _, err = rans_golden
    for name, blob, throws in err:
        with pytest.raises(RuntimeError) as e:
            orans.uncompress(blob)
        if 'Incorrect input length' in throws:
            assert 'Incorrect input length' in str(e.value), name
        if 'Unknown rANS order' in throws:
            assert 'Unknown rANS order' in str(e.value), name


def test_small_inputs_fall_back_to_order0(rans_golden):
    # RANS.compress [RANS.java:63-64]: n < 4 -> order 0
    for n in (1, 2, 3):
        blob = orans.compress(b'ACG'[:n], 1)
        assert blob[0] == 0
    assert orans.compress(b'', 1) == b''
    assert orans.uncompress(b'') == b''


@pytest.mark.parametrize('kind', sorted(synth.STREAM_KINDS))
@pytest.mark.parametrize('order', [0, 1])
def test_oracle_round_trip_streams(kind, order):
    for n in (1000, 65536 + 3, 300000):
        d = synth.STREAM_KINDS[kind](n, 100 + n % 97)
        blob = orans.compress(d, order)
        assert blob[0] == order
        assert int.from_bytes(blob[1:5], 'little') == len(blob) - 9
        assert int.from_bytes(blob[5:9], 'little') == n
        assert orans.uncompress(blob) == d.tobytes()


def test_frequencies_sum_to_4095():
    # the reference normalises to TOTFREQ - 1 (fsum++ before the fix-up, [Frequencies.java:131-135])
    d = synth.qs_stream(50000, 3)
    blob = orans.compress(d, 0)
    # parse the order-0 table
    p = 9
    total = 0
    rle = 0
    j = blob[p]; p += 1
    while True:
        f = blob[p]; p += 1
        if f >= 128:
            f = ((f & 127) << 8) | blob[p]; p += 1
        total += f
        if rle == 0 and blob[p] == j + 1:
            j = blob[p]; rle = blob[p + 1]; p += 2
        elif rle:
            rle -= 1; j += 1
        else:
            j = blob[p]; p += 1
        if j == 0:
            break
    assert total == 4095


def test_bound():
    for n in (0, 1, 100, 65536, 1 << 20, 8 << 20):
        assert orans.bound(n) == int(1.05 * n + 257 * 257 * 3 + 4)
