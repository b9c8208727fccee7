"""Regenerate tests/golden/rans_jvm.json by EXECUTING the reference jar's own bytecode
(htsjdk.samtools.cram.io.ExternalCompression.rans / unrans inside
/root/reference/cramtools-3.0.jar) with oracle/minijvm.py.  TEST INFRASTRUCTURE ONLY.

Run in the build container (needs /root/reference; takes a few minutes):

    python oracle/make_golden.py

The committed JSON is what the tests read; nothing at test time touches the jar.
"""
import base64
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

from minijvm import MiniJVM, JArray, JavaThrow, REFERENCE_JAR  # noqa: E402
from cramtools_b200 import synth  # noqa: E402

EC = 'htsjdk/samtools/cram/io/ExternalCompression'


def cases():
    rng = np.random.default_rng(20161)
    out = []
    for n in (1, 2, 3, 4, 5, 6, 7, 8, 9, 13, 31, 64, 257):
        out.append(('rand4sym_n%d' % n, bytes(rng.choice(np.frombuffer(b'ACGT', np.uint8), size=n))))
    out += [
        ('AAAA', b'AAAA'),
        ('abracadabra', b'abracadabra'),
        ('ACGTx8', b'ACGT' * 8),
        ('zeros_100', bytes(100)),
        ('ff_37', b'\xff' * 37),
        ('all256_once', bytes(range(256))),
        ('all256_x3_plus1', bytes(range(256)) * 3 + b'\x07'),
        ('two_syms_skewed', bytes((rng.random(3001) < 0.003).astype(np.uint8) * 200)),
        ('runs_of_consecutive_syms', bytes([10, 11, 12, 13, 20, 30, 31, 255, 254, 0, 1] * 23)),
        ('random_bytes_2051', bytes(rng.integers(0, 256, size=2051, dtype=np.uint8))),
        ('random_bytes_5000', bytes(rng.integers(0, 256, size=5000, dtype=np.uint8))),
        ('qs_like_20003', synth.qs_stream(20003, 11).tobytes()),
        ('ba_like_6002', synth.ba_stream(6002, 12).tobytes()),
        ('rn_like_7001', synth.rn_stream(7001, 13).tobytes()),
        ('itf8_like_4000', synth.itf8_stream(4000, 14).tobytes()),
        ('geometric_9000', bytes(np.minimum(rng.geometric(0.05, size=9000), 255).astype(np.uint8))),
    ]
    return out


def main():
    vm = MiniJVM([REFERENCE_JAR])

    def rans(data, order):
        return bytes(vm.call(EC, 'rans', '([BI)[B', [JArray('B', bytearray(data)), order]).a)

    def unrans(blob):
        return bytes(vm.call(EC, 'unrans', '([B)[B', [JArray('B', bytearray(blob))]).a)

    vectors = []
    t0 = time.time()
    for name, data in cases():
        for order in (0, 1):
            blob = rans(data, order)
            back = unrans(blob)
            assert back == data, name
            vectors.append({'name': name, 'order': order,
                            'input': base64.b64encode(data).decode(),
                            'output': base64.b64encode(blob).decode()})
            print('%-28s o%d  %6d -> %6d   (%.0fs)' % (name, order, len(data), len(blob), time.time() - t0), flush=True)
    # empty input [RANS.java:60-61]
    vectors.append({'name': 'empty', 'order': 1, 'input': '', 'output': base64.b64encode(rans(b'', 1)).decode()})

    # error behaviour of RANS.uncompress [RANS.java:29-55]
    errors = []
    good = rans(b'abracadabra', 0)
    for name, blob in (('truncated_by_1', good[:-1]), ('extended_by_1', good + b'\0'),
                       ('bad_order_2', bytes([2]) + good[1:])):
        try:
            unrans(blob)
            msg = None
        except JavaThrow as t:
            msg = '%s: %s' % (t.obj.cls, t.obj.fields.get('message'))
        errors.append({'name': name, 'blob': base64.b64encode(blob).decode(), 'throws': msg})
        print('error case', name, '->', msg)

    out = {'generator': 'oracle/make_golden.py (mini-JVM run of cramtools-3.0.jar, htsjdk 2.4.1 rANS bytecode)',
           'entry_point': 'htsjdk.samtools.cram.io.ExternalCompression.rans(byte[], int) / unrans(byte[])',
           'vectors': vectors, 'errors': errors}
    path = os.path.join(os.path.dirname(HERE), 'tests', 'golden', 'rans_jvm.json')
    with open(path, 'w') as f:
        json.dump(out, f, indent=0)
    print('wrote', path, 'bytecodes executed:', vm.nbytecodes)


if __name__ == '__main__':
    main()
