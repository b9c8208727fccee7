import json
import base64
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def rans_golden():
    """(input, order, output) triples produced by running the reference jar's bytecode (oracle/make_golden.py)"""
    with open(os.path.join(ROOT, 'tests', 'golden', 'rans_jvm.json')) as f:
        g = json.load(f)
    vec = [(v['name'], v['order'], base64.b64decode(v['input']), base64.b64decode(v['output'])) for v in g['vectors']]
    err = [(e['name'], base64.b64decode(e['blob']), e['throws']) for e in g['errors']]
    return vec, err


@pytest.fixture(scope='session')
def gpu_ctx():
    import cramtools_b200
    return cramtools_b200.default_context(0)
