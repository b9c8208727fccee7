"""cramtools_b200 -- B200-native CRAM slice codec behind cramtools' own plug-in points.

The product is libcramgpu.so (hand-written CUDA for sm_100a behind the C ABI of include/cramgpu.h).
This package is the host-side mirror of the reference's Java interfaces for the hot path, used by the
tests and the benchmark exactly the way the Java shim of INTEGRATION.md uses the library.
"""
from ._lib import Context, CramGpuError, default_context  # noqa: F401
from .external_compression import ExternalCompression, ExternalCompressor, rans_bound  # noqa: F401
from . import records  # noqa: F401,E402  (registers the record-feature entry points)
from .records import QualityScorePreservation  # noqa: F401,E402
