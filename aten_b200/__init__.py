"""aten_b200 -- B200-native (sm_100a) implementation of nackdai/aten's path-tracing hot path.

The product is ``libaten_b200.so`` (hand-written CUDA + a C ABI, include/aten_b200.h); this package
is the Python host side: ``api.PathTracing`` mirrors the reference's ``idaten::PathTracing`` and
``scene_io`` reads the flat scene blobs.  See DESIGN.md / INTEGRATION.md.
"""
from .api import AtnError, PathTracing, load_library  # noqa: F401
from . import scene_io  # noqa: F401
