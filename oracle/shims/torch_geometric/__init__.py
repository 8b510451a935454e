"""TEST INFRASTRUCTURE ONLY — minimal stand-in for the parts of `torch_geometric` that the
reference's `dgmc/` package and its tests import (dgmc.py:4-5, rel.py:4, spline.py:4,
gin.py:3, data.py:6, test_dgmc.py:2).  Used by `oracle/make_golden.py` (in the build
container, where /root/reference exists) to run the unmodified reference.  Semantics are
restated from PyG's documented behaviour; see each submodule.  Never imported by the product.
"""
from . import nn, utils, data  # noqa: F401
