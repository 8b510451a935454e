"""Drop-in for the reference's `dgmc` package (dgmc/__init__.py:1-11), B200 kernels underneath."""
from . import native  # noqa: F401  (fails loudly if libdgmc_b200.so is missing)
from . import models, utils  # noqa: F401
from .models.dgmc import DGMC

__version__ = '1.0.0'

__all__ = ['dgmc', 'DGMC', '__version__']
