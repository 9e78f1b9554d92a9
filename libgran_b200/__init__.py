"""libgran_b200 -- B200-native (sm_100a) implementation of libgran's per-timestep DEM hot path.

The compute path is the hand-written CUDA library ``libgran_b200/liblgb200.so`` behind the C ABI of
``include/lgb200/lgb200.h``; this package is only its Python host-side mirror (used by the tests and
``bench.py``).  There is no CPU fallback: importing :mod:`libgran_b200.system` raises if the CUDA
library has not been built (``python -c "import __graft_entry__ as g; g.build()"``).
"""
__version__ = "0.1.0"
