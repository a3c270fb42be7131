"""pycsou_b200: B200-native implementation of pycsou's iterative proximal-solver hot path.

Same module layout and public names as the reference for that path (``pycsou_b200.core``, ``.func``, ``.linop``,
``.math.prox``, ``.opt.proxalgs``); arrays are CUDA tensors, kernels are hand-written sm_100a CUDA behind a C ABI
(``include/pcsb200.h``, ``libpcsb200.so``).  There is no CPU fallback.
"""
__version__ = '0.1.0'
