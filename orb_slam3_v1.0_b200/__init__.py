"""B200-native ORB front-end (extractor + Hamming matching + stereo row search).

Product code = csrc/ (hand-written sm_100a CUDA behind the C ABI of include/orbx.h) plus a
thin ctypes binding (`orbx`) used by tests and bench.py.  There is no CPU fallback: every
entry point raises if liborbx.so or a CUDA device is missing."""
from . import synth  # noqa: F401
from . import orbx  # noqa: F401
from . import sharding  # noqa: F401
