"""vvc-b200: B200 (sm_100a) VVC reconstruction + in-loop-filter engine behind the VVCDSPContext surface.

The product is the C-ABI shared library built from ffmpeg_b200/csrc (see include/vvc_b200.h);
this package only holds the build recipe and a thin ctypes binding used by tests and bench.py.
"""
