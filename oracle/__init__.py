"""TEST INFRASTRUCTURE ONLY.

oracle/ holds the checker for the CUDA engine: (1) `_ref/libvvcref.so`, the reference's own C code
for the hot path compiled in place by oracle/Makefile, and (2) `port/`, a plain-C restatement of the
same algorithms.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; the product (ffmpeg_b200/, include/) never does.
"""
