"""c-raytracer_b200: B200-native trace loop for wojciech-graj/C-Raytracer.

The product is the C ABI in include/crt.h implemented by csrc/libcrt.so
(hand-written sm_100a CUDA) and the C host in src/.  This package is only the
Python binding used by tests/ and bench.py.  The directory name is not a valid
module name; import it with `tests/_load.py` / `bench.py`'s loader as
`c_raytracer_b200`.
"""
