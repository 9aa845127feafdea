"""pynama_b200 -- B200-native implementation of the FEM hot path of ibejarano/Pynama.

Sub-packages mirror the reference's layout (`src/{elements,domain,matrices,solver,cases,common}`) so that drivers
and tests written against the reference read the same:  Python on the host, hand-written sm_100a CUDA behind the
C ABI in include/pynama_b200.h for everything that touches an element, a matrix or a vector.
"""
__version__ = "0.1.0"
