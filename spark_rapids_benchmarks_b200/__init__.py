"""B200-native NDS query hot path.

Host-side mirror of the reference harness surface (nds_power / PysparkBenchReport /
nds_validate / nds_schema) over a C-ABI CUDA engine (csrc/, include/nds_b200.h).
There is no CPU fallback: importing `engine` without the built shared library, or
executing a plan with an unsupported node, raises.
"""
__version__ = "0.1.0"
