"""simglucose_b200 -- B200-native batched engine for the simglucose (UVa/Padova 2008) closed loop.

The compute path is the CUDA library behind include/simglucose_b200.h (csrc/); this package is the
host-side mirror of the reference's Python surface plus the vectorised batch environment.
"""
__version__ = '0.1.0'
