"""B200-native drop-in for SynapseML's LightGBM data-parallel training path.

Layout:
  csrc/       CUDA kernels for sm_100a and the C-ABI runtime (built into lib/lib_lightgbm.so by build.py)
  lightgbm/   host-side mirror of the reference's Scala lightgbm module above that C ABI
  data/       synthetic inputs of the BASELINE.json shapes
"""
__version__ = "0.1.0"
