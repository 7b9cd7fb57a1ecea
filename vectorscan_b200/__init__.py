"""vectorscan_b200 -- B200-native block-mode scanning runtime for Vectorscan databases.

  vectorscan_b200.gpu     ctypes mirror of the runtime C ABI (include/hs_gpu.h): alloc_scratch / scan / scan_vector
  vectorscan_b200.litdb   literal-set database writer (include/hs_litdb.h), stand-in for hs_compile_lit_multi
"""
__all__ = ["gpu", "litdb"]
