"""B200-native neighbourhood-PCA hot path of YangSiri/pointCloudProcessing.

The product is `libpcp.so` (CUDA sm_100a kernels behind the C-ABI of include/pcp.h);
`capi` is its ctypes binding, `synth` the synthetic clouds of the benchmark configs,
`sharding` the one-process-per-GPU tile/frame partitioning.  No CPU fallback exists.
"""
from . import capi, synth  # noqa: F401
from .capi import CENTER_MEAN, CENTER_QUERY, FEATURE_DTYPE, Pcp, PcpError  # noqa: F401
