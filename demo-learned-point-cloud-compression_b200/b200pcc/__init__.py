"""b200pcc — B200-native (sm_100a) learned point-cloud codec hot path behind the reference's
CompressionPipeline.compress() / DecompressionPipeline.decompress() surface.

Importing this package loads libpcc_b200.so (the C ABI of include/pcc_b200.h) and fails loudly if
it has not been built.  There is no CPU fallback.
"""
from . import _lib  # noqa: F401  (raises ImportError when the CUDA library is missing)
from .sparse import SparseTensor, CoordMap, sort_points, sort_tensor  # noqa: F401
from .model import ColorModel  # noqa: F401
from .pipeline import CompressionPipeline, DecompressionPipeline, load_model  # noqa: F401
from .single import CompressionPipelineSingle, DecompressionPipelineSingle  # noqa: F401

__all__ = ["SparseTensor", "CoordMap", "ColorModel", "CompressionPipeline", "DecompressionPipeline",
           "CompressionPipelineSingle", "DecompressionPipelineSingle", "load_model", "sort_points", "sort_tensor"]
