"""ocsmesh_b200 -- B200-native raster hot path of OCSMesh.

Drop-in mirror of the reference's public names on this path
(``ocsmesh/__init__.py:22-29``): ``Raster``, ``Hfun``, ``Mesh``, ``MeshData``.
All arithmetic runs in hand-written CUDA kernels (sm_100a) behind the C ABI of
``libocsb200.so`` (``include/ocsb200.h``); there is no CPU fallback.
"""
from .raster import (Affine, CRS, Raster, Window, get_iter_windows,
                     raster_from_numpy)
from .hfun import Hfun, HfunCollector, HfunRaster
from .mesh import EuclideanMesh2D, Mesh, MeshData

__all__ = [
    "Affine", "CRS", "Raster", "Window", "get_iter_windows",
    "raster_from_numpy", "Hfun", "HfunCollector", "HfunRaster",
    "EuclideanMesh2D", "Mesh", "MeshData",
]
__version__ = "0.1.0"
