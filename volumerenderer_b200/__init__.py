"""volumerenderer_b200 -- headless B200 (sm_100a) implementation of VolumeRenderer's many-light hot path.

The package is a thin host layer over libvr.so (C ABI in include/vr.h, kernels in csrc/). It has no CPU fallback:
importing works without a GPU (grid fixtures are host code), creating a VolumeRenderer needs a CUDA device.
"""
from ._lib import ALGO, LG_KEEP_ORIGIN, LG_RESET_ORIGIN, LIGHT_DTYPE, VrOptions, VrTile, VrUbo, VrLibraryMissing, load  # noqa: F401
from .renderer import (HEADLESS_PATH, VolumeRenderer, VrError, build_cloud_grid, build_dense_grid,  # noqa: F401
                       build_fog_sphere_grid, load_nvdb_grid, save_nvdb_grid, write_pfm, write_ppm)
from .group import RendererGroup  # noqa: F401

__all__ = ["VolumeRenderer", "VrError", "VrUbo", "VrTile", "ALGO", "LIGHT_DTYPE", "LG_RESET_ORIGIN", "LG_KEEP_ORIGIN",
           "build_cloud_grid", "build_fog_sphere_grid", "build_dense_grid", "load", "VrLibraryMissing",
           "load_nvdb_grid", "save_nvdb_grid", "write_ppm", "write_pfm", "HEADLESS_PATH", "RendererGroup"]
