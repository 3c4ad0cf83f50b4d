"""gstex_b200 -- B200-native (sm_100a) textured 2D-Gaussian-splatting rasterizer hot path.

Drop-in for the `gstex_cuda` submodule of nvnhat95/GStex on the path
project -> bounds -> tile count -> scan -> key emit -> radix sort -> tile ranges -> forward -> backward.
The Python layer mirrors the reference's entry points; all compute is hand-written CUDA behind the
C ABI declared in `include/gstex_b200.h` (no CPU path, no fallback).
"""
from .activations import activations  # noqa: F401
from .charts import build_charts, chart_texel_size, resample_texture  # noqa: F401
from .get_aabb_2d import get_aabb_2d, get_num_tiles_hit_2d, project_bounds, project_points  # noqa: F401
from .loss import photometric_loss  # noqa: F401
from .sh import deg_from_sh, num_sh_bases, spherical_harmonics  # noqa: F401
from .texture import texture_gaussians, pad_texture  # noqa: F401
from .texture_edit import texture_edit  # noqa: F401
from .texture_sample import texture_sample  # noqa: F401
from .utils import (begin_binning, bin_and_sort_gaussians, bin_gaussians_to_tiles, finish_binning, compute_cumulative_intersects, get_tile_bin_edges,  # noqa: F401
                    map_gaussian_to_intersects, sort_isects)

__all__ = [
    "project_points", "get_aabb_2d", "get_num_tiles_hit_2d", "project_bounds", "compute_cumulative_intersects",
    "map_gaussian_to_intersects", "sort_isects", "get_tile_bin_edges", "bin_and_sort_gaussians", "bin_gaussians_to_tiles", "begin_binning", "finish_binning",
    "texture_gaussians", "pad_texture", "activations", "build_charts", "chart_texel_size", "resample_texture", "spherical_harmonics", "num_sh_bases", "deg_from_sh", "texture_sample", "texture_edit", "photometric_loss",
]
