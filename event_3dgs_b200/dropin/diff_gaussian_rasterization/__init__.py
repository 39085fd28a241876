"""Drop-in for the reference's `diff_gaussian_rasterization` package: put
event_3dgs_b200/dropin on PYTHONPATH (after the repo root) and the reference's
gaussian_renderer / train.py import this instead of the CUDA extension they ship."""
from event_3dgs_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                        _RasterizeGaussians, cpu_deep_copy_tuple, rasterize_gaussians)
