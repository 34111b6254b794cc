"""vegs_b200 — B200 (sm_100a) implementation of VEGS's Gaussian-splatting hot path.

Operator modules (same interface as the reference's extensions):
    vegs_b200.diff_gaussian_rasterization   GaussianRasterizationSettings, GaussianRasterizer
    vegs_b200.simple_knn._C                 distCUDA2

`vegs_b200.install()` registers them under the reference's import names so that
`from diff_gaussian_rasterization import ...` and `from simple_knn._C import distCUDA2`
(gaussian_renderer/__init__.py:14, scene/gaussian_model.py:21) resolve to this package unchanged.
"""
from __future__ import annotations

import sys

__version__ = "0.1.0"


def install() -> None:
    """Alias this package's operator modules to the reference's top-level module names."""
    from . import diff_gaussian_rasterization as dgr
    from . import simple_knn as sk
    from .simple_knn import _C as skc
    sys.modules["diff_gaussian_rasterization"] = dgr
    sys.modules["simple_knn"] = sk
    sys.modules["simple_knn._C"] = skc
