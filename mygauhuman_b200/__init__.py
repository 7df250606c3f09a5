"""mygauhuman_b200 -- B200-native (sm_100a) implementation of GauHuman's render
hot path: LBS deformation, the tile-based differentiable Gaussian rasteriser and
simple-knn ``distCUDA2``, behind the reference's own operator API.

Sub-packages mirror the reference's extension packages:
``mygauhuman_b200.diff_gaussian_rasterization`` and ``mygauhuman_b200.simple_knn``.
``install()`` registers them under the reference's top-level names so the
reference's ``gaussian_renderer.render()`` / ``baking.py`` / ``scene/gaussian_model.py``
import them unchanged.
"""
import sys

__version__ = "0.1.0"


def install():
    """Alias the drop-in packages as ``diff_gaussian_rasterization`` and ``simple_knn``."""
    from . import diff_gaussian_rasterization as dgr
    from . import simple_knn as sknn
    sys.modules["diff_gaussian_rasterization"] = dgr
    sys.modules["diff_gaussian_rasterization._C"] = dgr._C
    sys.modules["simple_knn"] = sknn
    sys.modules["simple_knn._C"] = sknn._C
    return dgr, sknn
