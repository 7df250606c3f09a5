"""Interface-compatible stand-in for the reference package ``simple_knn``
(/root/reference/submodules/simple-knn): ``from simple_knn._C import distCUDA2``."""
from . import _C  # noqa: F401
