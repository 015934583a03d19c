"""Drop-in mirror of the reference's ``simple_knn`` package (knn/ext.cpp:16): ``simple_knn._C.distCUDA2``."""
from . import _C  # noqa: F401
