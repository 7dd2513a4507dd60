"""`simple_knn._C.distCUDA2` replacement (submodule source absent from the reference checkout)."""
from .._C import dist_cuda2 as distCUDA2  # noqa: F401
