"""Mirror of the `simple_knn` package the reference imports (scene/gaussian_model.py:22:
`from simple_knn._C import distCUDA2`)."""
from . import _C  # noqa: F401
