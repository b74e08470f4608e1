"""Drop-in for `simple_knn._C` (submodules/simple-knn/ext.cpp:15-17): MonoGS's
gaussian_splatting/scene/gaussian_model.py:18 does `from simple_knn._C import distCUDA2`."""
from monogs_b200.ops import distCUDA2  # noqa: F401
