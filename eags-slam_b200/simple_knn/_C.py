"""``simple_knn._C`` -- the one symbol the reference imports (``src/entities/gaussian_model.py:17``)."""
from eags_raster.neighbors import mean_dist2_3nn as distCUDA2  # noqa: F401
