"""Device-side pieces of the insightlib outputs that sit directly on the classification path (density grids)."""
from .density import allreduce_grids, density_grid_device, grid_shape  # noqa: F401
