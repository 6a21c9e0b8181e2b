"""Device-side input pipeline: the tail of the reference's data loading on the GPU."""

from .device_input import FLAIR_RASTERS, DeviceRasterPipeline, RasterSpec, draw_augmentation  # noqa: F401
