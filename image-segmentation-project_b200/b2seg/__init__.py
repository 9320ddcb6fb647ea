"""b2seg — B200-native (sm_100a) kernels for the segmentation training step, Python host layer."""
from ._lib import B2SegError, lib, LIB_PATH  # noqa: F401
