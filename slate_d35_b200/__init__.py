"""B200-native Level-3 tile-update engine for SLATE (sm_100a)."""
from ._lib import lib, check, TileDesc, TrsmDesc, LIB_PATH  # noqa: F401
