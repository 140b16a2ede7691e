"""acai_omr_b200: B200-native (sm_100a) drop-in for the ViTOMR encoder/decoder hot path of
jsnchon/acai-omr.  `acai_omr_b200.models.models` / `.models.kv_caching` mirror the reference's
`acai_omr.models` API; the arithmetic lives in libacai_b200.so (C ABI: include/acai_b200.h)."""
from ._lib import AcaiError, LIB_PATH  # noqa: F401

__all__ = ["AcaiError", "LIB_PATH"]
