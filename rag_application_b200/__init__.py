"""B200-native hybrid retrieval (dense top-k -> BM25 -> RRF) behind the reference's
Qdrant call surface.  See DESIGN.md; the C ABI is include/brag.h."""
from .engine import Engine, to_storage, TIE_ID, TIE_FIRST_SEEN  # noqa: F401

__all__ = ["Engine", "to_storage", "TIE_ID", "TIE_FIRST_SEEN"]
