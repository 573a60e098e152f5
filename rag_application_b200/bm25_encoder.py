"""fastembed-shaped sparse encoder: the object the reference builds at
`embedding_handler.py:41` (`SparseTextEmbedding(model_name="Qdrant/bm25")`) and calls at `:123`
(`list(self.sparse_model.embed([text]))[0]` -> `.indices`, `.values`).

Text normalisation uses fastembed's regular expressions (Unicode-aware, done here in Python); stop words,
Snowball English stemming, murmur3 ids and the TF weights run in the C++ library (`csrc/bm25_encode.cu`,
`brag_bm25_encode`).  fastembed is not installed in this image, so parity is unpinned: the known-answer
tests use the published Porter2 / MurmurHash3 vectors."""
from __future__ import annotations

import ctypes as C
import re
from dataclasses import dataclass
from typing import Iterable, Iterator, List, Union

import numpy as np

from . import _native

_NON_WORD_KEEP_SPACE = re.compile(r"[^\w\s]", flags=re.UNICODE)
_NON_WORD = re.compile(r"[^\w]", flags=re.UNICODE)
_SPACES = re.compile(r"\s+")


@dataclass
class SparseEmbedding:
    """Same attribute names as fastembed's SparseEmbedding."""
    values: np.ndarray
    indices: np.ndarray

    def as_object(self):
        return {"values": self.values, "indices": self.indices}

    def as_dict(self):
        return {int(i): float(v) for i, v in zip(self.indices, self.values)}


def tokenize(text: str) -> List[str]:
    """fastembed: remove_non_alphanumeric, then SimpleTokenizer.tokenize."""
    text = _NON_WORD_KEEP_SPACE.sub(" ", text)
    text = _NON_WORD.sub(" ", text.lower())
    return _SPACES.sub(" ", text).strip().split()


class SparseTextEmbedding:
    """`SparseTextEmbedding(model_name="Qdrant/bm25")` with fastembed's defaults k=1.2, b=0.75, avg_len=256."""

    def __init__(self, model_name: str = "Qdrant/bm25", k: float = 1.2, b: float = 0.75, avg_len: float = 256.0,
                 token_max_length: int = 40, **_ignored):
        if model_name.lower() != "qdrant/bm25":
            raise ValueError(f"only Qdrant/bm25 is implemented, not {model_name}")
        self.k, self.b, self.avg_len, self.token_max_length = float(k), float(b), float(avg_len), int(token_max_length)
        self._lib = _native.load()

    def _encode(self, text: str, query_mode: bool) -> SparseEmbedding:
        toks = tokenize(text)
        enc = [t.encode("utf-8") for t in toks]
        off = np.zeros(len(enc) + 1, dtype=np.int32)
        if enc:
            off[1:] = np.cumsum([len(t) for t in enc])
        ids = np.empty(max(1, len(enc)), dtype=np.int64)
        vals = np.empty(max(1, len(enc)), dtype=np.float64)
        n = C.c_int32(0)
        rc = self._lib.brag_bm25_encode(b"".join(enc), off.ctypes.data, len(enc), self.k, self.b, self.avg_len,
                                        self.token_max_length, int(query_mode), ids.ctypes.data, vals.ctypes.data,
                                        C.addressof(n))
        if rc != 0:
            raise _native.BragError("brag_bm25_encode failed")
        return SparseEmbedding(values=vals[: n.value].copy(), indices=ids[: n.value].copy())

    def embed(self, documents: Union[str, Iterable[str]], **_ignored) -> Iterator[SparseEmbedding]:
        if isinstance(documents, str):
            documents = [documents]
        for d in documents:
            yield self._encode(d, False)

    def query_embed(self, query: Union[str, Iterable[str]], **_ignored) -> Iterator[SparseEmbedding]:
        if isinstance(query, str):
            query = [query]
        for q in query:
            yield self._encode(q, True)

    def stem(self, word: str) -> str:
        buf = C.create_string_buffer(4 * len(word) + 16)
        if self._lib.brag_bm25_stem(word.encode("utf-8"), buf, len(buf)) != 0:
            raise _native.BragError("brag_bm25_stem failed")
        return buf.value.decode("utf-8")

    def token_id(self, stem: str) -> int:
        b = stem.encode("utf-8")
        h = self._lib.brag_murmur3_32(b, len(b), 0)
        h = h - (1 << 32) if h >= (1 << 31) else h
        return abs(h)
