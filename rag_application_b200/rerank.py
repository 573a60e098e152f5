"""Rerank seam of the reference.

`QdrantHandler.rerank_with_colbert(query, documents, results, max_tokens)` (qdrant_handler.py:388-412) asks
`reranker.client.rerank_documents(query, documents, max_tokens)` (huggingface.py:172-189) for an index order and
re-orders `results` with it, returning `results` unchanged when the order is empty or anything raises.  The encoder
(tokenizer + transformer) needs model weights and stays where it is; everything after it -- pooling, scoring, ordering
-- runs in `brag_rerank_meanpool` on the GPU.  `B200Reranker` takes the encoder as a callable
`encode(texts) -> float32 [len(texts), T, d]` (last hidden states, padded as the tokenizer pads them).
"""
from typing import Any, Callable, Dict, List, Sequence


class B200Reranker:
    def __init__(self, engine, encode: Callable[[Sequence[str]], Any]):
        self.engine = engine
        self.encode = encode

    def rerank_documents(self, query: str, documents: List[str], max_tokens: int) -> List[int]:
        """Same signature and result as the reference's `rerank_documents`: document indices, best first."""
        if not documents:
            return []
        # (the reference's `len(doc_tokens) > 8000` guard, huggingface.py:177-182, measures the number of fields the
        # tokenizer returned and never fires; `max_tokens` is accepted for signature parity and unused, as there)
        q = self.encode([query])[0]
        d = self.encode(documents)
        order, _ = self.engine.rerank_meanpool(q, d)
        return [int(i) for i in (order.tolist() if hasattr(order, "tolist") else order)]

    def rerank_with_colbert(self, query: str, documents: List[str], results: List[Dict], max_tokens: int) -> List[Dict]:
        """qdrant_handler.py:388-412: reorder `results`; on an empty order or any failure return them unchanged."""
        try:
            ranked = self.rerank_documents(query, documents, max_tokens)
            if not ranked:
                return results
            return [results[i] for i in ranked]
        except Exception:
            return results
