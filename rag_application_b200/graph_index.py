"""Vector-index procedure the reference delegates to Neo4j, on the engine's dense scan.

The reference creates cosine vector indexes over node / relationship embeddings
(app/core/graph_db/neo4j/neo4j_handler.py:37-55: 256-d; IndexerAPI/src/core/storage/neo4j_handler.py:70-73) and
searches them with

    CALL db.index.vector.queryNodes($index_name, $top_k, $embedding) YIELD node, score
    WITH node, score WHERE node.user_id = $user_id AND ...            (filter AFTER the top-k)
    RETURN ... ORDER BY score DESC LIMIT $top_k

(app/core/graph_db/neo4j/neo4j_search.py:33-41, IndexerAPI/.../neo4j_handler.py:809-870).  `queryNodes` is one more
consumer of the exact dense top-k kernel (row a1): this class keeps one dense space per index and answers the
procedure call with Neo4j's score convention for cosine indexes, (1 + cos) / 2.  The Cypher engine, the graph store
and the driver are out of scope; the property filter is applied to the k rows the index returned, as the query text
does, so fewer than k rows can come back.  (Neo4j's index is approximate; this one is exact.)
"""
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np


class B200GraphVectorIndexes:
    def __init__(self, engine, storage: str = "f32"):
        self.engine = engine
        self.storage = storage
        self.dims: Dict[str, int] = {}
        self.props: Dict[str, List[Dict[str, Any]]] = {}

    @staticmethod
    def _space(name: str) -> str:
        return f"graph\x1f{name}"

    def create_vector_index(self, name: str, dimensions: int, similarity_function: str = "cosine"):
        if similarity_function != "cosine":
            raise ValueError("the reference only creates cosine vector indexes")
        if name not in self.dims:                         # CREATE VECTOR INDEX ... IF NOT EXISTS
            self.dims[name] = int(dimensions)
            self.props[name] = []

    def set_nodes(self, name: str, embeddings, properties: Sequence[Dict[str, Any]]):
        """(Re)load the indexed nodes: embeddings [n, dimensions], one property map per node."""
        x = np.asarray(embeddings, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.dims[name] or len(properties) != len(x):
            raise ValueError("embeddings must be [n, dimensions] with one property map per row")
        self.props[name] = [dict(p) for p in properties]
        if len(x) == 0:
            self.engine.dense_drop(self._space(name))
            return
        norm = np.sqrt((x.astype(np.float64) ** 2).sum(axis=1))
        x = (x / np.maximum(norm, 1e-12)[:, None]).astype(np.float32)
        self.engine.dense_set(self._space(name), x, self.storage)

    def query_nodes(self, name: str, k: int, embedding) -> List[Tuple[Dict[str, Any], float]]:
        """db.index.vector.queryNodes(name, k, embedding): the k nearest indexed nodes, best first."""
        props = self.props[name]
        if not props:
            return []
        q = np.asarray(embedding, dtype=np.float32).reshape(1, -1)
        if q.shape[1] != self.dims[name]:
            raise ValueError("embedding has the wrong dimension for this index")
        q = q / max(float(np.sqrt((q.astype(np.float64) ** 2).sum())), 1e-12)
        k = max(1, min(int(k), len(props), 512))
        s, i, c = self.engine.dense_topk(self._space(name), q.astype(np.float32), k)
        s, i, c = np.asarray(s), np.asarray(i), np.asarray(c)
        return [(props[int(i[0, j])], (1.0 + float(s[0, j])) / 2.0) for j in range(int(c[0]))]

    def search(self, name: str, k: int, embedding, where: Optional[Dict[str, Any]] = None) -> List[Dict[str, Any]]:
        """One `_search_*` query of the reference: top-k of the index, then the property filter, score descending."""
        rows = []
        for node, score in self.query_nodes(name, k, embedding):
            if where and any(v is not None and node.get(key) != v for key, v in where.items()):
                continue
            rows.append({**node, "score": score})
        rows.sort(key=lambda r: r["score"], reverse=True)
        return rows[:max(1, int(k))]

    def search_across_spaces(self, names: Sequence[str], embedding, top_k: int,
                             where: Optional[Dict[str, Any]] = None) -> List[Dict[str, Any]]:
        """IndexerAPI neo4j_handler.py:809-827: every index searched with the same limit, one merged ranking."""
        limit = max(1, int(top_k))
        merged: List[Dict[str, Any]] = []
        for name in names:
            merged.extend(self.search(name, limit, embedding, where))
        merged.sort(key=lambda item: item["score"], reverse=True)
        return merged[:limit]
