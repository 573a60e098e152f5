"""`AsyncQdrantClient`-shaped front end over the B200 engine: the seven methods the reference's
vector-store adapter calls on `qdrant_session.client` (qdrant_handler.py:44,55,94,190,363,467,474),
with the same keyword names and result shapes, so `QdrantHandler` runs unmodified on top of it
(INTEGRATION.md shows the one-line change in qdrant_session.py:21).

Semantics follow qdrant-client local mode as restated in SURVEY.md 8c: COSINE spaces store
L2-normalised float32 rows and normalise the query; sparse scoring is an overlap-only dot product
with optional Modifier.IDF; a Prefetch with children re-scores the union of its children's ids; a
FusionQuery(RRF) fuses them with constant 2 and default limit 10; ranking is (score desc, insertion
row asc).  Points live on the host until the first query after a mutation, then every named vector
space of the collection is (re)uploaded to HBM in one go.

All numeric work goes through rag_application_b200.Engine (the C ABI); there is no CPU path here.
`engine_factory` exists so the host logic can be unit-tested against the oracle-backed stand-in in
tests/_oracle_engine.py."""
from __future__ import annotations

import asyncio
import math
import threading
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import models as M

EPS32 = float(np.finfo(np.float32).eps)
DEFAULT_LIMIT = 10          # qdrant's default `limit` (also used by a fusion prefetch that omits it)
_STORAGE = {None: "f32", "float32": "f32", "float16": "f16", "bfloat16": "bf16", "f32": "f32", "f16": "f16", "bf16": "bf16",
            "i8": "i8", "int8": "i8"}


def _prep_rows(x: np.ndarray, dist: str, storage: str) -> np.ndarray:
    """Rows as the engine stores them: L2-normalised for COSINE spaces, except int8 storage, which keeps the integer
    values the caller sent (the reference's `quantized` vector, qdrant_handler.py:144-146) and scores by their cosine
    in the engine (include/brag.h: BRAG_I8)."""
    if storage == "i8":
        r = np.rint(x)
        if not (np.array_equal(r, x) and (x.size == 0 or (r.min() >= -128 and r.max() <= 127))):
            raise ValueError("a vector stored as int8 must hold integer values in [-128, 127]")
        return np.asarray(x, np.float32)
    return _normalize_rows(x) if dist == "cosine" else x


def _get(obj, name, default=None):
    """Attribute or mapping access: accepts our dataclasses, real qdrant models and plain dicts."""
    if obj is None:
        return default
    if isinstance(obj, dict):
        return obj.get(name, default)
    return getattr(obj, name, default)


def _enum_value(x):
    return getattr(x, "value", x)


def _normalize_rows(x: np.ndarray) -> np.ndarray:
    x = np.array(x, dtype=np.float32, copy=True)
    n = np.linalg.norm(x, axis=-1, keepdims=True).astype(np.float32)
    return (x / np.where(n != 0.0, n, np.float32(EPS32))).astype(np.float32)


# ------------------------------------------------------------------------------ payload filters
_UNSUPPORTED_COND = ("values_count", "geo_bounding_box", "geo_radius", "geo_polygon", "is_empty", "is_null", "nested")
_UNSUPPORTED_MATCH = ("text", "except", "except_", "phrase")


def _match_column(col: List[Any], cond) -> np.ndarray:
    match, rng = _get(cond, "match"), _get(cond, "range")
    # anything this compiler does not evaluate is refused, never silently matched (the reference builds
    # Filter(**filters) from an arbitrary dict, qdrant_handler.py:297)
    for f in _UNSUPPORTED_COND:
        if _get(cond, f) is not None:
            raise NotImplementedError(f"filter condition {f!r} is not supported")
    if match is None and rng is None:
        raise NotImplementedError(f"field condition on {_get(cond, 'key')!r} carries neither match nor range")
    if match is not None:
        for f in _UNSUPPORTED_MATCH:
            if _get(match, f) is not None:
                raise NotImplementedError(f"match.{f} is not supported (match.value / match.any only)")
        if _get(match, "any") is None and _get(match, "value") is None:
            raise NotImplementedError("match carries neither value nor any")
    out = np.ones(len(col), dtype=bool)
    if match is not None:
        if _get(match, "any") is not None:
            allowed = set(_get(match, "any"))
            out &= np.fromiter((_contains(v, allowed) for v in col), bool, len(col))
        else:
            val = _get(match, "value")
            out &= np.fromiter((_equals(v, val) for v in col), bool, len(col))
    if rng is not None:
        lo, loe, hi, hie = _get(rng, "gt"), _get(rng, "gte"), _get(rng, "lt"), _get(rng, "lte")

        def ok(v):
            if not isinstance(v, (int, float)) or isinstance(v, bool):
                return False
            return ((lo is None or v > lo) and (loe is None or v >= loe) and
                    (hi is None or v < hi) and (hie is None or v <= hie))
        out &= np.fromiter((ok(v) for v in col), bool, len(col))
    return out


def _equals(v, val) -> bool:
    if isinstance(v, (list, tuple)):
        return any(e == val for e in v)      # qdrant: array payloads match if any element matches
    return v == val


def _contains(v, allowed) -> bool:
    if isinstance(v, (list, tuple)):
        return any(e in allowed for e in v)
    try:
        return v in allowed
    except TypeError:
        return False


def _payload_column(payloads: List[Optional[dict]], key: str) -> List[Any]:
    parts = key.split(".")
    col = []
    for p in payloads:
        v: Any = p
        for k in parts:
            v = v.get(k) if isinstance(v, dict) else None
        col.append(v)
    return col


def compile_filter(flt, payloads: List[Optional[dict]], ids: List[Any]) -> np.ndarray:
    """Filter(must / must_not / should) -> bool[n].  Conditions: FieldCondition(match value / any,
    range), HasIdCondition, nested Filter; dict spellings of the same are accepted."""
    n = len(payloads)
    if flt is None:
        return np.ones(n, dtype=bool)

    def cond_mask(c) -> np.ndarray:
        if _get(c, "must") is not None or _get(c, "must_not") is not None or _get(c, "should") is not None:
            return compile_filter(c, payloads, ids)
        if _get(c, "has_id") is not None:
            want = set(_get(c, "has_id"))
            return np.fromiter((i in want for i in ids), bool, n)
        key = _get(c, "key")
        if key is None:
            raise ValueError(f"unsupported filter condition: {c!r}")
        return _match_column(_payload_column(payloads, key), c)

    def as_list(x):
        if x is None:
            return []
        return list(x) if isinstance(x, (list, tuple)) else [x]

    out = np.ones(n, dtype=bool)
    for c in as_list(_get(flt, "must")):
        out &= cond_mask(c)
    for c in as_list(_get(flt, "must_not")):
        out &= ~cond_mask(c)
    should = as_list(_get(flt, "should"))
    if should:
        any_ = np.zeros(n, dtype=bool)
        for c in should:
            any_ |= cond_mask(c)
        out &= any_
    return out


# ------------------------------------------------------------------------------ collection (host side)
class _Collection:
    def __init__(self, name: str, vectors_config: Dict[str, Any], sparse_config: Optional[Dict[str, Any]],
                 default_storage: Optional[str], storage_overrides: Optional[Dict[str, str]] = None):
        self.name = name
        self.dense_cfg: Dict[str, Tuple[int, str, str]] = {}     # vector name -> (size, distance, storage)
        for vname, vp in (vectors_config or {}).items():
            dist = str(_enum_value(_get(vp, "distance", M.Distance.COSINE))).lower()
            if dist not in ("cosine", "dot"):
                raise ValueError(f"distance {dist!r} of vector {vname!r} is not supported (Cosine / Dot only)")
            storage = _STORAGE[_enum_value(_get(vp, "datatype"))] if _get(vp, "datatype") else (default_storage or "f32")
            storage = _STORAGE[(storage_overrides or {}).get(vname, storage)]
            if storage == "i8" and (dist != "cosine" or int(_get(vp, "size")) % 128 != 0):
                raise ValueError(f"vector {vname!r}: int8 storage needs a Cosine space whose size is a multiple of 128")
            self.dense_cfg[vname] = (int(_get(vp, "size")), dist, storage)
        self.sparse_cfg: Dict[str, str] = {}                      # vector name -> "none" | "idf"
        for vname, sp in (sparse_config or {}).items():
            self.sparse_cfg[vname] = str(_enum_value(_get(sp, "modifier") or "none")).lower()
        self.ids: List[Any] = []
        self.id2row: Dict[Any, int] = {}
        self.payloads: List[Optional[dict]] = []
        self.dense_rows: Dict[str, List[Optional[np.ndarray]]] = {v: [] for v in self.dense_cfg}
        self.sparse_rows: Dict[str, List[Optional[Tuple[np.ndarray, np.ndarray]]]] = {v: [] for v in self.sparse_cfg}
        self.version = 0            # bumped by every mutation
        self.committed = -1         # version present in HBM
        # incremental upload: rows already in HBM per dense space, and which of them were overwritten
        self.dense_in_hbm: Dict[str, int] = {}
        self.dense_dirty: Dict[str, set] = {v: set() for v in self.dense_cfg}
        # sparse spaces: rows in HBM as of the last commit; True once a row below that count changed its vector
        self.sparse_in_hbm: Dict[str, int] = {}
        self.sparse_dirty: Dict[str, bool] = {v: False for v in self.sparse_cfg}
        # filled by commit()
        self.term_ids: Dict[str, np.ndarray] = {}     # sparse space -> sorted external indices (row = position)
        self.df: Dict[str, np.ndarray] = {}
        self.n_sparse: Dict[str, int] = {}            # points that have that sparse vector (IDF's N)
        self.present: Dict[str, Optional[np.ndarray]] = {}   # dense space -> bool[n] when some points lack it
        self.mask_cache: Dict[Any, int] = {}

    def __len__(self):
        return len(self.ids)

    def upsert(self, points: Iterable[Any]):
        parsed = []                                   # validate everything before touching the collection
        for p in points:
            pid = _get(p, "id")
            vec = _get(p, "vector") or {}
            if not isinstance(vec, dict):
                if list(self.dense_cfg) == [""]:
                    vec = {"": vec}
                else:
                    raise ValueError("collection has named vectors: pass vector={name: ...}")
            dense, sparse = {}, {}
            for vname, v in vec.items():
                if vname in self.dense_cfg:
                    a = np.asarray(v, dtype=np.float32)
                    if a.shape != (self.dense_cfg[vname][0],):
                        raise ValueError(f"vector {vname!r}: expected dim {self.dense_cfg[vname][0]}, got {a.shape}")
                    dense[vname] = a
                elif vname in self.sparse_cfg:
                    idx = np.asarray(_get(v, "indices"), dtype=np.int64)
                    val = np.asarray(_get(v, "values"), dtype=np.float32)
                    if idx.shape != val.shape or idx.ndim != 1:
                        raise ValueError(f"sparse vector {vname!r}: indices/values length mismatch")
                    if len(np.unique(idx)) != len(idx):
                        raise ValueError(f"sparse vector {vname!r}: indices must be unique")
                    sparse[vname] = (idx, val)
                else:
                    raise ValueError(f"Not existing vector name error: {vname}")
            payload = _get(p, "payload")
            parsed.append((pid, dense, sparse, dict(payload) if payload is not None else None))
        for pid, dense, sparse, payload in parsed:
            row = self.id2row.get(pid)
            if row is None:
                row = len(self.ids)
                self.id2row[pid] = row
                self.ids.append(pid)
                self.payloads.append(None)
                for lst in self.dense_rows.values():
                    lst.append(None)
                for lst in self.sparse_rows.values():
                    lst.append(None)
            self.payloads[row] = payload
            # an upsert replaces the whole point (qdrant): named vectors the new point does not carry are gone
            for vname, lst in self.dense_rows.items():
                if vname not in dense and lst[row] is not None:
                    lst[row] = None
                    self.dense_in_hbm[vname] = 0              # presence changed: re-upload the space at the next commit
            for vname, lst in self.sparse_rows.items():
                if vname not in sparse and lst[row] is not None:
                    lst[row] = None
                    self.sparse_dirty[vname] = True
            for vname, a in dense.items():
                self.dense_rows[vname][row] = a
                if row < self.dense_in_hbm.get(vname, 0):
                    self.dense_dirty[vname].add(row)
            for vname, sv in sparse.items():
                old = self.sparse_rows[vname][row]
                if row < self.sparse_in_hbm.get(vname, 0) and not (
                        old is not None and np.array_equal(old[0], sv[0]) and np.array_equal(old[1], sv[1])):
                    self.sparse_dirty[vname] = True
                self.sparse_rows[vname][row] = sv
        self.version += 1

    # -------------------------------------------------------------- upload to HBM
    def space(self, vname: str) -> str:
        return f"{self.name}\x1f{vname}"

    @staticmethod
    def _coo(rows, which):
        idx = np.concatenate([rows[i][0] for i in which])
        val = np.concatenate([rows[i][1] for i in which])
        doc = np.concatenate([np.full(len(rows[i][0]), i, dtype=np.int64) for i in which])
        return idx, val, doc

    @staticmethod
    def _term_major(trow, doc, val, n_terms):
        o = np.lexsort((doc, trow))
        trow, doc, val = trow[o], doc[o], val[o]
        ptr = np.zeros(n_terms + 1, dtype=np.int64)
        np.add.at(ptr, trow + 1, 1)
        return np.cumsum(ptr), doc, val

    def commit(self, engine):
        if self.committed == self.version:
            return
        n = len(self.ids)
        self.mask_cache.clear()
        for vname, (size, dist, storage) in self.dense_cfg.items():
            rows = self.dense_rows[vname]
            have = np.fromiter((r is not None for r in rows), bool, n)
            if n == 0 or not have.any():
                engine.dense_drop(self.space(vname))
                self.present[vname] = np.zeros(n, dtype=bool)
                self.dense_in_hbm[vname] = 0
                self.dense_dirty[vname].clear()
                continue
            n_old = self.dense_in_hbm.get(vname, 0)
            if 0 < n_old <= n and have.all() and self.present.get(vname, False) is None:
                # every row has the vector, before and now: append the new rows and overwrite the changed ones in
                # place (brag_dense_append / brag_dense_update) instead of re-uploading the space
                if n > n_old:
                    xa = np.stack(rows[n_old:]).astype(np.float32)
                    engine.dense_append(self.space(vname), _prep_rows(xa, dist, storage))
                dirty = sorted(self.dense_dirty[vname])
                if dirty:
                    xu = np.stack([rows[i] for i in dirty]).astype(np.float32)
                    engine.dense_update(self.space(vname), dirty, _prep_rows(xu, dist, storage))
            else:
                x = np.zeros((n, size), dtype=np.float32)
                for i, r in enumerate(rows):
                    if r is not None:
                        x[i] = r
                engine.dense_set(self.space(vname), _prep_rows(x, dist, storage), storage)
            self.present[vname] = None if have.all() else have
            self.dense_in_hbm[vname] = n
            self.dense_dirty[vname].clear()
        for vname in self.sparse_cfg:
            rows = self.sparse_rows[vname]
            have = [i for i, r in enumerate(rows) if r is not None]
            self.n_sparse[vname] = len(have)
            n_old = self.sparse_in_hbm.get(vname, 0)
            dirty, self.sparse_dirty[vname] = self.sparse_dirty[vname], False
            self.sparse_in_hbm[vname] = n if have else 0
            if not have:
                engine.sparse_drop(self.space(vname))
                self.term_ids[vname] = np.zeros(0, np.int64)
                self.df[vname] = np.zeros(0, np.int64)
                continue
            if 0 < n_old <= n and not dirty and len(self.term_ids.get(vname, ())) > 0:
                # only new rows since the last commit: splice their postings into the index in HBM
                # (brag_sparse_append) instead of rebuilding and re-uploading the whole CSR
                new = [i for i in have if i >= n_old]
                if not new:
                    continue
                idx, val, doc = self._coo(rows, new)
                old_terms = self.term_ids[vname]
                terms = np.union1d(old_terms, idx)                 # ascending external index == ascending term row
                row_map = np.searchsorted(terms, old_terms)
                ptr, doc, val = self._term_major(np.searchsorted(terms, idx), doc, val, len(terms))
                engine.sparse_append(self.space(vname), ptr, doc.astype(np.uint32), val.astype(np.float32),
                                     row_map.astype(np.int32), n)
                df = np.diff(ptr)
                df[row_map] += self.df[vname]
                self.term_ids[vname], self.df[vname] = terms, df
                continue
            idx, val, doc = self._coo(rows, have)
            terms, trow = np.unique(idx, return_inverse=True)     # ascending external index == ascending term row
            ptr, doc, val = self._term_major(trow, doc, val, len(terms))
            engine.sparse_set_csr(self.space(vname), ptr, doc.astype(np.uint32), val.astype(np.float32), n)
            self.term_ids[vname] = terms
            self.df[vname] = np.diff(ptr)
        self.committed = self.version


# ------------------------------------------------------------------------------ synchronous core
class B200QdrantCore:
    """Synchronous implementation; `B200QdrantClient` below wraps it in the async surface the
    reference awaits.  One engine per client: one GPU (`device`), or every collection row-sharded over several
    (`devices=[0, 1, ...]`, multi.py); spaces are named '<collection>\\x1f<vector>'."""

    MAX_MASKS = 64

    def __init__(self, device: int = 0, storage: Optional[str] = None, engine_factory: Optional[Callable] = None,
                 rrf_tie: str = "id", filter_propagation: str = "server", devices: Optional[Sequence[int]] = None,
                 shard_block_rows: int = 1 << 16, storage_overrides: Optional[Dict[str, str]] = None):
        # storage_overrides: vector name -> "f32" | "f16" | "bf16" | "i8" for every collection of this client, e.g.
        # {"quantized": "i8"}: the reference's int8 vector (qdrant_handler.py:144-146) kept as int8 rows and scanned
        # by the kind::i8 tensor-core kernel instead of as one more float space
        self.storage_overrides = dict(storage_overrides or {})
        if filter_propagation not in ("server", "local"):
            raise ValueError("filter_propagation must be 'server' or 'local'")
        self.filter_propagation = filter_propagation
        if engine_factory is None:
            from .engine import Engine      # fails loudly if libbrag.so / a B200 is missing
            engine_factory = Engine
        if devices is not None and len(devices) > 1:
            # every collection row-sharded over several GPUs behind the same engine contract (multi.py)
            from .multi import MultiEngine
            self.engine = MultiEngine(list(devices), engine_factory, shard_block_rows)
        else:
            self.engine = engine_factory(devices[0] if devices else device)
        self.default_storage = storage
        self.rrf_tie = 0 if rrf_tie == "id" else 1
        self.collections: Dict[str, _Collection] = {}
        self._lock = threading.RLock()
        self._next_mask = 1
        self._mask_owner: Dict[int, Tuple[str, Any]] = {}
        self._qcache: Dict[Tuple[int, Any], np.ndarray] = {}

    def close(self):
        self.engine.close()

    # ---------------------------------------------------------------- persistence
    def save(self, path: str):
        """Write every collection to the directory `path`: the engine's flat index file(s) (brag_save: rows, postings
        and the built scan indexes, one file per GPU) and the host-side tables (ids, payloads, vocabularies) -- what
        Qdrant keeps on its volume (docker-compose.yml:30-31).  `load` reopens them without re-uploading or
        rebuilding anything (qdrant_handler.py:94-100 finds the user's collection again at start-up)."""
        import os
        import pickle
        with self._lock:
            os.makedirs(path, exist_ok=True)
            for col in self.collections.values():
                col.commit(self.engine)
            self.engine.save(os.path.join(path, "index.brag"))
            state = {"format": 1, "collections": self.collections, "engine": self.engine.host_state()}
            tmp = os.path.join(path, "host.pkl.tmp")
            with open(tmp, "wb") as f:
                pickle.dump(state, f, protocol=4)
            os.replace(tmp, os.path.join(path, "host.pkl"))

    def load(self, path: str):
        import os
        import pickle
        with self._lock:
            with open(os.path.join(path, "host.pkl"), "rb") as f:
                state = pickle.load(f)
            if state.get("format") != 1:
                raise ValueError("unknown collection file format")
            self.engine.load(os.path.join(path, "index.brag"), state["engine"])
            for name, col in state["collections"].items():
                col.mask_cache = {}
                self.collections[name] = col
            self._mask_owner.clear()

    # ---------------------------------------------------------------- collection management
    def get_collections(self) -> M.CollectionsResponse:
        with self._lock:
            return M.CollectionsResponse([M.CollectionDescription(n) for n in self.collections])

    def collection_exists(self, collection_name: str) -> bool:
        return collection_name in self.collections

    def create_collection(self, collection_name: str, vectors_config=None, sparse_vectors_config=None,
                          optimizers_config=None, on_disk_payload=None, **_ignored) -> bool:
        with self._lock:
            if collection_name in self.collections:
                raise ValueError(f"Collection {collection_name} already exists")
            if vectors_config is not None and not isinstance(vectors_config, dict):
                vectors_config = {"": vectors_config}
            self.collections[collection_name] = _Collection(collection_name, vectors_config or {},
                                                            sparse_vectors_config, self.default_storage,
                                                            self.storage_overrides)
            return True

    def delete_collection(self, collection_name: str, **_ignored) -> bool:
        with self._lock:
            col = self.collections.pop(collection_name, None)
            if col is None:
                return False
            for v in col.dense_cfg:
                self.engine.dense_drop(col.space(v))
            for v in col.sparse_cfg:
                self.engine.sparse_drop(col.space(v))
            return True

    def _col(self, name: str) -> _Collection:
        col = self.collections.get(name)
        if col is None:
            raise ValueError(f"Collection {name} not found")
        return col

    def upsert(self, collection_name: str, points, wait: bool = True, **_ignored) -> M.UpdateResult:
        with self._lock:
            self._col(collection_name).upsert(points)
            return M.UpdateResult()

    def get_collection(self, collection_name: str) -> M.CollectionInfo:
        with self._lock:
            col = self._col(collection_name)
            return M.CollectionInfo(points_count=len(col), vectors_count=len(col) * (len(col.dense_cfg) + len(col.sparse_cfg)))

    def count(self, collection_name: str, count_filter=None, exact: bool = True, **_ignored) -> M.CountResult:
        with self._lock:
            col = self._col(collection_name)
            if count_filter is None:
                return M.CountResult(len(col))
            return M.CountResult(int(compile_filter(count_filter, col.payloads, col.ids).sum()))

    # ---------------------------------------------------------------- masks
    def _mask_id(self, col: _Collection, mask: Optional[np.ndarray], key) -> int:
        """Upload a bool[n] mask (cached per collection version under `key`); None -> 0 (no filter)."""
        if mask is None:
            return 0
        if key is not None and key in col.mask_cache:
            return col.mask_cache[key]
        mid = self._next_mask
        self._next_mask = self._next_mask % self.MAX_MASKS + 1
        old = self._mask_owner.pop(mid, None)
        if old is not None:
            oc = self.collections.get(old[0])
            if oc is not None:
                oc.mask_cache.pop(old[1], None)
        self.engine.mask_set(mid, mask)
        if key is not None:
            col.mask_cache[key] = mid
            self._mask_owner[mid] = (col.name, key)
        return mid

    @staticmethod
    def _filter_key(flt):
        return None if flt is None else repr(flt)

    # ---------------------------------------------------------------- plan executor (one query)
    def _sparse_query(self, col: _Collection, using: str, q) -> Tuple[np.ndarray, np.ndarray]:
        """External (hashed) indices -> ascending term rows + float64 values (IDF applied if configured)."""
        idx = np.asarray(_get(q, "indices"), dtype=np.int64)
        val = np.asarray(_get(q, "values"), dtype=np.float64)
        terms = col.term_ids.get(using, np.zeros(0, np.int64))
        if len(terms) == 0 or len(idx) == 0:
            return np.zeros(0, np.int32), np.zeros(0, np.float64)
        pos = np.searchsorted(terms, idx)
        ok = (pos < len(terms)) & (terms[np.minimum(pos, len(terms) - 1)] == idx)
        pos, val = pos[ok], val[ok]
        o = np.argsort(pos, kind="stable")
        pos, val = pos[o], val[o]
        if col.sparse_cfg[using] == "idf":
            n = col.n_sparse[using]
            df = col.df[using][pos].astype(np.float64)
            val = np.array([float(v) * math.log(1.0 + (n - d + 0.5) / (d + 0.5)) for v, d in zip(val, df)], np.float64)
        return pos.astype(np.int32), val

    ENGINE_MAX_K = 512          # longest list one brag_dense_topk / brag_sparse_topk call returns
    ENGINE_MAX_K_I8 = 416       # ... from an int8 space (tensor-core scan only, include/brag.h)
    ENGINE_MAX_CAND = 8192      # candidates (and results) of one brag_dense_rescore call

    def _paged(self, col: _Collection, limit: int, mask: Optional[np.ndarray], key, call, max_k: Optional[int] = None):
        """One ranked list of `limit` entries for one query.  The reference's prefetch limits are free parameters (an
        LLM picks them per query, hybrid_search_workflow.py:28-66) and may exceed the engine's list length: longer
        lists are served in pages, each page a scan under the filter minus the rows already returned.  Every page is
        the best of what is left under (score desc, id asc), so the concatenation is exactly the long list."""
        max_k = max_k or self.ENGINE_MAX_K
        if limit <= max_k:
            return call(limit, self._mask_id(col, mask, key))
        left = np.ones(len(col), dtype=bool) if mask is None else mask.copy()
        ss, ii = [], []
        got = 0
        while got < limit:
            want = min(max_k, limit - got)
            s, i, c = call(want, self._mask_id(col, left, None))
            c0 = int(np.asarray(c)[0])
            if c0 == 0:
                break
            s0, i0 = np.asarray(s)[0, :c0], np.asarray(i)[0, :c0]
            ss.append(s0)
            ii.append(i0)
            got += c0
            if c0 < want:
                break
            left[i0] = False
        if not ss:
            return np.zeros((1, 0), np.float32), np.zeros((1, 0), np.int64), np.zeros(1, np.int32)
        s, i = np.concatenate(ss)[None, :], np.concatenate(ii)[None, :]
        return s, i, np.array([s.shape[1]], np.int32)

    def _run(self, col: _Collection, query, using, limit, prefetch, flt, inherited=(None, None)
             ) -> Tuple[np.ndarray, np.ndarray]:
        """Bottom-up evaluation of one Prefetch / top-level request -> (scores, rows), best first.
        inherited = (mask, cache key) of the ancestors' filters: a Qdrant SERVER merges the filter of a request into
        every prefetch below it (the reference talks to a server and passes query_filter at the root,
        qdrant_handler.py:363-372), so with filter_propagation="server" every leaf scan already runs under it;
        "local" keeps qdrant-client local mode's behaviour (a filter acts only on the node that carries it)."""
        n = len(col)
        limit = DEFAULT_LIMIT if limit is None else int(limit)
        empty = (np.zeros(0, np.float64), np.zeros(0, np.int64))
        fmask = None if flt is None else compile_filter(flt, col.payloads, col.ids)
        fkey = self._filter_key(flt)
        if inherited[0] is not None:
            fmask = inherited[0] if fmask is None else (fmask & inherited[0])
            fkey = (inherited[1], fkey)
        down = (fmask, fkey) if self.filter_propagation == "server" else (None, None)
        if prefetch is None:
            kids = []
        else:
            plist = prefetch if isinstance(prefetch, (list, tuple)) else [prefetch]
            kids = [self._run(col, _get(p, "query"), _get(p, "using"), _get(p, "limit"), _get(p, "prefetch"),
                              _get(p, "filter"), down) for p in plist]

        fusion = _get(query, "fusion")
        if fusion is not None:
            if str(_enum_value(fusion)).lower() != "rrf":
                raise NotImplementedError("only Fusion.RRF is implemented")
            if not kids:
                raise ValueError("a fusion query needs prefetches")
            kmax = max(1, max(len(r) for _, r in kids))
            ids = np.full((len(kids), 1, kmax), -1, np.int64)
            cnt = np.zeros((len(kids), 1), np.int32)
            for l, (_, r) in enumerate(kids):
                ids[l, 0, :len(r)] = r
                cnt[l, 0] = len(r)
            s, i, c = self.engine.rrf(ids, cnt, limit, self.rrf_tie)
            s, i = np.asarray(s)[0, :int(c[0])], np.asarray(i)[0, :int(c[0])]
            if fmask is not None and len(i):
                keep = fmask[i]
                s, i = s[keep], i[keep]
            return s.astype(np.float64), i.astype(np.int64)

        if query is None:
            raise NotImplementedError("query=None (scroll-like requests) is not part of the hot path")
        if n == 0 or limit <= 0:
            return empty
        cand = None
        if kids:
            cand = np.unique(np.concatenate([r for _, r in kids])) if any(len(r) for _, r in kids) else np.zeros(0, np.int64)
            if len(cand) == 0:
                return empty

        if _get(query, "indices") is not None:                       # sparse query
            if using not in col.sparse_cfg:
                raise ValueError(f"Not existing vector name error: {using}")
            terms, vals = self._sparse_query(col, using, query)
            if len(terms) == 0 or col.n_sparse.get(using, 0) == 0:
                return empty
            mask, key = fmask, fkey
            if cand is not None:
                cm = np.zeros(n, dtype=bool)
                cm[cand] = True
                mask, key = (cm if mask is None else (cm & mask)), None
            q_ptr = np.array([0, len(terms)], np.int32)
            s, i, c = self._paged(col, min(limit, n), mask, key,
                                  lambda kk, mid: self.engine.sparse_topk(col.space(using), q_ptr, terms, vals, kk, mid))
        else:                                                        # dense query
            using = "" if using is None else using
            if using not in col.dense_cfg:
                raise ValueError(f"Not existing vector name error: {using}")
            size, dist, storage = col.dense_cfg[using]
            # the reference passes the SAME Python list to several nodes of one request (qdrant_handler.py:330,346,363:
            # `dense_vector` three times): converting and normalising a 768-float list costs ~40 us, so once per request
            ck = (id(query), using)
            q = self._qcache.get(ck)
            if q is None:
                q = np.asarray(query, dtype=np.float32).reshape(1, -1)
                if q.shape[1] != size:
                    raise ValueError(f"query for {using!r}: expected dim {size}, got {q.shape[1]}")
                q = _prep_rows(q, dist, storage)
                self._qcache[ck] = q
            max_k = self.ENGINE_MAX_K_I8 if storage == "i8" else self.ENGINE_MAX_K
            present = col.present.get(using)
            if present is not None and not present.any():
                return empty
            mask, key = fmask, fkey
            if present is not None:
                mask, key = (present if mask is None else (present & mask)), (("present", using, key))
            if cand is not None:
                mid = self._mask_id(col, mask, key)
                if len(cand) <= self.ENGINE_MAX_CAND:
                    s, i, c = self.engine.dense_rescore(col.space(using), q, cand.reshape(1, -1), min(limit, len(cand)), mid)
                else:
                    # more candidates than one brag_dense_rescore call takes: score them in slices (every slice
                    # returns all of its passing candidates with exact scores) and order the union by (score desc, id)
                    ss, ii = [], []
                    for a in range(0, len(cand), self.ENGINE_MAX_CAND):
                        part = cand[a:a + self.ENGINE_MAX_CAND]
                        s, i, c = self.engine.dense_rescore(col.space(using), q, part.reshape(1, -1), len(part), mid)
                        c0 = int(np.asarray(c)[0])
                        ss.append(np.asarray(s)[0, :c0])
                        ii.append(np.asarray(i)[0, :c0])
                    s, i = np.concatenate(ss), np.concatenate(ii)
                    o = np.lexsort((i, -s.astype(np.float64)))[:limit]
                    s, i, c = s[o][None, :], i[o][None, :], np.array([len(o)], np.int32)
            else:
                s, i, c = self._paged(col, min(limit, n), mask, key,
                                      lambda kk, mid: self.engine.dense_topk(col.space(using), q, kk, mid), max_k)
        c0 = int(np.asarray(c)[0])
        return np.asarray(s)[0, :c0].astype(np.float64), np.asarray(i)[0, :c0].astype(np.int64)

    def _points(self, col: _Collection, scores, rows, with_payload) -> List[M.ScoredPoint]:
        out = []
        for s, r in zip(scores, rows):
            p = col.payloads[int(r)] if with_payload else None
            out.append(M.ScoredPoint(id=col.ids[int(r)], score=float(s), payload=dict(p) if p is not None else None))
        return out

    def query_points(self, collection_name: str, query=None, using: Optional[str] = None, prefetch=None,
                     query_filter=None, search_params=None, limit: int = DEFAULT_LIMIT, offset: Optional[int] = None,
                     with_payload=True, with_vectors=False, score_threshold: Optional[float] = None,
                     **_ignored) -> M.QueryResponse:
        """search_params (hnsw_ef / exact) are accepted and ignored: every scan here is exact."""
        with self._lock:
            col = self._col(collection_name)
            col.commit(self.engine)
            off = int(offset or 0)
            self._qcache = {}                  # prepared dense queries of this request (the tree keeps its lists alive)
            try:
                s, r = self._run(col, query, using, int(limit) + off, prefetch, query_filter)
            finally:
                self._qcache = {}
            if score_threshold is not None:
                keep = s >= score_threshold
                s, r = s[keep], r[keep]
            return M.QueryResponse(points=self._points(col, s[off:], r[off:], bool(with_payload)))

    # ---------------------------------------------------------------- batched north_star shape
    def hybrid_search_batch(self, collection_name: str, dense_queries, sparse_queries: Sequence[Any], limit: int = 10,
                            using_dense: str = "dense", using_sparse: str = "sparse", query_filter=None,
                            with_payload: bool = False) -> List[List[M.ScoredPoint]]:
        """dense top-`limit` + sparse top-`limit` + RRF(limit) for a whole batch of queries in one engine
        call (brag_hybrid_topk): the throughput path benchmarked in bench.py."""
        with self._lock:
            col = self._col(collection_name)
            col.commit(self.engine)
            size, dist, storage = col.dense_cfg[using_dense]
            Q = _prep_rows(np.asarray(dense_queries, dtype=np.float32).reshape(len(sparse_queries), size), dist, storage)
            terms, vals, ptr = [], [], [0]
            for sq in sparse_queries:
                t, v = self._sparse_query(col, using_sparse, sq)
                terms.append(t)
                vals.append(v)
                ptr.append(ptr[-1] + len(t))
            fmask = None if query_filter is None else compile_filter(query_filter, col.payloads, col.ids)
            fkey = self._filter_key(query_filter)
            mid = self._mask_id(col, fmask, fkey)
            k = min(int(limit), len(col))
            if k > self.ENGINE_MAX_K:
                raise ValueError(f"hybrid_search_batch: limit {k} exceeds the engine's {self.ENGINE_MAX_K}-entry lists; "
                                 "use query_points (paged) for longer lists")
            q_ptr = np.asarray(ptr, np.int32)
            q_terms = np.concatenate(terms) if terms else np.zeros(0, np.int32)
            q_vals = np.concatenate(vals) if vals else np.zeros(0, np.float64)
            present = col.present.get(using_dense)
            if present is None:
                s, i, c = self.engine.hybrid_topk(col.space(using_dense), col.space(using_sparse), Q, q_ptr, q_terms,
                                                  q_vals, k, mid)
            else:
                # some points lack the dense vector (zero rows in HBM): they stay out of the dense list only, so the
                # two stages run under different masks and are fused by brag_rrf
                dmid = self._mask_id(col, present if fmask is None else (fmask & present), ("present", using_dense, fkey))
                _, di, dc = self.engine.dense_topk(col.space(using_dense), Q, k, dmid)
                _, si, sc = self.engine.sparse_topk(col.space(using_sparse), q_ptr, q_terms, q_vals, k, mid)
                s, i, c = self.engine.rrf(np.stack([np.asarray(di), np.asarray(si)]),
                                          np.stack([np.asarray(dc), np.asarray(sc)]), k, 0)
            s, i, c = np.asarray(s), np.asarray(i), np.asarray(c)
            return [self._points(col, s[b, :int(c[b])], i[b, :int(c[b])], with_payload) for b in range(len(sparse_queries))]


# ------------------------------------------------------------------------------ async surface
class B200QdrantClient:
    """Drop-in for `AsyncQdrantClient(url=...)` at qdrant_session.py:21.  Every method is a coroutine;
    GPU work runs in a worker thread so the single uvicorn event loop is never blocked, and
    mutations / queries of one client are serialised (interleaved awaits are safe)."""

    def __init__(self, device: int = 0, storage: Optional[str] = None, engine_factory: Optional[Callable] = None,
                 rrf_tie: str = "id", filter_propagation: str = "server", devices: Optional[Sequence[int]] = None,
                 shard_block_rows: int = 1 << 16, storage_overrides: Optional[Dict[str, str]] = None, **_ignored):
        self._core = B200QdrantCore(device=device, storage=storage, engine_factory=engine_factory, rrf_tie=rrf_tie,
                                    filter_propagation=filter_propagation, devices=devices,
                                    shard_block_rows=shard_block_rows, storage_overrides=storage_overrides)

    @property
    def core(self) -> B200QdrantCore:
        return self._core

    async def _call(self, fn, *a, **kw):
        return await asyncio.to_thread(fn, *a, **kw)

    async def get_collections(self):
        return await self._call(self._core.get_collections)

    async def collection_exists(self, collection_name: str):
        return await self._call(self._core.collection_exists, collection_name)

    async def create_collection(self, collection_name: str, **kw):
        return await self._call(self._core.create_collection, collection_name, **kw)

    async def delete_collection(self, collection_name: str, **kw):
        return await self._call(self._core.delete_collection, collection_name, **kw)

    async def upsert(self, collection_name: str, points, **kw):
        return await self._call(self._core.upsert, collection_name, points, **kw)

    async def query_points(self, collection_name: str, **kw):
        return await self._call(self._core.query_points, collection_name, **kw)

    async def count(self, collection_name: str, **kw):
        return await self._call(self._core.count, collection_name, **kw)

    async def get_collection(self, collection_name: str):
        return await self._call(self._core.get_collection, collection_name)

    async def hybrid_search_batch(self, collection_name: str, *a, **kw):
        return await self._call(self._core.hybrid_search_batch, collection_name, *a, **kw)

    async def save(self, path: str):
        return await self._call(self._core.save, path)

    async def load(self, path: str):
        return await self._call(self._core.load, path)

    async def close(self):
        await self._call(self._core.close)
