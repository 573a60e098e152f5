"""Duck-typed stand-ins for the `qdrant_client.http.models` objects the reference builds
(qdrant_handler.py:3-5): same class and field names, nothing else.  `qdrant_client` is not a
dependency of this package; when it IS installed the client accepts its real model objects as well,
because everything below is read through attributes / mapping keys only."""
from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence, Union


class Distance(str, enum.Enum):
    COSINE = "Cosine"
    DOT = "Dot"
    EUCLID = "Euclid"       # accepted by the schema, rejected at query time (the reference never uses it)
    MANHATTAN = "Manhattan"


class Fusion(str, enum.Enum):
    RRF = "rrf"
    DBSF = "dbsf"           # not implemented (the reference only uses RRF)


class Modifier(str, enum.Enum):
    NONE = "none"
    IDF = "idf"


@dataclass
class VectorParams:
    size: int
    distance: Distance = Distance.COSINE
    on_disk: Optional[bool] = None
    datatype: Optional[str] = None      # "float32" (default) | "float16" | "bfloat16": HBM storage type


@dataclass
class SparseIndexParams:
    on_disk: Optional[bool] = None
    full_scan_threshold: Optional[int] = None


@dataclass
class SparseVectorParams:
    index: Optional[SparseIndexParams] = None
    modifier: Optional[Modifier] = None


@dataclass
class OptimizersConfigDiff:
    memmap_threshold: Optional[int] = None
    indexing_threshold: Optional[int] = None


@dataclass
class SparseVector:
    indices: List[int]
    values: List[float]


@dataclass
class PointStruct:
    id: Union[int, str]
    vector: Dict[str, Any]
    payload: Optional[Dict[str, Any]] = None


@dataclass
class FusionQuery:
    fusion: Fusion = Fusion.RRF


@dataclass
class MatchValue:
    value: Any


@dataclass
class MatchAny:
    any: Sequence[Any]


@dataclass
class Range:
    gt: Optional[float] = None
    gte: Optional[float] = None
    lt: Optional[float] = None
    lte: Optional[float] = None


@dataclass
class FieldCondition:
    key: str
    match: Optional[Union[MatchValue, MatchAny, Dict[str, Any]]] = None
    range: Optional[Union[Range, Dict[str, Any]]] = None


@dataclass
class HasIdCondition:
    has_id: Sequence[Union[int, str]]


@dataclass
class Filter:
    must: Optional[List[Any]] = None
    must_not: Optional[List[Any]] = None
    should: Optional[List[Any]] = None


@dataclass
class Prefetch:
    prefetch: Optional[Union["Prefetch", List["Prefetch"]]] = None
    query: Any = None
    using: Optional[str] = None
    limit: Optional[int] = None
    filter: Optional[Filter] = None
    params: Optional[Any] = None
    score_threshold: Optional[float] = None


@dataclass
class ScoredPoint:
    id: Union[int, str]
    score: float
    payload: Optional[Dict[str, Any]] = None
    vector: Optional[Any] = None
    version: int = 0


@dataclass
class QueryResponse:
    points: List[ScoredPoint] = field(default_factory=list)


@dataclass
class CollectionDescription:
    name: str


@dataclass
class CollectionsResponse:
    collections: List[CollectionDescription] = field(default_factory=list)


@dataclass
class CountResult:
    count: int


@dataclass
class CollectionInfo:
    points_count: int
    vectors_count: Optional[int] = None
    status: str = "green"
    config: Optional[Any] = None


@dataclass
class UpdateResult:
    operation_id: int = 0
    status: str = "completed"


class _Namespace:
    """`from qdrant_client.http.models import models` style access: models.Distance, models.FusionQuery..."""


models = _Namespace()
for _n, _v in list(globals().items()):
    if isinstance(_v, type) and not _n.startswith("_"):
        setattr(models, _n, _v)
