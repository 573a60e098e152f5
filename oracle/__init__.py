"""CPU oracle for the hybrid-retrieval hot path.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the arithmetic of this path lives in the un-vendored, un-pinned
third-party dependency ``qdrant-client[fastembed]`` (reference ``requirements.txt:3``),
which is not importable in the build container; the reference's own tests pin no
numerical result for it.  ``oracle.py`` therefore restates the
published qdrant-client *local mode* algorithm and anchors on the reference's own call
sites (``app/core/vector_store/qdrant/qdrant_handler.py:296-372``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / reference
legs may import this package.  The product (``rag_application_b200``) never does.
"""
