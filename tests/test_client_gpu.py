"""GPU: the reference's own request sequence (create_collection -> upsert -> hybrid_search's nested
prefetch tree -> count) through B200QdrantClient on the real engine, against the known-answer vectors."""
import pytest

from _client_cases import run_reference_flow

pytestmark = pytest.mark.gpu


def test_reference_flow_on_b200():
    run_reference_flow(engine_factory=None)
