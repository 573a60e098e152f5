"""the fastembed-shaped BM25 encoder behind `embedding_handler.py:41,123-138`.
fastembed is not installed here (parity unpinned), so the known answers are the published ones: the Snowball
English sample vocabulary, MurmurHash3 x86_32 vectors (mmh3.hash), and the TF formula worked by hand."""
import numpy as np
import pytest

from rag_application_b200.bm25_encoder import SparseTextEmbedding, tokenize


@pytest.fixture(scope="module")
def enc():
    return SparseTextEmbedding(model_name="Qdrant/bm25")


SNOWBALL_SAMPLE = {  # word -> stem, from the Snowball English stemmer's published sample vocabulary
    "consign": "consign", "consigned": "consign", "consigning": "consign", "consignment": "consign",
    "consist": "consist", "consisted": "consist", "consistency": "consist", "consistent": "consist",
    "consistently": "consist", "consisting": "consist", "consists": "consist", "consolation": "consol",
    "consolations": "consol", "consolatory": "consolatori", "console": "consol", "consoled": "consol",
    "consoles": "consol", "consolidate": "consolid", "consolidated": "consolid", "consolidating": "consolid",
    "consoling": "consol", "consolingly": "consol", "consols": "consol", "consonant": "conson", "consort": "consort",
    "consorted": "consort", "consorting": "consort", "conspicuous": "conspicu", "conspicuously": "conspicu",
    "conspiracy": "conspiraci", "conspirator": "conspir", "conspirators": "conspir", "conspire": "conspir",
    "conspired": "conspir", "conspiring": "conspir", "constable": "constabl", "constables": "constabl",
    "constance": "constanc", "constancy": "constanc", "constant": "constant",
    "knack": "knack", "knackeries": "knackeri", "knacks": "knack", "knag": "knag", "knave": "knave",
    "knaves": "knave", "knavish": "knavish", "kneaded": "knead", "kneading": "knead", "knee": "knee",
    "kneel": "kneel", "kneeled": "kneel", "kneeling": "kneel", "kneels": "kneel", "knees": "knee", "knell": "knell",
    "knelt": "knelt", "knew": "knew", "knick": "knick", "knif": "knif", "knife": "knife", "knight": "knight",
    "knightly": "knight", "knights": "knight", "knit": "knit", "knits": "knit", "knitted": "knit",
    "knitting": "knit", "knives": "knive", "knob": "knob", "knobs": "knob", "knock": "knock", "knocked": "knock",
    "knocker": "knocker", "knockers": "knocker", "knocking": "knock", "knocks": "knock", "knopp": "knopp",
    "knot": "knot", "knots": "knot",
}

PORTER2_RULES = {  # one word per rule of the algorithm description
    "caresses": "caress", "ponies": "poni", "ties": "tie", "cries": "cri", "gas": "gas", "gaps": "gap",
    "kiwis": "kiwi", "this": "this", "feed": "feed", "agreed": "agre", "plastered": "plaster", "bled": "bled",
    "motoring": "motor", "sing": "sing", "conflated": "conflat", "troubled": "troubl", "sized": "size",
    "hopping": "hop", "tanned": "tan", "falling": "fall", "hissing": "hiss", "fizzed": "fizz", "failing": "fail",
    "filing": "file", "happy": "happi", "cry": "cri", "by": "by", "say": "say", "relational": "relat",
    "conditional": "condit", "rational": "ration", "digitizer": "digit", "operator": "oper", "feudalism": "feudal",
    "decisiveness": "decis", "hopefulness": "hope", "callousness": "callous", "triplicate": "triplic",
    "formative": "format", "formalize": "formal", "electrical": "electr", "hopeful": "hope", "goodness": "good",
    "revival": "reviv", "allowance": "allow", "inference": "infer", "airliner": "airlin", "gyroscopic": "gyroscop",
    "adjustable": "adjust", "defensible": "defens", "irritant": "irrit", "replacement": "replac",
    "adjustment": "adjust", "dependent": "depend", "adoption": "adopt", "communism": "communism",
    "activate": "activ", "homologous": "homolog", "effective": "effect", "bowdlerize": "bowdler", "probate": "probat",
    "rate": "rate", "cease": "ceas", "controll": "control", "roll": "roll", "generously": "generous",
    "generate": "generat", "generic": "generic", "communication": "communic", "arsenal": "arsenal",
    # special words
    "skis": "ski", "skies": "sky", "dying": "die", "lying": "lie", "tying": "tie", "idly": "idl", "gently": "gentl",
    "ugly": "ugli", "early": "earli", "only": "onli", "singly": "singl", "sky": "sky", "news": "news",
    "howe": "howe", "atlas": "atlas", "cosmos": "cosmos", "bias": "bias", "andes": "andes", "inning": "inning",
    "outing": "outing", "canning": "canning", "herring": "herring", "earring": "earring", "proceed": "proceed",
    "exceed": "exceed", "succeed": "succeed", "youth": "youth", "boyish": "boyish", "flying": "fli",
}


def test_snowball_english_known_answers(enc):
    for word, stem in {**SNOWBALL_SAMPLE, **PORTER2_RULES}.items():
        assert enc.stem(word) == stem, word


def test_murmur3_known_answers(enc):
    # mmh3.hash(s) (signed 32-bit, seed 0): "" -> 0, "foo" -> -156908512, "hello" -> 613153351,
    # "hello world" -> 1586663183, "The quick brown fox jumps over the lazy dog" -> 776992547
    for s, h in (("", 0), ("foo", -156908512), ("hello", 613153351), ("hello world", 1586663183),
                 ("The quick brown fox jumps over the lazy dog", 776992547)):
        assert enc.token_id(s) == abs(h)
    lib = enc._lib
    assert lib.brag_murmur3_32(b"foo", 3, 0) == (-156908512) % (1 << 32)
    assert lib.brag_murmur3_32(b"", 0, 1) == 0x514E28B7          # reference vector of the algorithm (seed 1)
    assert lib.brag_murmur3_32(b"\xff\xff\xff\xff", 4, 0) == 0x76293B50


def test_tokenize_matches_fastembed_regexes():
    assert tokenize("Hello, World!  It's a naïve_test—42.") == ["hello", "world", "it", "s", "a", "naïve_test", "42"]
    assert tokenize("") == [] and tokenize(" \n\t ") == []


def test_embed_document_hand_computed(enc):
    text = "Hello, World! The quick brown foxes are running quickly; hello again."
    e = list(enc.embed([text]))[0]
    # kept tokens (stop words the/are/again dropped): hello world quick brown fox run quick hello -> len 8
    stems = ["hello", "world", "quick", "brown", "fox", "run"]
    tf = [2, 1, 2, 1, 1, 1]
    assert e.indices.tolist() == [enc.token_id(s) for s in stems]        # first-occurrence order
    k, b, avg, n = 1.2, 0.75, 256.0, 8
    want = [c * (k + 1) / (c + k * (1 - b + b * n / avg)) for c in tf]
    np.testing.assert_allclose(e.values, want, rtol=0, atol=1e-15)
    assert e.indices.dtype == np.int64 and e.values.dtype == np.float64
    # the reference also encodes QUERIES with .embed (embedding_handler.py:123); query_embed gives all-ones
    q = list(enc.query_embed("running foxes"))[0]
    assert q.indices.tolist() == [enc.token_id("run"), enc.token_id("fox")] and q.values.tolist() == [1.0, 1.0]


def test_filters_and_edge_cases(enc):
    assert len(list(enc.embed([""]))[0].indices) == 0
    assert len(list(enc.embed(["the and of to is"]))[0].indices) == 0          # only stop words
    long_tok = "x" * 41
    e = list(enc.embed([f"{long_tok} {'z' * 40} _"]))[0]                         # > 40 chars dropped, "_" is punctuation
    assert e.indices.tolist() == [enc.token_id("z" * 40)]
    a, b2 = list(enc.embed(["cats", "cat cat cat"]))
    assert a.indices.tolist() == b2.indices.tolist() and b2.values[0] > a.values[0]


def test_round_trip_into_the_client_shapes(enc):
    """The encoder output drops into the reference's SparseVector(indices=..., values=...) construction."""
    from rag_application_b200 import models
    e = list(enc.embed(["GPU kernels scan posting lists"]))[0]
    sv = models.SparseVector(indices=e.indices.tolist(), values=e.values.tolist())
    assert len(sv.indices) == len(sv.values) == 5
