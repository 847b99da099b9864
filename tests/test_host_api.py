"""Host-side mirror of the reference's Python API (no GPU needed): signatures, option dicts,
vocabulary / query id assignment, window tables, sharding."""

import numpy as np
import pytest

from vectorian_b200 import alignment, gaps, _cabi
from vectorian_b200.alignment import GlobalAlignment, SemiGlobalAlignment, LocalAlignment
from vectorian_b200.corpus import Corpus, Document, whitespace_nlp
from vectorian_b200.embedding import KeyedVectors, Vectors
from vectorian_b200.index import BruteForceIndex, Index, PreparedQuery, Query
from vectorian_b200.session import Session, Partition
from vectorian_b200.sim.span import OptimizedSpanSim
from vectorian_b200.sim.token import EmbeddingTokenSim
from vectorian_b200.sim.vector import CosineSim, PNormDistance, ModifiedVectorSim
from vectorian_b200.sim.kernel import Scale, DistanceToSimilarity


def tiny_session():
	docs = [
		Document.from_sentences([["the", "cat", "sat"], ["on", "the", "mat"]]),
		Document.from_sentences([["a", "dog", "barked"], ["the", "end"]]),
	]
	words = ["the", "cat", "sat", "on", "mat", "a", "dog", "barked", "end", "feline"]
	rng = np.random.default_rng(0)
	emb = KeyedVectors("toy", (words, rng.standard_normal((len(words), 8)).astype(np.float32)))
	return Session(Corpus(docs), [emb]), emb


def test_alignment_to_args_matches_reference_shape():
	# vectorian/alignment.py:31-38, 96-97, 129-130, 186-187
	for cls, loc in ((GlobalAlignment, _cabi.VX_GLOBAL), (SemiGlobalAlignment, _cabi.VX_SEMIGLOBAL), (LocalAlignment, _cabi.VX_LOCAL)):
		g = gaps.LinearGapCost(0.2)
		a = cls(gap=g).to_args(None)
		assert a == {'algorithm': 'pyalign', 'options': {'locality': loc, 'gap_cost': g}}
	assert isinstance(LocalAlignment().gap, gaps.ConstantGapCost)        # coalesce_default_gap
	with pytest.raises(ValueError):
		LocalAlignment(gap={'x': gaps.LinearGapCost(1)})
	with pytest.raises(NotImplementedError):
		alignment.WordMoversDistance()


def test_solver_selection():
	# pyalign/algorithm/factory.h:930-965
	assert gaps.resolve_gap_costs(None, 10, 5)['gap_kind'] == _cabi.VX_GAP_LINEAR
	assert gaps.resolve_gap_costs(gaps.ConstantGapCost(0), 10, 5)['gap_kind'] == _cabi.VX_GAP_LINEAR
	assert gaps.resolve_gap_costs(gaps.LinearGapCost(0.3), 10, 5)['lin'] == (0.3, 0.3)
	r = gaps.resolve_gap_costs(gaps.AffineGapCost(u=0.5, v=0.1), 10, 5)
	assert r['gap_kind'] == _cabi.VX_GAP_AFFINE and r['affine_s'] == (0.5, 0.1)
	r = gaps.resolve_gap_costs(gaps.smooth_gap_cost(5), 10, 5)
	assert r['gap_kind'] == _cabi.VX_GAP_GENERAL and len(r['w_s']) == 11 and len(r['w_t']) == 6
	assert r['w_s'][0] == 0 and abs(r['w_s'][5] - 0.5) < 1e-6
	r = gaps.resolve_gap_costs({'s': gaps.LinearGapCost(0.2), 't': gaps.ConstantGapCost(0.3)}, 4, 3)
	assert r['gap_kind'] == _cabi.VX_GAP_GENERAL
	np.testing.assert_allclose(r['w_s'], [0, 0.2, 0.4, 0.6, 0.8], rtol=1e-6)


def test_vectors_normalized():
	# vectorian/embedding/vectors.py:71-86: zero rows stay zero, others unit length
	v = Vectors(np.array([[3, 4], [0, 0], [1, 0]], dtype=np.float32))
	np.testing.assert_allclose(v.magnitudes, [5, 0, 1])
	np.testing.assert_allclose(v.normalized, [[0.6, 0.8], [0, 0], [1, 0]], rtol=1e-6)
	out = np.zeros((3, 3), dtype=np.float32)
	CosineSim()(v, v, out)
	np.testing.assert_allclose(np.diag(out), [1, 0, 1], rtol=1e-6)
	assert CosineSim().on_device and not PNormDistance().on_device


def test_session_vocabulary_and_query_ids():
	session, emb = tiny_session()
	# ids in insertion order while documents load (cpp/vocabulary.h:90-105)
	assert session.vocab.tokens.strings == ["the", "cat", "sat", "on", "mat", "a", "dog", "barked", "end"]
	assert session.documents[1].token_id.tolist() == [5, 6, 7, 0, 8]
	index = session.partition("sentence").index(session.default_metric(), nlp=whitespace_nlp)
	assert isinstance(index, BruteForceIndex) and isinstance(index, Index)
	q = index.make_query("the feline sat, feline!", n=3, min_score=0.1)
	assert q.options["max_matches"] == 3 and q.options["min_score"] == 0.1
	assert q.options["partition"] == {'level': 'sentence', 'window_size': 1, 'window_step': 1}
	assert q.options["metric"]["metric"] == 'alignment-isolated'
	p = q.prepare(whitespace_nlp)
	assert isinstance(p, PreparedQuery)
	assert p.token_str == ["the", "feline", "sat", ",", "feline", "!"]
	# query-only tokens get ids |V|, |V|+1, ... (IncrementalLexicon, cpp/vocabulary.h:152-157)
	assert p.token_id.tolist() == [0, 9, 2, 10, 9, 11]
	assert [t.text for t in p.span] == ["the", "feline", "sat", ",", "feline", "!"]


def test_span_sim_args_and_errors():
	session, emb = tiny_session()
	ts = EmbeddingTokenSim(emb, CosineSim())
	assert ts.to_args(None) == {'name': 'toy-cosine', 'embedding': 'toy', 'metric': ts.similarity}
	with pytest.raises(TypeError):
		OptimizedSpanSim("cosine")
	with pytest.raises(TypeError):
		session.partition("sentence").index("not a SpanSim")
	sim = OptimizedSpanSim(ts, tag_weights={'NN': 2.0}, pos_mismatch_penalty=0.5, similarity_threshold=0.1)
	index = session.partition("sentence").index(sim)
	a = sim.to_args(index)
	assert a['metric'] == 'alignment-tag-weighted' and a['pos_mismatch_penalty'] == 0.5 and a['tag_weights'] == {'NN': 2.0}
	with pytest.raises(ValueError):
		OptimizedSpanSim(ts, bogus=1).to_args(index)


def test_partition_and_windows():
	session, _ = tiny_session()
	p = session.partition("sentence", 2)
	assert isinstance(p, Partition) and p.window_step == 2 and p.contiguous
	assert session.partition("sentence", 1, 2).contiguous is False
	assert p.slice_id_to_slice(3) == ("sentence", 6, 2)
	assert session.max_len("sentence", 1) == 3 and session.max_len("sentence", 2) == 6
	assert session.max_len("token", 3) == 3
	d = session.documents[0]
	assert d.n_spans(session.partition("sentence")) == 2
	assert d.span(session.partition("sentence"), 1).text == "on the mat"
	assert d.span_info(session.partition("sentence"), 1) == {'start': 3, 'end': 6}


def test_token_similarity_modifiers_host():
	"""vectorian/sim/modifier.py: names and the host-side combination of operand matrices."""
	from vectorian_b200.sim.modifier import (MixedTokenSimilarity, MaximumTokenSimilarity, MinimumTokenSimilarity,
		UnaryTokenSimilarityModifier)
	from vectorian_b200.sim.kernel import Scale, Bias
	from vectorian_b200.sim.token import TokenSim

	class Named(TokenSim):
		def __init__(self, name):
			self._n = name
		name = property(lambda self: self._n)
		embeddings = property(lambda self: [self._n + "-emb"])

	a, b = Named("a"), Named("b")
	rng = np.random.default_rng(0)
	sa, sb = rng.random((6, 3), dtype=np.float32), rng.random((6, 3), dtype=np.float32)
	ops = [{"similarity": sa}, {"similarity": sb}]

	mixed = MixedTokenSimilarity([a, b], [1, 3])
	out = {"similarity": np.zeros_like(sa)}
	mixed(ops, out)
	np.testing.assert_allclose(out["similarity"], 0.25 * sa + 0.75 * sb, rtol=1e-6)
	assert mixed.name == "(0.25 * a + 0.75 * b)" and mixed.is_modifier and mixed.embeddings == ["a-emb", "b-emb"]

	for cls, nm in ((MaximumTokenSimilarity, "maximum"), (MinimumTokenSimilarity, "minimum")):
		m = cls([a, b])
		out = {"similarity": np.zeros_like(sa)}
		m(ops, out)
		np.testing.assert_array_equal(out["similarity"], np.maximum(sa, sb))   # the reference uses argmax for both
		assert m.name == f"{nm}(a, b)"

	u = UnaryTokenSimilarityModifier(a, [Scale(0.5), Bias(0.1)])
	out = {"similarity": np.zeros_like(sa)}
	u(ops[:1], out)
	np.testing.assert_allclose(out["similarity"], sa * 0.5 + 0.1, rtol=1e-6)
	assert u.operands == [a] and "a" in u.name
	with pytest.raises(ValueError):
		MixedTokenSimilarity([a, b], [1])


def test_token_filter_tables():
	"""pos_filter / tag_filter (cpp/query.cpp:220-229, cpp/query.h:22-53, cpp/slice/static.h:365-416): bit masks over
	the session's pos / tag ids; the filtered token table is, window by window, the list of passing tokens."""
	docs = [
		Document.from_sentences([["the", "cat", "sat"], ["on", "the", "mat"], ["end"]],
			pos=[["DET", "NOUN", "VERB"], ["ADP", "DET", "NOUN"], ["NOUN"]],
			tag=[["DT", "NN", "VBD"], ["IN", "DT", "NN"], ["NN"]]),
		Document.from_sentences([["a", "dog"], ["the"]], pos=[["DET", "NOUN"], ["DET"]], tag=[["DT", "NN"], ["DT"]]),
	]
	session = Session(Corpus(docs), [])
	assert session.token_filter({}) is None
	assert session.token_filter({"pos_filter": []}) is None
	with pytest.raises(RuntimeError, match="illegal value FOO for pos_filter"):
		session.token_filter({"pos_filter": ["FOO"]})
	tf = session.token_filter({"pos_filter": ["DET"], "tag_filter": ["VBD"]})
	assert tf == (1 << session.vocab.pos.get("DET"), 1 << session.vocab.tag.get("VBD"))
	part = session.partition("sentence")
	tb = session._filtered_tables(part, tf)
	strings = session.vocab.tokens.strings
	words = [strings[i] for i in tb["token_id"]]
	assert words == ["cat", "on", "mat", "end", "dog"]
	np.testing.assert_array_equal(tb["doc_tok_off"], [0, 4, 5])
	np.testing.assert_array_equal(tb["doc_span_off"], [0, 3, 5])
	np.testing.assert_array_equal(tb["span_start"], [0, 1, 3, 0, 1])
	np.testing.assert_array_equal(tb["span_end"], [1, 3, 4, 1, 1])      # the last sentence of doc 1 is emptied by the filter
	plain = session._filtered_tables(part, None)
	assert len(plain["token_id"]) == 10 and plain["keep"] == [None, None]
	# per window: exactly FilteredSliceFactory::create_slice's compaction
	for d, doc in enumerate(session.documents):
		keep = session.filter_pass(doc.token_pos, doc.token_tag, tf)
		s0, e0 = doc.span_table("sentence")
		for k in range(len(s0)):
			want = [int(doc.token_id[i]) for i in range(s0[k], e0[k]) if keep[i]]
			lo = tb["doc_tok_off"][d] + tb["span_start"][tb["doc_span_off"][d] + k]
			hi = tb["doc_tok_off"][d] + tb["span_end"][tb["doc_span_off"][d] + k]
			assert list(tb["token_id"][lo:hi]) == want


def test_host_token_matrix_mixes_static_and_contextual():
	"""Per-token matrices of a CONTEXTUAL modifier (cpp/query.cpp:156-215, cpp/metric/static.cpp:154-189,
	cpp/metric/contextual.cpp:26-63, cpp/metric/modifier.cpp:18-74), host side only: a static operand contributes the rows
	S[id(token)], a contextual one clip(VectorSim(X, Q)); a token filter selects the rows of the filtered token table."""
	from vectorian_b200.embedding import ContextualEmbedding
	from vectorian_b200.sim.modifier import MixedTokenSimilarity
	rng = np.random.default_rng(1)
	sents = [["the", "cat", "sat"], ["on", "the", "mat"]]
	pos = [["DET", "NOUN", "VERB"], ["ADP", "DET", "NOUN"]]
	x = rng.standard_normal((6, 5)).astype(np.float32)
	doc = Document.from_sentences(sents, pos=pos, contextual_embeddings={"ctx": x})
	words = ["the", "cat", "sat", "on", "mat"]
	emb = KeyedVectors("toy", (words, rng.standard_normal((5, 8)).astype(np.float32)))
	table = {}
	ctx = ContextualEmbedding("ctx", lambda toks: np.array([table.setdefault(t, rng.standard_normal(5).astype(np.float32)) for t in toks]), 5)
	session = Session(Corpus([doc]), [emb, ctx])
	a = EmbeddingTokenSim(emb, PNormDistance(p=2))          # evaluated on the host (no device needed)
	b = EmbeddingTokenSim(ctx, CosineSim())
	mixed = MixedTokenSimilarity([a, b], [1, 1])
	index = session.partition("sentence").index(OptimizedSpanSim(mixed, LocalAlignment()), nlp=whitespace_nlp)
	assert index._is_contextual(mixed) and index._is_contextual(b) and not index._is_contextual(a)
	p = index.make_query("cat mat").prepare(whitespace_nlp)
	sa = index._host_matrix(p, a)
	assert sa.shape == (session.vocab.size, 2)
	got = index._host_token_matrix(p, mixed, None)
	q = np.array([table["cat"], table["mat"]])
	sb = np.clip((x / np.linalg.norm(x, axis=1, keepdims=True)) @ (q / np.linalg.norm(q, axis=1, keepdims=True)).T, 0, 1)
	ids = np.concatenate([d.token_id for d in session.documents])
	np.testing.assert_allclose(got, (sa[ids] + sb) / 2, rtol=1e-6, atol=1e-6)
	tf = session.token_filter({"pos_filter": ["DET"]})
	keep = np.array([t != "DET" for s in pos for t in s])
	got_f = index._host_token_matrix(p, mixed, tf)
	np.testing.assert_allclose(got_f, got[keep], rtol=0, atol=0)
