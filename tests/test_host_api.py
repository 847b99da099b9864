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
