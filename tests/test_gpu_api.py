"""The drop-in Python API (Session / Partition / Index.find / match objects) on the GPU against the
oracle driven with the same token tables, the same similarity matrix and the same options."""

import numpy as np
import pytest

from vectorian_b200 import _cabi, gaps
from vectorian_b200.alignment import GlobalAlignment, SemiGlobalAlignment, LocalAlignment
from vectorian_b200.corpus import Corpus, Document, whitespace_nlp
from vectorian_b200.embedding import KeyedVectors, ContextualEmbedding
from vectorian_b200.session import Session
from vectorian_b200.sim.span import OptimizedSpanSim
from vectorian_b200.sim.token import EmbeddingTokenSim
from vectorian_b200.sim.vector import CosineSim, PNormDistance, ModifiedVectorSim
from vectorian_b200.sim.kernel import Scale, DistanceToSimilarity
from tests import helpers

pytestmark = pytest.mark.gpu

POS = ["NOUN", "VERB", "ADJ", "ADP", "DET"]
TAG = ["NN", "VB", "JJ", "IN", "DT"]


def make_session(n_docs=7, sent_per_doc=60, vocab=400, dim=16, seed=3, contextual_dim=0):
	rng = np.random.default_rng(seed)
	words = [f"w{i}" for i in range(vocab)]
	cdf = np.cumsum(1.0 / np.arange(1, vocab + 1))
	cdf /= cdf[-1]
	docs = []
	for d in range(n_docs):
		sents, pos, tag = [], [], []
		for _ in range(sent_per_doc):
			n = int(np.clip(np.rint(rng.normal(12, 5)), 1, 40))
			ids = np.minimum(np.searchsorted(cdf, rng.random(n)), vocab - 1)
			sents.append([words[i] for i in ids])
			k = rng.integers(0, len(POS), size=n)
			pos.append([POS[i] for i in k])
			tag.append([TAG[i] for i in k])
		ctx = {}
		if contextual_dim:
			n_tok = sum(len(s) for s in sents)
			ctx = {"ctx": rng.standard_normal((n_tok, contextual_dim)).astype(np.float32)}
		docs.append(Document.from_sentences(sents, pos=pos, tag=tag, contextual_embeddings=ctx))
	centroids = rng.standard_normal((30, dim)).astype(np.float32)
	matrix = centroids[rng.integers(0, 30, size=vocab)] + 0.3 * rng.standard_normal((vocab, dim)).astype(np.float32)
	emb = KeyedVectors("toy", (words, matrix))
	embeddings = [emb]
	if contextual_dim:
		crng = np.random.default_rng(seed + 100)
		table = {}
		def encode(tokens):
			return np.array([table.setdefault(t, crng.standard_normal(contextual_dim).astype(np.float32)) for t in tokens])
		embeddings.append(ContextualEmbedding("ctx", encode, contextual_dim))
	return Session(Corpus(docs), embeddings), embeddings


def oracle_corpus(oracle, session, partition, boost=None):
	docs = session.documents
	tok = np.concatenate([d.token_id for d in docs])
	pos = np.concatenate([d.token_pos for d in docs])
	starts, ends, span_off, tok_off = [], [], [0], [0]
	for d in docs:
		s, e = d.span_table(partition.level)
		starts.append(s)
		ends.append(e)
		span_off.append(span_off[-1] + len(s))
		tok_off.append(tok_off[-1] + d.n_tokens)
	return oracle.OracleCorpus(tok, np.array(tok_off), np.concatenate(starts), np.concatenate(ends), np.array(span_off),
		partition.window_size, partition.window_step, token_pos=pos, boost=boost)


def check_find(oracle, session, emb, index, text, n, min_score, s=None, qkw=None, contextual=False, oc=None, options=None):
	from vectorian_b200 import engine
	options = options or {}
	result = index.find(text, n=n, min_score=min_score, options=options)
	query = index.make_query(text, n=n, min_score=min_score, options=options)
	p = query.prepare(whitespace_nlp)
	t = p.n_tokens
	partition = index.partition
	if oc is None:
		oc = oracle_corpus(oracle, session, partition)
	if s is None:
		de, vv = session.device_embedding(emb)
		q_vec = session.to_encoder(emb).encode_tokens(p.token_str).unmodified
		q_id = np.where(p.token_id < session.vocab.size, p.token_id, -1).astype(np.int32)
		s = engine.DeviceSim.static(de, q_id, q_vec).download()     # the same bits vx_search consumed
	a = query.options["metric"]["alignment"]["options"]
	ap, keep = helpers.align_params(a["locality"], a["gap_cost"], session.device_corpus(partition).max_len, t)
	qp, keep2 = _cabi.make_query_params(max_matches=n, min_score=min_score, **(qkw or {}))
	o_all = oc.score_all(s, ap, qp, n_threads=2, contextual=contextual)
	ref = oc.search(s, ap, qp, n_threads=2, contextual=contextual)
	wd, wsl, _, _ = oc.windows()
	hits = np.zeros((len(result),), dtype=_cabi.HIT_DTYPE)
	for i, m in enumerate(result):
		hits[i] = (m.score, m.score_max, m.prepared_doc.id, m.slice_id)
	target = np.array([m.flow["target"] for m in result], dtype=np.int16).reshape(len(result), t)
	dist = np.array([m.flow["dist"] for m in result], dtype=np.float32).reshape(len(result), t)
	helpers.assert_hits_match((hits, target, dist), ref, o_all, wd, wsl)
	return result, p


QUERY = "w3 w17 w5 w120 w1 w9 unknownword w2"


@pytest.mark.parametrize("alignment", [
	LocalAlignment(gap=gaps.smooth_gap_cost(5)), GlobalAlignment(gap=gaps.LinearGapCost(0.2)),
	SemiGlobalAlignment(gap=gaps.AffineGapCost(u=0.3, v=0.1)), LocalAlignment()])
def test_find_matches_oracle(oracle, alignment):
	session, (emb,) = make_session()
	metric = OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), alignment)
	index = session.partition("sentence").index(metric, nlp=whitespace_nlp)
	result, p = check_find(oracle, session, emb, index, QUERY, n=25, min_score=0.05)
	assert len(result) > 0 and result.duration > 0
	m = result[0]
	assert m.metric == "toy-cosine" and m.level == "word"
	assert m.flow["type"] == "injective" and m.flow["target"].dtype == np.int16
	# match objects: regions / omitted / to_json are consistent with the flow
	matched = [j for j, u in enumerate(m.flow["target"]) if u >= 0]
	regions = m.regions(5)
	assert sum(1 for r in regions if r.match) == len(matched)
	assert len(m.omitted) == p.n_tokens - len(matched)
	span = m.doc_span
	for r in regions:
		if r.match:
			e = r.match.edges[0]
			assert span[int(m.flow["target"][e.t.index])].text == r.s
	js = m.to_json()
	assert js["slice"] == m.slice_id and js["score"] == m.score and len(js["regions"]) == len(regions)


def test_windows_and_token_level(oracle):
	session, (emb,) = make_session(n_docs=3, sent_per_doc=40)
	metric = OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), LocalAlignment(gap=gaps.LinearGapCost(0.1)))
	for partition in (session.partition("sentence", 3, 2), session.partition("token", 9, 4)):
		index = partition.index(metric, nlp=whitespace_nlp)
		check_find(oracle, session, emb, index, QUERY, n=10, min_score=0.0)


def test_tag_weighted_and_host_vector_sim(oracle):
	session, (emb,) = make_session()
	ts = EmbeddingTokenSim(emb, CosineSim())
	tag_weights = {"NN": 2.0, "VB": 0.5}
	metric = OptimizedSpanSim(ts, LocalAlignment(gap=gaps.LinearGapCost(0.2)), tag_weights=tag_weights,
		pos_mismatch_penalty=0.3, similarity_threshold=0.2)
	index = session.partition("sentence").index(metric, nlp=whitespace_nlp)
	p = index.make_query(QUERY).prepare(whitespace_nlp)
	qkw = dict(t_weights=np.array([tag_weights.get(x, 1.0) for x in p.token_tag], dtype=np.float32),
		t_pos=p.token_pos, pos_mismatch_penalty=0.3, similarity_threshold=0.2)
	check_find(oracle, session, emb, index, QUERY, n=15, min_score=0.01, qkw=qkw)

	# a VectorSim the device does not evaluate itself: computed on the host, uploaded
	pn = ModifiedVectorSim(PNormDistance(p=2), Scale(0.25), DistanceToSimilarity())
	metric = OptimizedSpanSim(EmbeddingTokenSim(emb, pn), LocalAlignment(gap=gaps.smooth_gap_cost(3)))
	index = session.partition("sentence").index(metric, nlp=whitespace_nlp)
	de, vv = session.device_embedding(emb)
	qv = session.to_encoder(emb).encode_tokens(p.token_str)
	s = np.zeros((session.vocab.size, p.n_tokens), dtype=np.float32)
	pn(vv, qv, s)
	for j, i in enumerate(p.token_id):
		if i < session.vocab.size:
			s[i, j] = 1.0
	np.clip(s, 0, 1, out=s)
	check_find(oracle, session, emb, index, QUERY, n=15, min_score=0.01, s=s)


def test_contextual_embedding(oracle):
	session, (emb, ctx) = make_session(n_docs=3, sent_per_doc=30, contextual_dim=24)
	metric = OptimizedSpanSim(EmbeddingTokenSim(ctx, CosineSim()), GlobalAlignment(gap=gaps.LinearGapCost(0.1)))
	index = session.partition("sentence").index(metric, nlp=whitespace_nlp)
	text = "w3 w17 w5 w120 w1"
	p = index.make_query(text).prepare(whitespace_nlp)
	x = np.concatenate([d.doc.contextual_embeddings["ctx"] for d in session.documents])
	q_vec = ctx.create_encoder(session).encode([[text[t['start']:t['end']] for t in p.tokens]])[0]
	s = oracle.sim_contextual(x, q_vec)
	result = index.find(text, n=10, min_score=-10.0)
	oc = oracle_corpus(oracle, session, index.partition)
	ap, keep = helpers.align_params(_cabi.VX_GLOBAL, gaps.LinearGapCost(0.1), session.device_corpus(index.partition).max_len, p.n_tokens)
	qp, _ = _cabi.make_query_params(max_matches=10, min_score=-10.0)
	ref = oc.search(s, ap, qp, n_threads=2, contextual=True)
	got = np.array([m.score for m in result], dtype=np.float32)
	np.testing.assert_allclose(got, ref[0]["score"], rtol=1e-5, atol=1e-6)


def test_empty_query_and_errors():
	session, (emb,) = make_session(n_docs=2, sent_per_doc=5)
	index = session.partition("sentence").index(session.default_metric(), nlp=whitespace_nlp)
	assert len(index.find("   ")) == 0                       # vectorian/index.py:533-534
	with pytest.raises(ValueError):
		index.find("w1 w2", options={"bogus_option": 1})     # cpp/query.cpp:59-68


def test_sweep_equals_individual_finds():
	"""index.sweep (parameter sweep over alignment settings, one device call) returns what index.find returns
	for every (query, index) pair."""
	from vectorian_b200.index import sweep
	session, (emb,) = make_session(n_docs=5, sent_per_doc=50)
	token_sim = EmbeddingTokenSim(emb, CosineSim())
	partition = session.partition("sentence")
	alignments = [LocalAlignment(gap=gaps.LinearGapCost(u)) for u in (0.0, 0.1, 0.2, 0.5)]
	alignments += [GlobalAlignment(gap=gaps.LinearGapCost(u)) for u in (0.0, 0.1, 0.2, 0.5)]
	alignments += [LocalAlignment(gap=gaps.smooth_gap_cost(5)), SemiGlobalAlignment(gap=gaps.LinearGapCost(0.3))]
	indexes = [partition.index(OptimizedSpanSim(token_sim, a), nlp=whitespace_nlp) for a in alignments]
	texts = [QUERY, "w1 w2 w3", "", "w40 w7 w7 w300 w12"]
	got = sweep(indexes, texts, n=15, min_score=0.01)
	assert len(got) == len(texts) and all(len(row) == len(indexes) for row in got)
	for text, row in zip(texts, got):
		for index, result in zip(indexes, row):
			ref = index.find(text, n=15, min_score=0.01, disable_progress=True)
			assert len(result) == len(ref)
			for m, r in zip(result, ref):
				assert (m.score, m.prepared_doc is r.prepared_doc, m.slice_id) == (r.score, True, r.slice_id)
				np.testing.assert_array_equal(m.flow["target"], r.flow["target"])
				assert m.to_json() == r.to_json()


def test_token_similarity_modifiers(oracle):
	"""Mixed / maximum / unary token-similarity modifiers (vectorian/sim/modifier.py): operand matrices from the
	device, combined on the host as the reference does, searched on the device; compared with the oracle fed with the
	independently combined matrix."""
	from vectorian_b200 import engine
	from vectorian_b200.sim.modifier import MixedTokenSimilarity, MaximumTokenSimilarity, UnaryTokenSimilarityModifier
	from vectorian_b200.sim.kernel import Power, Bias
	session, (emb,) = make_session()
	ts_cos = EmbeddingTokenSim(emb, CosineSim())
	pn = ModifiedVectorSim(PNormDistance(p=2), Scale(0.25), DistanceToSimilarity())
	ts_pn = EmbeddingTokenSim(emb, pn)
	partition = session.partition("sentence")
	p = partition.index(OptimizedSpanSim(ts_cos, LocalAlignment()), nlp=whitespace_nlp).make_query(QUERY).prepare(whitespace_nlp)
	de, vv = session.device_embedding(emb)
	qv = session.to_encoder(emb).encode_tokens(p.token_str)
	q_id = np.where(p.token_id < session.vocab.size, p.token_id, -1).astype(np.int32)
	s_cos = engine.DeviceSim.static(de, q_id, qv.unmodified).download()
	s_pn = np.zeros_like(s_cos)
	pn(vv, qv, s_pn)
	for j, i in enumerate(q_id):
		if i >= 0:
			s_pn[i, j] = 1.0
	np.clip(s_pn, 0, 1, out=s_pn)

	cases = [
		(MixedTokenSimilarity([ts_cos, ts_pn], [3, 1]), np.average([s_cos, s_pn], axis=0, weights=[3, 1]).astype(np.float32),
			"(0.75 * cosine[toy] + 0.25 * "),
		(MaximumTokenSimilarity([ts_cos, ts_pn]), np.maximum(s_cos, s_pn), "maximum(cosine[toy], "),
		(UnaryTokenSimilarityModifier(ts_cos, [Power(2.0), Bias(0.05)]), (s_cos ** 2.0 + 0.05).astype(np.float32), None),
	]
	for token_sim, s, name_prefix in cases:
		assert token_sim.is_modifier
		if name_prefix:
			assert token_sim.name.startswith(name_prefix)
		index = partition.index(OptimizedSpanSim(token_sim, LocalAlignment(gap=gaps.LinearGapCost(0.15))), nlp=whitespace_nlp)
		result, _ = check_find(oracle, session, emb, index, QUERY, n=12, min_score=0.01, s=s)
		assert len(result) > 0 and result[0].metric == token_sim.name


def test_token_filters(oracle):
	"""pos_filter / tag_filter query options: the document tokens that do not pass are dropped from every window before
	the alignment (FilteredSliceFactory, cpp/slice/static.h:365-416; TokenFilter::pass, cpp/query.h:22-28).  The oracle
	runs on a corpus filtered independently here, token by token."""
	session, (emb,) = make_session(n_docs=5, sent_per_doc=50)
	options = {"pos_filter": ["DET", "ADP"], "tag_filter": ["JJ"]}
	drop_pos = {session.vocab.pos.get(x) for x in options["pos_filter"]}
	drop_tag = {session.vocab.tag.get(x) for x in options["tag_filter"]}
	tok, pos, starts, ends, span_off, tok_off = [], [], [], [], [0], [0]
	index_maps = {}
	for d in session.documents:
		s0, e0 = d.span_table("sentence")
		n_doc_tok = 0
		for k in range(len(s0)):
			keep = [i for i in range(s0[k], e0[k]) if d.token_pos[i] not in drop_pos and d.token_tag[i] not in drop_tag]
			index_maps[(d.id, k)] = [i - s0[k] for i in keep]
			starts.append(n_doc_tok)
			n_doc_tok += len(keep)
			ends.append(n_doc_tok)
			tok.extend(int(d.token_id[i]) for i in keep)
			pos.extend(int(d.token_pos[i]) for i in keep)
		span_off.append(span_off[-1] + len(s0))
		tok_off.append(tok_off[-1] + n_doc_tok)
	assert len(tok) < sum(d.n_tokens for d in session.documents) * 0.7
	for alignment, part_args in ((LocalAlignment(gap=gaps.smooth_gap_cost(5)), ("sentence",)),
			(LocalAlignment(gap=gaps.AffineGapCost(u=0.2, v=0.1)), ("sentence", 3, 2))):
		partition = session.partition(*part_args)
		oc = oracle.OracleCorpus(np.array(tok, np.int32), np.array(tok_off), np.array(starts, np.int32), np.array(ends, np.int32),
			np.array(span_off), partition.window_size, partition.window_step, token_pos=np.array(pos, np.int8))
		metric = OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), alignment)
		index = partition.index(metric, nlp=whitespace_nlp)
		result, p = check_find(oracle, session, emb, index, QUERY, n=20, min_score=0.05, oc=oc, options=options)
		assert len(result) > 0
		# unfiltered search on the same index afterwards: a different device corpus, different hits
		plain = index.find(QUERY, n=20, min_score=0.05)
		check_find(oracle, session, emb, index, QUERY, n=20, min_score=0.05)
		if len(part_args) == 1:
			assert [(m.prepared_doc.id, m.slice_id, m.score) for m in plain] != [(m.prepared_doc.id, m.slice_id, m.score) for m in result]
			# regions(): matched document tokens are reported at their UNFILTERED positions (cpp/match/flow.cpp:49-61, 95-99),
			# exactly as the reference's Flow::py_regions does it -- including its grouping of edges at flow.cpp:143, which
			# compares a raw (filtered) target with a mapped one (oracle/render.py restates it; tests/test_render_oracle.py)
			from oracle import render
			for m in result[:5]:
				doc = m.prepared_doc.doc
				s0, s1 = m.prepared_doc.span_table("sentence")
				imap = index_maps[(m.prepared_doc.id, m.slice_id)]
				token_at, slice_len = int(s0[m.slice_id]), int(s1[m.slice_id]) - int(s0[m.slice_id])
				keep = [i in set(int(x) for x in imap) for i in range(slice_len)]
				s_tokens = [{'idx': int(doc.token_idx[i]), 'len': int(doc.token_len[i]), 'pos': doc.token_pos[i]} for i in range(doc.n_tokens)]
				t_tokens = [{'idx': t['start'], 'len': t['end'] - t['start'], 'pos': t.get('pos', 'X')} for t in m.query.tokens]
				edges = [(int(u), j, 1.0, float(m.flow["dist"][j])) for j, u in enumerate(m.flow["target"]) if u >= 0]
				gap_s, gap_t = m._gap_fns
				want = render.render(doc.text_str, m.query.text_str,
					render.py_regions(s_tokens, t_tokens, token_at, slice_len, edges, 3, filter_pass=keep, gap_cost_s=gap_s, gap_cost_t=gap_t), m.metric)
				got = [(r.s, None if r.match is None else (r.match.pos_s, [((e.t.text, e.t.index, e.t.pos), e.flow, e.distance, e.metric) for e in r.match.edges]),
					r.gap_penalty) for r in m.regions(3)]
				assert got == want
				# every matched region shows a document token that passed the filter
				passed = {doc.token_text[token_at + int(i)] for i in imap}
				assert all(r.s in passed for r in m.regions(3) if r.match)
	with pytest.raises(RuntimeError):
		index.find(QUERY, options={"pos_filter": ["NOPE"]})


def test_contextual_operands_in_modifiers(oracle):
	"""A modifier with a contextual operand is a CONTEXTUAL metric (cpp/query.cpp:156-215): every operand delivers a
	per-token matrix [n_tokens x T] -- contextual vectors through their VectorSim + clip (cpp/metric/contextual.cpp:26-63),
	static embeddings as the rows S[id(token)] (cpp/metric/static.cpp:154-189) -- and the modifier combines them
	(cpp/metric/modifier.cpp:18-74).  The oracle gets matrices assembled independently here; same for a contextual
	embedding under a VectorSim the device does not evaluate itself, and for a token filter on top."""
	from vectorian_b200 import engine
	from vectorian_b200.embedding import Vectors
	from vectorian_b200.sim.modifier import (MixedTokenSimilarity, MaximumTokenSimilarity, MinimumTokenSimilarity,
		UnaryTokenSimilarityModifier)
	from vectorian_b200.sim.kernel import Power, Threshold
	session, (emb, ctx) = make_session(n_docs=3, sent_per_doc=30, contextual_dim=24)
	a = EmbeddingTokenSim(emb, CosineSim())
	b = EmbeddingTokenSim(ctx, CosineSim())
	pn = ModifiedVectorSim(PNormDistance(p=2), Scale(0.1), DistanceToSimilarity())
	partition = session.partition("sentence")
	text = "w3 w17 w5 w120 w1 w9"
	probe = partition.index(OptimizedSpanSim(a, LocalAlignment()), nlp=whitespace_nlp)
	p = probe.make_query(text).prepare(whitespace_nlp)
	de, vv = session.device_embedding(emb)
	q_id = np.where(p.token_id < session.vocab.size, p.token_id, -1).astype(np.int32)
	qv = session.to_encoder(emb).encode_tokens(p.token_str)
	sim = engine.DeviceSim.static(de, q_id, qv.unmodified)
	s_static = sim.download()
	sim.close()
	q_ctx = session.to_encoder(ctx).encode([[p.text_str[t['start']:t['end']] for t in p.tokens]])[0]

	def matrices(keep):
		tok = np.concatenate([d.token_id for d in session.documents])[keep]
		x = np.concatenate([d.doc.contextual_embeddings["ctx"] for d in session.documents])[keep]
		sa = s_static[tok]
		sb = np.zeros((x.shape[0], q_ctx.shape[0]), dtype=np.float32)
		CosineSim()(Vectors(x), Vectors(q_ctx), sb)             # the host's VectorSim, as the reference's callback
		np.clip(sb, 0, 1, out=sb)
		np.testing.assert_allclose(sb, oracle.sim_contextual(x, q_ctx), rtol=0, atol=1e-5)
		sp = np.zeros_like(sb)
		pn(Vectors(x), Vectors(q_ctx), sp)
		np.clip(sp, 0, 1, out=sp)
		return sa, sb, sp

	n_tok = sum(d.n_tokens for d in session.documents)
	sa, sb, sp = matrices(np.ones(n_tok, dtype=bool))
	cases = [
		(MixedTokenSimilarity([a, b], [1, 3]), np.average([sa, sb], axis=0, weights=[1, 3]).astype(np.float32)),
		(MaximumTokenSimilarity([a, b]), np.maximum(sa, sb)),
		(UnaryTokenSimilarityModifier(MinimumTokenSimilarity([b, a]), [Power(2.0), Threshold(0.1)]), None),
		(EmbeddingTokenSim(ctx, pn), sp),
	]
	for token_sim, s in cases:
		if s is None:
			stack = np.array([sb, sa])
			s = np.take_along_axis(stack, np.argmax(stack, axis=0)[None], axis=0)[0]   # sim/modifier.py:87-93 (argmax for both)
			s = np.power(np.maximum(s, 0), 2.0)
			s = np.maximum(0, s - 0.1)
			s[s > 0] += 0.1                                     # sim/kernel.py Threshold
		for alignment in (LocalAlignment(gap=gaps.LinearGapCost(0.15)), GlobalAlignment(gap=gaps.smooth_gap_cost(4))):
			index = partition.index(OptimizedSpanSim(token_sim, alignment), nlp=whitespace_nlp)
			result, _ = check_find(oracle, session, emb, index, text, n=15, min_score=-1.0, s=np.asarray(s, np.float32), contextual=True)
			assert len(result) == 15
			assert result[0].metric == (token_sim.name if token_sim.is_modifier else token_sim.to_args(index)['name'])

	# with a token filter: rows of the filtered token table
	options = {"pos_filter": ["DET"]}
	det = session.vocab.pos.get("DET")
	keep = np.concatenate([d.token_pos for d in session.documents]) != det
	sa, sb, sp = matrices(keep)
	tok, pos, starts, ends, span_off, tok_off = [], [], [], [], [0], [0]
	for d in session.documents:
		k = d.token_pos != det
		cum = np.concatenate([[0], np.cumsum(k)])
		s0, e0 = d.span_table("sentence")
		tok.append(d.token_id[k]); pos.append(d.token_pos[k]); starts.append(cum[s0]); ends.append(cum[e0])
		span_off.append(span_off[-1] + len(s0)); tok_off.append(tok_off[-1] + int(k.sum()))
	oc = oracle.OracleCorpus(np.concatenate(tok), np.array(tok_off), np.concatenate(starts), np.concatenate(ends),
		np.array(span_off), 1, 1, token_pos=np.concatenate(pos))
	index = partition.index(OptimizedSpanSim(MixedTokenSimilarity([a, b], [1, 1]), LocalAlignment(gap=gaps.LinearGapCost(0.15))), nlp=whitespace_nlp)
	check_find(oracle, session, emb, index, text, n=15, min_score=0.0, s=np.average([sa, sb], axis=0, weights=[1, 1]).astype(np.float32), contextual=True,
		oc=oc, options=options)


def test_long_query_through_find(oracle):
	"""A 21-token query through Index.find: the similarity kernel takes 16 query columns per launch, so S is assembled from
	two device launches and the windows are scored by the exact per-window kernel; same answers as the oracle."""
	from vectorian_b200 import engine
	session, (emb,) = make_session(n_docs=4, sent_per_doc=40)
	text = " ".join(f"w{i}" for i in (3, 17, 5, 120, 1, 9, 2, 44, 8, 13, 21, 34, 55, 89, 144, 233, 377, 6, 10, 15, 28))
	metric = OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), LocalAlignment(gap=gaps.LinearGapCost(0.2)))
	index = session.partition("sentence").index(metric, nlp=whitespace_nlp)
	p = index.make_query(text).prepare(whitespace_nlp)
	assert p.n_tokens == 21
	de, vv = session.device_embedding(emb)
	q_vec = session.to_encoder(emb).encode_tokens(p.token_str).unmodified
	q_id = np.where(p.token_id < session.vocab.size, p.token_id, -1).astype(np.int32)
	s = index._device_static_matrix(de, q_id, q_vec)
	assert s.shape == (session.vocab.size, 21)
	np.testing.assert_allclose(s, oracle.sim_static(vv.unmodified, q_id, q_vec), rtol=0, atol=1e-5)
	result, _ = check_find(oracle, session, emb, index, text, n=10, min_score=0.01, s=s)
	assert len(result) == 10


def test_debug_hook_reports_matrix_and_every_slice():
	"""The `debug` option (cpp/query.cpp:73-74): one "static_similarity_matrix" call, then one "alignment" call per slice
	with {slice, similarity, flow, score} (cpp/metric/alignment.h:144-173) -- the flows of the slices that make the result
	equal the result's, the raw score is score * score_max."""
	session, (emb,) = make_session(n_docs=3, sent_per_doc=25)
	index = session.partition("sentence").index(
		OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), LocalAlignment(gap=gaps.smooth_gap_cost(5))), nlp=whitespace_nlp)
	calls = []
	result = index.find("w3 w17 w5 w9", n=10, min_score=0.0, debug=lambda name, data: calls.append((name, data)))
	names = [c[0] for c in calls]
	assert names[0] == "static_similarity_matrix" and names.count("static_similarity_matrix") == 1
	m = calls[0][1]
	assert m["similarity"].shape == (session.vocab.size, 4) and m["columns"]() == ["w3", "w17", "w5", "w9"]
	assert len(m["rows"]()) == session.vocab.size
	per_slice = [c[1] for c in calls if c[0] == "alignment"]
	assert len(per_slice) == 3 * 25          # every slice of every document, slice ids ascending per document
	assert [d["slice"] for d in per_slice[:25]] == list(range(25))
	by_key = {}
	doc = -1
	for d in per_slice:
		if d["slice"] == 0:
			doc += 1
		by_key[(doc, d["slice"])] = d
	assert len(result) > 0
	for match in result:
		d = by_key[(match.prepared_doc.id, match.slice_id)]
		np.testing.assert_array_equal(d["flow"]["target"], match.flow["target"])
		assert d["similarity"].shape[1] == 4 and d["similarity"].shape[0] >= 1
		np.testing.assert_allclose(d["score"], match.score * match.score_max, rtol=1e-6)
	big = session.partition("sentence").index(
		OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), LocalAlignment()), nlp=whitespace_nlp)
	big.DEBUG_MAX_WINDOWS = 10
	with pytest.raises(NotImplementedError, match="debug hook"):
		big.find("w3 w17", debug=lambda name, data: None)


def test_window_emptied_by_a_token_filter_is_not_reported():
	"""Documented deviation: the token filter is a property of the packed corpus, so a window whose
	tokens are ALL filtered has no tokens left and is never reported -- the reference aligns the empty window, scores it 0
	and admits it when min_score < 0 (cpp/slice/static.h:365-416).  With min_score >= 0 (every reference caller: Index.find
	passes 0.0) the results are identical; this test pins the documented behaviour for min_score < 0."""
	sents = [["w1", "w2", "w3"], ["w4", "w5"], ["w6", "w1", "w2"]]
	pos = [["NOUN", "VERB", "NOUN"], ["DET", "DET"], ["NOUN", "DET", "VERB"]]
	tag = [["NN", "VB", "NN"], ["DT", "DT"], ["NN", "DT", "VB"]]
	words = [f"w{i}" for i in range(1, 8)]
	rng = np.random.default_rng(0)
	emb = KeyedVectors("toy", (words, rng.standard_normal((len(words), 8)).astype(np.float32)))
	session = Session(Corpus([Document.from_sentences(sents, pos=pos, tag=tag)]), [emb])
	index = session.partition("sentence").index(
		OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), LocalAlignment()), nlp=whitespace_nlp)
	plain = index.find("w1 w2", n=10, min_score=-1.0)
	assert sorted(m.slice_id for m in plain) == [0, 1, 2]
	filtered = index.find("w1 w2", n=10, min_score=-1.0, options={"pos_filter": ["DET"]})
	assert sorted(m.slice_id for m in filtered) == [0, 2]          # slice 1 has only DET tokens
	assert [m.slice_id for m in index.find("w1 w2", n=10, min_score=0.0, options={"pos_filter": ["DET"]}) if m.slice_id == 1] == []
