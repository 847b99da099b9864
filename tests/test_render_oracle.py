"""Result rendering: CoreMatch.regions / omitted / to_json against oracle/render.py, the cited restatement
of Flow::py_regions / py_omitted (reference: vectorian/core/cpp/match/flow.cpp:8-188, region.h:8-157,
vectorian/index.py:249-292, 329-369).  Runs without a GPU: matches are built from synthetic alignments."""

import numpy as np
import pytest

from oracle import ref, render
from vectorian_b200 import gaps
from vectorian_b200._cabi import HIT_DTYPE
from vectorian_b200.alignment import LocalAlignment
from vectorian_b200.corpus import Corpus, Document, whitespace_nlp
from vectorian_b200.embedding import KeyedVectors
from vectorian_b200.index import CoreMatch, PartitionData, _gap_cost_fn
from vectorian_b200.session import Session
from vectorian_b200.sim.span import OptimizedSpanSim
from vectorian_b200.sim.token import EmbeddingTokenSim
from vectorian_b200.sim.vector import CosineSim

POS = ["NOUN", "VERB", "ADJ", "ADP", "DET"]
TAG = ["NN", "VB", "JJ", "IN", "DT"]


def make_session(seed=5, n_docs=3, sent_per_doc=12, vocab=60):
	rng = np.random.default_rng(seed)
	words = [f"w{i}" + "x" * int(rng.integers(0, 4)) for i in range(vocab)]      # tokens of different lengths
	docs = []
	for _ in range(n_docs):
		sents, pos, tag = [], [], []
		for _ in range(sent_per_doc):
			n = int(rng.integers(1, 14))
			ids = rng.integers(0, vocab, size=n)
			sents.append([words[i] for i in ids])
			k = rng.integers(0, len(POS), size=n)
			pos.append([POS[i] for i in k])
			tag.append([TAG[i] for i in k])
		docs.append(Document.from_sentences(sents, pos=pos, tag=tag))
	emb = KeyedVectors("toy", (words, rng.standard_normal((vocab, 8)).astype(np.float32)))
	return Session(Corpus(docs), [emb]), emb, words


def check_against_reference_flow(session, pdoc, doc, p_query, token_at, slice_len, target, dist, context, token_filter, gap_fns, want_regions, want_omitted):
	"""oracle/render.py against the reference's OWN Flow::py_regions / py_omitted (oracle/ref/flow_driver.cpp: the unmodified
	match/flow.cpp, match/match.h, match/region.h): slices, penalties, matched flags, edges, POS of every region."""
	pos_id = lambda name: POS.index(name) if name in POS else 99
	n = doc.n_tokens
	t_tok = p_query.tokens
	gap_s = [float(gap_fns[0](k)) for k in range(n + 2)]
	gap_t = [float(gap_fns[1](k)) for k in range(len(t_tok) + 2)]
	got, got_omitted = ref.regions(
		[int(doc.token_idx[i]) for i in range(n)], [int(doc.token_len[i]) for i in range(n)],
		[int(pdoc.token_pos[i]) for i in range(n)], [int(pdoc.token_tag[i]) for i in range(n)],
		[t['start'] for t in t_tok], [t['end'] - t['start'] for t in t_tok], [pos_id(t.get('pos', 'X')) for t in t_tok],
		token_at, slice_len, target, dist, context, token_filter, gap_s, gap_t)
	assert len(got) == len(want_regions)
	for g, w in zip(got, want_regions):
		assert g['s'] == tuple(w['s']) and g['matched'] == w['matched']
		assert np.float32(g['gap_penalty']) == np.float32(w['gap_penalty'])
		if g['matched']:
			assert g['pos_s'] == session._vocab.pos.get(w['pos_s'])
			assert len(g['edges']) == len(w['edges'])
			for ge, we in zip(g['edges'], w['edges']):
				assert ge['t_slice'] == tuple(we['t_slice']) and ge['t_index'] == we['t_index']
				assert np.float32(ge['flow']) == np.float32(we['flow']) and np.float32(ge['distance']) == np.float32(we['distance'])
				assert ge['t_pos'] == pos_id(we['t_pos'])
	assert got_omitted == [tuple(x) for x in want_omitted]
	return 1


def random_target(rng, n_s, n_t, p_match):
	"""A monotone injective alignment: target[j] strictly increasing over the matched query tokens, -1 elsewhere."""
	target = np.full((n_t,), -1, dtype=np.int16)
	u = -1
	for j in range(n_t):
		if rng.random() < p_match and u + 1 < n_s:
			u = int(rng.integers(u + 1, min(n_s, u + 4)))
			target[j] = u
	return target


@pytest.mark.parametrize("window,step,filtered,gap", [
	(1, 1, False, gaps.LinearGapCost(0.2)), (1, 1, True, gaps.smooth_gap_cost(3)), (3, 2, False, gaps.smooth_gap_cost(5)),
	(2, 1, True, gaps.LinearGapCost(0.0)), (1, 1, False, None)])
def test_regions_omitted_to_json_equal_the_reference_restatement(window, step, filtered, gap):
	session, emb, words = make_session()
	partition = session.partition("sentence", window, step)
	index = partition.index(OptimizedSpanSim(EmbeddingTokenSim(emb, CosineSim()), LocalAlignment(gap=gap)), nlp=whitespace_nlp)
	rng = np.random.default_rng(window * 100 + step * 10 + int(filtered))
	options = {"pos_filter": ["DET", "ADP"]} if filtered else {}
	n_cases = n_grouped = n_filtered_changes = n_ref_checked = 0
	for case in range(120):
		n_t = int(rng.integers(1, 8))
		text = " ".join(words[i] for i in rng.integers(0, len(words), size=n_t))
		query = index.make_query(text, options=options)
		p_query = query.prepare(whitespace_nlp)
		token_filter = session.token_filter(query.options)
		d = int(rng.integers(0, len(session.documents)))
		pdoc = session.documents[d]
		doc = pdoc.doc
		n_spans = len(doc.spans["sentence"]["start"]) if isinstance(doc.spans["sentence"], dict) else pdoc.n_spans(partition)
		slice_id = int(rng.integers(0, (n_spans + step - 1) // step)) * step
		span = pdoc.span(partition, slice_id)
		token_at, slice_len = span.start, len(span)
		if slice_len < 1:
			continue
		keep = None
		if token_filter is not None:
			w = slice(token_at, token_at + slice_len)
			keep = session.filter_pass(pdoc.token_pos[w], pdoc.token_tag[w], token_filter)
			n_s = int(keep.sum())
			n_filtered_changes += int(n_s != slice_len)
		else:
			n_s = slice_len
		target = random_target(rng, n_s, n_t, 0.0 if case % 17 == 0 else 0.7) if n_s > 0 else np.full((n_t,), -1, np.int16)
		dist = np.where(target >= 0, rng.random(n_t), 1.0).astype(np.float32)
		resolved = gaps.resolve_gap_costs(gap, 64, n_t)
		gap_fns = (_gap_cost_fn(resolved, 's'), _gap_cost_fn(resolved, 't'))
		hit = np.zeros((), dtype=HIT_DTYPE)
		hit["score"], hit["score_max"], hit["doc"], hit["slice"] = 0.5, n_t, d, slice_id
		context = int(rng.integers(0, 12))
		match = CoreMatch(index, p_query, hit, target, dist, "toy-cosine", gap_fns, token_filter)

		s_tokens = [{'idx': int(doc.token_idx[i]), 'len': int(doc.token_len[i]), 'pos': doc.token_pos[i]} for i in range(doc.n_tokens)]
		t_tokens = [{'idx': t['start'], 'len': t['end'] - t['start'], 'pos': t.get('pos', 'X')} for t in p_query.tokens]
		edges = [(int(u), j, 1.0, float(dist[j])) for j, u in enumerate(target) if u >= 0]
		want_regions = render.py_regions(s_tokens, t_tokens, token_at, slice_len, edges, context,
			filter_pass=None if keep is None else [bool(x) for x in keep], gap_cost_s=gap_fns[0], gap_cost_t=gap_fns[1])
		if ref.flow_available():
			n_ref_checked += check_against_reference_flow(session, pdoc, doc, p_query, token_at, slice_len, target, dist, context, token_filter, gap_fns,
				want_regions, render.py_omitted(t_tokens, target))
		want = render.render(doc.text_str, p_query.text_str, want_regions, "toy-cosine")
		got = [(r.s, None if r.match is None else (r.match.pos_s, [((e.t.text, e.t.index, e.t.pos), e.flow, e.distance, e.metric) for e in r.match.edges]),
			r.gap_penalty) for r in match.regions(context)]
		assert got == want, (case, target, token_at, slice_len)
		want_omitted = [p_query.text_str[slice(*s)] for s in render.py_omitted(t_tokens, target)]
		assert match.omitted == want_omitted
		location = pdoc.span_info(PartitionData(**query.options["partition"]), slice_id)
		assert match.to_json(context) == render.to_json(slice_id, location, match.score, "toy-cosine", want, want_omitted, "word")
		n_grouped += sum(1 for r in want_regions if r['matched'] and len(r['edges']) > 1)
		n_cases += 1
	assert n_cases > 80
	assert n_ref_checked == n_cases or not ref.flow_available()
	if filtered:
		assert n_filtered_changes > 20          # the filter really removed tokens
		assert n_grouped > 0                    # ... and the reference's raw-vs-mapped grouping (flow.cpp:143) occurred
