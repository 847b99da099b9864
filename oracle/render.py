"""TEST INFRASTRUCTURE -- CPU restatement of the reference's result rendering, used only by tests/ as the checker of
vectorian_b200.index.CoreMatch.regions / omitted / to_json.  Never imported by the product.

Follows, statement by statement:
  Flow<Index>::py_regions          /root/reference/vectorian/core/cpp/match/flow.cpp:8-167
  Flow<Index>::py_omitted          /root/reference/vectorian/core/cpp/match/flow.cpp:169-188
  InjectiveFlow::to_edges (order)  the edges handed in are (target, source, flow, distance) of the matched query tokens
  Region / MatchedRegion           /root/reference/vectorian/core/cpp/match/region.h:8-157
  Slice::to_py                     /root/reference/vectorian/core/cpp/common.h:54-56      (idx, idx + len)
  CoreMatch.regions / omitted      /root/reference/vectorian/index.py:329-369
  Match.to_json                    /root/reference/vectorian/index.py:249-292

Parity pinned how: no reference test or fixture covers rendering.  This file is a literal transcription of the control flow
above into plain Python over plain data; `tests/test_render_oracle.py` drives the product and this checker with the same
inputs AND checks this file against the reference's own code: `oracle/ref/flow_driver.cpp` compiles the unmodified
match/flow.cpp, match/match.h and match/region.h (include guards of the pybind11 / xtensor headers pre-defined, stand-in
types for the few names they touch) and returns the regions / omitted slices the reference builds for the same match.

One reference behaviour worth knowing (kept, because parity is the bar): with a token filter the grouping loop at
flow.cpp:143 compares the RAW (filtered) target of the next edge with the MAPPED target of the current one, so two
unrelated edges can land in one MatchedRegion.  The product reproduces it.
"""


def py_regions(s_tokens, t_tokens, token_at, slice_len, edges, window_size, filter_pass=None, gap_cost_s=None, gap_cost_t=None):
	"""s_tokens / t_tokens: lists of dicts with 'idx', 'len' (character offsets) and 'pos'.
	edges: list of (target, source, flow, distance), target relative to the (filtered) window.
	filter_pass: None or a list of bools over the window's slice_len tokens (TokenFilter::pass, cpp/query.h:22-28).
	Returns a list of dicts: {'s': (start, end), 'gap_penalty': p, 'matched': bool, 'pos_s': ..., 'edges': [...]}."""
	regions = []

	# flow.cpp:22-30
	if len(edges) < 1:
		s0 = s_tokens[token_at]
		s1 = s_tokens[min(token_at + slice_len, len(s_tokens) - 1)]
		regions.append({'s': (s0['idx'], s0['idx'] + (s1['idx'] - s0['idx'])), 'gap_penalty': 0.0, 'matched': False})
		return regions

	# flow.cpp:32-41: by target ascending, equal targets by flow descending (stable for the rest)
	edges = sorted(edges, key=lambda e: (e[0], -e[2]))

	# flow.cpp:43-45
	match_0 = edges[0][0]
	last_anchor = max(0, token_at + match_0 - window_size)
	last_matched = False

	# flow.cpp:49-61
	index_map = []
	if filter_pass is not None:
		k = 0
		index_map = [0] * slice_len
		for i in range(slice_len):
			index_map[k] = i
			if filter_pass[i]:
				k += 1

	edge_index = 0
	last_source = -1

	# flow.cpp:89-155
	while edge_index < len(edges):
		edge = edges[edge_index]
		target_filtered = edge[0]
		assert target_filtered >= 0
		target = target_filtered if not index_map else index_map[target_filtered]
		s = s_tokens[token_at + target]

		idx0 = s_tokens[last_anchor]['idx']
		if s['idx'] > idx0:
			if last_matched:
				p = gap_cost_s(token_at + target - last_anchor)
			else:
				p = 0.0
			regions.append({'s': (idx0, idx0 + (s['idx'] - idx0)), 'gap_penalty': p, 'matched': False})

		region_edges = []
		while True:
			source = edges[edge_index][1]
			if last_source >= 0:
				p = gap_cost_t(source - last_source - 1)
				if p > 0.0:
					a = s_tokens[last_anchor]['idx']
					regions.append({'s': (a, a + 0), 'gap_penalty': p, 'matched': False})
			last_source = source
			t = t_tokens[source]
			region_edges.append({
				't_slice': (t['idx'], t['idx'] + t['len']), 't_index': source, 't_pos': t['pos'],
				'flow': edges[edge_index][2], 'distance': edges[edge_index][3]})
			edge_index += 1
			if not (edge_index < len(edges) and edges[edge_index][0] == target):   # RAW target vs MAPPED target (flow.cpp:143)
				break

		regions.append({'s': (s['idx'], s['idx'] + s['len']), 'gap_penalty': 0.0, 'matched': True,
			'pos_s': s['pos'], 'edges': region_edges})

		last_anchor = token_at + target + 1
		last_matched = True

	# flow.cpp:157-164
	up_to = min(last_anchor + window_size, len(s_tokens) - 1)
	if up_to > last_anchor:
		idx0 = s_tokens[last_anchor]['idx']
		regions.append({'s': (idx0, idx0 + (s_tokens[up_to]['idx'] - idx0)), 'gap_penalty': 0.0, 'matched': False})
	return regions


def py_omitted(t_tokens, target):
	"""flow.cpp:169-188: (idx, idx + len) of every query token without a partner."""
	return [(t_tokens[i]['idx'], t_tokens[i]['idx'] + t_tokens[i]['len']) for i in range(len(target)) if target[i] < 0]


def render(s_text, t_text, regions, metric):
	"""CoreMatch.regions (vectorian/index.py:335-369) over the dicts above: Region / TokenMatch / TokenMatchEdge /
	TokenMatchT as plain tuples (s, match, gap_penalty) with match = (pos_s, [((text, index, pos), flow, distance, metric)])."""
	out = []
	for r in regions:
		s = s_text[slice(*r['s'])]
		if r['matched']:
			edges = [((t_text[slice(*e['t_slice'])], e['t_index'], e['t_pos']), e['flow'], e['distance'], metric) for e in r['edges']]
			out.append((s, (r['pos_s'], edges), r['gap_penalty']))
		else:
			out.append((s, None, r['gap_penalty']))
	return out


def to_json(slice_id, location, score, metric, rendered, omitted, level):
	"""Match.to_json (vectorian/index.py:249-292) over render()'s tuples."""
	regions = []
	for s, match, gap_penalty in rendered:
		if match:
			edges = [{'t': {'text': t[0], 'index': t[1], 'pos': t[2]}, 'flow': flow, 'distance': distance, 'metric': m}
				for t, flow, distance, m in match[1]]
			regions.append(dict(s=s, pos_s=match[0], edges=edges))
		else:
			regions.append(dict(s=s, gap_penalty=gap_penalty))
	return dict(slice=slice_id, location=location, score=score, metric=metric, regions=regions, omitted=omitted, level=level)
