"""Query / Match / Index / BruteForceIndex with the reference's signatures (vectorian/index.py).

``BruteForceIndex._find`` is where the reference fans out ``core.Document.find`` over a thread pool
(index.py:530-560); here it is one ``vx_sim_*`` + one ``vx_search`` on the device corpus of the
partition and, with several ranks, one gather of the per-rank top-k.
"""

import collections
import functools
import logging
import time

import numpy as np

from . import _cabi, engine, gaps
from .corpus import whitespace_nlp, Span
from .distributed import gather_topk
from .embedding import Vectors


class Query:
	"""vectorian/index.py:26-54"""

	def __init__(self, index, vocab, text, options):
		self._index = index
		self._vocab = vocab
		self._text = text
		self._options = options

	@property
	def index(self):
		return self._index

	@property
	def text(self):
		return self._text

	@property
	def options(self):
		return self._options

	def prepare(self, nlp):
		return PreparedQuery(self, self._vocab, nlp)


class _QueryDoc:
	"""The query seen as a one-span document (what Token / Span objects need)."""

	def __init__(self, text, tokens, strs):
		self.text_str = text
		self.token_idx = np.array([t['start'] for t in tokens], dtype=np.int32)
		self.token_len = np.array([t['end'] - t['start'] for t in tokens], dtype=np.int32)
		self.token_str = strs


class PreparedQuery:
	"""vectorian/index.py:56-169: tokenise, filter, normalise, map to vocabulary ids.

	``nlp`` is any callable text -> list of token dicts with 'start', 'end', 'pos', 'tag' (the
	``doc.to_json()["tokens"]`` of a spaCy pipeline); objects with ``to_json`` are accepted too."""

	def __init__(self, query, vocab, nlp):
		self._query = query
		self._vocab = vocab
		text = query.text
		doc = (nlp or whitespace_nlp)(text)
		tokens = doc.to_json()["tokens"] if hasattr(doc, "to_json") else list(doc)
		for attr in ('pos', 'tag'):
			f = query.options.get(f'{attr}_filter')
			if f:
				s = set(f)
				tokens = [t for t in tokens if t.get(attr) not in s]
		session = query.index.session
		norm = session.normalization
		kept, strs = [], []
		for t in tokens:
			s = norm.token_to_str(text[t['start']:t['end']])
			if s:                                  # TokenTable.add drops empty tokens (corpus/document.py:63-100)
				kept.append(t)
				strs.append(s)
		self._tokens = kept
		self._token_str = strs
		# ids: corpus vocabulary first, then per-query ids |V|, |V|+1, ... (IncrementalLexicon,
		# vectorian/core/cpp/vocabulary.h:152-157)
		extra = {}
		ids = []
		for s in strs:
			i = vocab.tokens.get(s)
			if i < 0:
				i = extra.setdefault(s, vocab.size + len(extra))
			ids.append(i)
		self._token_id = np.array(ids, dtype=np.int32)
		self._token_pos = np.array([vocab.pos.get(t.get('pos', 'X')) for t in kept], dtype=np.int8)
		self._token_tag = [t.get('tag', 'X') for t in kept]
		self._doc = _QueryDoc(text, kept, strs)

	@property
	def index(self):
		return self._query.index

	@property
	def text_str(self):
		return self._query.text

	@property
	def options(self):
		return self._query.options

	@property
	def n_tokens(self):
		return len(self._tokens)

	@property
	def token_str(self):
		return self._token_str

	@property
	def token_id(self):
		return self._token_id

	@property
	def token_pos(self):
		return self._token_pos

	@property
	def token_tag(self):
		return self._token_tag

	@property
	def tokens(self):
		return self._tokens

	@property
	def span(self):
		return Span(self._doc, 0, self.n_tokens)

	def __getitem__(self, i):
		return self.span[i]

	def __len__(self):
		return self.n_tokens


Region = collections.namedtuple('Region', ['s', 'match', 'gap_penalty'])
TokenMatch = collections.namedtuple('TokenMatch', ['pos_s', 'edges'])
TokenMatchEdge = collections.namedtuple('TokenMatchEdge', ['t', 'flow', 'distance', 'metric'])
TokenMatchT = collections.namedtuple('TokenMatchT', ['text', 'index', 'pos'])
PartitionData = collections.namedtuple('PartitionData', ['level', 'window_size', 'window_step'])


def _gap_cost_fn(resolved, side):
	"""gap_cost_s / gap_cost_t of the solver in use (cpp/metric/alignment.h:296-304)."""
	kind = resolved['gap_kind']
	k = 0 if side == 's' else 1
	if kind == _cabi.VX_GAP_LINEAR:
		u = resolved['lin'][k]
		return lambda n: float(np.float32(u) * np.float32(n))
	if kind == _cabi.VX_GAP_AFFINE:
		open0, ext = resolved['affine_s' if side == 's' else 'affine_t']
		return lambda n: 0.0 if n == 0 else float(np.float32(ext) * np.float32(n) + np.float32(open0))
	w = resolved['w_s' if side == 's' else 'w_t']
	return lambda n: float(w[min(int(n), len(w) - 1)])


class Match:
	"""vectorian/index.py:192-292"""

	@property
	def partition(self):
		return self.index.partition

	@property
	def doc_span(self):
		return self.prepared_doc.span(self.query.index.partition, self.slice_id)

	@property
	def doc(self):
		return self.prepared_doc.doc

	@property
	def slice(self):
		return self.partition.slice_id_to_slice(self.slice_id)

	@property
	def flow(self):
		return None

	def to_json(self, context_size=10):
		regions = []
		partition = self.query.options["partition"]
		span_info = self.prepared_doc.span_info(PartitionData(**partition), self.slice_id)
		for region in self.regions(context_size):
			if region.match:
				edges = [{
					't': {'text': e.t.text, 'index': e.t.index, 'pos': e.t.pos},
					'flow': e.flow, 'distance': e.distance, 'metric': e.metric} for e in region.match.edges]
				regions.append(dict(s=region.s, pos_s=region.match.pos_s, edges=edges))
			else:
				regions.append(dict(s=region.s, gap_penalty=region.gap_penalty))
		return dict(
			slice=self.slice_id, location=span_info, score=self.score, metric=self.metric,
			regions=regions, omitted=self.omitted, level=self.level)


class CoreMatch(Match):
	"""A match produced by the device search; same getters as vectorian/index.py:295-379 over
	Match / InjectiveFlow (vectorian/core/cpp/match/match.h:275-425, cpp/match/flow.cpp:190-216)."""

	def __init__(self, index, query, hit, target, dist, metric_name, gap_fns, token_filter=None):
		self._index = index
		self._session = index.session
		self._query = query
		self._score = float(hit["score"])
		self._score_max = float(hit["score_max"])
		self._doc_id = int(hit["doc"])
		self._slice_id = int(hit["slice"])
		self._target = np.array(target, dtype=np.int16)
		self._dist = np.array(dist, dtype=np.float32)
		self._metric = metric_name
		self._gap_fns = gap_fns
		self._token_filter = token_filter
		self._level = "word"

	@property
	def index(self):
		return self._index

	@property
	def query(self):
		return self._query

	@property
	def prepared_doc(self):
		return self._session.documents[self._doc_id]

	@property
	def slice_id(self):
		return self._slice_id

	@property
	def score(self):
		return self._score

	@property
	def score_max(self):
		return self._score_max

	@property
	def metric(self):
		return self._metric

	@property
	def level(self):
		return self._level

	@property
	def flow(self):
		"""InjectiveFlow::to_py (cpp/match/flow.cpp:190-216)."""
		matched = self._target >= 0
		return {
			'type': 'injective',
			'target': self._target,
			'flow': matched.astype(np.float32),
			'dist': self._dist,
		}

	@property
	def omitted(self):
		"""Query tokens without a partner (Flow::py_omitted, cpp/match/flow.cpp:169-188)."""
		text = self._query.text_str
		return [text[t['start']:t['end']] for t, u in zip(self._query.tokens, self._target) if u < 0]

	def regions(self, context_size=10):
		"""Matched / unmatched character regions of the window (Flow::py_regions, cpp/match/flow.cpp:8-167)."""
		pdoc = self.prepared_doc
		doc = pdoc.doc
		s_text = doc.text_str
		t_text = self._query.text_str
		part = self._index.partition
		span = pdoc.span(part, self._slice_id)
		token_at, n_tok = span.start, doc.n_tokens
		idx, ln = doc.token_idx, doc.token_len
		# InjectiveFlow::to_edges + the sort of flow.cpp:32-41 (by target; all flows of an injective alignment are 1)
		edges = sorted(((int(u), j) for j, u in enumerate(self._target) if u >= 0), key=lambda e: e[0])
		index_map = None
		if self._token_filter is not None:
			# targets count the tokens that pass the filter: back to window positions (cpp/match/flow.cpp:49-61, 95-99)
			w = slice(token_at, token_at + len(span))
			index_map = np.flatnonzero(self._session.filter_pass(pdoc.token_pos[w], pdoc.token_tag[w], self._token_filter))
		regions = []
		if not edges:
			s0 = int(idx[token_at])
			s1 = int(idx[min(token_at + len(span), n_tok - 1)])
			return [Region(s=s_text[s0:s1], match=None, gap_penalty=0.0)]
		gap_s, gap_t = self._gap_fns
		last_anchor = max(0, token_at + edges[0][0] - context_size)
		last_matched = False
		last_source = -1
		i = 0
		while i < len(edges):
			target = edges[i][0]
			if index_map is not None:
				target = int(index_map[target])
			s_i = token_at + target
			idx0 = int(idx[last_anchor])
			if int(idx[s_i]) > idx0:
				p = gap_s(s_i - last_anchor) if last_matched else 0.0
				regions.append(Region(s=s_text[idx0:int(idx[s_i])], match=None, gap_penalty=p))
			region_edges = []
			while True:
				source = edges[i][1]
				if last_source >= 0:
					p = gap_t(source - last_source - 1)
					if p > 0.0:
						regions.append(Region(s="", match=None, gap_penalty=p))
				last_source = source
				t = self._query.tokens[source]
				region_edges.append(TokenMatchEdge(
					t=TokenMatchT(text=t_text[t['start']:t['end']], index=source, pos=t.get('pos', 'X')),
					flow=1.0, distance=float(self._dist[source]), metric=self._metric))
				i += 1
				# flow.cpp:143 groups the following edges whose RAW target equals this edge's MAPPED target; without a
				# token filter the two coincide and an injective flow never repeats a target.  Kept as the reference has it.
				if not (i < len(edges) and edges[i][0] == target):
					break
			regions.append(Region(
				s=s_text[int(idx[s_i]):int(idx[s_i]) + int(ln[s_i])],
				match=TokenMatch(pos_s=doc.token_pos[s_i], edges=region_edges), gap_penalty=0.0))
			last_anchor = s_i + 1
			last_matched = True
		up_to = min(last_anchor + context_size, n_tok - 1)
		if up_to > last_anchor:
			idx0 = int(idx[last_anchor])
			regions.append(Region(s=s_text[idx0:int(idx[up_to])], match=None, gap_penalty=0.0))
		return regions


class Index:
	"""vectorian/index.py:430-501"""

	def __init__(self, partition, sim):
		self._partition = partition
		self._sim = sim
		if not partition.contiguous:
			logging.warning("the used partition is non-contiguous, you will miss parts of the content.")

	@property
	def partition(self):
		return self._partition

	@property
	def session(self):
		return self._partition.session

	@property
	def sim(self):
		return self._sim

	def describe(self):
		data = {'partition': self._partition.to_args(), 'metric': self._sim.to_args(self)}
		try:
			import yaml
			print(yaml.dump(data))
		except ImportError:
			print(data)

	def make_query(self, text, n=10, min_score=0.0, debug=None, options: dict = dict()):
		options = options.copy()
		options["max_matches"] = n
		options["min_score"] = min_score
		if debug is not None:
			options["debug"] = debug
		options["partition"] = self._partition.to_args()
		if self._sim is not None:
			sim_args = self._sim.to_args(self)
			if sim_args:
				options["metric"] = sim_args
		return Query(self, self._partition.session.vocab, text, options)

	def find(self, text, n=10, min_score=0.0, debug=None, disable_progress=False,
			run_task=None, make_result=None, options: dict = dict()):
		start_time = time.time()
		query = self.make_query(text, n=n, min_score=min_score, debug=debug, options=options)
		session = self._partition.session
		if make_result is None:
			make_result = session.make_result
		if run_task is None:
			run_task = functools.partial(session.on_progress, disable_progress=disable_progress)
		matches = run_task(lambda progress: self._find(query, progress=progress))
		return make_result(self, matches, duration=time.time() - start_time)


# options core.Query.initialize accepts (vectorian/core/cpp/query.cpp:40-57)
_KNOWN_OPTIONS = frozenset([
	"metric", "pos_filter", "tag_filter", "submatch_weight", "bidirectional", "max_matches",
	"min_score", "partition", "debug", "slice_strategy"])


class BruteForceIndex(Index):
	"""vectorian/index.py:508-560"""

	def __init__(self, *args, nlp=None, saliency=None, **kwargs):
		super().__init__(*args, **kwargs)
		self._nlp = nlp
		self._saliency = saliency   # callable(prepared_doc, partition) -> float32[n_spans], the Booster values
		self.last_timing = None

	def _set_boost(self, dc):
		"""Booster values (cpp/document.h:175-187) by span.  Device corpora are shared by all indexes of a partition
		(one per token filter), so the corpus remembers whose values it carries."""
		owner = getattr(dc, "boost_owner", None)
		if self._saliency is None:
			if owner is not None:
				dc.set_boost(None)
				dc.boost_owner = None
				dc.boost_values = None
			return
		if owner is self:
			return
		session = self.session
		lo, hi = session.shard[2:]
		boost = [np.asarray(self._saliency(d, self.partition), dtype=np.float32) for d in session.documents[lo:hi]]
		dc.boost_values = np.concatenate(boost) if boost else np.zeros((0,), np.float32)
		dc.set_boost(dc.boost_values)
		dc.boost_owner = self

	@staticmethod
	def _device_static_matrix(de, q_id, q_vec):
		"""S [V x T] of a static embedding, formed on the device and brought to the host.  The similarity kernel takes
		up to 16 query columns per launch; longer queries go through it in groups of 16 (every column of S depends on
		its own query token only, the identity fix-up included)."""
		parts = []
		for lo in range(0, len(q_id), engine.MAX_SIM_COLUMNS):
			sim = engine.DeviceSim.static(de, q_id[lo:lo + engine.MAX_SIM_COLUMNS], q_vec[lo:lo + engine.MAX_SIM_COLUMNS])
			parts.append(sim.download())
			sim.close()
		return np.ascontiguousarray(np.hstack(parts)) if parts else np.zeros((de.shape[0], 0), dtype=np.float32)

	def _host_matrix(self, p_query, token_sim):
		"""[V x T] similarity matrix of a static token similarity in host memory: identity fix-up and clip applied
		(cpp/metric/static.cpp:58-75); modifiers combine their operands' matrices (cpp/metric/modifier.cpp:18-74)."""
		session = self.session
		if token_sim.is_modifier:
			operands = [{"similarity": self._host_matrix(p_query, op)} for op in token_sim.operands]
			if not operands:
				raise ValueError("a token similarity modifier needs at least one operand")
			out = {"similarity": np.zeros_like(operands[0]["similarity"])}
			token_sim(operands, out)
			return out["similarity"]
		embedding = token_sim.embedding
		vector_sim = token_sim.similarity
		encoder = session.to_encoder(embedding)
		if encoder.is_contextual:
			raise ValueError("a contextual embedding has no [V x T] matrix (use _host_token_matrix)")
		q_vectors = encoder.encode_tokens(p_query.token_str)
		q_id = np.where(p_query.token_id < session.vocab.size, p_query.token_id, -1).astype(np.int32)
		if vector_sim.on_device:
			de, _ = session.device_embedding(embedding)
			return self._device_static_matrix(de, q_id, q_vectors.unmodified)
		# any other VectorSim: evaluated on the host exactly as the reference's callback does
		s = np.zeros((session.vocab.size, p_query.n_tokens), dtype=np.float32)
		vector_sim(session.vocab_vectors(embedding), q_vectors, s)
		for j, i in enumerate(q_id):
			if i >= 0:
				s[i, j] = 1.0
		np.clip(s, 0.0, 1.0, out=s)
		return s

	def _is_contextual(self, token_sim):
		"""Query::create_strategy (cpp/query.cpp:156-215): a modifier is CONTEXTUAL as soon as one operand is."""
		if token_sim.is_modifier:
			return any(self._is_contextual(op) for op in token_sim.operands)
		return self.session.to_encoder(token_sim.embedding).is_contextual

	def _query_vectors(self, p_query, encoder):
		q_vec = encoder.encode([[p_query.text_str[t['start']:t['end']] for t in p_query.tokens]])[0]
		if q_vec.shape[0] != p_query.n_tokens:
			raise ValueError("contextual encoder returned a wrong number of query vectors")
		return q_vec

	def _host_token_matrix(self, p_query, token_sim, token_filter):
		"""[n_tokens x T] per-token similarity matrix in host memory, rows in the token order of the device corpus:
		a contextual embedding through any VectorSim + clip (cpp/metric/contextual.cpp:26-63); a static embedding as
		the rows S[id(token)] of its matrix (cpp/metric/static.cpp:154-189); modifiers combine their operands
		(cpp/metric/modifier.cpp:18-74).  The reference builds these per document; the shard is one block here."""
		session = self.session
		if token_sim.is_modifier:
			operands = [{"similarity": self._host_token_matrix(p_query, op, token_filter)} for op in token_sim.operands]
			if not operands:
				raise ValueError("a token similarity modifier needs at least one operand")
			out = {"similarity": np.zeros_like(operands[0]["similarity"])}
			token_sim(operands, out)
			return out["similarity"]
		embedding = token_sim.embedding
		encoder = session.to_encoder(embedding)
		if encoder.is_contextual:
			from .embedding import Vectors
			x = session.contextual_vectors(embedding, token_filter)
			q_vec = np.asarray(self._query_vectors(p_query, encoder), dtype=np.float32)
			s = np.zeros((x.shape[0], p_query.n_tokens), dtype=np.float32)
			token_sim.similarity(Vectors(x), Vectors(q_vec), s)
			np.clip(s, 0.0, 1.0, out=s)
			return s
		ids = session.shard_token_ids(self.partition, token_filter)
		return np.ascontiguousarray(self._host_matrix(p_query, token_sim)[ids])

	def _similarity(self, p_query, metric_args, token_filter=None):
		"""The query's similarity matrix on the device (cpp/metric/static.cpp:9-78, contextual.cpp:26-63)."""
		session = self.session
		token_sim = metric_args['token_metric']
		if token_sim.is_modifier:
			dc = session.device_corpus(self.partition, token_filter)
			if self._is_contextual(token_sim):
				return dc, engine.DeviceSim.upload_tokens(dc, self._host_token_matrix(p_query, token_sim, token_filter))
			return dc, engine.DeviceSim.upload(np.ascontiguousarray(self._host_matrix(p_query, token_sim), dtype=np.float32))
		embedding = token_sim.embedding
		vector_sim = token_sim.similarity
		encoder = session.to_encoder(embedding)
		if encoder.is_contextual:
			if not vector_sim.on_device:
				# any other VectorSim over contextual vectors: evaluated on the host like the reference's callback
				dc = session.device_corpus(self.partition, token_filter)
				return dc, engine.DeviceSim.upload_tokens(dc, self._host_token_matrix(p_query, token_sim, token_filter))
			q_vec = self._query_vectors(p_query, encoder)
			dc = session.load_contextual(self.partition, embedding, token_filter)
			return dc, engine.DeviceSim.contextual(q_vec)
		dc = session.device_corpus(self.partition, token_filter)
		if vector_sim.on_device:
			de, _ = session.device_embedding(embedding)
			q_vectors = encoder.encode_tokens(p_query.token_str)
			q_id = np.where(p_query.token_id < session.vocab.size, p_query.token_id, -1).astype(np.int32)
			if p_query.n_tokens <= engine.MAX_SIM_COLUMNS:
				return dc, engine.DeviceSim.static(de, q_id, q_vectors.unmodified)
			# long queries: S assembled from 16-column device launches, searched by the exact per-window kernel
			return dc, engine.DeviceSim.upload(self._device_static_matrix(de, q_id, q_vectors.unmodified))
		return dc, engine.DeviceSim.upload(self._host_matrix(p_query, token_sim))

	@staticmethod
	def _metric_name(index, token_sim):
		# Match::metric is the name of the token metric (cpp/query.cpp:158-189)
		return token_sim.name if token_sim.is_modifier else token_sim.to_args(index)['name']

	def _align_params(self, metric_args, dc, t):
		alignment = metric_args['alignment']
		if alignment.get('algorithm') != 'pyalign':
			raise NotImplementedError(f"algorithm {alignment.get('algorithm')} is outside the B200 hot path")
		a_opt = alignment['options']
		resolved = gaps.resolve_gap_costs(a_opt.get('gap_cost'), max(dc.max_len, 1), t)
		gap_fns = (_gap_cost_fn(resolved, 's'), _gap_cost_fn(resolved, 't'))
		r = dict(resolved)
		kind = r.pop('gap_kind')
		ap, keep_a = _cabi.make_align_params(a_opt['locality'], kind, **r)
		return ap, keep_a, gap_fns

	@staticmethod
	def _query_params(options, metric_args, p_query):
		qkw = {}
		if metric_args['metric'] == 'alignment-tag-weighted':
			tw = metric_args['tag_weights']
			qkw = dict(
				t_weights=np.array([tw.get(tag, 1.0) for tag in p_query.token_tag], dtype=np.float32),
				t_pos=p_query.token_pos,
				pos_mismatch_penalty=metric_args.get('pos_mismatch_penalty', 0),
				similarity_threshold=metric_args.get('similarity_threshold', 0))
		k = int(options.get("max_matches", 100))
		qp, keep_q = _cabi.make_query_params(
			max_matches=k, min_score=options.get("min_score", 0.2),
			submatch_weight=options.get("submatch_weight", 0.0), **qkw)
		return qp, keep_q, k

	DEBUG_MAX_WINDOWS = 8192   # the traceback of every window comes from ONE search with max_matches = #windows

	def _call_debug_hook(self, hook, dc, sim, ap, metric_args, options, p_query, token_filter):
		"""The `debug` query option (cpp/query.cpp:73-74): the reference calls the hook once with the static similarity
		matrix ("static_similarity_matrix": similarity, rows, columns; cpp/metric/static.cpp:196-231, cpp/query.cpp:132-134)
		and then for EVERY slice with {"slice", "similarity" [len_s x len_t], "flow", "score" (raw aligner score)} under the
		name "alignment" (cpp/metric/alignment.h:144-173, 279-282, 472).  Here the per-slice data comes from one extra search
		with max_matches = number of windows and no score floor; windows are reported in document order, slice ascending.
		A debugging aid for small corpora: more than DEBUG_MAX_WINDOWS windows per shard are refused."""
		session = self.session
		s_matrix = None
		if sim.rows > 0:
			s_matrix = sim.download()
			if not metric_args['token_metric'].is_modifier:
				vocab = session.vocab
				hook("static_similarity_matrix", {
					"similarity": s_matrix,
					"rows": lambda: list(vocab.tokens.strings),
					"columns": lambda: list(p_query.token_str)})
		n_win = dc.n_windows
		if n_win > self.DEBUG_MAX_WINDOWS:
			raise NotImplementedError(
				f"the debug hook reports every slice; {n_win} windows exceed the {self.DEBUG_MAX_WINDOWS} a debug search can trace")
		if n_win == 0:
			return
		d_opts = dict(options, max_matches=n_win, min_score=-np.inf)
		qp_all, keep_all, _ = self._query_params(d_opts, metric_args, p_query)
		was_sharded = dc.sharded
		comm = getattr(dc, "_comm", None)
		if was_sharded:
			dc.set_shard(None, 0)       # the hook reports this rank's slices
		try:
			hits, target, dist = dc.search(sim, ap, qp_all)
		finally:
			if was_sharded:
				dc.set_shard(comm, session.shard[2])
		tb = session._filtered_tables(self.partition, token_filter)
		ids = tb["token_id"]
		boost = getattr(dc, "boost_values", None)
		size, step = self.partition.window_size, self.partition.window_step
		rows = {(int(h["doc"]), int(h["slice"])): i for i, h in enumerate(hits)}
		for d in range(len(tb["doc_tok_off"]) - 1):
			st = tb["span_start"][tb["doc_span_off"][d]:tb["doc_span_off"][d + 1]]
			en = tb["span_end"][tb["doc_span_off"][d]:tb["doc_span_off"][d + 1]]
			n = len(st)
			token_at = 0
			for slice_id in range(0, n, step):              # Spans::iterate, cpp/document.h:147-169
				len_s = int(en[min(slice_id + size - 1, n - 1)]) - int(st[slice_id])
				if len_s >= 1 and (d, slice_id) in rows:
					i = rows[(d, slice_id)]
					t0 = int(tb["doc_tok_off"][d]) + token_at
					sim_ij = s_matrix[ids[t0:t0 + len_s]] if s_matrix is not None else None
					t = sim.t
					b = float(boost[int(tb["doc_span_off"][d]) + slice_id]) if boost is not None else 1.0
					hook("alignment", {
						"slice": slice_id,
						"similarity": sim_ij,
						"flow": {"type": "injective", "target": target[i].copy(),
							"flow": (target[i] >= 0).astype(np.float32), "dist": dist[i].copy()},
						"score": float(hits["score"][i]) * float(hits["score_max"][i]) / (b if b else 1.0)})
				token_at += int(en[min(slice_id + step - 1, n - 1)]) - int(st[slice_id])

	def _find(self, query, n_threads=None, progress=None):
		p_query = query.prepare(self._nlp)
		if len(p_query) == 0:
			return []
		options = query.options
		for k in options.keys():
			if k not in _KNOWN_OPTIONS:
				raise ValueError(f"illegal option {k}")          # cpp/query.cpp:59-68
		metric_args = options.get("metric")
		if not metric_args:
			raise ValueError("no metric specified")
		session = self.session
		t = p_query.n_tokens

		token_filter = session.token_filter(options)
		dc, sim = self._similarity(p_query, metric_args, token_filter)
		self._set_boost(dc)

		ap, keep_a, gap_fns = self._align_params(metric_args, dc, t)
		qp, keep_q, k = self._query_params(options, metric_args, p_query)
		metric_name = self._metric_name(self, metric_args['token_metric'])

		if options.get("debug") is not None:
			self._call_debug_hook(options["debug"], dc, sim, ap, metric_args, options, p_query, token_filter)
		hits, target, dist = dc.search(sim, ap, qp)
		self.last_timing = dict(dc.last_timing(), sim_ms=sim.build_ms)
		sim.close()
		rank, world, doc_lo, _ = session.shard
		if dc.sharded:
			pass    # vx_search has already gathered and merged the ranks' rows on the device; document ids are global
		elif world > 1:
			hits, target, dist = gather_topk(hits, target, dist, k, doc_offset=doc_lo)   # process group without NCCL
		elif doc_lo:
			hits = hits.copy()
			hits["doc"] += doc_lo
		if progress:
			progress(1.0)
		return [CoreMatch(self, p_query, h, tg, ds, metric_name, gap_fns, token_filter) for h, tg, ds in zip(hits, target, dist)]



def sweep(indexes, texts, n=10, min_score=0.0, options: dict = dict()):
	"""Interactive parameter sweep (BASELINE.json config 5): every query text against every index of `indexes`
	-- BruteForceIndex objects over the SAME partition and token similarity that differ in their alignment
	(locality / gap costs) -- in one device call (vx_search_batch).  The reference runs
	`index.find(text, ...)` once per pair (vectorian/index.py:481-501); this returns the same Result objects as
	`[[index.find(text, n, min_score) for index in indexes] for text in texts]`.
	"""
	if not indexes or not texts:
		return [[] for _ in texts]
	first = indexes[0]
	session = first.session
	queries = [[index.make_query(text, n=n, min_score=min_score, options=options) for index in indexes] for text in texts]
	margs = [q.options.get("metric") for q in queries[0]]
	same = all(index.partition is first.partition and index._saliency is None for index in indexes) and all(
		m and m['metric'] == 'alignment-isolated' and m['token_metric'] is margs[0]['token_metric'] for m in margs)
	if not same:
		# different partitions / token metrics or tag weights: one find per pair
		return [[index.find(text, n=n, min_score=min_score, options=options) for index in indexes] for text in texts]
	start_time = time.time()
	prepared, sims, live = [], [], []
	dc = None
	token_filter = session.token_filter(queries[0][0].options)
	for qi, row in enumerate(queries):
		p_query = row[0].prepare(first._nlp)
		prepared.append(p_query)
		if len(p_query) > 0:
			dc, sim = first._similarity(p_query, margs[0], token_filter)
			sims.append(sim)
			live.append(qi)
	results = [[session.make_result(index, [], duration=0.0) for index in indexes] for _ in texts]
	if sims:
		t_max = max(p.n_tokens for p in prepared if len(p) > 0)
		aps, keeps, gap_fns = [], [], []
		for index, m in zip(indexes, margs):
			ap, keep, fns = index._align_params(m, dc, t_max)
			aps.append(ap)
			keeps.append(keep)
			gap_fns.append(fns)
		qp, keep_q, k = first._query_params(queries[0][0].options, margs[0], prepared[live[0]])
		first._set_boost(dc)
		rank, world, doc_lo, _ = session.shard
		if dc.sharded:
			# sharded session: ONE all-gather + ONE merge launch for all (query, setting) pairs inside vx_search_batch
			out = dc.search_batch(sims, aps, qp)
			doc_lo = 0
		elif world > 1:
			from .distributed import gather_topk_many   # process group without NCCL: host path
			flat = gather_topk_many(dc.search_batch(sims, aps, qp, raw=True), k, doc_offset=doc_lo)
			out = [[flat[q * len(aps) + a] for a in range(len(aps))] for q in range(len(sims))]
			doc_lo = 0
		else:
			out = dc.search_batch(sims, aps, qp)
		for sim in sims:
			sim.close()
		duration = (time.time() - start_time) / (len(sims) * len(indexes))
		for row, qi in zip(out, live):
			for ai, (index, (hits, target, dist)) in enumerate(zip(indexes, row)):
				if doc_lo:
					hits = hits.copy()
					hits["doc"] += doc_lo
				metric_name = index._metric_name(index, margs[ai]['token_metric'])
				matches = [CoreMatch(index, prepared[qi], h, tg, ds, metric_name, gap_fns[ai], token_filter)
					for h, tg, ds in zip(hits, target, dist)]
				results[qi][ai] = session.make_result(index, matches, duration=duration)
	return results
