"""Session / Partition / Result with the reference's signatures (vectorian/session.py:24-258).

A Session owns the vocabulary (token string -> id in insertion order while documents are loaded,
vectorian/core/cpp/vocabulary.h:90-105), the per-document token-id tables, the static embedding
matrices over that vocabulary and -- per partition -- the packed device corpus (vx_corpus_create).
With torch.distributed initialised, every rank keeps only its contiguous range of documents
(sentence shards; document ids stay global) and Index.find merges the per-rank top-k.
"""

import collections
import time

import numpy as np

from . import engine
from .corpus import Corpus, Span
from .embedding import TokenEmbedding, Vectors
from .sim.vector import CosineSim
from .sim.token import EmbeddingTokenSim
from .sim.span import SpanSim, OptimizedSpanSim


class Normalization:
	"""Token text normalisation applied to documents and queries alike (vectorian/normalization.py).
	The reference chains regex / unicode normalizers; here: optional lower-casing plus user callables."""

	def __init__(self, name="vanilla", lowercase=False, normalizers=None):
		self.name = name
		self._lowercase = lowercase
		self.normalizers = list(normalizers or [])

	def token_to_str(self, text):
		if self._lowercase:
			text = text.lower()
		for f in self.normalizers:
			text = f(text)
		return text


class VanillaNormalization(Normalization):
	def __init__(self):
		super().__init__("vanilla")


class Lexicon:
	"""String -> id in insertion order (cpp/vocabulary.h:31-143)."""

	def __init__(self):
		self._ids = {}
		self._strs = []

	def add(self, s):
		i = self._ids.get(s)
		if i is None:
			i = len(self._strs)
			self._ids[s] = i
			self._strs.append(s)
		return i

	def get(self, s, default=-1):
		return self._ids.get(s, default)

	def __len__(self):
		return len(self._strs)

	@property
	def strings(self):
		return self._strs


class Vocabulary:
	"""Token, POS and tag lexicons of a session (cpp/vocabulary.h:275-434)."""

	def __init__(self):
		self.tokens = Lexicon()
		self.pos = Lexicon()
		self.tag = Lexicon()

	@property
	def size(self):
		return len(self.tokens)


class Result:
	"""vectorian/session.py:24-55"""

	def __init__(self, index, matches, duration):
		self._index = index
		self._matches = matches
		self._duration = duration

	@property
	def index(self):
		return self._index

	@property
	def matches(self):
		return self._matches

	def __len__(self):
		return len(self._matches)

	def __iter__(self):
		return iter(self._matches)

	def __getitem__(self, i):
		return self._matches[i]

	def to_json(self, context_size=10):
		return [m.to_json(context_size) for m in self._matches]

	def limit_to(self, n):
		return type(self)(self._index, self._matches[:n], self._duration)

	@property
	def duration(self):
		return self._duration


class PreparedDocument:
	"""A document bound to a session: token ids / POS ids (the role of core.Document,
	corpus/document.py:626-779, cpp/document.cpp:30-69)."""

	def __init__(self, session, doc, doc_id):
		self._session = session
		self._doc = doc
		self._id = doc_id
		vocab = session.vocab
		norm = session.normalization
		strs = [norm.token_to_str(t) for t in doc.token_text]
		doc.token_str = strs
		self.token_id = np.array([vocab.tokens.add(s) for s in strs], dtype=np.int32)
		self.token_pos = np.array([vocab.pos.add(p) for p in doc.token_pos], dtype=np.int8)
		self.token_tag = np.array([vocab.tag.add(t) for t in doc.token_tag], dtype=np.int8)

	@property
	def id(self):
		return self._id

	@property
	def doc(self):
		return self._doc

	@property
	def session(self):
		return self._session

	@property
	def metadata(self):
		return self._doc.metadata

	@property
	def n_tokens(self):
		return self._doc.n_tokens

	@property
	def text_str(self):
		return self._doc.text_str

	def token(self, i):
		from .corpus import Token
		return Token(self._doc, i)

	def span_table(self, level):
		"""(start, end) int32 arrays of a partition level; 'token' = one span per token."""
		if level == "token":
			start = np.arange(self.n_tokens, dtype=np.int32)
			return start, start + 1
		table = self._doc.spans[level]
		return table['start'], table['end']

	def n_spans(self, partition):
		start, _ = self.span_table(partition.level)
		n = start.shape[0]
		while n > 0 and start[n - 1] >= self.n_tokens:
			n -= 1
		k = n // partition.window_step
		if k * partition.window_step < n:
			k += 1
		return k

	def span(self, partition, slice_id):
		"""Token span of the window that starts at span `slice_id` (cpp/document.h:96-99)."""
		start, end = self.span_table(partition.level)
		last = min(slice_id + partition.window_size - 1, len(start) - 1)
		return Span(self._doc, start[slice_id], end[last])

	def span_info(self, partition, slice_id):
		if partition.level == "token":
			return {'start': slice_id, 'end': slice_id + partition.window_size}
		start, end = self.span_table(partition.level)
		return {'start': int(start[slice_id]), 'end': int(end[slice_id])}


Slice = collections.namedtuple('Slice', ['level', 'start', 'end'])


class Partition:
	"""vectorian/session.py:83-146"""

	def __init__(self, session, level, window_size, window_step):
		self._session = session
		self._level = level
		self._window_size = window_size
		self._window_step = window_step

	@property
	def contiguous(self):
		return self._window_step <= self._window_size

	@property
	def session(self):
		return self._session

	@property
	def level(self):
		return self._level

	@property
	def window_size(self):
		return self._window_size

	@property
	def window_step(self):
		return self._window_step

	def to_args(self):
		return {'level': self._level, 'window_size': self._window_size, 'window_step': self._window_step}

	@property
	def cache_key(self):
		return self._level, self._window_size, self._window_step

	def max_len(self):
		return self._session.max_len(self._level, self._window_size)

	def index(self, metric, nlp=None, **kwargs):
		if not isinstance(metric, SpanSim):
			raise TypeError(metric)
		if nlp:
			kwargs = kwargs.copy()
			kwargs['nlp'] = nlp
		return metric.create_index(self, **kwargs)

	def slice_id_to_slice(self, slice_id):
		return Slice(self._level, self._window_step * slice_id, self._window_size)


class Session:
	"""vectorian/session.py:165-258.  `corpus` is a vectorian_b200.corpus.Corpus (or a list of Documents)."""

	def __init__(self, corpus, embeddings=None, normalization: Normalization = None):
		if not isinstance(corpus, Corpus):
			corpus = Corpus(corpus)
		self._corpus = corpus
		self._normalization = normalization or VanillaNormalization()
		embeddings = embeddings or []
		self._token_embeddings = tuple(x for x in embeddings if isinstance(x, TokenEmbedding))
		for embedding in self._token_embeddings:
			if embedding.is_contextual:
				for doc in corpus:
					if not doc.has_contextual_embedding(embedding.name):
						raise RuntimeError(
							f"doc {corpus.get_unique_id(doc)} misses contextual embedding {embedding.name}")
		self._embedding_encoders = collections.OrderedDict()
		for embedding in self._token_embeddings:
			encoder = embedding.create_encoder(self)
			self._embedding_encoders[encoder.name] = encoder

		self._vocab = Vocabulary()
		self._documents = [PreparedDocument(self, doc, i) for i, doc in enumerate(corpus)]

		# sentence shards: this rank's contiguous range of documents (ids stay global)
		self._rank, self._world = 0, 1
		try:
			import torch.distributed as dist
			if dist.is_available() and dist.is_initialized():
				self._rank, self._world = dist.get_rank(), dist.get_world_size()
		except ImportError:
			pass
		from .distributed import shard_documents
		weights = [d.n_tokens for d in self._documents]
		self._doc_lo, self._doc_hi = shard_documents(len(self._documents), self._world, self._rank, weights)

		self._vocab_vectors = {}       # encoder name -> Vectors over the vocabulary (host)
		self._device_embeddings = {}   # encoder name -> (engine.DeviceEmbedding, Vectors over the vocabulary)
		self._device_corpora = {}      # partition cache key -> engine.DeviceCorpus
		self._contextual_loaded = {}   # partition cache key -> embedding name currently on the device corpus

	# ---- reference surface ----

	@property
	def corpus(self):
		return self._corpus

	@property
	def vocab(self):
		return self._vocab

	@property
	def normalization(self):
		return self._normalization

	@property
	def normalizers(self):
		return self._normalization.normalizers

	def default_metric(self):
		embedding = self._token_embeddings[0]
		return OptimizedSpanSim(EmbeddingTokenSim(embedding, CosineSim()))

	@property
	def documents(self):
		return self._documents

	@property
	def encoders(self):
		return self._embedding_encoders

	def to_encoder(self, embedding):
		return self._embedding_encoders[embedding.name]

	def max_len(self, level, window_size):
		"""Longest window of the partition (Spans::max_len, cpp/document.h:116-133)."""
		best = 0
		for doc in self._documents:
			start, end = doc.span_table(level)
			n = len(start)
			for i in range(n):
				last = min(i + window_size - 1, n - 1)
				best = max(best, int(end[last]) - int(start[i]))
		return best

	def make_result(self, *args, **kwargs):
		return Result(*args, **kwargs)

	def on_progress(self, task, disable_progress=False):
		return task(None)

	def partition(self, level, window_size=1, window_step=None):
		if window_step is None:
			window_step = window_size
		return Partition(self, level, window_size, window_step)

	def index(self, *args, **kwargs):
		return self.partition("sentence").index(*args, **kwargs)

	def word_vec(self, embedding, token_or_tokens):
		encoder = self.to_encoder(embedding)
		get = lambda t: encoder.word_vec(t if isinstance(t, str) else t.text)
		if isinstance(token_or_tokens, (list, tuple)):
			return [get(x) for x in token_or_tokens]
		return get(token_or_tokens)

	def similarity(self, token_sim, a, b):
		"""Similarity of two tokens under a token similarity (vectorian/session.py:281-321)."""
		encoder = self.to_encoder(token_sim.embedding)
		va = Vectors([encoder.word_vec(a if isinstance(a, str) else a.text)])
		vb = Vectors([encoder.word_vec(b if isinstance(b, str) else b.text)])
		out = np.zeros((1, 1), dtype=np.float32)
		token_sim.similarity(va, vb, out)
		return out[0, 0]

	# ---- device state (what the reference keeps in core.Document / core.StaticEmbedding) ----

	@property
	def shard(self):
		"""(rank, world size, first document id, one past the last document id) of this process."""
		return self._rank, self._world, self._doc_lo, self._doc_hi

	def vocab_vectors(self, embedding):
		"""Static embedding vectors over the session vocabulary in host memory (cpp/vocabulary.h:439-497
		compile_embeddings), cached until the vocabulary grows."""
		encoder = self.to_encoder(embedding)
		vectors = self._vocab_vectors.get(encoder.name)
		if vectors is None or vectors.shape[0] != self._vocab.size:
			vectors = encoder.encode_tokens(self._vocab.tokens.strings)
			self._vocab_vectors[encoder.name] = vectors
		return vectors

	def device_embedding(self, embedding):
		"""Static embedding matrix over the session vocabulary, normalised on the device
		(cpp/vocabulary.h:439-497 compile_embeddings + embedding/vectors.py:71-86)."""
		encoder = self.to_encoder(embedding)
		vectors = self.vocab_vectors(embedding)
		entry = self._device_embeddings.get(encoder.name)
		if entry is None or entry[1] is not vectors:
			entry = (engine.DeviceEmbedding(vectors.unmodified), vectors)
			self._device_embeddings[encoder.name] = entry
		return entry

	def token_filter(self, options):
		"""The query's document-side token filter: (pos bit mask, tag bit mask), or None when the options filter
		nothing.  Query::make_token_filter + parse_filter_mask (cpp/query.cpp:220-257, cpp/query.h:8-53): bit i of a
		mask drops every document token whose pos / tag id is i; an unknown name is an error."""
		bits = []
		for attr, lexicon in (('pos', self._vocab.pos), ('tag', self._vocab.tag)):
			mask = 0
			for name in (options.get(f'{attr}_filter') or []):
				i = lexicon.get(name)
				if i < 0:
					raise RuntimeError(f"illegal value {name} for {attr}_filter")
				if i >= 64:
					raise RuntimeError(f"{attr}_filter: id {i} of {name} does not fit the 64-bit filter mask")   # uint64_t masks, cpp/query.h:13-14
				mask |= 1 << i
			bits.append(mask)
		return tuple(bits) if any(bits) else None

	@staticmethod
	def filter_pass(token_pos, token_tag, token_filter):
		"""TokenFilter::pass (cpp/query.h:22-28) over token columns: bool[n_tokens]."""
		pos_bits, tag_bits = token_filter
		drop = ((np.uint64(pos_bits) >> token_pos.astype(np.uint64)) | (np.uint64(tag_bits) >> token_tag.astype(np.uint64))) & np.uint64(1)
		return drop == 0

	def _filtered_tables(self, partition, token_filter):
		"""Token and span columns of this rank's documents for one partition; with a token filter the tokens that do
		not pass are removed and the span bounds are moved to the compacted coordinates -- what FilteredSliceFactory::
		create_slice does per window with its index map (cpp/slice/static.h:365-416), done once per (partition, filter)
		for the whole shard so that the DP kernels see dense windows again."""
		docs = self._documents[self._doc_lo:self._doc_hi]
		tok, pos, keeps, starts, ends, span_off, tok_off = [], [], [], [], [], [0], [0]
		for d in docs:
			s, e = d.span_table(partition.level)
			s = np.asarray(s, dtype=np.int32)
			e = np.asarray(e, dtype=np.int32)
			if token_filter is None:
				t_id, t_pos = d.token_id, d.token_pos
				keep = None
			else:
				keep = self.filter_pass(d.token_pos, d.token_tag, token_filter)
				cum = np.concatenate([[0], np.cumsum(keep)]).astype(np.int32)
				n = d.n_tokens
				s = cum[np.clip(s, 0, n)]
				e = cum[np.clip(e, 0, n)]
				t_id, t_pos = d.token_id[keep], d.token_pos[keep]
			tok.append(t_id)
			pos.append(t_pos)
			keeps.append(keep)
			starts.append(s)
			ends.append(e)
			span_off.append(span_off[-1] + len(s))
			tok_off.append(tok_off[-1] + len(t_id))
		cat = lambda xs, dt: np.concatenate(xs).astype(dt) if xs else np.zeros((0,), dtype=dt)
		return dict(token_id=cat(tok, np.int32), token_pos=cat(pos, np.int8), doc_tok_off=np.array(tok_off, dtype=np.int64),
			span_start=cat(starts, np.int32), span_end=cat(ends, np.int32), doc_span_off=np.array(span_off, dtype=np.int64),
			keep=keeps)

	def device_corpus(self, partition, token_filter=None):
		"""This rank's documents packed for one partition (vx_corpus_create), cached -- per token filter, so that
		queries that switch filters between calls (the reference's use case, cpp/query.h:9-11) pack once per filter."""
		key = (partition.cache_key, token_filter)
		dc = self._device_corpora.get(key)
		if dc is None:
			tb = self._filtered_tables(partition, token_filter)
			dc = engine.DeviceCorpus(
				tb["token_id"], tb["doc_tok_off"], tb["span_start"], tb["span_end"], tb["doc_span_off"],
				partition.window_size, partition.window_step, token_pos=tb["token_pos"])
			if self._world > 1:
				from .distributed import device_comm
				comm = device_comm()      # NCCL process groups: gather + merge inside vx_search (csrc/vx_shard.cu)
				if comm is not None:
					dc.set_shard(comm, self._doc_lo)
			self._device_corpora[key] = dc
		return dc

	def shard_token_ids(self, partition, token_filter=None):
		"""Token ids of this rank's documents in the token order of the device corpus (after the token filter)."""
		return self._filtered_tables(partition, token_filter)["token_id"]

	def contextual_vectors(self, embedding, token_filter=None):
		"""[n_tokens x d] per-token vectors of this rank's documents, rows aligned with shard_token_ids."""
		docs = self._documents[self._doc_lo:self._doc_hi]
		dim = embedding.dimension
		xs = [np.asarray(d.doc.contextual_embeddings[embedding.name], dtype=np.float32).reshape(-1, dim) for d in docs]
		for d, x in zip(docs, xs):
			if x.shape[0] != d.n_tokens:
				raise ValueError(f"contextual embedding of document {d.id} has {x.shape[0]} rows for {d.n_tokens} tokens")
		if token_filter is not None:
			xs = [x[self.filter_pass(d.token_pos, d.token_tag, token_filter)] for d, x in zip(docs, xs)]
		return np.concatenate(xs) if xs else np.zeros((0, dim), dtype=np.float32)

	def load_contextual(self, partition, embedding, token_filter=None):
		"""Make the per-token vectors of `embedding` resident next to the partition's device corpus."""
		key = (partition.cache_key, token_filter)
		dc = self.device_corpus(partition, token_filter)
		if self._contextual_loaded.get(key) != embedding.name:
			dc.set_contextual(self.contextual_vectors(embedding, token_filter))
			self._contextual_loaded[key] = embedding.name
		return dc


class LabSession(Session):
	"""The reference's notebook session (session.py:407-460) adds renderers and progress bars only (UI,
	out of scope); kept as an alias so that user code constructing it keeps working."""

	def __init__(self, *args, location_formatter=None, **kwargs):
		super().__init__(*args, **kwargs)
