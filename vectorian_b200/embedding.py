"""Embeddings: in-memory replacements of the reference's token embedding classes.

The reference loads vectors through gensim / spaCy / fastText (vectorian/embedding/token/*.py), none
of which is installable here and none of which is on the hot path; what the hot path needs is
``Vectors`` (embedding/vectors.py:56-86: unmodified / normalized / magnitudes), an encoder with
``word_vec`` / ``encode_tokens`` / ``dimension`` (embedding/token/keyed.py:283-311) and a contextual
embedding that supplies one vector per document token (embedding/token/contextual.py:93-160).
"""

import numpy as np


class Vectors:
	"""embedding/vectors.py:56-86"""

	def __init__(self, unmodified):
		self._unmodified = np.asarray(unmodified, dtype=np.float32)
		self._normalized = None
		self._magnitudes = None

	def close(self):
		pass

	@property
	def shape(self):
		return self._unmodified.shape

	@property
	def unmodified(self):
		return self._unmodified

	@property
	def magnitudes(self):
		if self._magnitudes is None:
			data = np.linalg.norm(self._unmodified, axis=1)
			np.nan_to_num(data, copy=False, nan=0)
			self._magnitudes = data
		return self._magnitudes

	@property
	def normalized(self):
		if self._normalized is None:
			with np.errstate(divide='ignore', invalid='ignore'):
				data = self._unmodified / self.magnitudes[:, np.newaxis]
			np.nan_to_num(data, copy=False, nan=0, posinf=0, neginf=0)
			self._normalized = data.astype(np.float32)
		return self._normalized


class TokenEmbedding:
	"""embedding/token/token.py:1-35"""

	@property
	def is_static(self):
		return False

	@property
	def is_contextual(self):
		return False

	def create_encoder(self, session):
		raise NotImplementedError()

	@property
	def name(self):
		raise NotImplementedError()


class StaticEmbedding(TokenEmbedding):
	@property
	def is_static(self):
		return True


class KeyedVectors(StaticEmbedding):
	"""A static embedding given as (words, matrix); signature of keyed.py:279-327 where ``wv`` may be
	a gensim-like object (``index_to_key`` / ``vectors``), a dict word -> vector, or (words, matrix).
	Unknown words get the zero vector (cosine 0 to everything)."""

	class Encoder:
		def __init__(self, embedding, name, words, matrix):
			self._embedding = embedding
			self._name = name
			self._index = {w: i for i, w in enumerate(words)}
			self._matrix = matrix

		@property
		def is_static(self):
			return True

		@property
		def is_contextual(self):
			return False

		@property
		def embedding(self):
			return self._embedding

		@property
		def name(self):
			return self._name

		@property
		def dimension(self):
			return self._matrix.shape[1]

		def word_vec(self, t):
			i = self._index.get(t)
			if i is None:
				return np.zeros((self.dimension,), dtype=np.float32)
			return self._matrix[i]

		def encode_tokens(self, tokens):
			data = np.zeros((len(tokens), self.dimension), dtype=np.float32)
			for i, t in enumerate(tokens):
				j = self._index.get(t)
				if j is not None:
					data[i, :] = self._matrix[j]
			return Vectors(data)

	def __init__(self, name, wv):
		self._name = name
		if isinstance(wv, dict):
			words = list(wv.keys())
			matrix = np.array([wv[w] for w in words], dtype=np.float32)
		elif isinstance(wv, (tuple, list)) and len(wv) == 2:
			words, matrix = list(wv[0]), np.asarray(wv[1], dtype=np.float32)
		else:
			words, matrix = list(wv.index_to_key), np.asarray(wv.vectors, dtype=np.float32)
		if matrix.ndim != 2 or matrix.shape[0] != len(words):
			raise ValueError("expected one vector per word")
		self._words = words
		self._matrix = np.ascontiguousarray(matrix)

	def create_encoder(self, session):
		return KeyedVectors.Encoder(self, self._name, self._words, self._matrix)

	@property
	def name(self):
		return self._name


class ContextualEmbedding(TokenEmbedding):
	"""Per-token vectors that depend on the context (contextual.py:121-160).

	``encode`` is a callable: list of token texts of one document (or query) -> array [n_tokens x d].
	In the reference this is a spaCy transformer pipeline, optionally PCA-compressed (768 -> 128)."""

	class Encoder:
		def __init__(self, embedding):
			self._embedding = embedding

		@property
		def is_static(self):
			return False

		@property
		def is_contextual(self):
			return True

		@property
		def embedding(self):
			return self._embedding

		@property
		def name(self):
			return self._embedding.name

		@property
		def dimension(self):
			return self._embedding.dimension

		def encode(self, token_lists):
			return [np.asarray(self._embedding._encode(toks), dtype=np.float32) for toks in token_lists]

	def __init__(self, name, encode, dimension):
		self._name = name
		self._encode = encode
		self._dimension = int(dimension)

	@property
	def is_contextual(self):
		return True

	def create_encoder(self, session):
		return ContextualEmbedding.Encoder(self)

	@property
	def name(self):
		return self._name

	@property
	def dimension(self):
		return self._dimension
