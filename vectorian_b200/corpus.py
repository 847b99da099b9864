"""In-memory documents and corpus.

The reference stores documents in h5py + sqlite (vectorian/corpus/corpus.py:245-425) after a spaCy
import (vectorian/importers.py); both are absent here and outside the hot path.
What the hot path consumes is the token table ``{id/str, idx, len, pos, tag}`` and the span tables
``{level: {start, end}}`` of each document (corpus/document.py:626-662 -> core.Document,
vectorian/core/cpp/document.cpp:30-69); these classes hold exactly that.
"""

import re

import numpy as np


def whitespace_nlp(text):
	"""Minimal stand-in for a spaCy pipeline: one token per run of word characters or punctuation mark.
	Returns the list of token dicts PreparedQuery reads from ``doc.to_json()["tokens"]``
	(vectorian/index.py:76): start, end, pos, tag."""
	tokens = []
	for m in re.finditer(r"\w+|[^\w\s]", text, flags=re.UNICODE):
		tokens.append({'start': m.start(), 'end': m.end(), 'pos': 'X', 'tag': 'X'})
	return tokens


class Token:
	"""corpus/document.py:541-573"""

	def __init__(self, doc, index):
		self._doc = doc
		self._index = index

	@property
	def doc(self):
		return self._doc

	@property
	def index(self):
		return self._index

	def to_slice(self):
		offset = int(self._doc.token_idx[self._index])
		return slice(offset, offset + int(self._doc.token_len[self._index]))

	@property
	def text(self):
		return self._doc.text_str[self.to_slice()]

	@property
	def normalized_text(self):
		return self._doc.token_str[self._index]


class Span:
	"""corpus/document.py:576-623"""

	def __init__(self, doc, start, end):
		self._doc = doc
		self._start = int(start)
		self._end = int(end)

	@property
	def start(self):
		return self._start

	@property
	def end(self):
		return self._end

	def __len__(self):
		return self._end - self._start

	def __getitem__(self, i):
		n = self._end - self._start
		if i < 0 or i >= n:
			raise IndexError(f'{i} not in [0, {n}[')
		return Token(self._doc, self._start + i)

	def __iter__(self):
		for i in range(len(self)):
			yield self[i]

	@property
	def text(self):
		if len(self) == 0:
			return ""
		i0 = int(self._doc.token_idx[self._start])
		last = self._end - 1
		i1 = int(self._doc.token_idx[last]) + int(self._doc.token_len[last])
		return self._doc.text_str[i0:i1]


class Document:
	"""One document: text, token table and span tables (+ optional contextual vectors).

	tokens : list of str (token texts, already in text order)
	spans  : {level: {'start': int[], 'end': int[]}} token offsets, e.g. level 'sentence'
	pos/tag: optional lists of str per token (universal POS / Penn tag)
	"""

	def __init__(self, tokens, spans, text=None, pos=None, tag=None, token_idx=None, token_len=None,
			unique_id=None, metadata=None, contextual_embeddings=None):
		self.token_text = list(tokens)
		n = len(self.token_text)
		if text is None:
			text = " ".join(self.token_text)
			lens = np.array([len(t) for t in self.token_text], dtype=np.int64)
			token_idx = np.concatenate([[0], np.cumsum(lens + 1)[:-1]]) if n else np.zeros(0, np.int64)
			token_len = lens
		elif token_idx is None or token_len is None:
			raise ValueError("token_idx and token_len are needed when text is given")
		self.text_str = text
		self.token_idx = np.asarray(token_idx, dtype=np.int32)
		self.token_len = np.asarray(token_len, dtype=np.int32)
		self.token_pos = list(pos) if pos is not None else ['X'] * n
		self.token_tag = list(tag) if tag is not None else ['X'] * n
		if not (len(self.token_pos) == len(self.token_tag) == len(self.token_idx) == n):
			raise ValueError("token columns differ in length")
		self.spans = {}
		for level, table in spans.items():
			start = np.asarray(table['start'], dtype=np.int32)
			end = np.asarray(table['end'], dtype=np.int32)
			if start.shape != end.shape:
				raise ValueError(f"span table '{level}' is ragged")
			self.spans[level] = {'start': start, 'end': end}
		self.unique_id = unique_id
		self.metadata = metadata or {}
		self.contextual_embeddings = dict(contextual_embeddings or {})
		self.token_str = None   # normalised token strings, set by the session

	@staticmethod
	def from_sentences(sentences, pos=None, tag=None, **kwargs):
		"""sentences: list of list of str."""
		tokens = [t for s in sentences for t in s]
		lens = np.array([len(s) for s in sentences], dtype=np.int64)
		end = np.cumsum(lens)
		start = end - lens
		flat = lambda x: [t for s in x for t in s] if x is not None else None
		return Document(tokens, {'sentence': {'start': start, 'end': end}}, pos=flat(pos), tag=flat(tag), **kwargs)

	@property
	def n_tokens(self):
		return len(self.token_text)

	def has_contextual_embedding(self, name):
		return name in self.contextual_embeddings


class Corpus:
	"""An ordered collection of documents; the document id is the corpus order index
	(vectorian/corpus/corpus.py:370-381)."""

	def __init__(self, docs=None):
		self._docs = list(docs or [])

	def add_doc(self, doc):
		self._docs.append(doc)

	@property
	def docs(self):
		return self._docs

	def __len__(self):
		return len(self._docs)

	def __iter__(self):
		return iter(self._docs)

	def __getitem__(self, i):
		return self._docs[i]

	def get_doc_index(self, doc):
		return self._docs.index(doc)

	def get_unique_id(self, doc):
		return doc.unique_id if doc.unique_id is not None else self.get_doc_index(doc)
