"""Corpus storage: the reference's HDF5 layout read through an h5py-like mapping, and the packed snapshot."""

import json

import numpy as np
import pytest

from vectorian_b200 import storage
from vectorian_b200.corpus import Corpus, Document
from vectorian_b200.session import Session


class Group(dict):
	"""Stand-in for an h5py group: a mapping with `attrs` (h5py itself is not installable here)."""

	def __init__(self, *args, attrs=None, **kwargs):
		super().__init__(*args, **kwargs)
		self.attrs = attrs or {}


def reference_layout():
	"""Two documents written the way vectorian/corpus writes them: token tables with ALL tokens in corpus.h5 (span tables
	count those), the flavor keeps only the unmasked tokens (corpus/corpus.py:121-182)."""
	text_a = "The cat ,  sat on a mat . Done"
	#          0   1   2  3   4  5 6   7 8
	toks_a = ["The", "cat", ",", "sat", "on", "a", "mat", ".", "Done"]
	start_a = [0, 4, 8, 11, 15, 18, 20, 24, 26]
	mask_a = [True, True, False, True, True, True, True, False, True]          # punctuation dropped by the flavor
	pos_a = ["DET", "NOUN", "PUNCT", "VERB", "ADP", "DET", "NOUN", "PUNCT", "VERB"]
	text_b = "dogs bark"
	toks_b = ["dogs", "bark"]
	mappings = {"str": [], "pos": [], "tag": []}

	def ids(attr, values):
		out = []
		for v in values:
			if v not in mappings[attr]:
				mappings[attr].append(v)
			out.append(mappings[attr].index(v))
		return np.array(out, dtype=np.uint8)

	def flavor_doc(toks, starts, mask, pos):
		keep = [i for i, m in enumerate(mask) if m]
		return Group({
			"token_mask": np.array(mask, dtype=bool),
			"span": np.array([[starts[i], len(toks[i])] for i in keep], dtype=np.uint32).reshape(-1, 2),
			"str": ids("str", [toks[i].lower() for i in keep]),
			"pos": ids("pos", [pos[i] for i in keep]),
			"tag": ids("tag", ["T" + pos[i][:2] for i in keep])})

	flavor = Group({"documents": Group({
		"doc-a": flavor_doc(toks_a, start_a, mask_a, pos_a),
		"doc-b": flavor_doc(toks_b, [0, 5], [True, True], ["NOUN", "VERB"])})})
	flavor["mappings"] = Group({k: np.array([s.encode("utf8") for s in v], dtype=object) for k, v in mappings.items()})
	corpus = Group({"documents": Group({
		"doc-a": Group({"spans": Group({"sentence": Group({"start": np.array([0, 8]), "end": np.array([8, 9])})})},
			attrs={"metadata": json.dumps({"title": "A", "author": "x"})}),
		"doc-b": Group({"spans": Group({"sentence": Group({"start": np.array([0]), "end": np.array([2])})})},
			attrs={"metadata": json.dumps({"title": "B"})})})})
	texts = {"doc-a": text_a, "doc-b": text_b}
	return corpus, flavor, texts, np.array(mask_a)


def test_documents_from_reference_layout():
	corpus_root, flavor_root, texts, mask_a = reference_layout()
	rng = np.random.default_rng(0)
	x_a = rng.standard_normal((9, 4)).astype(np.float32)
	docs = storage.documents_from_h5_layout(corpus_root, flavor_root, texts.__getitem__,
		vectors_of=lambda uid: {"ctx": x_a} if uid == "doc-a" else {})
	assert [d.unique_id for d in docs] == ["doc-a", "doc-b"]
	a, b = docs
	assert a.token_text == ["the", "cat", "sat", "on", "a", "mat", "done"]          # masked tokens are gone
	assert a.token_pos == ["DET", "NOUN", "VERB", "ADP", "DET", "NOUN", "VERB"] and a.token_tag[0] == "TDE"
	# spans re-indexed by the prefix sum of the mask (corpus/document.py:641-649): tokens 0..8 -> 0..6, 8..9 -> 6..7
	np.testing.assert_array_equal(a.spans["sentence"]["start"], [0, 6])
	np.testing.assert_array_equal(a.spans["sentence"]["end"], [6, 7])
	assert [a.text_str[i:i + n] for i, n in zip(a.token_idx, a.token_len)] == ["The", "cat", "sat", "on", "a", "mat", "Done"]
	assert a.metadata == {"title": "A", "author": "x"} and b.metadata == {"title": "B"}
	np.testing.assert_array_equal(a.contextual_embeddings["ctx"], x_a[mask_a])
	assert b.token_text == ["dogs", "bark"]
	# a Session binds them like any other corpus
	session = Session(Corpus(docs), [])
	assert session.vocab.size == 9
	start, end = session.documents[0].span_table("sentence")
	assert list(end - start) == [6, 1]
	with pytest.raises(ValueError):
		storage.documents_from_h5_layout(corpus_root, flavor_root, texts.__getitem__, vectors_of=lambda uid: {"ctx": x_a[:5]})


def test_open_reference_corpus_needs_h5py(tmp_path):
	try:
		import h5py  # noqa: F401
	except ImportError:
		with pytest.raises(ImportError, match="h5py"):
			storage.open_reference_corpus(str(tmp_path), "default")


def test_snapshot_round_trip(tmp_path):
	corpus_root, flavor_root, texts, _ = reference_layout()
	rng = np.random.default_rng(1)
	docs = storage.documents_from_h5_layout(corpus_root, flavor_root, texts.__getitem__,
		vectors_of=lambda uid: {"ctx": rng.standard_normal((9 if uid == "doc-a" else 2, 3)).astype(np.float32)})
	docs.append(Document.from_sentences([["ünï", "cödé"], []], unique_id=None, contextual_embeddings={"ctx": np.ones((2, 3), np.float32)}))
	path = storage.save_snapshot(Corpus(docs), tmp_path / "snap")
	loaded = storage.load_snapshot(path)
	assert len(loaded) == 3
	for d0, d1 in zip(docs, loaded):
		assert d0.token_text == d1.token_text and d0.token_pos == d1.token_pos and d0.token_tag == d1.token_tag
		assert d0.text_str == d1.text_str and d0.unique_id == d1.unique_id and d0.metadata == d1.metadata
		np.testing.assert_array_equal(d0.token_idx, d1.token_idx)
		np.testing.assert_array_equal(d0.token_len, d1.token_len)
		assert d0.spans.keys() == d1.spans.keys()
		for level in d0.spans:
			np.testing.assert_array_equal(d0.spans[level]["start"], d1.spans[level]["start"])
			np.testing.assert_array_equal(d0.spans[level]["end"], d1.spans[level]["end"])
		np.testing.assert_array_equal(d0.contextual_embeddings["ctx"], d1.contextual_embeddings["ctx"])
	s0, s1 = Session(Corpus(docs), []), Session(loaded, [])
	for p0, p1 in zip(s0.documents, s1.documents):
		np.testing.assert_array_equal(p0.token_id, p1.token_id)
		np.testing.assert_array_equal(p0.token_pos, p1.token_pos)
	with pytest.raises(ValueError):
		np.savez(tmp_path / "other.npz", format=np.array("something else"))
		storage.load_snapshot(tmp_path / "other.npz")
