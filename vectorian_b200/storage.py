"""Corpus storage: the reference's on-disk layout -> in-memory token tables, and a packed one-shot snapshot.

The reference keeps a corpus as a directory (vectorian/corpus/corpus.py:245-330):

  corpus.h5                documents/<unique_id>/  attrs["metadata"] (JSON)
                                                   tokens/{start,end,pos,tag,...}       corpus/document.py:196-212
                                                   spans/<level>/{start,end}            corpus/document.py:298-306
  corpus.db                sqlite: text(unique_id, content, content_hash)              corpus/corpus.py:266-276
  flavors/<name>.h5        mappings/{str,pos,tag}   string tables                       corpus/corpus.py:55-66, 225-233
                           documents/<unique_id>/{token_mask, span[n,2] = (idx, len), str, pos, tag}   corpus.py:121-182

and binds a document to a session by dropping the masked tokens and re-indexing the span tables with the prefix sum of
the mask (PreparedDocument, corpus/document.py:626-662).  `documents_from_h5_layout` restates exactly that over any
h5py-like mapping (h5py.File objects, or nested dicts in the tests -- h5py is not installable in the build image, so
`open_reference_corpus`, which opens the real files, raises ImportError there); the result feeds `Session`, whose
`device_corpus` packs it once into the device-resident tiles (vx_corpus_create).

`save_snapshot` / `load_snapshot` are this repo's own one-shot format: one .npz with the concatenated token columns of
all documents (no pickle, no per-document files), so a large corpus is re-opened without the HDF5 / sqlite stack.
"""

import json
import os

import numpy as np

from .corpus import Corpus, Document


def _as_str(x):
	if isinstance(x, bytes):
		return x.decode("utf8")
	return str(x)


def _mappings(flavor_root):
	"""FlavorCache.__init__ (corpus/corpus.py:55-66): id -> string tables of a flavor."""
	group = flavor_root["mappings"]
	return {k: [_as_str(v) for v in np.asarray(group[k]).tolist()] for k in group.keys()}


def documents_from_h5_layout(corpus_root, flavor_root, text_of, vectors_of=None):
	"""Documents of a reference corpus, in the order of corpus_root["documents"].

	corpus_root / flavor_root : h5py.File-like mappings with the layout above
	text_of(unique_id)        : the document text (corpus.db or documents/<uid>/document.txt)
	vectors_of(unique_id)     : optional {embedding name: [n_tokens_unmasked x d]} contextual vectors; rows are masked
	                            like the tokens (MaskedVectorsRef, corpus/document.py:651-652)
	"""
	mappings = _mappings(flavor_root)
	docs = []
	for unique_id in corpus_root["documents"].keys():
		group = corpus_root["documents"][unique_id]
		record = flavor_root["documents"][unique_id]
		token_mask = np.asarray(record["token_mask"], dtype=bool)
		span = np.asarray(record["span"])
		if span.size and np.any(span[:, 1] > np.iinfo(np.uint8).max):
			raise RuntimeError(f"token len > {np.iinfo(np.uint8).max} is not supported")      # corpus.py:38-40
		unmap = lambda attr: [mappings[attr][int(i)] for i in np.asarray(record[attr]).tolist()]
		strs, pos, tag = unmap("str"), unmap("pos"), unmap("tag")
		n = len(strs)
		if not (span.shape[0] == n == int(token_mask.sum())):
			raise ValueError(f"document {unique_id}: flavor columns and token mask differ in length")
		# span tables count unmasked tokens: re-index with the prefix sum of the mask (corpus/document.py:641-649)
		reindex = np.cumsum(np.concatenate(([False], token_mask)), dtype=np.int32)
		spans = {}
		for level in group["spans"].keys():
			table = group["spans"][level]
			spans[level] = {
				"start": reindex[np.asarray(table["start"], dtype=np.int64)],
				"end": reindex[np.asarray(table["end"], dtype=np.int64)]}
		attrs = getattr(group, "attrs", {})
		metadata = json.loads(_as_str(attrs["metadata"])) if "metadata" in attrs else {}
		contextual = {}
		if vectors_of is not None:
			for name, x in (vectors_of(unique_id) or {}).items():
				x = np.asarray(x, dtype=np.float32)
				if x.shape[0] != token_mask.shape[0]:
					raise ValueError(f"document {unique_id}: embedding {name} has {x.shape[0]} rows for {token_mask.shape[0]} tokens")
				contextual[name] = x[token_mask]
		docs.append(Document(
			strs, spans, text=text_of(unique_id), pos=pos, tag=tag,
			token_idx=span[:, 0] if n else np.zeros(0, np.int32), token_len=span[:, 1] if n else np.zeros(0, np.int32),
			unique_id=unique_id, metadata=metadata, contextual_embeddings=contextual))
	return docs


def open_reference_corpus(path, flavor, vectors_of=None):
	"""Corpus from a reference corpus directory (corpus.h5 + corpus.db + flavors/<flavor>.h5).  Needs h5py."""
	try:
		import h5py
	except ImportError as err:
		raise ImportError("reading the reference's corpus.h5 needs h5py, which is not installed; "
			"use Document / Corpus with in-memory token tables or load_snapshot") from err
	import sqlite3
	db = sqlite3.connect(os.path.join(path, "corpus.db"))

	def text_of(unique_id):
		row = db.execute("SELECT content FROM text WHERE unique_id = ?", (unique_id,)).fetchone()
		if row is not None:
			return row[0]
		with open(os.path.join(path, "documents", unique_id, "document.txt"), encoding="utf8") as f:   # corpus/document.py:309-316
			return f.read()

	with h5py.File(os.path.join(path, "corpus.h5"), "r") as corpus_root, \
			h5py.File(os.path.join(path, "flavors", flavor + ".h5"), "r") as flavor_root:
		docs = documents_from_h5_layout(corpus_root, flavor_root, text_of, vectors_of)
	db.close()
	return Corpus(docs)


# ---------------------------------------------------------------------------
# packed snapshot
# ---------------------------------------------------------------------------

_FORMAT = "vectorian_b200.corpus.v1"


def _string_table(strings):
	"""(ids, unique strings as one utf-8 blob + offsets)"""
	index = {}
	ids = np.empty((len(strings),), dtype=np.int32)
	for i, s in enumerate(strings):
		ids[i] = index.setdefault(s, len(index))
	blobs = [s.encode("utf8") for s in index.keys()]
	offsets = np.concatenate([[0], np.cumsum([len(b) for b in blobs])]).astype(np.int64)
	return ids, np.frombuffer(b"".join(blobs), dtype=np.uint8), offsets


def _strings_of(blob, offsets):
	raw = blob.tobytes()
	return [raw[offsets[i]:offsets[i + 1]].decode("utf8") for i in range(len(offsets) - 1)]


def save_snapshot(corpus, path):
	"""All documents of `corpus` in one .npz: token columns concatenated, string columns as id + table."""
	docs = list(corpus)
	tok_off = np.concatenate([[0], np.cumsum([d.n_tokens for d in docs])]).astype(np.int64)
	cat = lambda xs, dt: np.concatenate(xs).astype(dt) if xs else np.zeros((0,), dtype=dt)
	arrays = {"format": np.array(_FORMAT), "doc_tok_off": tok_off}
	for name, column in (("text", lambda d: d.token_text), ("pos", lambda d: d.token_pos), ("tag", lambda d: d.token_tag)):
		ids, blob, offsets = _string_table([s for d in docs for s in column(d)])
		arrays[f"token_{name}_id"], arrays[f"token_{name}_blob"], arrays[f"token_{name}_off"] = ids, blob, offsets
	arrays["token_idx"] = cat([d.token_idx for d in docs], np.int32)
	arrays["token_len"] = cat([d.token_len for d in docs], np.int32)
	texts = [d.text_str.encode("utf8") for d in docs]
	arrays["text_blob"] = np.frombuffer(b"".join(texts), dtype=np.uint8)
	arrays["text_off"] = np.concatenate([[0], np.cumsum([len(t) for t in texts])]).astype(np.int64)
	arrays["meta"] = np.array(json.dumps([{"unique_id": d.unique_id, "metadata": d.metadata} for d in docs]))
	levels = sorted({level for d in docs for level in d.spans})
	arrays["levels"] = np.array(json.dumps(levels))
	for li, level in enumerate(levels):
		counts = [len(d.spans[level]["start"]) if level in d.spans else -1 for d in docs]
		arrays[f"span{li}_count"] = np.array(counts, dtype=np.int64)
		arrays[f"span{li}_start"] = cat([d.spans[level]["start"] for d in docs if level in d.spans], np.int32)
		arrays[f"span{li}_end"] = cat([d.spans[level]["end"] for d in docs if level in d.spans], np.int32)
	names = sorted({name for d in docs for name in d.contextual_embeddings})
	arrays["embeddings"] = np.array(json.dumps(names))
	for ei, name in enumerate(names):
		if not all(name in d.contextual_embeddings for d in docs):
			raise ValueError(f"contextual embedding {name} is missing from some documents")
		arrays[f"ctx{ei}"] = np.concatenate([np.asarray(d.contextual_embeddings[name], dtype=np.float32) for d in docs])
	if not str(path).endswith(".npz"):
		path = str(path) + ".npz"
	np.savez(path, **arrays)
	return path


def load_snapshot(path):
	"""Corpus from a file written by save_snapshot."""
	with np.load(path, allow_pickle=False) as z:
		if str(z["format"]) != _FORMAT:
			raise ValueError(f"{path} is not a {_FORMAT} snapshot")
		tok_off = z["doc_tok_off"]
		columns = {}
		for name in ("text", "pos", "tag"):
			table = _strings_of(z[f"token_{name}_blob"], z[f"token_{name}_off"])
			columns[name] = [table[i] for i in z[f"token_{name}_id"].tolist()]
		token_idx, token_len = z["token_idx"], z["token_len"]
		text_raw, text_off = z["text_blob"].tobytes(), z["text_off"]
		meta = json.loads(str(z["meta"]))
		levels = json.loads(str(z["levels"]))
		span_tables = []
		for li, level in enumerate(levels):
			counts = z[f"span{li}_count"]
			off = np.concatenate([[0], np.cumsum(np.maximum(counts, 0))])
			span_tables.append((level, counts, off, z[f"span{li}_start"], z[f"span{li}_end"]))
		names = json.loads(str(z["embeddings"]))
		ctx = [z[f"ctx{ei}"] for ei in range(len(names))]
		docs = []
		for i, m in enumerate(meta):
			lo, hi = int(tok_off[i]), int(tok_off[i + 1])
			spans = {}
			for level, counts, off, start, end in span_tables:
				if counts[i] >= 0:
					spans[level] = {"start": start[off[i]:off[i + 1]], "end": end[off[i]:off[i + 1]]}
			docs.append(Document(
				columns["text"][lo:hi], spans, text=text_raw[text_off[i]:text_off[i + 1]].decode("utf8"),
				pos=columns["pos"][lo:hi], tag=columns["tag"][lo:hi],
				token_idx=token_idx[lo:hi], token_len=token_len[lo:hi],
				unique_id=m["unique_id"], metadata=m["metadata"],
				contextual_embeddings={name: x[lo:hi] for name, x in zip(names, ctx)}))
	return Corpus(docs)
