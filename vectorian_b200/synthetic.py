"""Synthetic corpora, embeddings and queries of the shapes BASELINE.json names.

Zipf(1.0) token ids over V (id 0 most frequent), sentence lengths
clip(round(N(25, 8)), 3, 64), 1000 sentences per document, E ~ N(0, 1) (optionally
clustered so that alignments are non-trivial), queries of T distinct Zipf tokens.
Seeds: corpus 1234, embedding 4321, queries 99.
"""

import numpy as np

CORPUS_SEED, EMBEDDING_SEED, QUERY_SEED = 1234, 4321, 99


def zipf_cdf(vocab_size, s=1.0):
	p = 1.0 / np.arange(1, vocab_size + 1, dtype=np.float64) ** s
	cdf = np.cumsum(p)
	return cdf / cdf[-1]


def sentence_lengths(rng, n, mean=25.0, sd=8.0, lo=3, hi=64):
	return np.clip(np.rint(rng.normal(mean, sd, size=n)), lo, hi).astype(np.int32)


def make_corpus(n_sentences, vocab_size, seed=CORPUS_SEED, sentences_per_doc=1000, with_pos=False, n_pos=17):
	"""Returns a dict with the reference-shaped arrays (numpy, host)."""
	rng = np.random.default_rng(seed)
	lens = sentence_lengths(rng, n_sentences)
	n_tok = int(lens.sum())
	cdf = zipf_cdf(vocab_size)
	token_id = np.searchsorted(cdf, rng.random(n_tok), side="left").astype(np.int32)
	np.minimum(token_id, vocab_size - 1, out=token_id)
	n_doc = (n_sentences + sentences_per_doc - 1) // sentences_per_doc
	doc_span_off = np.minimum(np.arange(n_doc + 1, dtype=np.int64) * sentences_per_doc, n_sentences)
	ends = np.cumsum(lens.astype(np.int64))
	starts = ends - lens
	doc_tok_off = np.concatenate([starts[doc_span_off[:-1]], [n_tok]]).astype(np.int64) if n_sentences else np.zeros(1, np.int64)
	doc_of_sentence = np.repeat(np.arange(n_doc), np.diff(doc_span_off))
	span_start = (starts - doc_tok_off[doc_of_sentence]).astype(np.int32)
	span_end = (ends - doc_tok_off[doc_of_sentence]).astype(np.int32)
	out = dict(token_id=token_id, doc_tok_off=doc_tok_off, span_start=span_start, span_end=span_end,
		doc_span_off=doc_span_off, lens=lens, n_doc=n_doc, vocab_size=vocab_size)
	if with_pos:
		out["token_pos"] = rng.integers(0, n_pos, size=n_tok).astype(np.int8)
	return out


def make_embedding(vocab_size, dim, seed=EMBEDDING_SEED, clustered=True, n_centroids=2000, noise=0.3):
	rng = np.random.default_rng(seed)
	if not clustered:
		return rng.standard_normal((vocab_size, dim), dtype=np.float32)
	centroids = rng.standard_normal((n_centroids, dim), dtype=np.float32)
	assign = rng.integers(0, n_centroids, size=vocab_size)
	return centroids[assign] + noise * rng.standard_normal((vocab_size, dim), dtype=np.float32)


def make_query(t, vocab_size, seed=QUERY_SEED):
	"""T distinct token ids drawn from the corpus' Zipf distribution."""
	rng = np.random.default_rng(seed)
	cdf = zipf_cdf(vocab_size)
	ids = []
	while len(ids) < t:
		x = int(min(np.searchsorted(cdf, rng.random()), vocab_size - 1))
		if x not in ids:
			ids.append(x)
	return np.array(ids, dtype=np.int32)


def make_corpus_device(n_sentences, vocab_size, seed=CORPUS_SEED, sentences_per_doc=1000, device="cuda"):
	"""Same distribution generated on the GPU with torch (for the 50M-sentence bench corpus).

	Returns (token_id CUDA int32 tensor, host span tables dict).  The token stream differs from
	make_corpus (different RNG); parity checks on it read a slice back to the host.
	"""
	import torch
	g = torch.Generator(device=device)
	g.manual_seed(seed)
	lens = torch.clamp(torch.round(torch.normal(25.0, 8.0, size=(n_sentences,), generator=g, device=device)), 3, 64).to(torch.int32)
	n_tok = int(lens.sum().item())
	cdf = torch.from_numpy(zipf_cdf(vocab_size)).to(device)
	token_id = torch.empty((n_tok,), dtype=torch.int32, device=device)
	chunk = 1 << 26
	for i in range(0, n_tok, chunk):
		n = min(chunk, n_tok - i)
		u = torch.rand((n,), generator=g, device=device, dtype=torch.float64)
		token_id[i:i + n] = torch.clamp(torch.searchsorted(cdf, u), max=vocab_size - 1).to(torch.int32)
	lens_h = lens.cpu().numpy()
	n_doc = (n_sentences + sentences_per_doc - 1) // sentences_per_doc
	doc_span_off = np.minimum(np.arange(n_doc + 1, dtype=np.int64) * sentences_per_doc, n_sentences)
	ends = np.cumsum(lens_h.astype(np.int64))
	starts = ends - lens_h
	doc_tok_off = np.concatenate([starts[doc_span_off[:-1]], [n_tok]]).astype(np.int64)
	doc_of_sentence = np.repeat(np.arange(n_doc), np.diff(doc_span_off))
	span_start = (starts - doc_tok_off[doc_of_sentence]).astype(np.int32)
	span_end = (ends - doc_tok_off[doc_of_sentence]).astype(np.int32)
	return token_id, dict(doc_tok_off=doc_tok_off, span_start=span_start, span_end=span_end,
		doc_span_off=doc_span_off, lens=lens_h, n_doc=n_doc, vocab_size=vocab_size)
