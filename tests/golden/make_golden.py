#!/usr/bin/env python
"""Generates tests/golden/reference_py.npz by IMPORTING the reference's own Python modules.

Run in the build container only (needs /root/reference, which does not travel to the GPU box):
    python tests/golden/make_golden.py
The reference package cannot be imported as a whole here (h5py 2.10, spaCy, gensim, cached_property
and the compiled vectorian_core are absent), so the pure-numpy modules on the hot path are loaded
file by file with two stub modules in sys.modules:
    pyalign/gaps.py                       -> GapCost.costs(n) tables           (pyalign/gaps.py:89-292)
    vectorian/embedding/vectors.py        -> Vectors.normalized / magnitudes   (embedding/vectors.py:56-86)
    vectorian/sim/kernel.py, sim/vector.py-> CosineSim / PNormDistance / ModifiedVectorSim / ... .compute (sim/vector.py:63-200, sim/kernel.py)
Inputs are seeded; the outputs are what the reference's code returned, unmodified.
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def load(name, path):
	spec = importlib.util.spec_from_file_location(name, path)
	mod = importlib.util.module_from_spec(spec)
	sys.modules[name] = mod
	spec.loader.exec_module(mod)
	return mod


def main():
	import functools
	stub = types.ModuleType("cached_property")
	stub.cached_property = functools.cached_property
	sys.modules["cached_property"] = stub
	sys.modules["h5py"] = types.ModuleType("h5py")
	for pkg in ("vectorian", "vectorian.sim", "vectorian.embedding"):
		m = types.ModuleType(pkg)
		m.__path__ = []
		sys.modules[pkg] = m

	gaps = load("ref_gaps", f"{REF}/vectorian/core/lib/pyalign/pyalign/gaps.py")
	vectors = load("vectorian.embedding.vectors", f"{REF}/vectorian/embedding/vectors.py")
	kern = load("vectorian.sim.kernel", f"{REF}/vectorian/sim/kernel.py")
	vsim = load("vectorian.sim.vector", f"{REF}/vectorian/sim/vector.py")

	out = {}
	n = 24
	specs = {
		"constant_0": gaps.ConstantGapCost(0), "constant_0.25": gaps.ConstantGapCost(0.25),
		"linear_0.2": gaps.LinearGapCost(0.2), "affine_0.3_0.1": gaps.AffineGapCost(u=0.3, v=0.1),
		"affine_open1_ext0.25": gaps.AffineGapCost(1, 0.25),
		"log_0.2_0.1": gaps.LogarithmicGapCost(0.2, 0.1), "exp_2_0.2": gaps.ExponentialGapCost(2, 0.2),
		"smooth_5": gaps.smooth_gap_cost(5), "smooth_0": gaps.smooth_gap_cost(0),
		"user_sqrt": gaps.UserFuncGapCost(lambda k: 0.1 * np.sqrt(k)),
	}
	for name, g in specs.items():
		out[f"gap/{name}"] = np.asarray(g.costs(n), dtype=np.float32)
		sc = g.to_special_case()
		out[f"gapspecial/{name}"] = np.array(
			[sc.get("linear", np.nan)] + list(sc.get("affine", (np.nan, np.nan))), dtype=np.float64)

	rng = np.random.default_rng(2024)
	e = rng.standard_normal((60, 20)).astype(np.float32)
	e[7, :] = 0            # vanishing vector
	e[11, :] *= 1e-7       # below the 100 * eps magnitude cut
	q = rng.standard_normal((6, 20)).astype(np.float32)
	ve, vq = vectors.Vectors(e.copy()), vectors.Vectors(q.copy())
	out["vec/e"], out["vec/q"] = e, q
	out["vec/e_normalized"] = np.asarray(ve.normalized)
	out["vec/e_magnitudes"] = np.asarray(ve.magnitudes)
	for name, s in (("cosine", vsim.CosineSim()), ("pnorm2", vsim.ModifiedVectorSim(vsim.PNormDistance(p=2), kern.Scale(0.25), kern.DistanceToSimilarity())),
			("sqrtcos", vsim.ImprovedSqrtCosineSim()), ("cos_thr", vsim.ModifiedVectorSim(vsim.CosineSim(), kern.Bias(0.1), kern.Power(2), kern.Threshold(0.2), kern.RadialBasis(0.5))), ("fuzzy_jaccard", vsim.FuzzyJaccardSim())):
		res = np.zeros((e.shape[0], q.shape[0]), dtype=np.float32)
		if name == "fuzzy_jaccard":
			s(vectors.Vectors(np.abs(e) + 0.01), vectors.Vectors(np.abs(q) + 0.01), res)
		else:
			s(ve, vq, res)
		out[f"sim/{name}"] = res
		out[f"simname/{name}"] = np.array(s.name)
	# token-similarity modifiers (vectorian/sim/modifier.py:22-117) on two operand matrices: the operator __call__ as the
	# reference's ModifiedSimilarityMatrixFactory invokes it (cpp/metric/modifier.cpp:18-74)
	token = load("vectorian.sim.token", f"{REF}/vectorian/sim/token.py")
	mod = load("vectorian.sim.modifier", f"{REF}/vectorian/sim/modifier.py")

	class Operand(token.TokenSim):
		def __init__(self, name):
			self._n = name

		@property
		def name(self):
			return self._n

	a, b = Operand("A"), Operand("B")
	m_a = rng.random((40, 6)).astype(np.float32)
	m_b = rng.random((40, 6)).astype(np.float32)
	out["mod/a"], out["mod/b"] = m_a, m_b
	mods = {
		"mixed_1_3": mod.MixedTokenSimilarity([a, b], [1, 3]),
		"maximum": mod.MaximumTokenSimilarity([a, b]),
		"minimum": mod.MinimumTokenSimilarity([a, b]),
		"unary_pow_thr": mod.UnaryTokenSimilarityModifier(a, [kern.Power(2.0), kern.Threshold(0.3)]),
		"unary_bias_scale": mod.UnaryTokenSimilarityModifier(b, [kern.Bias(-0.2), kern.Scale(1.5)]),
	}
	for name, m in mods.items():
		res = {"similarity": np.zeros_like(m_a)}
		operands = [{"similarity": m_a.copy()}, {"similarity": m_b.copy()}] if len(m.operands) == 2 else \
			[{"similarity": (m_a if m.operands[0] is a else m_b).copy()}]
		m(operands, res)
		out[f"mod/{name}"] = res["similarity"]
		out[f"modname/{name}"] = np.array(m.name)
	np.savez_compressed(os.path.join(HERE, "reference_py.npz"), **out)
	print("wrote", len(out), "arrays")


if __name__ == "__main__":
	main()
