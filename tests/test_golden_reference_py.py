"""Golden vectors produced by the reference's OWN Python modules (tests/golden/make_golden.py imports
pyalign/gaps.py, vectorian/embedding/vectors.py, vectorian/sim/{kernel,vector}.py from /root/reference)
against this repo's host classes and against the C oracle."""

import os

import numpy as np
import pytest

from vectorian_b200 import gaps
from vectorian_b200.embedding import Vectors
from vectorian_b200.sim import vector as vsim, kernel as kern

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_py.npz"))

GAP_SPECS = {
	"constant_0": lambda: gaps.ConstantGapCost(0), "constant_0.25": lambda: gaps.ConstantGapCost(0.25),
	"linear_0.2": lambda: gaps.LinearGapCost(0.2), "affine_0.3_0.1": lambda: gaps.AffineGapCost(u=0.3, v=0.1),
	"affine_open1_ext0.25": lambda: gaps.AffineGapCost(1, 0.25),
	"log_0.2_0.1": lambda: gaps.LogarithmicGapCost(0.2, 0.1), "exp_2_0.2": lambda: gaps.ExponentialGapCost(2, 0.2),
	"smooth_5": lambda: gaps.smooth_gap_cost(5), "smooth_0": lambda: gaps.smooth_gap_cost(0),
	"user_sqrt": lambda: gaps.UserFuncGapCost(lambda k: 0.1 * np.sqrt(k)),
}


@pytest.mark.parametrize("name", list(GAP_SPECS.keys()))
def test_gap_cost_tables_bit_exact(name):
	g = GAP_SPECS[name]()
	want = G[f"gap/{name}"]
	got = np.asarray(g.costs(len(want)), dtype=np.float32)
	np.testing.assert_array_equal(got, want)
	sc = g.to_special_case()
	got_sc = np.array([sc.get("linear", np.nan)] + list(sc.get("affine", (np.nan, np.nan))), dtype=np.float64)
	np.testing.assert_array_equal(got_sc, G[f"gapspecial/{name}"])


def test_vectors_normalized_and_oracle():
	e = G["vec/e"]
	v = Vectors(e.copy())
	np.testing.assert_array_equal(v.magnitudes, G["vec/e_magnitudes"])
	np.testing.assert_array_equal(v.normalized, G["vec/e_normalized"])
	from oracle import oracle
	np.testing.assert_allclose(oracle.normalize_rows(e), G["vec/e_normalized"], rtol=5e-7, atol=1e-12)   # summation order of the norm: 1 ulp
	assert np.all(oracle.normalize_rows(e)[7] == 0)


def test_vector_sims_match_reference():
	e, q = G["vec/e"], G["vec/q"]
	ve, vq = Vectors(e.copy()), Vectors(q.copy())
	sims = {
		"cosine": vsim.CosineSim(),
		"pnorm2": vsim.ModifiedVectorSim(vsim.PNormDistance(p=2), kern.Scale(0.25), kern.DistanceToSimilarity()),
		"sqrtcos": vsim.ImprovedSqrtCosineSim(),
		"cos_thr": vsim.ModifiedVectorSim(vsim.CosineSim(), kern.Bias(0.1), kern.Power(2), kern.Threshold(0.2), kern.RadialBasis(0.5)),
		"fuzzy_jaccard": vsim.FuzzyJaccardSim(),
	}
	for name, s in sims.items():
		out = np.zeros((e.shape[0], q.shape[0]), dtype=np.float32)
		if name == "fuzzy_jaccard":
			s(Vectors(np.abs(e) + 0.01), Vectors(np.abs(q) + 0.01), out)
		else:
			s(ve, vq, out)
		np.testing.assert_allclose(out, G[f"sim/{name}"], rtol=1e-6, atol=1e-7, err_msg=name)
		assert s.name == str(G[f"simname/{name}"])


def test_oracle_similarity_matrix_against_reference_cosine():
	"""S of the oracle = clip01(reference CosineSim) with the identity fix-up (cpp/metric/static.cpp:58-75)."""
	from oracle import oracle
	e, q = G["vec/e"], G["vec/q"]
	q_id = np.array([3, 7, -1, 20, 59, 11], dtype=np.int32)
	want = G["sim/cosine"].copy()
	for j, i in enumerate(q_id):
		if i >= 0:
			want[i, j] = 1.0
	want = np.clip(want, 0, 1)
	got = oracle.sim_static(e, q_id, q)
	np.testing.assert_allclose(got, want, rtol=1e-5, atol=2e-7)


def test_token_similarity_modifiers_match_reference():
	"""vectorian_b200.sim.modifier against the outputs of the reference's own operators (vectorian/sim/modifier.py:22-117,
	invoked like ModifiedSimilarityMatrixFactory::create does, cpp/metric/modifier.cpp:18-74): bit-identical matrices and
	the same metric names (Match::metric)."""
	from vectorian_b200.sim import modifier as mod
	from vectorian_b200.sim.token import TokenSim

	class Operand(TokenSim):
		def __init__(self, name):
			self._n = name

		@property
		def name(self):
			return self._n

	a, b = Operand("A"), Operand("B")
	m_a, m_b = G["mod/a"], G["mod/b"]
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
		np.testing.assert_array_equal(res["similarity"], G[f"mod/{name}"], err_msg=name)
		assert m.name == str(G[f"modname/{name}"])
