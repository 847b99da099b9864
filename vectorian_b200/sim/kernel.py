"""Unary operators that turn a vector distance / similarity into the similarity the alignment
consumes; same classes and call signatures as vectorian/sim/kernel.py:6-97.  They run on the host
over the [V x T] matrix (once per query), exactly where the reference runs them."""

import numpy as np


class UnaryOperator:
	def kernel(self, data):
		raise NotImplementedError()

	def name(self, operand):
		raise NotImplementedError()


class RadialBasis(UnaryOperator):
	def __init__(self, gamma):
		self._gamma = gamma

	def kernel(self, data):
		data[:, :] = np.exp(-self._gamma * np.power(data, 2))

	def name(self, operand):
		return f'radialbasis({operand}, {self._gamma})'


class DistanceToSimilarity(UnaryOperator):
	def kernel(self, data):
		data[:, :] = np.maximum(0, 1 - data)

	def name(self, operand):
		return f'(1 - {operand})'


class Bias(UnaryOperator):
	def __init__(self, bias):
		self._bias = bias

	def kernel(self, data):
		data += self._bias

	def name(self, operand):
		return f'({operand} + {self._bias})'


class Scale(UnaryOperator):
	def __init__(self, scale):
		self._scale = scale

	def kernel(self, data):
		data *= self._scale

	def name(self, operand):
		return f'({operand} * {self._scale})'


class Power(UnaryOperator):
	def __init__(self, exp):
		self._exp = exp

	def kernel(self, data):
		data[:, :] = np.power(np.maximum(data, 0), self._exp)

	def name(self, operand):
		return f'({operand} ** {self._exp})'


class Threshold(UnaryOperator):
	def __init__(self, threshold):
		self._threshold = threshold

	def kernel(self, data):
		x = np.maximum(0, data - self._threshold)
		x[x > 0] += self._threshold
		data[:, :] = x

	def name(self, operand):
		return f'threshold({operand}, {self._threshold})'


class Kernel:
	def __init__(self, operators):
		self._operators = operators

	def __call__(self, data):
		for x in self._operators:
			x.kernel(data)

	def name(self, operand):
		name = operand
		for x in self._operators:
			name = x.name(name)
		return name
