"""The C-ABI library loads on a CPU-only box and exports every symbol include/vectorian_b200.h declares."""

import ctypes
import os
import re

import pytest

from vectorian_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
	text = open(os.path.join(ROOT, "include", "vectorian_b200.h")).read()
	text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
	return sorted(set(re.findall(r"\b(vx_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree():
	assert declared_symbols() == sorted(_cabi.EXPORTED_SYMBOLS)


def test_library_exports_every_symbol():
	if not os.path.exists(_cabi.LIB_PATH):
		import __graft_entry__
		__graft_entry__.build()
	lib = ctypes.CDLL(_cabi.LIB_PATH)
	for name in declared_symbols():
		assert hasattr(lib, name), name


def test_no_cpu_fallback():
	"""Without a GPU the product path must fail loudly (no compute call is made here)."""
	import torch
	if torch.cuda.is_available():
		pytest.skip("GPU present")
	lib = _cabi.load_library()
	assert lib.vx_init(0) != 0
	assert b"no CUDA device" in lib.vx_last_error() or b"CPU fallback" in lib.vx_last_error()
	from vectorian_b200 import engine
	with pytest.raises(_cabi.VxError):
		engine.DeviceEmbedding([[1.0, 2.0]])
