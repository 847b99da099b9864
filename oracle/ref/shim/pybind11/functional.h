// stand-in: pyalign/algorithm/common.h includes pybind11, solver.h does not use it
