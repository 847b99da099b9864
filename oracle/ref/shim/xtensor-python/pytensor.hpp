// the solver does not use pytensor; vectorian/core/lib/pyalign/pyalign/algorithm/common.h includes it
#include "../xt_shim.hpp"
