#include "../xt_shim.hpp"
