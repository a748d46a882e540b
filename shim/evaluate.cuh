// phantom-fhe "evaluate.cuh" stand-in (dotprod_test.cu:12): multiply, relinearize_inplace, mod_switch_to_next_inplace,
// rotate_inplace, add_inplace live in shim/phantom_bgv.h.
#pragma once
#include "phantom.h"
