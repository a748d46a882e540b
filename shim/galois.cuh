// phantom-fhe "galois.cuh" stand-in (dotprod_test.cu:14): PhantomGaloisKey lives in shim/phantom_bgv.h.
#pragma once
#include "phantom.h"
