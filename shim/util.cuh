// phantom-fhe "util.cuh" stand-in (dotprod_test.cu:16): PhantomBatchEncoder / PhantomPlaintext / PhantomCiphertext live in
// shim/phantom_bgv.h, the stream / pointer helpers in shim/phantom.h.
#pragma once
#include "phantom.h"
