// Drop-in for the reference header <gravomg/multigrid.h>: same include path, same namespace, same
// signatures. All declarations live in b200_facade.h; the implementations call the CUDA library
// through the C ABI (include/gravomg_b200.h).
#pragma once
#include "b200_facade.h"
