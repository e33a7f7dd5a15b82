// oracle/ref/zero_malloc.h -- TEST INFRASTRUCTURE, force-included when compiling the
// reference's reference_executor.cpp only.  The reference reads memory it never
// initialised (cg.cpp:41,69 x_padded; demos/spmv.cpp:78 y), which works by allocator
// luck; the oracle makes those reads deterministic by handing out zeroed memory.
// The standard headers are pulled in first so the macro cannot touch their declarations.
#pragma once
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <stdlib.h>
#include <typeinfo>
#define malloc(n) calloc(1, (n))
