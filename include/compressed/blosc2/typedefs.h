// compressed/blosc2/typedefs.h -- context_ptr / context_raw_ptr (blosc2/typedefs.h) are declared in wrapper.h.
#pragma once
#include "wrapper.h"
