// compressed/blosc2/util.h -- map_error_code (blosc2/util.h:16-19) is declared in wrapper.h.
#pragma once
#include "wrapper.h"
