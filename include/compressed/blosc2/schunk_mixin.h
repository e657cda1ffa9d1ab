// compressed/blosc2/schunk_mixin.h -- the reference shares the size / index checks of its two super-chunk classes through
// a mixin (blosc2/schunk_mixin.h:20-250); with one super-chunk class they are members of blosc2::schunk<T>.
#pragma once
#include "schunk.h"
