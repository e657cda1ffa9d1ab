// compressed/containers/chunk_span.h -- the reference keeps container::chunk_span<T> here (containers/chunk_span.h:20-110);
// in this tree it is declared in channel.h.
#pragma once
#include "../channel.h"
