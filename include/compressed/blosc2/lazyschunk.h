// compressed/blosc2/lazyschunk.h -- the reference's lazy_schunk<T> (blosc2/lazyschunk.h:47-340) is the lazy_chunk
// alternative of this tree's single blosc2::schunk<T> (a super-chunk whose chunks are either compressed bytes or a fill
// value); the name is kept as an alias.
#pragma once
#include "schunk.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace blosc2
	{
		template <typename T>
		using lazy_schunk = schunk<T>;
	}
}
