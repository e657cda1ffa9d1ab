// compressed/enums.h -- codec enum of the reference API (compressed_image/include/compressed/enums.h:14-20).
// The OpenImageIO TypeDesc mapping of the reference is part of its file-read path and is not provided.
#pragma once
#include "macros.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace enums
	{
		enum class codec
		{
			blosclz,
			lz4,
			lz4hc,   // decodes on the GPU (LZ4 block format); encoding is not part of the b2g engine
			zstd,    // decodes on the GPU; encoding is not part of the b2g engine
		};
	}
}
