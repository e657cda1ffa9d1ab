// compressed/fwd.h -- forward declarations (fwd.h:1-14).
#pragma once

#include "macros.h"
#include "constants.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	template <typename T>
	struct channel;

	template <typename T>
	struct image;
}
