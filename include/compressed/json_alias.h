// compressed/json_alias.h -- json_ordered, the type of image<T>::metadata() (json_alias.h:1-12).  nlohmann/json.hpp comes
// from the including project, as in the reference (thirdparty/json); without it on the include path the metadata
// accessors of image<T> are left out (COMPRESSED_IMAGE_HAVE_JSON undefined) and everything else still compiles.
#pragma once

#include "macros.h"

#if __has_include("nlohmann/json.hpp")
#include "nlohmann/json.hpp"
#define COMPRESSED_IMAGE_HAVE_JSON 1

namespace NAMESPACE_COMPRESSED_IMAGE
{
	using json_ordered = nlohmann::ordered_json;
}
#endif
