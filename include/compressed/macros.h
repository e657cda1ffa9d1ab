// compressed/macros.h -- same namespace macro as the reference (compressed_image/include/compressed/macros.h).
#pragma once
#define NAMESPACE_COMPRESSED_IMAGE compressed
