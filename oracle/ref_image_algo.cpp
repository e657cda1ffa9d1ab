// oracle/ref_image_algo.cpp -- TEST INFRASTRUCTURE ONLY.
// Flat entry points over the UNMODIFIED compressed::image_algo::interleave / deinterleave of the reference
// (compressed_image/include/compressed/image_algo.h:34-65, :84-111; header-only, included where it lies under
// /root/reference), so that the numpy restatement the interleave tests check the kernels against is itself pinned.
#include <cstdint>
#include <ranges>      // the reference header uses std::views without including it
#include <span>
#include <stdexcept>
#include <vector>

#include "compressed/image_algo.h"

namespace {
template <typename T>
int run(bool to_interleaved, int nch, void* const* planes, void* inter, int64_t n) {
  try {
    if (to_interleaved) {
      std::vector<std::span<const T>> spans;
      for (int c = 0; c < nch; c++) spans.emplace_back(static_cast<const T*>(planes[c]), static_cast<size_t>(n));
      compressed::image_algo::interleave<T>(std::span<T>(static_cast<T*>(inter), static_cast<size_t>(n) * nch), spans);
    } else {
      std::vector<std::span<T>> spans;
      for (int c = 0; c < nch; c++) spans.emplace_back(static_cast<T*>(planes[c]), static_cast<size_t>(n));
      compressed::image_algo::deinterleave<T>(std::span<const T>(static_cast<const T*>(inter), static_cast<size_t>(n) * nch), spans);
    }
  } catch (const std::exception&) {
    return -1;
  }
  return 0;
}
int dispatch(bool to_interleaved, int typesize, int nch, void* const* planes, void* inter, int64_t n) {
  switch (typesize) {
    case 1: return run<uint8_t>(to_interleaved, nch, planes, inter, n);
    case 2: return run<uint16_t>(to_interleaved, nch, planes, inter, n);
    case 4: return run<uint32_t>(to_interleaved, nch, planes, inter, n);
    case 8: return run<uint64_t>(to_interleaved, nch, planes, inter, n);
    default: return -2;
  }
}
}  // namespace

extern "C" {
int ref_interleave(int typesize, int nch, void* const* planes, void* inter, int64_t n) {
  return dispatch(true, typesize, nch, planes, inter, n);
}
int ref_deinterleave(int typesize, int nch, const void* inter, void* const* planes, int64_t n) {
  return dispatch(false, typesize, nch, planes, const_cast<void*>(inter), n);
}
}
