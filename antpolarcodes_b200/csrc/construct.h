// Host-side code construction (frozen-set selection).  Same results as the reference's
// PolarCode::Construction::frozen_bits (src/polarcode/construction/constructor.cpp:41-63).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace polar {
// type: contains "be" -> beta expansion, "5g" -> 3GPP reliability table, anything else -> Bhattacharyya
// (case-insensitive, like the reference).  Returns false and sets err on invalid arguments.
bool construct_frozen_bits(int N, int K, float designSnr, const std::string& type, std::vector<uint32_t>& out,
                           std::string& err);
} // namespace polar
