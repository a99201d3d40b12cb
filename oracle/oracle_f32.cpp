/* TEST INFRASTRUCTURE ONLY -- fp32 decoders of the CPU oracle (placeholder: parity unpinned until written). */
#include "oracle.h"
#include "oracle_common.h"
namespace oracle {
int fastssc_f32_decode(int, const uint8_t*, int, int, const float*, uint8_t*) { return -3; }
int scl_f32_decode(int, int, const uint8_t*, int, int, const float*, uint8_t*) { return -3; }
}
