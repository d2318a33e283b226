// Helpers shared by the translation units of libfmpi_host.so (not part of the ABI).
#pragma once

#include <cstddef>
#include <cstdint>

namespace fmpi_host_detail {

// Records a thread-local message for fmpi_last_error() and returns 1.
int fail(const char* fmt, ...) __attribute__((format(printf, 1, 2)));

double now_s();
size_t dtype_size(int dtype);

// out[i] = op over r = 0..world-1 of slot_r[i], i in [lo, hi); slot_r = slots + r * stride.
// Folded in slot order; 16-bit floats accumulate in fp32 and round once.
void reduce_slots(int dtype, const uint8_t* slots, size_t stride, int world, uint8_t* out,
                  size_t lo, size_t hi, int op);

}  // namespace fmpi_host_detail
