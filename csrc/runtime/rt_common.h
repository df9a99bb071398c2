// Shared bits of the host runtime library.
#pragma once
#include <stdint.h>

#include <memory>
#include <stdexcept>
#include <string>

#define PDB_RT_EXPORT extern "C" __attribute__((visibility("default")))

namespace pdbrt {
void set_error(const char* fmt, ...);
const char* last_error();
}  // namespace pdbrt
