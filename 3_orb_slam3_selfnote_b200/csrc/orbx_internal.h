// Internal helpers shared by the translation units of liborbx.so.
#pragma once
#include <stdarg.h>

namespace orbx {
void set_error(const char* fmt, ...);
}
