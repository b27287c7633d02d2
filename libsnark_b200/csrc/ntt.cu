#include "runtime.hpp"
extern "C" int b200_ntt(int, int, void *, size_t, const void *) { return b200::rt::set_error(B200_ERR_ARG, "ntt: not built yet"); }
