// qe_internal.h -- what the translation units of libqegpu.so share beyond the public C ABI
#ifndef QE_INTERNAL_H
#define QE_INTERNAL_H

#include <string>

// records the message for qe_last_error() on the calling thread and returns `code`
int qe_fail(int code, const std::string &msg);

#endif
