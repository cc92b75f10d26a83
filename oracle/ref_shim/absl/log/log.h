// Minimal stand-in for absl/log/log.h (test infrastructure only).
#ifndef REF_SHIM_ABSL_LOG_LOG_H_
#define REF_SHIM_ABSL_LOG_LOG_H_
#include <iostream>
#define LOG(level) std::cerr
#endif
