// context.hpp -- internal: the facade's shared dgnn_ctx per CUDA device (a ctx is single-threaded; the facade
// serialises constructors on it with a mutex, results are copied out before the lock is released).
#pragma once

#include "dgnn_cabi.h"

#include <mutex>
#include <string>

namespace defect_gnn::graph::detail {

struct DeviceCtx {
    dgnn_ctx* ctx = nullptr;
    std::mutex mu;
};

// Lazily creates the context; throws std::runtime_error when no B200 is usable (there is no CPU fallback).
DeviceCtx& device_ctx(int device = 0);

// Translate a C-ABI status into std::runtime_error carrying dgnn_last_error().
[[noreturn]] void throw_status(dgnn_ctx* ctx, int code, const char* where);

}  // namespace defect_gnn::graph::detail
