// context.hpp -- internal: the facade's dgnn_ctx of the CALLING THREAD on a CUDA device.  A ctx is single-threaded, so
// every thread that constructs NeighborList / CrystalGraph objects gets its own (created on first use, destroyed when
// the thread ends): concurrent per-structure constructions -- the OpenMP-over-files upgrade of the reference's
// preprocess_betti.cpp:37 loop -- run side by side instead of queueing on one mutex.
#pragma once

#include "dgnn_cabi.h"

#include <string>

namespace defect_gnn::graph::detail {

struct DeviceCtx {
    dgnn_ctx* ctx = nullptr;
};

// The device the facade's per-structure classes use on this thread: graph::set_device(), else the environment variable
// DGNN_DEVICE, else 0.
int current_device();
void set_current_device(int device);

// Lazily creates the calling thread's context; throws std::runtime_error when no B200 is usable (no CPU fallback).
DeviceCtx& device_ctx(int device = -1);

// Translate a C-ABI status into std::runtime_error carrying dgnn_last_error().
[[noreturn]] void throw_status(dgnn_ctx* ctx, int code, const char* where);

}  // namespace defect_gnn::graph::detail
