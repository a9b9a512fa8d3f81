#include "context.hpp"

#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>

namespace defect_gnn::graph::detail {

namespace {
struct ThreadCtxs {
    std::map<int, std::unique_ptr<DeviceCtx>> by_device;
    ~ThreadCtxs() {
        for (auto& kv : by_device)
            if (kv.second && kv.second->ctx) dgnn_destroy(kv.second->ctx);
    }
};
thread_local ThreadCtxs t_ctxs;
thread_local int t_device = -1;

int env_device() {
    const char* e = std::getenv("DGNN_DEVICE");
    if (!e || !*e) return 0;
    char* end = nullptr;
    const long v = std::strtol(e, &end, 10);
    return (end && *end == '\0' && v >= 0 && v < 1024) ? static_cast<int>(v) : 0;
}
}  // namespace

int current_device() { return t_device >= 0 ? t_device : env_device(); }
void set_current_device(int device) { t_device = device; }

DeviceCtx& device_ctx(int device) {
    if (device < 0) device = current_device();
    auto it = t_ctxs.by_device.find(device);
    if (it != t_ctxs.by_device.end()) return *it->second;
    auto dc = std::make_unique<DeviceCtx>();
    const int rc = dgnn_create(device, 0, &dc->ctx);
    if (rc != DGNN_OK)
        throw std::runtime_error(std::string("defect_gnn::graph: cannot create the B200 context on device ") +
                                 std::to_string(device) + " (" + dgnn_strerror(rc) + "); this build has no CPU fallback");
    DeviceCtx& ref = *dc;
    t_ctxs.by_device.emplace(device, std::move(dc));
    return ref;
}

void throw_status(dgnn_ctx* ctx, int code, const char* where) {
    std::string msg = std::string("defect_gnn::graph::") + where + ": " + dgnn_strerror(code);
    const char* detail = ctx ? dgnn_last_error(ctx) : nullptr;
    if (detail && *detail) msg += std::string(" -- ") + detail;
    if (code == DGNN_ERR_INVALID_ARG) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
}

}  // namespace defect_gnn::graph::detail
