#include "context.hpp"

#include <map>
#include <memory>
#include <stdexcept>

namespace defect_gnn::graph::detail {

namespace {
struct Registry {
    std::mutex mu;
    std::map<int, std::unique_ptr<DeviceCtx>> by_device;
    ~Registry() {
        for (auto& kv : by_device)
            if (kv.second && kv.second->ctx) dgnn_destroy(kv.second->ctx);
    }
};
Registry& registry() {
    static Registry r;
    return r;
}
}  // namespace

DeviceCtx& device_ctx(int device) {
    Registry& r = registry();
    std::lock_guard<std::mutex> lock(r.mu);
    auto it = r.by_device.find(device);
    if (it != r.by_device.end()) return *it->second;
    auto dc = std::make_unique<DeviceCtx>();
    const int rc = dgnn_create(device, 0, &dc->ctx);
    if (rc != DGNN_OK)
        throw std::runtime_error(std::string("defect_gnn::graph: cannot create the B200 context (") +
                                 dgnn_strerror(rc) + "); this build has no CPU fallback");
    DeviceCtx& ref = *dc;
    r.by_device.emplace(device, std::move(dc));
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
