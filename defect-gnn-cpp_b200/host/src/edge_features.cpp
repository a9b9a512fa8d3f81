// graph::gaussian_rbf on the B200 path (reference: src/graph/edge_features.cpp:7-24).
#include "graph/edge_features.hpp"

#include "context.hpp"

#include <cmath>

namespace defect_gnn::graph {

Eigen::VectorXd gaussian_rbf(double distance, double r_cutoff, double dr) {
    const int n_bins = static_cast<int>(std::floor(r_cutoff / dr));  // edge_features.cpp:10
    Eigen::VectorXd out;
    out.resize(n_bins > 0 ? n_bins : 0);
    if (n_bins <= 0) return out;  // the reference returns an empty vector (g.resize(0), edge_features.cpp:15-16)
    detail::DeviceCtx& dc = detail::device_ctx();
    int32_t k = 0;
    const int rc = dgnn_rbf_expand_f64(dc.ctx, &distance, 1, r_cutoff, dr, out.data(), &k);
    if (rc != DGNN_OK) detail::throw_status(dc.ctx, rc, "gaussian_rbf");
    return out;
}

}  // namespace defect_gnn::graph
