// graph/edge_features.hpp -- drop-in for the reference's include/graph/edge_features.hpp:14.
#pragma once

#include <Eigen/Dense>

namespace defect_gnn::graph {

// Gaussian radial basis expansion of one distance: K = floor(r_cutoff / dr) bins, sigma = r_cutoff / 3,
//   g[k] = exp(-0.5 * (k*dr - distance)^2 / sigma^2) / (sigma * sqrt(2*pi))
// (reference src/graph/edge_features.cpp:7-24).  Evaluated on the GPU in fp64 (dgnn_rbf_expand_f64); one launch per
// call, so legacy per-edge use is slow -- CrystalGraph and build_batch expand all edges in one launch instead.
Eigen::VectorXd gaussian_rbf(double distance, double r_cutoff, double dr);

}  // namespace defect_gnn::graph
