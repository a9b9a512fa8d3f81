// graph/neighbor_list.hpp -- drop-in for the reference's include/graph/neighbor_list.hpp:11-24 (public part).
// Same names, argument meaning and defaults; the nanoflann/KD-tree private members are gone: construction runs the
// B200 kernels through the C ABI (include/dgnn_cabi.h) and materialises the rows on the host for legacy callers
// (WasmAPI::build_graph, src/viz/wasm_bindings.cpp:131-136; compute_structure_betti_features,
// src/topology/betti_features.cpp:107).  Row order: ascending distance with the canonical tie-break
// (distance, idx, n_a, n_b, n_c) -- the reference's std::sort leaves ties unspecified (neighbor_list.cpp:56-58).
#pragma once

#include "crystal/structure.hpp"

#include <Eigen/Dense>

#include <array>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace defect_gnn::graph {

struct Neighbor {
    size_t idx;
    double distance;
    Eigen::Vector3d displacement;  // image position - query position (neighbor_list.cpp:51)
};

class NeighborList {
public:
    explicit NeighborList(const crystal::Structure& structure,
                          double r_cutoff = 10.0,
                          size_t max_neighbors = 20,
                          double epsilon = 1e-10);

    // Unchecked like the reference (neighbor_list.cpp:23-25); safe for concurrent const reads.
    [[nodiscard]] const std::vector<Neighbor>& neighbors(size_t atom_idx) const;

    // ---- additions (not in the reference) ----
    // periodic image (n_a, n_b, n_c) of every neighbour, same row layout as neighbors()
    [[nodiscard]] const std::vector<std::array<int8_t, 3>>& images(size_t atom_idx) const;
    [[nodiscard]] size_t num_edges() const { return num_edges_; }
    [[nodiscard]] int64_t near_cutoff_pairs() const { return n_near_cutoff_; }  // |d - r_c| <= 1e-6 A
    [[nodiscard]] int64_t tie_rows() const { return n_tie_rows_; }              // rows truncated through a tie
    // device pointer to the row-ordered fp64 distances while they are still the current result of context `ctx`
    // (the constructing thread's dgnn_ctx), else nullptr
    [[nodiscard]] const double* device_distances(const void* ctx) const;

private:
    double r_cutoff_;
    size_t max_neighbors_;
    double epsilon_;
    size_t num_edges_ = 0;
    int64_t n_near_cutoff_ = 0;
    int64_t n_tie_rows_ = 0;
    std::vector<std::vector<Neighbor>> neighbor_lists_;
    std::vector<std::vector<std::array<int8_t, 3>>> images_;
    const void* device_ctx_ = nullptr;
    const double* device_dist_ = nullptr;
    uint64_t device_serial_ = 0;
};

// ---- additions (not in the reference): CUDA device the per-structure classes use on the calling thread.  Default:
// environment variable DGNN_DEVICE, else 0.  Every thread has its own context, so constructions on different threads
// (and devices) run concurrently.
void set_device(int device);
[[nodiscard]] int current_device();

}  // namespace defect_gnn::graph
