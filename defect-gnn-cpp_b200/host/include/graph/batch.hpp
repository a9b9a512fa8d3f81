// graph/batch.hpp -- NEW surface (not in the reference, which builds one Structure per constructor call): the batched
// entry point the B200 path is designed around.  Structures are repacked once into SoA coordinate / lattice buffers
// (pinned host memory), one dgnn_build call constructs every graph of the batch on the device, and the result is a
// view of device-resident CSR + fp32 feature buffers; only counts / offsets come back to the host unless asked.
#pragma once

#include "crystal/structure.hpp"

#include <Eigen/Dense>

#include <cstddef>
#include <cstdint>
#include <map>
#include <memory>
#include <vector>

namespace defect_gnn::graph {

struct BatchParams {
    double r_cutoff = 10.0;         // NeighborList default (neighbor_list.hpp:20)
    size_t max_neighbors = 20;      // NeighborList default; SIZE_MAX = uncapped (betti_features.cpp:107)
    double epsilon = 1e-10;
    double rbf_r_cutoff = 10.0;     // CrystalGraph default (crystal_graph.hpp:18)
    double dr = 0.1;                // CrystalGraph default; <= 0: no features
    int device = 0;
};

// SoA repack of crystal::Structure objects (include/crystal/structure.hpp:12-38): Cartesian positions exactly as
// stored (never wrapped), lattice rows a, b, c.
struct SoABatch {
    std::vector<int64_t> atom_offsets;  // [S+1]
    std::vector<double> x, y, z;        // [N]
    std::vector<double> lattice;        // [9 S] row-major
    std::vector<int> element;           // [N] Atom::element (the key CrystalGraph looks the embedding up with)
    void add(const crystal::Structure& s);
    [[nodiscard]] int64_t n_structs() const { return (int64_t)atom_offsets.size() - 1; }
    [[nodiscard]] int64_t n_atoms() const { return atom_offsets.empty() ? 0 : atom_offsets.back(); }
    SoABatch() : atom_offsets{0} {}
};

// The same layout in page-locked host memory (dgnn_host_alloc): inputs that cross PCIe asynchronously, written in
// place by io::load_vasp_batch_pinned.  Move-only; frees its buffers on destruction.
struct PinnedSoABatch {
    int64_t n_structs = 0, n_atoms = 0;
    int64_t* atom_offsets = nullptr;  // [S+1]
    double *x = nullptr, *y = nullptr, *z = nullptr;  // [N]
    double* lattice = nullptr;        // [9 S]
    int* element = nullptr;           // [N]
    PinnedSoABatch() = default;
    PinnedSoABatch(int64_t n_structs, int64_t n_atoms);  // throws std::runtime_error if pinned memory is unavailable
    ~PinnedSoABatch();
    PinnedSoABatch(PinnedSoABatch&& o) noexcept;
    PinnedSoABatch& operator=(PinnedSoABatch&& o) noexcept;
    PinnedSoABatch(const PinnedSoABatch&) = delete;
    PinnedSoABatch& operator=(const PinnedSoABatch&) = delete;
};

// Device-resident result of one batched build.  Pointers are DEVICE pointers owned by the builder; they stay valid
// until the next build() on the same builder or its destruction.
struct BatchedGraphs {
    int64_t n_structs = 0, n_atoms = 0, n_edges = 0;
    int32_t K = 0, max_degree = 0;
    const int64_t* row_ptr = nullptr;              // [N+1] CSR by global source atom
    const int32_t* dst = nullptr;                  // [E] neighbour atom, local to its structure
    const int32_t* image = nullptr;                // [E] int8 n_a | n_b << 8 | n_c << 16
    const double* dist = nullptr;                  // [E]
    const float* rbf = nullptr;                    // [E*K] row-major fp32
    const int64_t* struct_edge_offsets = nullptr;  // [S+1]
    int64_t n_near_cutoff = 0, n_tie_rows = 0;
    std::vector<int64_t> host_struct_edge_offsets;  // copied back: the only D2H of the hot path
};

// The training consumer's batch (the planned train::BatchedGraph, PROJECT_ROADMAP.md:892-904) left on the device with
// the dtypes of the Python consumer (utilities.py:69-71).  DEVICE pointers owned by the builder, valid until its next
// build() / collate().
struct CollatedBatch {
    int64_t num_nodes = 0, num_edges = 0;
    int num_graphs = 0;
    int32_t F = 0, K = 0;
    const float* x = nullptr;             // [num_nodes, F]   node_features rows (crystal_graph.cpp:19-21), fp32
    const int64_t* edge_index = nullptr;  // [2, num_edges]   global node ids (per-structure offsets applied)
    const float* edge_attr = nullptr;     // [num_edges, K]   = BatchedGraphs::rbf
    const int64_t* batch = nullptr;       // [num_nodes]      graph membership
};

// The viewer's flat getters (src/viz/wasm_bindings.cpp:206-294) for every edge of the batch, host vectors.
struct EdgeViews {
    std::vector<int> sources, targets;        // atom indices local to the structure
    std::vector<float> distances;             // static_cast<float>(Neighbor::distance)
    std::vector<float> displacements;         // 3 per edge, static_cast<float>(Neighbor::displacement)
};

class BatchBuilder {
public:
    explicit BatchBuilder(int device = 0);
    ~BatchBuilder();
    BatchBuilder(const BatchBuilder&) = delete;
    BatchBuilder& operator=(const BatchBuilder&) = delete;

    // Throws std::runtime_error with the C-ABI error text on failure; there is no CPU fallback.
    BatchedGraphs build(const SoABatch& batch, const BatchParams& params);
    BatchedGraphs build(const PinnedSoABatch& batch, const BatchParams& params);  // asynchronous chunked upload
    BatchedGraphs build(const std::vector<const crystal::Structure*>& structures, const BatchParams& params);

    // Node features gathered on the device: x[i] = atom_embeddings.at(node_keys[i]); throws std::out_of_range where
    // std::map::at would (crystal_graph.cpp:20).  node_keys: one per atom of the batch (SoABatch::element).
    CollatedBatch collate(const BatchedGraphs& g, const std::vector<int>& node_keys,
                          const std::map<int, Eigen::VectorXd>& atom_embeddings);
    void copy_collated_to_host(const CollatedBatch& c, float* x, int64_t* edge_index, int64_t* batch) const;
    EdgeViews edge_views(const BatchedGraphs& g, bool with_displacements = true);

    // Host copies of the last result (any pointer may be null).
    void copy_edges_to_host(const BatchedGraphs& g, int64_t* row_ptr, int32_t* dst, int32_t* image, double* dist,
                            float* rbf) const;

private:
    struct Impl;
    std::unique_ptr<Impl> impl_;
};

}  // namespace defect_gnn::graph
