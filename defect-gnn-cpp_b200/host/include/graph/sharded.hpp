// graph/sharded.hpp -- NEW surface: one batch over several GPUs of a box.  Every structure's graph depends
// only on that structure, so the batch is cut into contiguous structure ranges, one per device, balanced by the
// predicted edge count  N_s * (N_s / V_s) * 4/3 pi r_c^3 ; every device has its own BatchBuilder (context, arena,
// stream) driven by its own host thread.  There is NO collective on the data path: the only exchange is the host-side
// scan of the per-shard edge counts into global edge offsets.  (bench.py does the same with one PROCESS per GPU under
// torchrun; this is the in-process form for a C++ caller.)
#pragma once

#include "graph/batch.hpp"

#include <memory>
#include <vector>

namespace defect_gnn::graph {

// per-structure edge estimate used for balancing
std::vector<double> predicted_edges(const SoABatch& batch, double r_cutoff);

// contiguous ranges [bounds[r], bounds[r+1]) of structures with near-equal total weight; bounds.size() = n_shards + 1
std::vector<int64_t> partition_structures(const std::vector<double>& weights, int n_shards);

struct ShardedGraphs {
    std::vector<int64_t> struct_bounds;          // [n_shards + 1] structure range of every shard
    std::vector<BatchedGraphs> shards;           // device-resident results, shard r lives on devices[r]
    std::vector<int64_t> global_edge_offsets;    // [S + 1] first edge of every structure in the concatenated edge list
    [[nodiscard]] int64_t n_edges() const { return global_edge_offsets.empty() ? 0 : global_edge_offsets.back(); }
};

class ShardedBuilder {
public:
    // one shard per entry; a device may appear more than once (separate contexts)
    explicit ShardedBuilder(const std::vector<int>& devices);
    ~ShardedBuilder();
    ShardedBuilder(const ShardedBuilder&) = delete;
    ShardedBuilder& operator=(const ShardedBuilder&) = delete;

    ShardedGraphs build(const SoABatch& batch, const BatchParams& params);
    [[nodiscard]] int n_shards() const { return static_cast<int>(builders_.size()); }
    BatchBuilder& builder(int shard) { return *builders_[static_cast<size_t>(shard)]; }

private:
    std::vector<int> devices_;
    std::vector<std::unique_ptr<BatchBuilder>> builders_;
};

}  // namespace defect_gnn::graph
