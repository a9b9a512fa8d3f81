// graph::ShardedBuilder -- structure-sharded multi-GPU build inside one process (see graph/sharded.hpp).
#include "graph/sharded.hpp"

#include <cmath>
#include <exception>
#include <stdexcept>
#include <thread>

namespace defect_gnn::graph {

std::vector<double> predicted_edges(const SoABatch& batch, double r_cutoff) {
    const int64_t S = batch.n_structs();
    std::vector<double> w(static_cast<size_t>(S));
    const double sphere = 4.0 / 3.0 * 3.14159265358979323846 * r_cutoff * r_cutoff * r_cutoff;
    for (int64_t s = 0; s < S; ++s) {
        const double* L = batch.lattice.data() + 9 * s;
        double vol = std::fabs(L[0] * (L[4] * L[8] - L[5] * L[7]) + L[1] * (L[5] * L[6] - L[3] * L[8]) +
                               L[2] * (L[3] * L[7] - L[4] * L[6]));
        if (!(vol > 0.0)) vol = 1.0;
        const double n = static_cast<double>(batch.atom_offsets[static_cast<size_t>(s) + 1] -
                                             batch.atom_offsets[static_cast<size_t>(s)]);
        w[static_cast<size_t>(s)] = n * (n / vol) * sphere;
    }
    return w;
}

std::vector<int64_t> partition_structures(const std::vector<double>& weights, int n_shards) {
    if (n_shards <= 0) throw std::invalid_argument("partition_structures: n_shards must be positive");
    const auto S = static_cast<int64_t>(weights.size());
    std::vector<double> cum(static_cast<size_t>(S) + 1, 0.0);
    for (int64_t s = 0; s < S; ++s) {
        if (!(weights[static_cast<size_t>(s)] >= 0.0)) throw std::invalid_argument("partition_structures: negative weight");
        cum[static_cast<size_t>(s) + 1] = cum[static_cast<size_t>(s)] + weights[static_cast<size_t>(s)];
    }
    std::vector<int64_t> bounds{0};
    for (int r = 1; r < n_shards; ++r) {
        const double target = cum.back() * r / n_shards;
        int64_t cut = bounds.back();
        while (cut < S && cum[static_cast<size_t>(cut)] < target) ++cut;  // first cut point at or beyond the target
        if (cut > bounds.back() &&
            std::fabs(cum[static_cast<size_t>(cut) - 1] - target) <= std::fabs(cum[static_cast<size_t>(cut)] - target))
            --cut;  // the closer of the two neighbouring cut points
        bounds.push_back(cut);
    }
    bounds.push_back(S);
    return bounds;
}

ShardedBuilder::ShardedBuilder(const std::vector<int>& devices) : devices_(devices) {
    if (devices.empty()) throw std::invalid_argument("ShardedBuilder: no devices");
    for (int d : devices) builders_.emplace_back(new BatchBuilder(d));
}

ShardedBuilder::~ShardedBuilder() = default;

ShardedGraphs ShardedBuilder::build(const SoABatch& batch, const BatchParams& params) {
    ShardedGraphs out;
    const int R = n_shards();
    out.struct_bounds = partition_structures(predicted_edges(batch, params.r_cutoff), R);
    out.shards.resize(static_cast<size_t>(R));
    std::vector<std::exception_ptr> errors(static_cast<size_t>(R));
    auto work = [&](int r) {
        try {
            const int64_t s0 = out.struct_bounds[static_cast<size_t>(r)], s1 = out.struct_bounds[static_cast<size_t>(r) + 1];
            // the shard's slice of the SoA batch (offsets rebased to 0)
            SoABatch part;
            const int64_t a0 = batch.atom_offsets[static_cast<size_t>(s0)], a1 = batch.atom_offsets[static_cast<size_t>(s1)];
            part.atom_offsets.resize(static_cast<size_t>(s1 - s0) + 1);
            for (int64_t s = s0; s <= s1; ++s)
                part.atom_offsets[static_cast<size_t>(s - s0)] = batch.atom_offsets[static_cast<size_t>(s)] - a0;
            part.x.assign(batch.x.begin() + a0, batch.x.begin() + a1);
            part.y.assign(batch.y.begin() + a0, batch.y.begin() + a1);
            part.z.assign(batch.z.begin() + a0, batch.z.begin() + a1);
            if (!batch.element.empty()) part.element.assign(batch.element.begin() + a0, batch.element.begin() + a1);
            part.lattice.assign(batch.lattice.begin() + 9 * s0, batch.lattice.begin() + 9 * s1);
            BatchParams p = params;
            p.device = devices_[static_cast<size_t>(r)];
            out.shards[static_cast<size_t>(r)] = builders_[static_cast<size_t>(r)]->build(part, p);
        } catch (...) {
            errors[static_cast<size_t>(r)] = std::current_exception();
        }
    };
    std::vector<std::thread> pool;
    for (int r = 1; r < R; ++r) pool.emplace_back(work, r);
    work(0);
    for (auto& t : pool) t.join();
    for (auto& e : errors)
        if (e) std::rethrow_exception(e);
    // the only cross-shard step: per-structure edge counts -> global edge offsets (host scan)
    out.global_edge_offsets.assign(1, 0);
    for (int r = 0; r < R; ++r) {
        const std::vector<int64_t>& seo = out.shards[static_cast<size_t>(r)].host_struct_edge_offsets;
        for (size_t s = 0; s + 1 < seo.size(); ++s)
            out.global_edge_offsets.push_back(out.global_edge_offsets.back() + (seo[s + 1] - seo[s]));
    }
    return out;
}

}  // namespace defect_gnn::graph
