// graph::NeighborList on the B200 path.  Replaces the reference's src/graph/neighbor_list.cpp:14-94 (image cloud ->
// nanoflann KD-tree -> radius search -> sort -> truncate) by one dgnn_build call; the rows are then materialised on
// the host in the reference's own types.  Neighbor::displacement is re-derived here as
//   (pos_j + ((n_a*row0 + n_b*row1) + n_c*row2)) - pos_i
// in the reference's operand order (neighbor_list.cpp:82-87,51), so it is bit-identical without shipping 24 bytes
// per edge from the device.  Build this file without FMA contraction (-ffp-contract=off).
#include "graph/neighbor_list.hpp"

#include "context.hpp"

#include <limits>

namespace defect_gnn::graph {

NeighborList::NeighborList(const crystal::Structure& structure, double r_cutoff, size_t max_neighbors,
                           double epsilon)
    : r_cutoff_(r_cutoff), max_neighbors_(max_neighbors), epsilon_(epsilon) {
    const size_t n = structure.num_atoms();
    neighbor_lists_.resize(n);
    images_.resize(n);
    // SoA repack of the structure
    std::vector<double> x(n), y(n), z(n);
    for (size_t i = 0; i < n; ++i) {
        const Eigen::Vector3d& p = structure.atoms()[i].position;
        x[i] = p[0];
        y[i] = p[1];
        z[i] = p[2];
    }
    double lat[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) lat[3 * r + c] = structure.lattice()(r, c);
    const int64_t offsets[2] = {0, static_cast<int64_t>(n)};

    dgnn_batch b{};
    b.n_structs = 1;
    b.atom_offsets = offsets;
    b.x = x.data();
    b.y = y.data();
    b.z = z.data();
    b.lattice = lat;
    b.r_cutoff = r_cutoff_;
    b.max_neighbors = max_neighbors_ >= std::numeric_limits<uint64_t>::max() ? DGNN_UNCAPPED
                                                                             : static_cast<uint64_t>(max_neighbors_);
    b.epsilon = epsilon_;
    b.rbf_r_cutoff = r_cutoff_;
    b.dr = 0.0;
    b.flags = DGNN_SKIP_RBF;  // the neighbour list never asks for features

    std::vector<int64_t> row_ptr(n + 1, 0);
    std::vector<int32_t> dst, image;
    std::vector<double> dist;
    {
        detail::DeviceCtx& dc = detail::device_ctx();  // this thread's context on the selected device
        dgnn_graphs g{};
        int rc = dgnn_build(dc.ctx, &b, &g);
        if (rc != DGNN_OK) detail::throw_status(dc.ctx, rc, "NeighborList");
        dst.resize(static_cast<size_t>(g.n_edges));
        image.resize(static_cast<size_t>(g.n_edges));
        dist.resize(static_cast<size_t>(g.n_edges));
        rc = dgnn_copy_to_host(dc.ctx, &g, row_ptr.data(), dst.data(), image.data(), dist.data(), nullptr, nullptr);
        if (rc != DGNN_OK) detail::throw_status(dc.ctx, rc, "NeighborList");
        num_edges_ = static_cast<size_t>(g.n_edges);
        n_near_cutoff_ = g.n_near_cutoff;
        n_tie_rows_ = g.n_tie_rows;
        // the device-resident result stays current until this thread's next build: CrystalGraph reuses the distances
        device_ctx_ = dc.ctx;
        device_dist_ = g.dist;
        device_serial_ = dgnn_build_serial(dc.ctx);
    }
    for (size_t i = 0; i < n; ++i) {
        const auto e0 = static_cast<size_t>(row_ptr[i]), e1 = static_cast<size_t>(row_ptr[i + 1]);
        std::vector<Neighbor>& row = neighbor_lists_[i];
        std::vector<std::array<int8_t, 3>>& img = images_[i];
        row.resize(e1 - e0);
        img.resize(e1 - e0);
        for (size_t e = e0; e < e1; ++e) {
            const auto j = static_cast<size_t>(dst[e]);
            const uint32_t packed = static_cast<uint32_t>(image[e]);
            const int na = static_cast<int8_t>(packed & 0xffu), nb = static_cast<int8_t>((packed >> 8) & 0xffu),
                      nc = static_cast<int8_t>((packed >> 16) & 0xffu);
            Neighbor& nb_out = row[e - e0];
            nb_out.idx = j;
            nb_out.distance = dist[e];
            for (int k = 0; k < 3; ++k) {
                const double off = (static_cast<double>(na) * lat[k] + static_cast<double>(nb) * lat[3 + k]) +
                                   static_cast<double>(nc) * lat[6 + k];
                const double pj = (k == 0 ? x[j] : k == 1 ? y[j] : z[j]) + off;
                const double pi = k == 0 ? x[i] : k == 1 ? y[i] : z[i];
                nb_out.displacement[k] = pj - pi;
            }
            img[e - e0] = {static_cast<int8_t>(na), static_cast<int8_t>(nb), static_cast<int8_t>(nc)};
        }
    }
}

const std::vector<Neighbor>& NeighborList::neighbors(size_t atom_idx) const { return neighbor_lists_[atom_idx]; }

const std::vector<std::array<int8_t, 3>>& NeighborList::images(size_t atom_idx) const { return images_[atom_idx]; }

const double* NeighborList::device_distances(const void* ctx) const {
    return (ctx != nullptr && ctx == device_ctx_ && dgnn_build_serial(static_cast<const dgnn_ctx*>(ctx)) == device_serial_)
               ? device_dist_
               : nullptr;
}

void set_device(int device) { detail::set_current_device(device); }
int current_device() { return detail::current_device(); }

}  // namespace defect_gnn::graph
