// graph::BatchBuilder -- batched, device-resident graph construction (new surface, see graph/batch.hpp).
#include "graph/batch.hpp"

#include "context.hpp"

#include <limits>
#include <utility>
#include <stdexcept>

namespace defect_gnn::graph {

void SoABatch::add(const crystal::Structure& s) {
    for (const crystal::Atom& a : s.atoms()) {
        x.push_back(a.position[0]);
        y.push_back(a.position[1]);
        z.push_back(a.position[2]);
        element.push_back(a.element);
    }
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) lattice.push_back(s.lattice()(r, c));
    atom_offsets.push_back(atom_offsets.back() + static_cast<int64_t>(s.num_atoms()));
}

struct BatchBuilder::Impl {
    dgnn_ctx* ctx = nullptr;
};

BatchBuilder::BatchBuilder(int device) : impl_(new Impl) {
    const int rc = dgnn_create(device, 0, &impl_->ctx);
    if (rc != DGNN_OK)
        throw std::runtime_error(std::string("BatchBuilder: cannot create the B200 context (") + dgnn_strerror(rc) +
                                 "); this build has no CPU fallback");
}

BatchBuilder::~BatchBuilder() {
    if (impl_ && impl_->ctx) dgnn_destroy(impl_->ctx);
}

namespace {
BatchedGraphs build_raw(dgnn_ctx* ctx, int64_t n_structs, const int64_t* atom_offsets, const double* x, const double* y,
                        const double* z, const double* lattice, const BatchParams& p) {
    dgnn_batch b{};
    b.n_structs = n_structs;
    b.atom_offsets = atom_offsets;
    b.x = x;
    b.y = y;
    b.z = z;
    b.lattice = lattice;
    b.r_cutoff = p.r_cutoff;
    b.max_neighbors = p.max_neighbors >= std::numeric_limits<uint64_t>::max() ? DGNN_UNCAPPED
                                                                             : static_cast<uint64_t>(p.max_neighbors);
    b.epsilon = p.epsilon;
    b.rbf_r_cutoff = p.rbf_r_cutoff;
    b.dr = p.dr;
    b.flags = p.dr > 0.0 ? 0u : DGNN_SKIP_RBF;
    dgnn_graphs g{};
    int rc = dgnn_build(ctx, &b, &g);
    if (rc != DGNN_OK) detail::throw_status(ctx, rc, "BatchBuilder::build");
    BatchedGraphs out;
    out.n_structs = g.n_structs;
    out.n_atoms = g.n_atoms;
    out.n_edges = g.n_edges;
    out.K = g.K;
    out.max_degree = g.max_degree;
    out.row_ptr = g.row_ptr;
    out.dst = g.dst;
    out.image = g.image;
    out.dist = g.dist;
    out.rbf = g.rbf;
    out.struct_edge_offsets = g.struct_edge_offsets;
    out.n_near_cutoff = g.n_near_cutoff;
    out.n_tie_rows = g.n_tie_rows;
    out.host_struct_edge_offsets.resize(static_cast<size_t>(g.n_structs) + 1);
    rc = dgnn_copy_to_host(ctx, &g, nullptr, nullptr, nullptr, nullptr, nullptr, out.host_struct_edge_offsets.data());
    if (rc != DGNN_OK) detail::throw_status(ctx, rc, "BatchBuilder::build");
    return out;
}
}  // namespace

BatchedGraphs BatchBuilder::build(const SoABatch& batch, const BatchParams& p) {
    return build_raw(impl_->ctx, batch.n_structs(), batch.atom_offsets.data(), batch.x.data(), batch.y.data(),
                     batch.z.data(), batch.lattice.data(), p);
}

BatchedGraphs BatchBuilder::build(const PinnedSoABatch& batch, const BatchParams& p) {
    return build_raw(impl_->ctx, batch.n_structs, batch.atom_offsets, batch.x, batch.y, batch.z, batch.lattice, p);
}

// ---- page-locked SoA buffers ------------------------------------------------------------------------------------------
namespace {
template <class T>
T* pinned(int64_t n) {
    void* p = nullptr;
    if (dgnn_host_alloc(static_cast<size_t>(n > 0 ? n : 1) * sizeof(T), &p) != DGNN_OK || !p)
        throw std::runtime_error("PinnedSoABatch: page-locked allocation failed (a CUDA device is required)");
    return static_cast<T*>(p);
}
}  // namespace

PinnedSoABatch::PinnedSoABatch(int64_t ns, int64_t na) : n_structs(ns), n_atoms(na) {
    try {
        atom_offsets = pinned<int64_t>(ns + 1);
        x = pinned<double>(na);
        y = pinned<double>(na);
        z = pinned<double>(na);
        lattice = pinned<double>(9 * ns);
        element = pinned<int>(na);
    } catch (...) {
        this->~PinnedSoABatch();
        throw;
    }
    atom_offsets[0] = 0;
}
PinnedSoABatch::~PinnedSoABatch() {
    for (void* p : {(void*)atom_offsets, (void*)x, (void*)y, (void*)z, (void*)lattice, (void*)element})
        if (p) dgnn_host_free(p);
    atom_offsets = nullptr;
    x = y = z = lattice = nullptr;
    element = nullptr;
}
PinnedSoABatch::PinnedSoABatch(PinnedSoABatch&& o) noexcept { *this = std::move(o); }
PinnedSoABatch& PinnedSoABatch::operator=(PinnedSoABatch&& o) noexcept {
    if (this != &o) {
        this->~PinnedSoABatch();
        n_structs = o.n_structs; n_atoms = o.n_atoms;
        atom_offsets = o.atom_offsets; x = o.x; y = o.y; z = o.z; lattice = o.lattice; element = o.element;
        o.atom_offsets = nullptr; o.x = o.y = o.z = o.lattice = nullptr; o.element = nullptr;
        o.n_structs = o.n_atoms = 0;
    }
    return *this;
}

BatchedGraphs BatchBuilder::build(const std::vector<const crystal::Structure*>& structures, const BatchParams& p) {
    SoABatch b;
    for (const crystal::Structure* s : structures) {
        if (!s) throw std::invalid_argument("BatchBuilder::build: null structure");
        b.add(*s);
    }
    return build(b, p);
}

namespace {
dgnn_graphs as_cabi(const BatchedGraphs& g) {
    dgnn_graphs v{};
    v.n_structs = g.n_structs;
    v.n_atoms = g.n_atoms;
    v.n_edges = g.n_edges;
    v.K = g.K;
    v.max_degree = g.max_degree;
    v.row_ptr = g.row_ptr;
    v.dst = g.dst;
    v.image = g.image;
    v.dist = g.dist;
    v.rbf = g.rbf;
    v.struct_edge_offsets = g.struct_edge_offsets;
    return v;
}
}  // namespace

CollatedBatch BatchBuilder::collate(const BatchedGraphs& g, const std::vector<int>& node_keys,
                                    const std::map<int, Eigen::VectorXd>& atom_embeddings) {
    if (static_cast<int64_t>(node_keys.size()) != g.n_atoms)
        throw std::invalid_argument("BatchBuilder::collate: one key per atom expected");
    std::vector<int32_t> keys;
    std::vector<double> table;
    int F = 0;
    for (const auto& kv : atom_embeddings) {
        if (keys.empty()) F = static_cast<int>(kv.second.size());
        if (static_cast<int>(kv.second.size()) != F)
            throw std::invalid_argument("BatchBuilder::collate: embeddings of different widths");
        keys.push_back(kv.first);
        for (int d = 0; d < F; ++d) table.push_back(kv.second[d]);
    }
    // map::at semantics up front: the exception type the reference's constructor lets escape
    for (int k : node_keys)
        if (atom_embeddings.find(k) == atom_embeddings.end()) throw std::out_of_range("map::at");
    const dgnn_graphs v = as_cabi(g);
    dgnn_collated c{};
    std::vector<int32_t> nk(node_keys.begin(), node_keys.end());
    const bool with_x = !keys.empty();
    const int rc = dgnn_collate(impl_->ctx, &v, with_x ? nk.data() : nullptr, with_x ? keys.data() : nullptr,
                                with_x ? table.data() : nullptr, static_cast<int32_t>(keys.size()), F, 0u, &c);
    if (rc != DGNN_OK) detail::throw_status(impl_->ctx, rc, "BatchBuilder::collate");
    CollatedBatch out;
    out.num_nodes = c.n_nodes;
    out.num_edges = c.n_edges;
    out.num_graphs = static_cast<int>(g.n_structs);
    out.F = c.F;
    out.K = c.K;
    out.x = c.x;
    out.edge_index = c.edge_index;
    out.edge_attr = c.edge_attr;
    out.batch = c.batch;
    return out;
}

void BatchBuilder::copy_collated_to_host(const CollatedBatch& c, float* x, int64_t* edge_index, int64_t* batch) const {
    dgnn_collated v{};
    v.n_nodes = c.num_nodes;
    v.n_edges = c.num_edges;
    v.F = c.F;
    v.K = c.K;
    v.x = c.x;
    v.edge_index = c.edge_index;
    v.batch = c.batch;
    v.edge_attr = c.edge_attr;
    const int rc = dgnn_copy_collated_to_host(impl_->ctx, &v, x, edge_index, batch);
    if (rc != DGNN_OK) detail::throw_status(impl_->ctx, rc, "BatchBuilder::copy_collated_to_host");
}

EdgeViews BatchBuilder::edge_views(const BatchedGraphs& g, bool with_displacements) {
    EdgeViews out;
    const auto e = static_cast<size_t>(g.n_edges);
    out.sources.resize(e);
    out.targets.resize(e);
    out.distances.resize(e);
    if (with_displacements) out.displacements.resize(3 * e);
    const dgnn_graphs v = as_cabi(g);
    const int rc = dgnn_edge_views(impl_->ctx, &v, out.sources.data(), out.targets.data(), out.distances.data(),
                                   with_displacements ? out.displacements.data() : nullptr);
    if (rc != DGNN_OK) detail::throw_status(impl_->ctx, rc, "BatchBuilder::edge_views");
    return out;
}

void BatchBuilder::copy_edges_to_host(const BatchedGraphs& g, int64_t* row_ptr, int32_t* dst, int32_t* image,
                                      double* dist, float* rbf) const {
    dgnn_graphs v{};
    v.n_structs = g.n_structs;
    v.n_atoms = g.n_atoms;
    v.n_edges = g.n_edges;
    v.K = g.K;
    v.row_ptr = g.row_ptr;
    v.dst = g.dst;
    v.image = g.image;
    v.dist = g.dist;
    v.rbf = g.rbf;
    v.struct_edge_offsets = g.struct_edge_offsets;
    const int rc = dgnn_copy_to_host(impl_->ctx, &v, row_ptr, dst, image, dist, rbf, nullptr);
    if (rc != DGNN_OK) detail::throw_status(impl_->ctx, rc, "BatchBuilder::copy_edges_to_host");
}

}  // namespace defect_gnn::graph
