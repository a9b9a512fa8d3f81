// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY.
//
// extern "C" driver over the reference's OWN classes, compiled from the reference's own sources where they lie
// (oracle/Makefile target `ref` -> oracle/_ref/libref_graph.so):
//     /root/reference/src/io/vasp_parser.cpp        io::parse_vasp
//     /root/reference/src/crystal/structure.cpp     crystal::Structure
//     /root/reference/src/graph/neighbor_list.cpp   graph::NeighborList
//     /root/reference/src/graph/edge_features.cpp   graph::gaussian_rbf
//     /root/reference/src/graph/crystal_graph.cpp   graph::CrystalGraph
// Nothing here re-implements the path: every result is read out of the reference's objects.  Only the two third-party
// headers are stand-ins (oracle/ref_shim/Eigen/Dense, oracle/ref_shim/nanoflann.hpp).  Used by tests/ (to pin the
// oracle restatement and the golden fixtures) and by bench.py --impl reference / cpu_baseline (kind
// "reference+shim").  The product library never links, loads or calls it.
#include "crystal/structure.hpp"
#include "graph/crystal_graph.hpp"
#include "graph/edge_features.hpp"
#include "graph/neighbor_list.hpp"
#include "io/vasp_parser.hpp"

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

using defect_gnn::crystal::Structure;
using defect_gnn::graph::CrystalGraph;
using defect_gnn::graph::Neighbor;
using defect_gnn::graph::NeighborList;
using defect_gnn::io::VASPStructure;

namespace {

VASPStructure make_vasp(int64_t n, const double* frac, const double* lattice, const int32_t* atom_types) {
    VASPStructure v;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) v.lattice(i, j) = lattice[3 * i + j];  // rows a, b, c (vasp_parser.cpp:32-38)
    v.frac_coords.resize(n, 3);
    v.atom_types.resize(static_cast<size_t>(n));
    int max_type = -1;
    for (int64_t i = 0; i < n; ++i) {
        for (int j = 0; j < 3; ++j) v.frac_coords(i, j) = frac[3 * i + j];
        v.atom_types[static_cast<size_t>(i)] = atom_types ? atom_types[i] : 0;
        max_type = std::max(max_type, v.atom_types[static_cast<size_t>(i)]);
    }
    v.counts.assign(static_cast<size_t>(max_type + 1), 0);
    for (int64_t i = 0; i < n; ++i) v.counts[static_cast<size_t>(v.atom_types[static_cast<size_t>(i)])]++;
    v.elements.assign(v.counts.size(), "X");
    return v;
}

struct RefList {
    std::unique_ptr<Structure> structure;
    std::unique_ptr<NeighborList> list;
    int64_t n_atoms = 0;
    int64_t n_edges = 0;
};

struct RefGraph {
    std::unique_ptr<CrystalGraph> graph;
};

thread_local std::string g_error;

}  // namespace

extern "C" {

const char* ref_last_error() { return g_error.c_str(); }

// crystal::Structure(VASPStructure) + graph::NeighborList(structure, r_cutoff, max_neighbors, epsilon)
void* ref_nl_create(int64_t n, const double* frac, const double* lattice, const int32_t* atom_types, double r_cutoff,
                    uint64_t max_neighbors, double epsilon) {
    try {
        auto h = std::make_unique<RefList>();
        h->structure = std::make_unique<Structure>(make_vasp(n, frac, lattice, atom_types));
        h->list = std::make_unique<NeighborList>(*h->structure, r_cutoff, static_cast<size_t>(max_neighbors), epsilon);
        h->n_atoms = static_cast<int64_t>(h->structure->num_atoms());
        for (int64_t i = 0; i < h->n_atoms; ++i)
            h->n_edges += static_cast<int64_t>(h->list->neighbors(static_cast<size_t>(i)).size());
        return h.release();
    } catch (const std::exception& e) {
        g_error = e.what();
        return nullptr;
    }
}

int64_t ref_nl_num_atoms(const void* h) { return static_cast<const RefList*>(h)->n_atoms; }
int64_t ref_nl_num_edges(const void* h) { return static_cast<const RefList*>(h)->n_edges; }

// rows in the reference's own order (std::sort by distance, ties as the reference left them)
void ref_nl_export(const void* hv, double* pos /*n*3*/, int64_t* row_ptr /*n+1*/, int64_t* idx /*E*/,
                   double* distance /*E*/, double* displacement /*E*3*/) {
    const RefList* h = static_cast<const RefList*>(hv);
    int64_t e = 0;
    for (int64_t i = 0; i < h->n_atoms; ++i) {
        if (pos)
            for (int k = 0; k < 3; ++k) pos[3 * i + k] = h->structure->atoms()[static_cast<size_t>(i)].position[k];
        if (row_ptr) row_ptr[i] = e;
        for (const Neighbor& nb : h->list->neighbors(static_cast<size_t>(i))) {
            if (idx) idx[e] = static_cast<int64_t>(nb.idx);
            if (distance) distance[e] = nb.distance;
            if (displacement)
                for (int k = 0; k < 3; ++k) displacement[3 * e + k] = nb.displacement[k];
            ++e;
        }
    }
    if (row_ptr) row_ptr[h->n_atoms] = e;
}

void ref_nl_free(void* h) { delete static_cast<RefList*>(h); }

// graph::gaussian_rbf(distance, r_cutoff, dr); returns K, writes K doubles if out != NULL
int ref_gaussian_rbf(double distance, double r_cutoff, double dr, double* out) {
    Eigen::VectorXd g = defect_gnn::graph::gaussian_rbf(distance, r_cutoff, dr);
    if (out)
        for (Eigen::Index k = 0; k < g.size(); ++k) out[k] = g[k];
    return static_cast<int>(g.size());
}

// graph::CrystalGraph(structure, neighbors, atom_embeddings, dims, r_cutoff, dr).  Returns NULL and sets
// ref_last_error() when the constructor throws (std::map::at on a missing element key, crystal_graph.cpp:20).
void* ref_graph_create(const void* hv, int n_keys, const int32_t* keys, const double* table /*n_keys*dims row-major*/,
                       int dims, double r_cutoff, double dr) {
    const RefList* h = static_cast<const RefList*>(hv);
    try {
        std::map<int, Eigen::VectorXd> emb;
        for (int k = 0; k < n_keys; ++k) {
            Eigen::VectorXd v(dims);
            for (int d = 0; d < dims; ++d) v[d] = table[static_cast<size_t>(k) * dims + d];
            emb[keys[k]] = v;
        }
        auto g = std::make_unique<RefGraph>();
        g->graph = std::make_unique<CrystalGraph>(*h->structure, *h->list, emb, dims, r_cutoff, dr);
        return g.release();
    } catch (const std::exception& e) {
        g_error = e.what();
        return nullptr;
    }
}

void ref_graph_dims(const void* gv, int64_t* num_nodes, int64_t* num_edges, int64_t* feat_dims, int64_t* K) {
    const CrystalGraph& g = *static_cast<const RefGraph*>(gv)->graph;
    *num_nodes = static_cast<int64_t>(g.num_nodes());
    *num_edges = static_cast<int64_t>(g.num_edges());
    *feat_dims = g.node_features().cols();
    *K = g.edge_attr().cols();
}

// raw storage of the three matrices, exactly as the reference holds them (column-major)
void ref_graph_export(const void* gv, double* node_features, int32_t* edge_index, double* edge_attr, double* target) {
    const CrystalGraph& g = *static_cast<const RefGraph*>(gv)->graph;
    if (node_features)
        std::memcpy(node_features, g.node_features().data(), sizeof(double) * static_cast<size_t>(g.node_features().size()));
    if (edge_index)
        std::memcpy(edge_index, g.edge_index().data(), sizeof(int) * static_cast<size_t>(g.edge_index().size()));
    if (edge_attr) std::memcpy(edge_attr, g.edge_attr().data(), sizeof(double) * static_cast<size_t>(g.edge_attr().size()));
    if (target) *target = g.target();
}

void ref_graph_free(void* g) { delete static_cast<RefGraph*>(g); }

// io::parse_vasp(path): returns the atom count (or -1), fills lattice[9] (rows a,b,c), and up to `cap` rows of frac /
// atom_types.
int64_t ref_parse_vasp(const char* path, double* lattice, double* frac, int32_t* atom_types, int64_t cap) {
    try {
        VASPStructure v = defect_gnn::io::parse_vasp(path);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) lattice[3 * i + j] = v.lattice(i, j);
        const int64_t n = v.frac_coords.rows();
        for (int64_t i = 0; i < n && i < cap; ++i) {
            for (int j = 0; j < 3; ++j) frac[3 * i + j] = v.frac_coords(i, j);
            atom_types[i] = v.atom_types[static_cast<size_t>(i)];
        }
        return n;
    } catch (const std::exception& e) {
        g_error = e.what();
        return -1;
    }
}

// Whole batch the way a caller of the reference would do it (one Structure + NeighborList + CrystalGraph per
// structure, preprocess_betti.cpp:37-79 loop shape), `threads` std::threads over structures; nothing retained.
typedef struct {
    int64_t n_edges;
    double checksum;
    double seconds;
    int32_t threads;
} ref_batch_result;

int ref_build_batch(int64_t n_structs, const int64_t* atom_offsets, const double* frac /*N*3*/,
                    const double* lattices /*S*9*/, const int32_t* atom_types /*N or NULL*/, double r_cutoff,
                    uint64_t max_neighbors, double epsilon, double rbf_r_cutoff, double dr, int embed_dims,
                    int threads, ref_batch_result* out) {
    unsigned nt = threads > 0 ? static_cast<unsigned>(threads) : std::max(1u, std::thread::hardware_concurrency());
    if (static_cast<int64_t>(nt) > n_structs) nt = static_cast<unsigned>(std::max<int64_t>(n_structs, 1));
    std::map<int, Eigen::VectorXd> emb;
    int max_type = 0;
    if (atom_types)
        for (int64_t i = 0; i < atom_offsets[n_structs]; ++i) max_type = std::max(max_type, atom_types[i]);
    for (int t = 0; t <= max_type; ++t) {
        Eigen::VectorXd v(embed_dims);
        for (int d = 0; d < embed_dims; ++d) v[d] = 0.01 * (t + 1) + 1e-3 * d;
        emb[t] = v;
    }
    std::atomic<int64_t> next{0};
    std::atomic<int> failed{0};
    std::vector<int64_t> edges(nt, 0);
    std::vector<double> sums(nt, 0.0);
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](unsigned tid) {
        try {
            for (;;) {
                const int64_t s = next.fetch_add(1);
                if (s >= n_structs) break;
                const int64_t a0 = atom_offsets[s], n = atom_offsets[s + 1] - a0;
                Structure st(make_vasp(n, frac + 3 * a0, lattices + 9 * s, atom_types ? atom_types + a0 : nullptr));
                NeighborList nl(st, r_cutoff, static_cast<size_t>(max_neighbors), epsilon);
                CrystalGraph g(st, nl, emb, embed_dims, rbf_r_cutoff, dr);
                edges[tid] += static_cast<int64_t>(g.num_edges());
                const Eigen::MatrixXd& ea = g.edge_attr();
                if (ea.size() > 0) sums[tid] += ea(0, 0) + ea(ea.rows() - 1, ea.cols() - 1);
            }
        } catch (const std::exception&) {
            failed.store(1);
        }
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < nt; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    auto t1 = std::chrono::steady_clock::now();
    out->n_edges = 0;
    out->checksum = 0;
    for (unsigned t = 0; t < nt; ++t) {
        out->n_edges += edges[t];
        out->checksum += sums[t];
    }
    out->seconds = std::chrono::duration<double>(t1 - t0).count();
    out->threads = static_cast<int32_t>(nt);
    return failed.load() ? -1 : 0;
}

}  // extern "C"
