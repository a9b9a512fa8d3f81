// facade_shim.cpp -- TEST code: a C-callable window onto the C++ facade (defect_gnn::graph::NeighborList,
// gaussian_rbf, CrystalGraph, BatchBuilder) so pytest can drive the reference-shaped API through ctypes and compare
// it with the oracle.  It constructs the objects exactly the way a reference caller would
// (VASPStructure -> crystal::Structure -> NeighborList -> CrystalGraph, ETL_PIPELINE_PLAN.md MAIN-PIPELINE).
#include "crystal/structure.hpp"
#include "graph/batch.hpp"
#include "io/vasp_batch.hpp"
#include "graph/crystal_graph.hpp"
#include "graph/edge_features.hpp"
#include "graph/neighbor_list.hpp"
#include "graph/sharded.hpp"

#include <cstring>
#include <exception>
#include <limits>
#include <map>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

using namespace defect_gnn;

namespace {
thread_local std::string g_err;

io::VASPStructure make_vasp(int64_t n, const double* frac, const double* lattice, const int* types) {
    io::VASPStructure v;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) v.lattice(r, c) = lattice[3 * r + c];
    v.frac_coords.resize(n, 3);
    int max_type = -1;
    for (int64_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) v.frac_coords(i, k) = frac[3 * i + k];
        v.atom_types.push_back(types ? types[i] : 0);
        if (v.atom_types.back() > max_type) max_type = v.atom_types.back();
    }
    v.counts.assign(static_cast<size_t>(max_type + 1), 0);
    for (int t : v.atom_types) v.counts[static_cast<size_t>(t)]++;
    v.elements.assign(v.counts.size(), "X");
    return v;
}

template <class F>
int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (const std::out_of_range& e) {
        g_err = e.what();
        return -20;
    } catch (const std::invalid_argument& e) {
        g_err = e.what();
        return -21;
    } catch (const std::exception& e) {
        g_err = e.what();
        return -22;
    }
}
}  // namespace

extern "C" {

const char* facade_last_error() { return g_err.c_str(); }

// Cartesian positions as crystal::Structure computes them (for feeding the oracle the identical inputs).
int facade_positions(int64_t n, const double* frac, const double* lattice, double* xyz /* n*3 */) {
    return guarded([&] {
        crystal::Structure s(make_vasp(n, frac, lattice, nullptr));
        for (int64_t i = 0; i < n; ++i)
            for (int k = 0; k < 3; ++k) xyz[3 * i + k] = s.atoms()[static_cast<size_t>(i)].position[k];
    });
}

int facade_neighbor_list(int64_t n, const double* frac, const double* lattice, double r_cutoff, uint64_t max_neighbors,
                         double epsilon, int use_defaults, int64_t edge_cap, int64_t* row_ptr, int32_t* dst,
                         double* dist, double* disp, int8_t* image, int64_t* n_edges, int64_t* near, int64_t* ties) {
    return guarded([&] {
        crystal::Structure s(make_vasp(n, frac, lattice, nullptr));
        const size_t cap = max_neighbors == std::numeric_limits<uint64_t>::max() ? std::numeric_limits<size_t>::max()
                                                                                  : static_cast<size_t>(max_neighbors);
        graph::NeighborList nl = use_defaults ? graph::NeighborList(s) : graph::NeighborList(s, r_cutoff, cap, epsilon);
        int64_t e = 0;
        for (int64_t i = 0; i < n; ++i) {
            row_ptr[i] = e;
            const auto& row = nl.neighbors(static_cast<size_t>(i));
            const auto& img = nl.images(static_cast<size_t>(i));
            for (size_t q = 0; q < row.size(); ++q, ++e) {
                if (e >= edge_cap) throw std::runtime_error("edge_cap too small");
                dst[e] = static_cast<int32_t>(row[q].idx);
                dist[e] = row[q].distance;
                for (int k = 0; k < 3; ++k) {
                    disp[3 * e + k] = row[q].displacement[k];
                    image[3 * e + k] = img[q][static_cast<size_t>(k)];
                }
            }
        }
        row_ptr[n] = e;
        *n_edges = e;
        *near = nl.near_cutoff_pairs();
        *ties = nl.tie_rows();
        if (static_cast<size_t>(e) != nl.num_edges()) throw std::runtime_error("num_edges mismatch");
    });
}

// Concurrent per-structure constructions (the OpenMP-over-files shape of preprocess_betti.cpp:37): n_threads threads each
// build Structure + NeighborList + CrystalGraph `reps` times on their own context; per thread the edge total and a
// checksum of edge_attr come back.  reuse[t] = how often CrystalGraph found the NeighborList's distances still in HBM.
int facade_concurrent(int64_t n, const double* frac, const double* lattice, int n_threads, int reps, int device,
                      int64_t* edges, double* checksum, int* reuse) {
    return guarded([&] {
        std::map<int, Eigen::VectorXd> emb;
        Eigen::VectorXd v(4);
        for (int d = 0; d < 4; ++d) v[d] = 0.5 * d;
        emb[0] = v;
        std::vector<std::string> errs(static_cast<size_t>(n_threads));
        std::vector<std::thread> pool;
        for (int t = 0; t < n_threads; ++t)
            pool.emplace_back([&, t] {
                try {
                    graph::set_device(device);
                    edges[t] = 0;
                    checksum[t] = 0.0;
                    reuse[t] = 0;
                    for (int r = 0; r < reps; ++r) {
                        crystal::Structure s(make_vasp(n, frac, lattice, nullptr));
                        graph::NeighborList nl(s, 10.0, 20, 1e-10);
                        const bool resident = nl.device_distances(nullptr) == nullptr;  // (null ctx never matches)
                        (void)resident;
                        graph::CrystalGraph g(s, nl, emb, 4, 10.0, 0.1);
                        edges[t] += static_cast<int64_t>(g.num_edges());
                        const Eigen::MatrixXd& ea = g.edge_attr();
                        double c = 0.0;
                        for (Eigen::Index i = 0; i < ea.rows(); ++i) c += ea(i, static_cast<Eigen::Index>(i % ea.cols()));
                        checksum[t] += c;
                        reuse[t] += graph::current_device() == device ? 1 : 0;
                    }
                } catch (const std::exception& e) {
                    errs[static_cast<size_t>(t)] = e.what();
                }
            });
        for (auto& th : pool) th.join();
        for (const auto& e : errs)
            if (!e.empty()) throw std::runtime_error(e);
    });
}

// emb_keys: which species indices get an embedding (value = key + 0.25 * dim); a missing key must throw out_of_range.
int facade_crystal_graph(int64_t n, const double* frac, const double* lattice, const int* types, double nl_cutoff,
                         uint64_t max_neighbors, double rbf_cutoff, double dr, int emb_dims, int n_emb_keys,
                         const int* emb_keys, int64_t edge_cap, int32_t* edge_index /* 2*E col-major */,
                         double* edge_attr /* E*K col-major */, double* node_features /* N*dims col-major */,
                         int64_t* n_edges, int32_t* K, int64_t* n_nodes) {
    return guarded([&] {
        crystal::Structure s(make_vasp(n, frac, lattice, types));
        const size_t cap = max_neighbors == std::numeric_limits<uint64_t>::max() ? std::numeric_limits<size_t>::max()
                                                                                  : static_cast<size_t>(max_neighbors);
        graph::NeighborList nl(s, nl_cutoff, cap);
        std::map<int, Eigen::VectorXd> emb;
        for (int q = 0; q < n_emb_keys; ++q) {
            Eigen::VectorXd v;
            v.resize(emb_dims);
            for (int d = 0; d < emb_dims; ++d) v[d] = emb_keys[q] + 0.25 * d;
            emb[emb_keys[q]] = v;
        }
        graph::CrystalGraph g(s, nl, emb, emb_dims, rbf_cutoff, dr);
        const auto E = static_cast<int64_t>(g.num_edges());
        if (E > edge_cap) throw std::runtime_error("edge_cap too small");
        *n_edges = E;
        *K = static_cast<int32_t>(g.edge_attr().cols());
        *n_nodes = static_cast<int64_t>(g.num_nodes());
        std::memcpy(edge_index, g.edge_index().data(), sizeof(int) * 2 * static_cast<size_t>(E));
        std::memcpy(edge_attr, g.edge_attr().data(), sizeof(double) * static_cast<size_t>(E) * static_cast<size_t>(*K));
        std::memcpy(node_features, g.node_features().data(),
                    sizeof(double) * static_cast<size_t>(n) * static_cast<size_t>(emb_dims));
        if (g.target() != 0.0) throw std::runtime_error("target must start at 0");
        g.set_target(1.5);
        if (g.target() != 1.5) throw std::runtime_error("set_target");
    });
}

int facade_gaussian_rbf(double d, double r_cutoff, double dr, double* out, int32_t* K) {
    return guarded([&] {
        Eigen::VectorXd v = graph::gaussian_rbf(d, r_cutoff, dr);
        *K = static_cast<int32_t>(v.size());
        for (int i = 0; i < *K; ++i) out[i] = v[i];
    });
}

// Batched surface: S structures given as concatenated frac coords; returns per-structure edge offsets and host copies.
int facade_build_batch(int64_t n_structs, const int64_t* atom_offsets, const double* frac, const double* lattices,
                       double r_cutoff, uint64_t max_neighbors, double rbf_cutoff, double dr, int64_t edge_cap,
                       int64_t* struct_edge_offsets, int64_t* row_ptr, int32_t* dst, int32_t* image, double* dist,
                       float* rbf, int64_t* n_edges, int32_t* K) {
    return guarded([&] {
        std::vector<crystal::Structure> structs;
        for (int64_t s = 0; s < n_structs; ++s)
            structs.emplace_back(make_vasp(atom_offsets[s + 1] - atom_offsets[s], frac + 3 * atom_offsets[s],
                                           lattices + 9 * s, nullptr));
        std::vector<const crystal::Structure*> ptrs;
        for (const auto& st : structs) ptrs.push_back(&st);
        graph::BatchParams p;
        p.r_cutoff = r_cutoff;
        p.max_neighbors = max_neighbors == std::numeric_limits<uint64_t>::max() ? std::numeric_limits<size_t>::max()
                                                                                 : static_cast<size_t>(max_neighbors);
        p.rbf_r_cutoff = rbf_cutoff;
        p.dr = dr;
        graph::BatchBuilder bb(0);
        graph::BatchedGraphs g = bb.build(ptrs, p);
        if (g.n_edges > edge_cap) throw std::runtime_error("edge_cap too small");
        *n_edges = g.n_edges;
        *K = g.K;
        for (int64_t s = 0; s <= n_structs; ++s) struct_edge_offsets[s] = g.host_struct_edge_offsets[static_cast<size_t>(s)];
        bb.copy_edges_to_host(g, row_ptr, dst, image, dist, g.K > 0 ? rbf : nullptr);
    });
}

// Batched collation + viewer arrays through the facade: embeddings value = key + 0.25 * dim (as above).
// missing_is_out_of_range: returns 2 when std::out_of_range escaped (the reference's map::at behaviour).
int facade_collate_batch(int64_t n_structs, const int64_t* atom_offsets, const double* frac, const double* lattices,
                         const int* types, double r_cutoff, uint64_t max_neighbors, double dr, int emb_dims,
                         int n_emb_keys, const int* emb_keys, int64_t edge_cap, float* x, int64_t* edge_index,
                         int64_t* batch, int* sources, int* targets, float* distances, float* displacements,
                         int64_t* n_edges, int32_t* num_graphs) {
    try {
        std::vector<crystal::Structure> structs;
        for (int64_t s = 0; s < n_structs; ++s)
            structs.emplace_back(make_vasp(atom_offsets[s + 1] - atom_offsets[s], frac + 3 * atom_offsets[s],
                                           lattices + 9 * s, types ? types + atom_offsets[s] : nullptr));
        graph::SoABatch soa;
        for (const auto& st : structs) soa.add(st);
        graph::BatchParams p;
        p.r_cutoff = r_cutoff;
        p.max_neighbors = max_neighbors == std::numeric_limits<uint64_t>::max() ? std::numeric_limits<size_t>::max()
                                                                                 : static_cast<size_t>(max_neighbors);
        p.rbf_r_cutoff = r_cutoff;
        p.dr = dr;
        graph::BatchBuilder bb(0);
        graph::BatchedGraphs g = bb.build(soa, p);
        if (g.n_edges > edge_cap) throw std::runtime_error("edge_cap too small");
        std::map<int, Eigen::VectorXd> emb;
        for (int q = 0; q < n_emb_keys; ++q) {
            Eigen::VectorXd v;
            v.resize(emb_dims);
            for (int d = 0; d < emb_dims; ++d) v[d] = emb_keys[q] + 0.25 * d;
            emb[emb_keys[q]] = v;
        }
        graph::CollatedBatch c = bb.collate(g, soa.element, emb);
        bb.copy_collated_to_host(c, x, edge_index, batch);
        graph::EdgeViews v = bb.edge_views(g);
        std::memcpy(sources, v.sources.data(), sizeof(int) * v.sources.size());
        std::memcpy(targets, v.targets.data(), sizeof(int) * v.targets.size());
        std::memcpy(distances, v.distances.data(), sizeof(float) * v.distances.size());
        std::memcpy(displacements, v.displacements.data(), sizeof(float) * v.displacements.size());
        *n_edges = g.n_edges;
        *num_graphs = c.num_graphs;
        return 0;
    } catch (const std::out_of_range& e) {
        g_err = e.what();
        return 2;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// io::parse_vasp of one file: lattice rows, fractional coordinates, species indices; returns the atom count via *n.
int facade_parse_vasp(const char* path, double* lattice, double* frac, int* types, int64_t cap, int64_t* n) {
    return guarded([&] {
        const io::VASPStructure v = io::parse_vasp(path);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) lattice[3 * i + j] = v.lattice(i, j);
        *n = v.frac_coords.rows();
        for (int64_t i = 0; i < *n && i < cap; ++i) {
            for (int j = 0; j < 3; ++j) frac[3 * i + j] = v.frac_coords(i, j);
            types[i] = v.atom_types[static_cast<size_t>(i)];
        }
    });
}

// Bulk ingest: newline-separated file paths -> SoA arrays (io::load_vasp_batch).
int facade_load_batch(const char* paths_nl, unsigned n_threads, int64_t atom_cap, int64_t struct_cap,
                      int64_t* atom_offsets, double* x, double* y, double* z, double* lattices, int* element,
                      int64_t* n_structs) {
    return guarded([&] {
        std::vector<std::string> paths;
        std::string cur;
        for (const char* q = paths_nl; *q; ++q) {
            if (*q == '\n') {
                if (!cur.empty()) paths.push_back(cur);
                cur.clear();
            } else {
                cur.push_back(*q);
            }
        }
        if (!cur.empty()) paths.push_back(cur);
        graph::SoABatch b = io::load_vasp_batch(paths, n_threads);
        if (b.n_structs() > struct_cap || b.n_atoms() > atom_cap) throw std::runtime_error("caps too small");
        *n_structs = b.n_structs();
        std::memcpy(atom_offsets, b.atom_offsets.data(), sizeof(int64_t) * b.atom_offsets.size());
        std::memcpy(x, b.x.data(), sizeof(double) * b.x.size());
        std::memcpy(y, b.y.data(), sizeof(double) * b.y.size());
        std::memcpy(z, b.z.data(), sizeof(double) * b.z.size());
        std::memcpy(lattices, b.lattice.data(), sizeof(double) * b.lattice.size());
        std::memcpy(element, b.element.data(), sizeof(int) * b.element.size());
    });
}

// Pinned bulk ingest + build from the pinned buffers: returns positions / offsets and the edge total of the build.
int facade_load_batch_pinned_build(const char* paths_nl, unsigned n_threads, int64_t atom_cap, int64_t struct_cap,
                                   int64_t* atom_offsets, double* x, double* y, double* z, double* lattices, int* element,
                                   int64_t* n_structs, double r_cutoff, uint64_t max_neighbors, int64_t* n_edges) {
    return guarded([&] {
        std::vector<std::string> paths;
        std::string cur;
        for (const char* q = paths_nl; *q; ++q) {
            if (*q == '\n') {
                if (!cur.empty()) paths.push_back(cur);
                cur.clear();
            } else {
                cur.push_back(*q);
            }
        }
        if (!cur.empty()) paths.push_back(cur);
        graph::PinnedSoABatch b = io::load_vasp_batch_pinned(paths, n_threads);
        if (b.n_structs > struct_cap || b.n_atoms > atom_cap) throw std::runtime_error("caps too small");
        *n_structs = b.n_structs;
        std::memcpy(atom_offsets, b.atom_offsets, sizeof(int64_t) * (b.n_structs + 1));
        std::memcpy(x, b.x, sizeof(double) * b.n_atoms);
        std::memcpy(y, b.y, sizeof(double) * b.n_atoms);
        std::memcpy(z, b.z, sizeof(double) * b.n_atoms);
        std::memcpy(lattices, b.lattice, sizeof(double) * 9 * b.n_structs);
        std::memcpy(element, b.element, sizeof(int) * b.n_atoms);
        graph::BatchBuilder builder(0);
        graph::BatchParams p;
        p.r_cutoff = r_cutoff;
        p.max_neighbors = max_neighbors == UINT64_MAX ? SIZE_MAX : static_cast<size_t>(max_neighbors);
        p.dr = 0.0;
        const graph::BatchedGraphs g = builder.build(b, p);
        *n_edges = g.n_edges;
    });
}

// host logic of the sharded builder (no GPU): contiguous structure ranges balanced by predicted edges
int facade_partition(int64_t n_structs, const int64_t* atom_offsets, const double* lattices, double r_cutoff,
                     int n_shards, int64_t* bounds, double* weights) {
    return guarded([&] {
        graph::SoABatch b;
        b.atom_offsets.assign(atom_offsets, atom_offsets + n_structs + 1);
        b.lattice.assign(lattices, lattices + 9 * n_structs);
        const std::vector<double> w = graph::predicted_edges(b, r_cutoff);
        const std::vector<int64_t> bd = graph::partition_structures(w, n_shards);
        std::memcpy(weights, w.data(), sizeof(double) * w.size());
        std::memcpy(bounds, bd.data(), sizeof(int64_t) * bd.size());
    });
}

// sharded build over `n_shards` contexts (device ids given; the same device may repeat), host copies concatenated
int facade_build_sharded(int64_t n_structs, const int64_t* atom_offsets, const double* frac, const double* lattices,
                         double r_cutoff, uint64_t max_neighbors, double dr, int n_shards, const int* devices,
                         int64_t edge_cap, int64_t* bounds, int64_t* global_edge_offsets, int32_t* dst, double* dist,
                         int64_t* n_edges) {
    return guarded([&] {
        graph::SoABatch soa;
        for (int64_t s = 0; s < n_structs; ++s)
            soa.add(crystal::Structure(make_vasp(atom_offsets[s + 1] - atom_offsets[s], frac + 3 * atom_offsets[s],
                                                 lattices + 9 * s, nullptr)));
        graph::BatchParams p;
        p.r_cutoff = r_cutoff;
        p.max_neighbors = max_neighbors == std::numeric_limits<uint64_t>::max() ? std::numeric_limits<size_t>::max()
                                                                                 : static_cast<size_t>(max_neighbors);
        p.rbf_r_cutoff = r_cutoff;
        p.dr = dr;
        graph::ShardedBuilder sb(std::vector<int>(devices, devices + n_shards));
        graph::ShardedGraphs g = sb.build(soa, p);
        if (g.n_edges() > edge_cap) throw std::runtime_error("edge_cap too small");
        *n_edges = g.n_edges();
        std::memcpy(bounds, g.struct_bounds.data(), sizeof(int64_t) * g.struct_bounds.size());
        std::memcpy(global_edge_offsets, g.global_edge_offsets.data(), sizeof(int64_t) * g.global_edge_offsets.size());
        int64_t e0 = 0;
        for (int r = 0; r < n_shards; ++r) {
            const graph::BatchedGraphs& sh = g.shards[static_cast<size_t>(r)];
            sb.builder(r).copy_edges_to_host(sh, nullptr, dst + e0, nullptr, dist + e0, nullptr);
            e0 += sh.n_edges;
        }
    });
}

}  // extern "C"
