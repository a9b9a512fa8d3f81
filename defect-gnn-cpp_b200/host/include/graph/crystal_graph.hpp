// graph/crystal_graph.hpp -- drop-in for the reference's include/graph/crystal_graph.hpp:12-41.
#pragma once

#include "crystal/structure.hpp"
#include "graph/neighbor_list.hpp"

#include <Eigen/Dense>

#include <map>

namespace defect_gnn::graph {

class CrystalGraph {
public:
    // Same semantics as the reference constructor (src/graph/crystal_graph.cpp:9-43): node features are gathered
    // with atom_embeddings.at(atom.element) (throws std::out_of_range on a missing key, like the reference);
    // edge_index is 2 x E (source ascending, then row order); edge_attr is E x floor(r_cutoff/dr), column-major
    // doubles, filled by ONE fp64 RBF launch over all edges instead of one gaussian_rbf call per edge.
    CrystalGraph(const crystal::Structure& structure,
                 const NeighborList& neighbors,
                 const std::map<int, Eigen::VectorXd>& atom_embeddings,
                 int atom_embedding_dims = 92,
                 double r_cutoff = 10,
                 double dr = 0.1);

    [[nodiscard]] const Eigen::MatrixXd& node_features() const;
    [[nodiscard]] const Eigen::MatrixXi& edge_index() const;
    [[nodiscard]] const Eigen::MatrixXd& edge_attr() const;

    [[nodiscard]] double target() const;
    void set_target(double y);

    void add_topo_features(const Eigen::MatrixXd& topo);  // empty in the reference too (crystal_graph.cpp:65-67)

    [[nodiscard]] size_t num_nodes() const;
    [[nodiscard]] size_t num_edges() const;

private:
    Eigen::MatrixXd node_features_;
    Eigen::MatrixXi edge_index_;
    Eigen::MatrixXd edge_attr_;
    double target_;
};

}  // namespace defect_gnn::graph
