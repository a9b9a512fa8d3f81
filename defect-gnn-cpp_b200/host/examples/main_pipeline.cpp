// main_pipeline.cpp -- BASELINE config 1 driver: one POSCAR -> crystal::Structure -> NeighborList -> CrystalGraph with
// the reference's default parameters (r_c = 10, max_neighbors = 20, dr = 0.1), i.e. the MAIN-PIPELINE sequence of the
// reference's design doc (ETL_PIPELINE_PLAN.md:721-745; src/main.cpp itself only prints the parsed inputs).
//   usage: main_pipeline <file.vasp> [r_cutoff max_neighbors dr]
#include "crystal/structure.hpp"
#include "graph/crystal_graph.hpp"
#include "graph/neighbor_list.hpp"
#include "io/vasp_parser.hpp"

#include <cstdio>
#include <cstdlib>
#include <exception>
#include <map>

int main(int argc, char** argv) {
    if (argc < 2) {
        std::fprintf(stderr, "usage: %s <file.vasp> [r_cutoff max_neighbors dr]\n", argv[0]);
        return 2;
    }
    try {
        using namespace defect_gnn;
        const double rc = argc > 2 ? std::atof(argv[2]) : 10.0;
        const size_t cap = argc > 3 ? static_cast<size_t>(std::atoll(argv[3])) : 20;
        const double dr = argc > 4 ? std::atof(argv[4]) : 0.1;
        io::VASPStructure vasp = io::parse_vasp(argv[1]);
        crystal::Structure structure(vasp);
        graph::NeighborList nl(structure, rc, cap);
        std::map<int, Eigen::VectorXd> emb;  // species index -> 92-dim embedding (stand-in values)
        for (size_t e = 0; e < vasp.counts.size(); ++e) {
            Eigen::VectorXd v;
            v.resize(92);
            for (int d = 0; d < 92; ++d) v[d] = 0.0;
            emb[static_cast<int>(e)] = v;
        }
        graph::CrystalGraph g(structure, nl, emb, 92, rc, dr);
        std::printf("%s: %zu atoms, %zu edges, edge_attr %ld x %ld, near-cutoff pairs %lld, tie rows %lld\n", argv[1],
                    g.num_nodes(), g.num_edges(), static_cast<long>(g.edge_attr().rows()),
                    static_cast<long>(g.edge_attr().cols()), static_cast<long long>(nl.near_cutoff_pairs()),
                    static_cast<long long>(nl.tie_rows()));
        if (g.num_edges() > 0)
            std::printf("edge 0: %d -> %d, d = %.15g, g[0] = %.15g\n", g.edge_index()(0, 0), g.edge_index()(1, 0),
                        nl.neighbors(0)[0].distance, g.edge_attr()(0, 0));
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
