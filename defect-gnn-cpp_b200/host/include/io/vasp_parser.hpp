// io/vasp_parser.hpp -- the INPUT types of the graph layer, interface-compatible with the reference's
// include/io/vasp_parser.hpp:10-18.  The POSCAR parser is out of scope for the B200 path (north_star: "unchanged");
// a maintainer integrating libdefect_gnn_graph keeps the reference's own io/ and crystal/ sources.
// This stand-alone copy of the interface exists so the facade can be built and tested without the reference tree.
#pragma once

#include <Eigen/Dense>

#include <string>
#include <vector>

namespace defect_gnn::io {

struct VASPStructure {
    Eigen::Matrix3d lattice;          // rows a, b, c in Angstrom (already multiplied by the POSCAR scale)
    std::vector<std::string> elements;
    std::vector<int> counts;
    Eigen::MatrixXd frac_coords;      // N x 3
    std::vector<int> atom_types;      // species index per atom
};

// POSCAR (VASP 5) text file -> VASPStructure; Direct or Cartesian coordinates.  Throws std::runtime_error.
VASPStructure parse_vasp(const std::string& filepath);

}  // namespace defect_gnn::io
