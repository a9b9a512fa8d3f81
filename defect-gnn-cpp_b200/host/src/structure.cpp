// crystal::Structure -- stand-alone implementation of the reference interface (include/crystal/structure.hpp:12-38,
// behaviour of src/crystal/structure.cpp:7-66): frac -> Cartesian with lattice^T, fractional coordinates kept as
// given, minimum-image displacement by rounding the fractional difference.
#include "crystal/structure.hpp"

#include <cmath>

namespace defect_gnn::crystal {

Structure::Structure(const io::VASPStructure& vasp) : lattice_(vasp.lattice), inv_lattice_(vasp.lattice.inverse()) {
    for (size_t e = 0; e < vasp.counts.size(); ++e) counts_[static_cast<int>(e)] = vasp.counts[e];
    const auto n = static_cast<size_t>(vasp.frac_coords.rows());
    atoms_.reserve(n);
    for (size_t i = 0; i < n; ++i) {
        Atom a;
        a.element = vasp.atom_types[i];
        a.frac_position = vasp.frac_coords.row(static_cast<Eigen::Index>(i));
        // the reference's own expression (structure.cpp:17): with real Eigen the evaluation order -- (e0 + e1) + e2,
        // vectorised redux over a contiguous column of the lattice -- comes from the expression type, so keep it
        a.position = lattice_.transpose() * a.frac_position;
        atoms_.push_back(a);
    }
}

const Eigen::Matrix3d& Structure::lattice() const { return lattice_; }
const std::vector<Atom>& Structure::atoms() const { return atoms_; }
size_t Structure::num_atoms() const { return atoms_.size(); }

Eigen::Vector3d Structure::displacement(size_t i, size_t j) const {
    Eigen::Vector3d d = atoms_[j].frac_position - atoms_[i].frac_position;
    for (int k = 0; k < 3; ++k) d[k] -= std::round(d[k]);
    return lattice_.transpose() * d;
}

double Structure::distance(size_t i, size_t j) const { return displacement(i, j).norm(); }

Eigen::MatrixXd Structure::compute_distance_matrix() const {
    const auto n = static_cast<Eigen::Index>(atoms_.size());
    Eigen::MatrixXd D(n, n);
    for (Eigen::Index i = 0; i < n; ++i) {
        D(i, i) = 0.0;
        for (Eigen::Index j = i + 1; j < n; ++j) {
            const double d = distance(static_cast<size_t>(i), static_cast<size_t>(j));
            D(i, j) = d;
            D(j, i) = d;
        }
    }
    return D;
}

int Structure::count(int element) const { return counts_.at(element); }

}  // namespace defect_gnn::crystal
