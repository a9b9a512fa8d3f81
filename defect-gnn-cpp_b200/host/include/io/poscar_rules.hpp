// io/poscar_rules.hpp -- the reference parser's semantics (src/io/vasp_parser.cpp:13-78) in one place, shared by the
// per-file parser (src/vasp_parser.cpp) and the bulk ingest (src/vasp_batch.cpp).  The north_star keeps the reference's
// io layer unchanged, so a file must give the same Structure through either parser here as through the reference's:
//
//   * line 8 selects Direct iff its FIRST character is 'd' or 'D' (vasp_parser.cpp:52) -- no whitespace skipping, and
//     anything else (Cartesian, Kartesisch, a blank line) is Cartesian;
//   * Cartesian coordinates are converted as  frac = cart * inverse(lattice)  with the lattice already multiplied by
//     the scale factor and the coordinates NOT (vasp_parser.cpp:72-75) -- VASP itself would scale them;
//   * lattice.inverse() and the N x 3 by 3 x 3 product are evaluated in Eigen 3.4's order on x86-64 (cofactor inverse,
//     InverseImpl.h; lazy coefficient product with slice-vectorised assignment), restated in oracle/ref_shim/Eigen/Dense
//     and held against oracle/_ref by tests/test_facade.py.
//
// One deliberate difference, listed in INTEGRATION.md section 1: a "Selective dynamics" line before the mode line is
// skipped (and the T/F flags after the coordinates ignored).  The reference does not know the keyword: it takes the
// line for the mode line, reads the first coordinate triple from the real mode line and leaves what the failed
// extraction does not set uninitialised -- there is no defined behaviour to reproduce.
#pragma once

#include <cstddef>
#include <cstdint>

namespace defect_gnn::io::rules {

// vasp_parser.cpp:52
inline bool is_direct(const char* b, const char* e) { return b < e && (*b == 'd' || *b == 'D'); }
inline bool is_selective(const char* b, const char* e) { return b < e && (*b == 's' || *b == 'S'); }

// Eigen::Matrix3d::inverse() of a row-major 3x3 given as L[3*i + j] (InverseImpl.h, compute_inverse_size3_helper):
// cofactor(i,j) = m(i1,j1) m(i2,j2) - m(i1,j2) m(i2,j1); det = (c00 m00 + c10 m10) + c20 m20; inv(i,j) = cofactor(j,i)/det
inline void inverse3(const double* L, double* inv) {
    auto m = [&](int i, int j) { return L[3 * i + j]; };
    auto cof = [&](int i, int j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return m(i1, j1) * m(i2, j2) - m(i1, j2) * m(i2, j1);
    };
    const double c0 = cof(0, 0), c1 = cof(1, 0), c2 = cof(2, 0);
    const double det = (c0 * m(0, 0) + c1 * m(1, 0)) + c2 * m(2, 0);
    const double id = 1.0 / det;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) inv[3 * i + j] = cof(j, i) * id;
}

// Row i of  frac = cart * inv  for an N-row coordinate matrix (vasp_parser.cpp:74).  Eigen evaluates the product
// column by column into an aligned temporary: rows inside the column's run of whole packets (two doubles) accumulate
// (e0 + e1) + e2, the scalar rows before / after it e0 + (e1 + e2); the run of column j starts at row (j * (N % 2)) % 2.
inline void cart_to_frac_row(int64_t i, int64_t N, const double* cart, const double* inv, double* frac) {
    const int64_t odd = N % 2;
    for (int j = 0; j < 3; ++j) {
        int64_t a0 = (j * odd) % 2;
        if (a0 > N) a0 = N;
        const int64_t a1 = a0 + ((N - a0) & ~int64_t(1));
        const double e0 = cart[0] * inv[j], e1 = cart[1] * inv[3 + j], e2 = cart[2] * inv[6 + j];
        frac[j] = (i >= a0 && i < a1) ? (e0 + e1) + e2 : e0 + (e1 + e2);
    }
}

}  // namespace defect_gnn::io::rules
