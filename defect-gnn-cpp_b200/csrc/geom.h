// geom.h -- per-structure geometry, uniform cell grid and stencil planning.  Host/device shared, so the same code
// can be exercised on the CPU (tests/cpu_harness) and inside the kernels.
//
// Replaces the sizing logic of the reference's image cloud (src/graph/neighbor_list.cpp:68-94): instead of
// replicating all atoms over the (2n+1)^3 cube and building a KD-tree, atoms' periodic images are binned into a
// uniform grid of parallelepiped cells (lattice vectors divided by g_k) and only the images that fall inside the
// halo of the occupied cells are materialised.  The reference cube |n_k| <= n is kept as a hard clip (rule R3).
#pragma once

#include <stdint.h>
#include <math.h>
#include <limits.h>

#if defined(__CUDACC__)
#define DG_HD __host__ __device__ __forceinline__
#else
#define DG_HD inline
#endif

// Unfused IEEE fp64 ops: the reference is compiled without FMA (CMakeLists.txt:33-37); every product and sum that
// reaches the accept test must round separately.
#if defined(__CUDA_ARCH__)
#define DG_MUL(a, b) __dmul_rn((a), (b))
#define DG_ADD(a, b) __dadd_rn((a), (b))
#define DG_SUB(a, b) __dsub_rn((a), (b))
#else
#define DG_MUL(a, b) ((a) * (b))
#define DG_ADD(a, b) ((a) + (b))
#define DG_SUB(a, b) ((a) - (b))
#endif

#define DG_MAX_REACH 15       // stencil reach limit per axis (tile dims <= 32)
#define DG_MAX_TILE_DIM 32
#define DG_RC_MARGIN 1e-6     // relative safety margin on r_c for candidate generation (never for acceptance)
#define DG_MAX_HOME_DIM 4096
// Candidate-generation radius: r_c plus a relative and an absolute margin (covers the 1e-6 A near-cutoff band and
// all rounding in the fractional-coordinate binning).  Never used for acceptance.
#define DG_RCM(rc) ((rc) * (1.0 + DG_RC_MARGIN) + 4e-6)
#define DG_MAX_HALO_CELLS (1 << 27)

enum {
    DG_ST_OK = 0,
    DG_ST_SKIP = 1,  // structure handled by the direct path: the cell-list pipeline ignores it
    DG_ST_SINGULAR = 4,
    DG_ST_IMAGE_RANGE = 5,
    DG_ST_GEOMETRY = 6,
    DG_ST_UNWRAPPED = 7,
};

struct DgGeom {
    double L[9];      // lattice rows a, b, c
    double Linv[9];   // frac_k = x*Linv[0+k] + y*Linv[3+k] + z*Linv[6+k]
    double cvec[9];   // grid-cell edge vectors: L[k] / g[k]
    double cellw[3];  // perpendicular width of a grid cell along lattice axis k
    double ext[3];    // Cartesian AABB width of one grid cell: sum_axis |cvec[axis][k]|
    int g[3];         // grid cells per lattice vector
    int R[3];         // stencil reach (cells)
    int cmin[3], cmax[3];  // occupied range of (unwrapped) source cell coordinates
    int hmin[3];      // halo origin = cmin - R
    int hd[3];        // halo dims = (cmax - cmin + 1) + 2R
    int bd[3];        // source-block dims (cells) of one work item
    int nb[3];        // blocks per axis
    int nimg;         // reference image range n (neighbor_list.cpp:68-72)
    int natoms;
    int status;
    int nitems;
    long long ncells;     // halo cells
    long long cell_base;  // offset into the batch-global cell arrays
    long long atom_base;
    long long item_base;
};

// lattice.row(k).norm() as Eigen evaluates it for a strided 3-vector: x^2 + (y^2 + z^2).
DG_HD double dg_row_norm(const double* r) {
    return sqrt(DG_ADD(DG_MUL(r[0], r[0]), DG_ADD(DG_MUL(r[1], r[1]), DG_MUL(r[2], r[2]))));
}

// neighbor_list.cpp:68-72.  Returns -1 when the result is not representable (degenerate lattice).
DG_HD int dg_num_images(const double* L, double r_cutoff) {
    double l0 = dg_row_norm(L), l1 = dg_row_norm(L + 3), l2 = dg_row_norm(L + 6);
    double lmin = l0 < l1 ? l0 : l1;
    lmin = lmin < l2 ? lmin : l2;
    if (!(lmin > 0.0)) return -1;
    double q = ceil(r_cutoff / lmin);
    if (!(q < 1.0e6)) return -1;
    return (int)q + 1;
}

// neighbor_list.cpp:82-84 -- one Cartesian component of n_a*row0 + n_b*row1 + n_c*row2, left to right.
DG_HD double dg_image_offset(const double* L, int k, int na, int nb, int nc) {
    return DG_ADD(DG_ADD(DG_MUL((double)na, L[k]), DG_MUL((double)nb, L[3 + k])), DG_MUL((double)nc, L[6 + k]));
}

DG_HD int dg_floordiv(int a, int b) {
    int q = a / b;
    if ((a % b != 0) && ((a < 0) != (b < 0))) --q;
    return q;
}
DG_HD int dg_ceildiv(int a, int b) { return -dg_floordiv(-a, b); }

// Geometry that does not depend on the atoms.  cell_frac: target cell width / r_cutoff.
DG_HD void dg_geom_init(DgGeom* G, const double* L, int natoms, double r_cutoff, double cell_frac) {
    for (int i = 0; i < 9; ++i) G->L[i] = L[i];
    G->natoms = natoms;
    G->status = DG_ST_OK;
    G->nitems = 0;
    G->ncells = 0;
    for (int k = 0; k < 3; ++k) {
        G->cmin[k] = INT_MAX;
        G->cmax[k] = INT_MIN;
        G->g[k] = 1;
        G->R[k] = 1;
        G->hmin[k] = 0;
        G->hd[k] = 0;
        G->bd[k] = 1;
        G->nb[k] = 0;
    }
    G->nimg = dg_num_images(L, r_cutoff);
    if (G->nimg < 0) {
        G->status = DG_ST_SINGULAR;
        return;
    }
    if (G->nimg > 127) {
        G->status = DG_ST_IMAGE_RANGE;
        return;
    }
    const double c00 = L[4] * L[8] - L[5] * L[7];
    const double c01 = L[5] * L[6] - L[3] * L[8];
    const double c02 = L[3] * L[7] - L[4] * L[6];
    const double det = L[0] * c00 + L[1] * c01 + L[2] * c02;
    if (det == 0.0 || !isfinite(det)) {
        G->status = DG_ST_SINGULAR;
        return;
    }
    const double id = 1.0 / det;
    G->Linv[0] = c00 * id;
    G->Linv[1] = (L[2] * L[7] - L[1] * L[8]) * id;
    G->Linv[2] = (L[1] * L[5] - L[2] * L[4]) * id;
    G->Linv[3] = c01 * id;
    G->Linv[4] = (L[0] * L[8] - L[2] * L[6]) * id;
    G->Linv[5] = (L[2] * L[3] - L[0] * L[5]) * id;
    G->Linv[6] = c02 * id;
    G->Linv[7] = (L[1] * L[6] - L[0] * L[7]) * id;
    G->Linv[8] = (L[0] * L[4] - L[1] * L[3]) * id;
    const double vol = fabs(det);
    double wt = cell_frac * r_cutoff;
    const double wdens = cbrt(vol / (2.0 * (double)(natoms > 0 ? natoms : 1)));
    if (wdens > wt) wt = wdens;
    const double rcm = DG_RCM(r_cutoff);
    for (int k = 0; k < 3; ++k) {
        const double cn = sqrt(G->Linv[k] * G->Linv[k] + G->Linv[3 + k] * G->Linv[3 + k] +
                               G->Linv[6 + k] * G->Linv[6 + k]);
        if (!(cn > 0.0) || !isfinite(cn)) {
            G->status = DG_ST_SINGULAR;
            return;
        }
        const double h = 1.0 / cn;  // perpendicular height of the unit cell along axis k
        // target reach R_t = ceil(rcm / wt) cells; the widest grid that keeps it: cell width h / g >= rcm / R_t.
        // (Dividing h by the target width instead lets a cell a hair below rcm / R_t push the reach to R_t + 1: 2.4x
        // the stencil volume for a 236 A cell at r_c = 8 with 4 A cells.)
        double rt = ceil(rcm / wt - 0.01);  // (the candidate margin in rcm must not add a cell to the reach)
        if (rt < 1.0) rt = 1.0;
        double gf = floor(h * rt / rcm);
        if (gf < 1.0) gf = 1.0;
        if (gf > 1024.0) gf = 1024.0;
        G->g[k] = (int)gf;
        G->cellw[k] = h / gf;
        const double rr = ceil(rcm / G->cellw[k]);
        if (!(rr <= (double)DG_MAX_REACH)) {
            G->status = DG_ST_GEOMETRY;
            return;
        }
        G->R[k] = (int)rr < 1 ? 1 : (int)rr;
    }
    for (int a = 0; a < 3; ++a)
        for (int k = 0; k < 3; ++k) G->cvec[3 * a + k] = L[3 * a + k] / (double)G->g[a];
    for (int k = 0; k < 3; ++k) G->ext[k] = fabs(G->cvec[k]) + fabs(G->cvec[3 + k]) + fabs(G->cvec[6 + k]);
}

// Unwrapped grid-cell coordinate of a Cartesian point along axis k: floor(frac_k * g_k).
DG_HD int dg_cell_coord(const DgGeom* G, int k, double x, double y, double z) {
    const double f = x * G->Linv[k] + y * G->Linv[3 + k] + z * G->Linv[6 + k];
    double u = floor(f * (double)G->g[k]);
    if (!(u > -1.0e9)) u = -1.0e9;  // also catches NaN
    if (u > 1.0e9) u = 1.0e9;
    return (int)u;
}

// After cmin/cmax are known: halo box, block decomposition, counts.
// tile_target: desired max image points per staged tile; src_target: desired max source atoms per item.
DG_HD void dg_geom_plan(DgGeom* G, int tile_target, int src_target) {
    if (G->status != DG_ST_OK) return;
    if (G->natoms <= 0) {
        G->nitems = 0;
        G->ncells = 0;
        return;
    }
    long long hc = 1;
    int D[3];
    for (int k = 0; k < 3; ++k) {
        const long long d = (long long)G->cmax[k] - (long long)G->cmin[k] + 1;
        if (d < 1 || d > DG_MAX_HOME_DIM) {
            G->status = DG_ST_UNWRAPPED;
            return;
        }
        D[k] = (int)d;
        G->hmin[k] = G->cmin[k] - G->R[k];
        G->hd[k] = D[k] + 2 * G->R[k];
        hc *= G->hd[k];
        if (hc > DG_MAX_HALO_CELLS) {
            G->status = DG_ST_UNWRAPPED;
            return;
        }
    }
    G->ncells = hc;
    const double rho = (double)G->natoms / ((double)D[0] * (double)D[1] * (double)D[2]);
    int bd[3];
    for (int k = 0; k < 3; ++k) {
        const int lim = DG_MAX_TILE_DIM - 2 * G->R[k];
        bd[k] = D[k] < lim ? D[k] : lim;
        if (bd[k] < 1) bd[k] = 1;
    }
    for (int it = 0; it < 64; ++it) {
        const double tpts = rho * (double)(bd[0] + 2 * G->R[0]) * (double)(bd[1] + 2 * G->R[1]) *
                            (double)(bd[2] + 2 * G->R[2]);
        const double spts = rho * (double)bd[0] * (double)bd[1] * (double)bd[2];
        if (tpts <= (double)tile_target && spts <= (double)src_target) break;
        int kmax = 0;
        if (bd[1] > bd[kmax]) kmax = 1;
        if (bd[2] > bd[kmax]) kmax = 2;
        if (bd[kmax] <= 1) break;
        bd[kmax] = (bd[kmax] + 1) / 2;
    }
    long long ni = 1;
    for (int k = 0; k < 3; ++k) {
        G->bd[k] = bd[k];
        G->nb[k] = (D[k] + bd[k] - 1) / bd[k];
        ni *= G->nb[k];
    }
    G->nitems = (int)ni;
}

// Can any point of grid cell (0,0,0) be within rcm of any point of grid cell (da,db,dc)?  Conservative.
DG_HD bool dg_cells_may_touch(const DgGeom* G, int da, int db, int dc, double rcm) {
    const int d[3] = {da, db, dc};
    for (int k = 0; k < 3; ++k) {
        const int ad = d[k] < 0 ? -d[k] : d[k];
        if (ad > 1 && (double)(ad - 1) * G->cellw[k] >= rcm) return false;
    }
    double s = 0.0;
    for (int k = 0; k < 3; ++k) {
        const double Dk = (double)da * G->cvec[k] + (double)db * G->cvec[3 + k] + (double)dc * G->cvec[6 + k];
        const double gap = fabs(Dk) - G->ext[k];
        if (gap > 0.0) s += gap * gap;
    }
    // shave the bound a little: it is computed in floating point
    return s * (1.0 - 1e-9) < rcm * rcm;
}

// Range [zlo, zhi] of dc that may touch for a given (da, db); false if the whole stencil row is empty.
DG_HD bool dg_stencil_row(const DgGeom* G, int da, int db, double rcm, int* zlo, int* zhi) {
    int lo = INT_MAX, hi = INT_MIN;
    for (int dc = -G->R[2]; dc <= G->R[2]; ++dc) {
        if (dg_cells_may_touch(G, da, db, dc, rcm)) {
            if (dc < lo) lo = dc;
            if (dc > hi) hi = dc;
        }
    }
    if (lo > hi) return false;
    *zlo = lo;
    *zhi = hi;
    return true;
}

// Valid reference-image range of an atom in (unwrapped) cell coordinate c along axis k: images n with
// |n| <= nimg whose cell c + n*g lies inside the halo box.  Returns count (0 if none).
DG_HD int dg_image_range(const DgGeom* G, int k, int c, int* nlo) {
    const int hmax = G->hmin[k] + G->hd[k] - 1;
    int lo = dg_ceildiv(G->hmin[k] - c, G->g[k]);
    int hi = dg_floordiv(hmax - c, G->g[k]);
    if (lo < -G->nimg) lo = -G->nimg;
    if (hi > G->nimg) hi = G->nimg;
    *nlo = lo;
    return hi >= lo ? hi - lo + 1 : 0;
}

// image packing of the ABI: int8 n_a | n_b << 8 | n_c << 16
DG_HD uint32_t dg_pack_image(int na, int nb, int nc) {
    return ((uint32_t)(na & 0xff)) | ((uint32_t)(nb & 0xff) << 8) | ((uint32_t)(nc & 0xff) << 16);
}
// monotone tie-break key of a packed image: (n_a, n_b, n_c) lexicographic as signed values
DG_HD uint32_t dg_image_sort_key(uint32_t packed) {
    const uint32_t b = packed ^ 0x00808080u;  // bias each int8 by +128
    return ((b & 0xffu) << 16) | (b & 0xff00u) | ((b >> 16) & 0xffu);
}

// =====================================================================================================================
// Direct (small-structure) path: no cell list.  For a source atom i and a candidate atom j only the periodic images
// n whose fractional offset can bring j within r_c of i are tested:  the projection of the displacement on the unit
// normal of lattice plane k is |f_j[k] + n_k - f_i[k]| * h_k (h_k = perpendicular height of the cell along axis k), and
// it never exceeds the distance, so  |f_j[k] + n_k - f_i[k]| < r_c / h_k  for every true neighbour.  The window is
// evaluated with the candidate radius DG_RCM(r_c) (margin >> rounding of the fractional coordinates) and then clipped
// to the reference cube |n_k| <= nimg (rule R3).  Acceptance itself always uses the reference arithmetic.
// =====================================================================================================================
enum { DG_MODE_CELL = 0, DG_MODE_ATOMPAR = 1, DG_MODE_IMGPAR = 2 };

struct DgSmall {
    double L[9];      // lattice rows a, b, c
    double Linv[9];   // frac_k = x*Linv[0+k] + y*Linv[3+k] + z*Linv[6+k]
    double w[3];      // DG_RCM(r_c) / h_k: half-width of the image window in fractional units
    long long atom_base;
    int natoms;
    int nimg;         // reference image range n (neighbor_list.cpp:68-72)
    int status;       // DG_ST_*
    int mode;         // DG_MODE_*
};

// direct_max_atoms: structures with more atoms go to the cell-list path.
DG_HD void dg_small_init(DgSmall* G, const double* L, long long atom_base, int natoms, double r_cutoff,
                         int direct_max_atoms) {
    for (int i = 0; i < 9; ++i) G->L[i] = L[i];
    for (int i = 0; i < 9; ++i) G->Linv[i] = 0.0;
    G->w[0] = G->w[1] = G->w[2] = 0.0;
    G->atom_base = atom_base;
    G->natoms = natoms;
    G->status = DG_ST_OK;
    G->mode = DG_MODE_ATOMPAR;
    G->nimg = dg_num_images(L, r_cutoff);
    if (G->nimg < 0) {
        G->status = DG_ST_SINGULAR;
        return;
    }
    if (G->nimg > 127) {
        G->status = DG_ST_IMAGE_RANGE;
        return;
    }
    const double c00 = L[4] * L[8] - L[5] * L[7];
    const double c01 = L[5] * L[6] - L[3] * L[8];
    const double c02 = L[3] * L[7] - L[4] * L[6];
    const double det = L[0] * c00 + L[1] * c01 + L[2] * c02;
    if (det == 0.0 || !isfinite(det)) {
        G->status = DG_ST_SINGULAR;
        return;
    }
    const double id = 1.0 / det;
    G->Linv[0] = c00 * id;
    G->Linv[1] = (L[2] * L[7] - L[1] * L[8]) * id;
    G->Linv[2] = (L[1] * L[5] - L[2] * L[4]) * id;
    G->Linv[3] = c01 * id;
    G->Linv[4] = (L[0] * L[8] - L[2] * L[6]) * id;
    G->Linv[5] = (L[2] * L[3] - L[0] * L[5]) * id;
    G->Linv[6] = c02 * id;
    G->Linv[7] = (L[1] * L[6] - L[0] * L[7]) * id;
    G->Linv[8] = (L[0] * L[4] - L[1] * L[3]) * id;
    const double rcm = DG_RCM(r_cutoff);
    double test = 1.0;  // estimated images per candidate pair
    for (int k = 0; k < 3; ++k) {
        const double cn = sqrt(G->Linv[k] * G->Linv[k] + G->Linv[3 + k] * G->Linv[3 + k] +
                               G->Linv[6 + k] * G->Linv[6 + k]);  // = 1 / h_k
        if (!(cn > 0.0) || !isfinite(cn)) {
            G->status = DG_ST_SINGULAR;
            return;
        }
        G->w[k] = rcm * cn;
        double t = 2.0 * G->w[k];
        const double tmax = (double)(2 * G->nimg + 1);
        if (t < 1.0) t = 1.0;
        if (t > tmax) t = tmax;
        test *= t;
    }
    if (natoms > direct_max_atoms) {
        G->mode = DG_MODE_CELL;
        return;
    }
    // lane efficiency of the two direct traversals: lanes over atoms (slot loop over images, ~60 % balance) versus
    // lanes over the images of one pair
    const double n = (double)(natoms > 0 ? natoms : 1);
    const double eff_atom = n / (ceil(n / 32.0) * 32.0) * (test > 1.5 ? 0.6 : 1.0);
    const double eff_img = test / (ceil(test / 32.0) * 32.0);
    G->mode = eff_img > eff_atom ? DG_MODE_IMGPAR : DG_MODE_ATOMPAR;
}

DG_HD double dg_frac_coord(const double* Linv, int k, double x, double y, double z) {
    return x * Linv[k] + y * Linv[3 + k] + z * Linv[6 + k];
}

// Images n of candidate j along one axis: integers in [t - w, t + w] with t = f_i - f_j, clipped to [-nimg, nimg].
// Returns the count (0 if none), *lo = first n.
DG_HD int dg_axis_window(double t, double w, int nimg, int* lo) {
    double a = ceil(t - w), b = floor(t + w);
    const double lim = (double)nimg;
    if (!(a >= -lim)) a = -lim;  // also catches NaN
    if (!(b <= lim)) b = lim;
    if (a > lim + 1.0) a = lim + 1.0;  // keep the int conversions defined for absurd inputs
    if (b < -lim - 1.0) b = -lim - 1.0;
    *lo = (int)a;
    const int c = (int)b - (int)a + 1;
    return c > 0 ? c : 0;
}
