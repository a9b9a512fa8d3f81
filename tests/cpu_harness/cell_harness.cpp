// cell_harness.cpp -- CPU emulation of the product's cell-list pipeline (planning, halo-pruned image cloud binning,
// work-item tiles, stencil-row traversal) built on the SAME geom.h the CUDA kernels include.  It exists so the
// geometry / stencil / halo logic can be checked against the oracle on a machine without a GPU.  It is test code:
// it is not shipped, not linked into libdgnn_b200.so, and it is not the oracle.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../defect-gnn-cpp_b200/csrc/geom.h"

namespace {
struct Pt {
    double x, y, z;
    int j;
    uint32_t n;
};
struct Edge {
    double dist;
    int dst;
    uint32_t n;
};
inline bool edge_less(const Edge& a, const Edge& b) {
    if (a.dist != b.dist) return a.dist < b.dist;
    if (a.dst != b.dst) return a.dst < b.dst;
    return dg_image_sort_key(a.n) < dg_image_sort_key(b.n);
}
}  // namespace

extern "C" {

// Returns 0 on success, DG_ST_* (>0) on geometry errors.  row_ptr has n+1 entries; dst/image/dist hold up to
// edge_cap edges (returns -1 if more are needed; n_edges_out is always the exact count).
int harness_neighbor_list(int64_t n, const double* x, const double* y, const double* z, const double* lattice,
                          double r_cutoff, uint64_t max_neighbors, double epsilon, double cell_frac,
                          int tile_target, int src_target, int64_t edge_cap, int64_t* row_ptr, int32_t* dst,
                          uint32_t* image, double* dist, int64_t* n_edges_out, int64_t* n_near_out,
                          int64_t* info /* [10]: g0 g1 g2 R0 R1 R2 nitems cloud_points candidates stencil_row_visits */) {
    DgGeom G;
    dg_geom_init(&G, lattice, (int)n, r_cutoff, cell_frac);
    G.atom_base = 0;
    G.cell_base = 0;
    G.item_base = 0;
    if (G.status != DG_ST_OK) return G.status;
    std::vector<int> acell(3 * (size_t)n);
    for (int64_t a = 0; a < n; ++a)
        for (int k = 0; k < 3; ++k) {
            const int c = dg_cell_coord(&G, k, x[a], y[a], z[a]);
            acell[3 * a + k] = c;
            G.cmin[k] = std::min(G.cmin[k], c);
            G.cmax[k] = std::max(G.cmax[k], c);
        }
    dg_geom_plan(&G, tile_target, src_target);
    if (G.status != DG_ST_OK) return G.status;
    const int64_t ncells = G.ncells;
    // ---- K1: count -> scan -> scatter ----
    std::vector<int64_t> cellstart((size_t)ncells + 1, 0);
    auto cell_id = [&](int X0, int X1, int X2) {
        return ((int64_t)(X0 - G.hmin[0]) * G.hd[1] + (X1 - G.hmin[1])) * G.hd[2] + (X2 - G.hmin[2]);
    };
    for (int pass = 0; pass < 2; ++pass) {
        std::vector<int64_t> fill;
        std::vector<Pt>* cloudp = nullptr;
        static thread_local std::vector<Pt> cloud;
        if (pass == 1) {
            for (int64_t c = 0, run = 0; c <= ncells; ++c) {
                const int64_t v = c < ncells ? cellstart[c] : 0;
                cellstart[c] = run;
                run += v;
            }
            cloud.assign((size_t)cellstart[ncells], Pt{});
            fill.assign((size_t)ncells, 0);
            cloudp = &cloud;
        }
        for (int64_t a = 0; a < n; ++a) {
            int nlo[3], cnt[3];
            for (int k = 0; k < 3; ++k) cnt[k] = dg_image_range(&G, k, acell[3 * a + k], &nlo[k]);
            for (int ia = 0; ia < cnt[0]; ++ia)
                for (int ib = 0; ib < cnt[1]; ++ib)
                    for (int ic = 0; ic < cnt[2]; ++ic) {
                        const int na = nlo[0] + ia, nb = nlo[1] + ib, nc = nlo[2] + ic;
                        const int64_t cid = cell_id(acell[3 * a] + na * G.g[0], acell[3 * a + 1] + nb * G.g[1],
                                                    acell[3 * a + 2] + nc * G.g[2]);
                        if (pass == 0) {
                            cellstart[cid]++;
                        } else {
                            Pt p;
                            p.x = x[a] + dg_image_offset(G.L, 0, na, nb, nc);
                            p.y = y[a] + dg_image_offset(G.L, 1, na, nb, nc);
                            p.z = z[a] + dg_image_offset(G.L, 2, na, nb, nc);
                            p.j = (int)a;
                            p.n = dg_pack_image(na, nb, nc);
                            (*cloudp)[(size_t)(cellstart[cid] + fill[cid]++)] = p;
                        }
                    }
        }
        if (pass == 1) {
            // ---- K2: items -> sources -> stencil rows ----
            const double r2 = r_cutoff * r_cutoff;
            const double rcm = DG_RCM(r_cutoff);
            std::vector<std::vector<Edge>> rows((size_t)n);
            int64_t near = 0, ncand = 0, nrowvis = 0;
            const int NBr = 2 * G.R[1] + 1, nst = (2 * G.R[0] + 1) * NBr;
            std::vector<int> zlo(nst), zhi(nst);
            for (int q = 0; q < nst; ++q) {
                int lo = 1, hi = 0;
                if (!dg_stencil_row(&G, q / NBr - G.R[0], q % NBr - G.R[1], rcm, &lo, &hi)) {
                    lo = 1;
                    hi = 0;
                }
                zlo[q] = lo;
                zhi[q] = hi;
            }
            for (int item = 0; item < G.nitems; ++item) {
                const int b2 = item % G.nb[2], b1 = (item / G.nb[2]) % G.nb[1], b0 = item / (G.nb[2] * G.nb[1]);
                const int bb[3] = {b0, b1, b2};
                int c0[3], c1[3], t0[3], td[3];
                for (int k = 0; k < 3; ++k) {
                    const int D = G.hd[k] - 2 * G.R[k];
                    c0[k] = G.R[k] + bb[k] * G.bd[k];
                    c1[k] = std::min(c0[k] + G.bd[k], G.R[k] + D);
                    t0[k] = c0[k] - G.R[k];
                    td[k] = (c1[k] - c0[k]) + 2 * G.R[k];
                    if (td[k] > DG_MAX_TILE_DIM) return 100;
                }
                for (int a0 = c0[0]; a0 < c1[0]; ++a0)
                    for (int a1 = c0[1]; a1 < c1[1]; ++a1)
                        for (int a2 = c0[2]; a2 < c1[2]; ++a2) {
                            const int ta = a0 - t0[0], tb = a1 - t0[1], tc = a2 - t0[2];
                            const int64_t gc = ((int64_t)a0 * G.hd[1] + a1) * G.hd[2] + a2;
                            for (int64_t ps = cellstart[gc]; ps < cellstart[gc + 1]; ++ps) {
                                const Pt& S = cloud[(size_t)ps];
                                if (S.n != 0u) continue;
                                std::vector<Edge>& row = rows[(size_t)S.j];
                                for (int q = 0; q < nst; ++q) {
                                    if (zlo[q] > zhi[q]) continue;
                                    const int ra = ta + (q / NBr - G.R[0]), rb = tb + (q % NBr - G.R[1]);
                                    if (ra < 0 || ra >= td[0] || rb < 0 || rb >= td[1]) return 101;
                                    if (tc + zlo[q] < 0 || tc + zhi[q] >= td[2]) return 102;
                                    ++nrowvis;
                                    const int64_t g0 = ((int64_t)(t0[0] + ra) * G.hd[1] + (t0[1] + rb)) * G.hd[2] +
                                                       (t0[2] + tc + zlo[q]);
                                    for (int64_t p = cellstart[g0]; p < cellstart[g0 + (zhi[q] - zlo[q]) + 1]; ++p) {
                                        const Pt& C = cloud[(size_t)p];
                                        ++ncand;
                                        const double dx = S.x - C.x, dy = S.y - C.y, dz = S.z - C.z;
                                        const double d2 = ((dx * dx) + (dy * dy)) + (dz * dz);
                                        if (std::fabs(std::sqrt(d2) - r_cutoff) <= 1e-6) ++near;
                                        if (!(d2 < r2)) continue;
                                        const double d = std::sqrt(d2);
                                        if (C.j == S.j && d < epsilon) continue;
                                        row.push_back(Edge{d, C.j, C.n});
                                    }
                                }
                            }
                        }
            }
            int64_t e = 0;
            bool fits = true;
            for (int64_t i = 0; i < n; ++i) {
                std::vector<Edge>& row = rows[(size_t)i];
                std::sort(row.begin(), row.end(), edge_less);
                if (row.size() > max_neighbors) row.resize((size_t)max_neighbors);
                row_ptr[i] = e;
                for (const Edge& ed : row) {
                    if (e < edge_cap) {
                        dst[e] = ed.dst;
                        image[e] = ed.n;
                        dist[e] = ed.dist;
                    } else {
                        fits = false;
                    }
                    ++e;
                }
            }
            row_ptr[n] = e;
            *n_edges_out = e;
            *n_near_out = near;
            if (info) {
                info[0] = G.g[0]; info[1] = G.g[1]; info[2] = G.g[2];
                info[3] = G.R[0]; info[4] = G.R[1]; info[5] = G.R[2];
                info[6] = G.nitems; info[7] = cellstart[ncells]; info[8] = ncand; info[9] = nrowvis;
            }
            return fits ? 0 : -1;
        }
    }
    return 0;
}


// CPU emulation of the direct (small-structure) traversal of dgnn_direct.cu: per source atom, per candidate atom, the
// fractional-coordinate image window (dg_small_init / dg_axis_window from geom.h), reference arithmetic for the accept
// test.  mode_out receives the traversal the classification picks; info[0] = candidates tested.
int harness_direct_neighbor_list(int64_t n, const double* x, const double* y, const double* z, const double* lattice,
                                 double r_cutoff, uint64_t max_neighbors, double epsilon, int64_t edge_cap,
                                 int64_t* row_ptr, int32_t* dst, uint32_t* image, double* dist, int64_t* n_edges_out,
                                 int64_t* n_near_out, int64_t* info /* [4]: candidates mode nimg - */) {
    DgSmall G;
    dg_small_init(&G, lattice, 0, (int)n, r_cutoff, 1 << 30);
    if (G.status != DG_ST_OK) return G.status;
    std::vector<double> fa((size_t)n), fb((size_t)n), fc((size_t)n);
    for (int64_t a = 0; a < n; ++a) {
        fa[a] = dg_frac_coord(G.Linv, 0, x[a], y[a], z[a]);
        fb[a] = dg_frac_coord(G.Linv, 1, x[a], y[a], z[a]);
        fc[a] = dg_frac_coord(G.Linv, 2, x[a], y[a], z[a]);
    }
    const double r2 = r_cutoff * r_cutoff;
    int64_t e = 0, near = 0, ncand = 0;
    bool fits = true;
    std::vector<Edge> row;
    for (int64_t i = 0; i < n; ++i) {
        row.clear();
        for (int64_t j = 0; j < n; ++j) {
            int lo0, lo1, lo2;
            const int c0 = dg_axis_window(fa[i] - fa[j], G.w[0], G.nimg, &lo0);
            const int c1 = dg_axis_window(fb[i] - fb[j], G.w[1], G.nimg, &lo1);
            const int c2 = dg_axis_window(fc[i] - fc[j], G.w[2], G.nimg, &lo2);
            for (int ia = 0; ia < c0; ++ia)
                for (int ib = 0; ib < c1; ++ib)
                    for (int ic = 0; ic < c2; ++ic) {
                        const int na = lo0 + ia, nb = lo1 + ib, nc = lo2 + ic;
                        ++ncand;
                        const double px = x[j] + dg_image_offset(G.L, 0, na, nb, nc);
                        const double py = y[j] + dg_image_offset(G.L, 1, na, nb, nc);
                        const double pz = z[j] + dg_image_offset(G.L, 2, na, nb, nc);
                        const double dx = x[i] - px, dy = y[i] - py, dz = z[i] - pz;
                        const double d2 = ((dx * dx) + (dy * dy)) + (dz * dz);
                        if (std::fabs(std::sqrt(d2) - r_cutoff) <= 1e-6) ++near;
                        if (!(d2 < r2)) continue;
                        const double d = std::sqrt(d2);
                        if (j == i && d < epsilon) continue;
                        row.push_back(Edge{d, (int)j, dg_pack_image(na, nb, nc)});
                    }
        }
        std::sort(row.begin(), row.end(), edge_less);
        if (row.size() > max_neighbors) row.resize((size_t)max_neighbors);
        row_ptr[i] = e;
        for (const Edge& ed : row) {
            if (e < edge_cap) {
                dst[e] = ed.dst;
                image[e] = ed.n;
                dist[e] = ed.dist;
            } else {
                fits = false;
            }
            ++e;
        }
    }
    row_ptr[n] = e;
    *n_edges_out = e;
    *n_near_out = near;
    if (info) {
        info[0] = ncand;
        info[1] = G.mode;
        info[2] = G.nimg;
        info[3] = 0;
    }
    return fits ? 0 : -1;
}

}  // extern "C"
