// dgnn_oracle.cpp -- CPU ORACLE (test infrastructure only; see dgnn_oracle.h for scope and parity status).
//
// Restates, without Eigen / nanoflann, the reference algorithm of
//   /root/reference/src/graph/neighbor_list.cpp:27-94, src/graph/edge_features.cpp:7-24,
//   src/graph/crystal_graph.cpp:23-40, src/crystal/structure.cpp:17.
// Third-party arithmetic that is NOT under /root/reference (un-vendored downloads, Makefile:157-186) is restated
// from the published upstream sources as recalled: nanoflann v1.5.5 (L2_Simple_Adaptor::evalMetric,
// RadiusResultSet::addPoint, KDTreeBaseClass::divideTree/middleSplit_/planeSplit, searchLevel) and Eigen 3.4.0
// (coefficient-wise expression evaluation order, redux unrolling order for 3-vectors).
//
// Build with -ffp-contract=off: the reference is compiled -O3 without -march (CMakeLists.txt:33-37), i.e. plain
// IEEE fp64, one rounding per operation, no FMA.

#include "dgnn_oracle.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

#include <atomic>
#include <thread>

namespace {

struct Vec3 {
    double v[3];
};

// ---- Eigen-order helpers ---------------------------------------------------------------------------------
// lattice.row(k).norm(): sqrt of cwiseAbs2().sum() on a strided 3-vector -> Eigen's scalar redux unroller splits
// [0,3) as [0,1) + [1,3), i.e. x^2 + (y^2 + z^2).
inline double row_norm(const double* r) {
    return std::sqrt(r[0] * r[0] + (r[1] * r[1] + r[2] * r[2]));
}

// neighbor_list.cpp:68-72
inline int num_images(const double* L, double r_cutoff) {
    double l_min = std::min({row_norm(L + 0), row_norm(L + 3), row_norm(L + 6)});
    return static_cast<int>(std::ceil(r_cutoff / l_min)) + 1;
}

// neighbor_list.cpp:82-84: (n_a*row0 + n_b*row1 + n_c*row2), coefficient-wise, left to right.
inline Vec3 image_offset(const double* L, int na, int nb, int nc) {
    Vec3 o;
    for (int k = 0; k < 3; ++k) {
        o.v[k] = (static_cast<double>(na) * L[0 + k] + static_cast<double>(nb) * L[3 + k]) +
                 static_cast<double>(nc) * L[6 + k];
    }
    return o;
}

// nanoflann L2_Simple_Adaptor::evalMetric: result = 0; for d: diff = a[d] - pt[d]; result += diff*diff.
inline double eval_metric(const double* q, const double* p) {
    double result = 0.0;
    for (int d = 0; d < 3; ++d) {
        const double diff = q[d] - p[d];
        result += diff * diff;
    }
    return result;
}

struct Cloud {  // neighbor_list.hpp:27-55 PointCloud + the image each point came from
    std::vector<Vec3> pos;
    std::vector<int32_t> orig;
    std::vector<int8_t> img;  // 3 per point
};

// neighbor_list.cpp:74-94 (a outermost, c innermost, then atoms in order)
Cloud make_cloud(int64_t n, const double* x, const double* y, const double* z, const double* L, int nimg) {
    Cloud c;
    for (int na = -nimg; na <= nimg; ++na)
        for (int nb = -nimg; nb <= nimg; ++nb)
            for (int nc = -nimg; nc <= nimg; ++nc) {
                Vec3 off = image_offset(L, na, nb, nc);
                for (int64_t i = 0; i < n; ++i) {
                    Vec3 p;
                    p.v[0] = x[i] + off.v[0];
                    p.v[1] = y[i] + off.v[1];
                    p.v[2] = z[i] + off.v[2];
                    c.pos.push_back(p);
                    c.orig.push_back(static_cast<int32_t>(i));
                    c.img.push_back(static_cast<int8_t>(na));
                    c.img.push_back(static_cast<int8_t>(nb));
                    c.img.push_back(static_cast<int8_t>(nc));
                }
            }
    return c;
}

struct Edge {
    double dist;
    int32_t dst;
    int8_t n[3];
    double disp[3];
};

inline bool canonical_less(const Edge& a, const Edge& b) {
    if (a.dist != b.dist) return a.dist < b.dist;
    if (a.dst != b.dst) return a.dst < b.dst;
    if (a.n[0] != b.n[0]) return a.n[0] < b.n[0];
    if (a.n[1] != b.n[1]) return a.n[1] < b.n[1];
    return a.n[2] < b.n[2];
}

inline bool near_cutoff(double d2, double r_cutoff) {
    return std::fabs(std::sqrt(d2) - r_cutoff) <= 1e-6;
}

// neighbor_list.cpp:44-54: one radius-search match -> Neighbor (or the excluded zero-distance self).
inline bool make_edge(int64_t i, const double* q, const Cloud& c, size_t m, double d2, double epsilon, Edge* e) {
    const int32_t orig = c.orig[m];
    const double dist = std::sqrt(d2);
    if (orig == i && dist < epsilon) return false;
    e->dist = dist;
    e->dst = orig;
    for (int k = 0; k < 3; ++k) {
        e->n[k] = c.img[3 * m + k];
        e->disp[k] = c.pos[m].v[k] - q[k];
    }
    return true;
}

// ---- KD-tree restating nanoflann v1.5.5 KDTreeSingleIndexAdaptor (leaf_max_size 10, neighbor_list.cpp:31-33) ----
struct KDTree {
    struct Interval {
        double low, high;
    };
    struct Node {
        // leaf: left,right index range into vAcc; inner: divfeat, divlow, divhigh
        int32_t child1 = -1, child2 = -1;
        int32_t left = 0, right = 0;
        int divfeat = -1;
        double divlow = 0, divhigh = 0;
    };
    const Cloud& cloud;
    std::vector<uint32_t> vAcc;
    std::vector<Node> nodes;
    Interval root_bbox[3];
    size_t leaf_max;

    explicit KDTree(const Cloud& c, size_t leaf) : cloud(c), leaf_max(leaf) {}

    inline double pt(uint32_t idx, int d) const { return cloud.pos[idx].v[d]; }

    void build() {
        const size_t n = cloud.pos.size();
        vAcc.resize(n);
        for (size_t i = 0; i < n; ++i) vAcc[i] = static_cast<uint32_t>(i);
        nodes.clear();
        if (n == 0) return;
        // computeBoundingBox (kdtree_get_bbox returns false -> scan, neighbor_list.hpp:35-38)
        for (int d = 0; d < 3; ++d) root_bbox[d].low = root_bbox[d].high = pt(0, d);
        for (size_t k = 1; k < n; ++k)
            for (int d = 0; d < 3; ++d) {
                const double v = pt(static_cast<uint32_t>(k), d);
                if (v < root_bbox[d].low) root_bbox[d].low = v;
                if (v > root_bbox[d].high) root_bbox[d].high = v;
            }
        nodes.reserve(2 * n / leaf_max + 16);
        Interval bbox[3] = {root_bbox[0], root_bbox[1], root_bbox[2]};
        divide(0, n, bbox);
    }

    int32_t divide(size_t left, size_t right, Interval* bbox) {
        const int32_t id = static_cast<int32_t>(nodes.size());
        nodes.emplace_back();
        if ((right - left) <= leaf_max) {
            nodes[id].left = static_cast<int32_t>(left);
            nodes[id].right = static_cast<int32_t>(right);
            for (int d = 0; d < 3; ++d) bbox[d].low = bbox[d].high = pt(vAcc[left], d);
            for (size_t k = left + 1; k < right; ++k)
                for (int d = 0; d < 3; ++d) {
                    const double v = pt(vAcc[k], d);
                    if (bbox[d].low > v) bbox[d].low = v;
                    if (bbox[d].high < v) bbox[d].high = v;
                }
            return id;
        }
        size_t idx;
        int cutfeat;
        double cutval;
        middle_split(left, right - left, idx, cutfeat, cutval, bbox);
        Interval lb[3] = {bbox[0], bbox[1], bbox[2]};
        lb[cutfeat].high = cutval;
        const int32_t c1 = divide(left, left + idx, lb);
        Interval rb[3] = {bbox[0], bbox[1], bbox[2]};
        rb[cutfeat].low = cutval;
        const int32_t c2 = divide(left + idx, right, rb);
        nodes[id].child1 = c1;
        nodes[id].child2 = c2;
        nodes[id].divfeat = cutfeat;
        nodes[id].divlow = lb[cutfeat].high;
        nodes[id].divhigh = rb[cutfeat].low;
        for (int d = 0; d < 3; ++d) {
            bbox[d].low = std::min(lb[d].low, rb[d].low);
            bbox[d].high = std::max(lb[d].high, rb[d].high);
        }
        return id;
    }

    void min_max(size_t ind, size_t count, int d, double& mn, double& mx) const {
        mn = mx = pt(vAcc[ind], d);
        for (size_t k = 1; k < count; ++k) {
            const double v = pt(vAcc[ind + k], d);
            if (v < mn) mn = v;
            if (v > mx) mx = v;
        }
    }

    void middle_split(size_t ind, size_t count, size_t& index, int& cutfeat, double& cutval,
                      const Interval* bbox) {
        const double EPS = 0.00001;
        double max_span = bbox[0].high - bbox[0].low;
        for (int d = 1; d < 3; ++d) max_span = std::max(max_span, bbox[d].high - bbox[d].low);
        double max_spread = -1;
        cutfeat = 0;
        double min_elem = 0, max_elem = 0;
        for (int d = 0; d < 3; ++d) {
            const double span = bbox[d].high - bbox[d].low;
            if (span > (1 - EPS) * max_span) {
                double mn, mx;
                min_max(ind, count, d, mn, mx);
                const double spread = mx - mn;
                if (spread > max_spread) {
                    cutfeat = d;
                    max_spread = spread;
                    min_elem = mn;
                    max_elem = mx;
                }
            }
        }
        const double split_val = (bbox[cutfeat].low + bbox[cutfeat].high) / 2;
        if (split_val < min_elem)
            cutval = min_elem;
        else if (split_val > max_elem)
            cutval = max_elem;
        else
            cutval = split_val;
        size_t lim1, lim2;
        plane_split(ind, count, cutfeat, cutval, lim1, lim2);
        if (lim1 > count / 2)
            index = lim1;
        else if (lim2 < count / 2)
            index = lim2;
        else
            index = count / 2;
    }

    void plane_split(size_t ind, size_t count, int cutfeat, double cutval, size_t& lim1, size_t& lim2) {
        size_t left = 0, right = count - 1;
        for (;;) {
            while (left <= right && pt(vAcc[ind + left], cutfeat) < cutval) ++left;
            while (right && left <= right && pt(vAcc[ind + right], cutfeat) >= cutval) --right;
            if (left > right || !right) break;
            std::swap(vAcc[ind + left], vAcc[ind + right]);
            ++left;
            --right;
        }
        lim1 = left;
        right = count - 1;
        for (;;) {
            while (left <= right && pt(vAcc[ind + left], cutfeat) <= cutval) ++left;
            while (right && left <= right && pt(vAcc[ind + right], cutfeat) > cutval) --right;
            if (left > right || !right) break;
            std::swap(vAcc[ind + left], vAcc[ind + right]);
            ++left;
            --right;
        }
        lim2 = left;
    }

    // radiusSearch(query, radius(=r^2), matches), default SearchParameters: eps = 0, sorted = true.
    void radius_search(const double* q, double radius, std::vector<std::pair<uint32_t, double>>& out) const {
        out.clear();
        if (nodes.empty()) return;
        double dists[3] = {0, 0, 0};
        double distsq = 0;
        for (int d = 0; d < 3; ++d) {  // computeInitialDistances
            if (q[d] < root_bbox[d].low) {
                dists[d] = (q[d] - root_bbox[d].low) * (q[d] - root_bbox[d].low);
                distsq += dists[d];
            }
            if (q[d] > root_bbox[d].high) {
                dists[d] = (q[d] - root_bbox[d].high) * (q[d] - root_bbox[d].high);
                distsq += dists[d];
            }
        }
        search_level(q, 0, distsq, dists, radius, out);
        std::sort(out.begin(), out.end(),
                  [](const std::pair<uint32_t, double>& a, const std::pair<uint32_t, double>& b) {
                      return a.second < b.second;
                  });
    }

    void search_level(const double* q, int32_t id, double mindist, double* dists, double radius,
                      std::vector<std::pair<uint32_t, double>>& out) const {
        const Node& nd = nodes[id];
        if (nd.child1 < 0) {
            for (int32_t k = nd.left; k < nd.right; ++k) {
                const uint32_t a = vAcc[k];
                const double dist = eval_metric(q, cloud.pos[a].v);
                if (dist < radius) {  // searchLevel: dist < worst_dist; RadiusResultSet::addPoint: dist < radius
                    out.emplace_back(a, dist);
                }
            }
            return;
        }
        const int idx = nd.divfeat;
        const double val = q[idx];
        const double diff1 = val - nd.divlow;
        const double diff2 = val - nd.divhigh;
        int32_t best, other;
        double cut_dist;
        if ((diff1 + diff2) < 0) {
            best = nd.child1;
            other = nd.child2;
            cut_dist = (val - nd.divhigh) * (val - nd.divhigh);
        } else {
            best = nd.child2;
            other = nd.child1;
            cut_dist = (val - nd.divlow) * (val - nd.divlow);
        }
        search_level(q, best, mindist, dists, radius, out);
        const double dst = dists[idx];
        mindist = mindist + cut_dist - dst;
        dists[idx] = cut_dist;
        if (mindist * 1.0 <= radius) {  // epsError = 1 + eps, eps = 0
            search_level(q, other, mindist, dists, radius, out);
        }
        dists[idx] = dst;
    }
};

// ---- per-structure neighbour search, three algorithms --------------------------------------------------------

struct RowSink {
    std::vector<std::vector<Edge>> rows;
    int64_t n_near = 0;
};

void search_brute(int64_t n, const double* x, const double* y, const double* z, const double* L, double r_cutoff,
                  double epsilon, int nimg, RowSink& sink) {
    const Cloud cloud = make_cloud(n, x, y, z, L, nimg);
    const double r2 = r_cutoff * r_cutoff;
    const size_t M = cloud.pos.size();
    for (int64_t i = 0; i < n; ++i) {
        const double q[3] = {x[i], y[i], z[i]};
        std::vector<Edge>& row = sink.rows[i];
        for (size_t m = 0; m < M; ++m) {
            const double d2 = eval_metric(q, cloud.pos[m].v);
            if (near_cutoff(d2, r_cutoff)) ++sink.n_near;
            if (d2 < r2) {
                Edge e;
                if (make_edge(i, q, cloud, m, d2, epsilon, &e)) row.push_back(e);
            }
        }
    }
}

void search_kdtree(int64_t n, const double* x, const double* y, const double* z, const double* L, double r_cutoff,
                   double epsilon, int nimg, RowSink& sink) {
    const Cloud cloud = make_cloud(n, x, y, z, L, nimg);
    KDTree tree(cloud, 10);
    tree.build();
    const double r2 = r_cutoff * r_cutoff;
    std::vector<std::pair<uint32_t, double>> matches;
    for (int64_t i = 0; i < n; ++i) {
        const double q[3] = {x[i], y[i], z[i]};
        tree.radius_search(q, r2, matches);
        std::vector<Edge>& row = sink.rows[i];
        for (const auto& m : matches) {
            if (near_cutoff(m.second, r_cutoff)) ++sink.n_near;  // inside side only: outside is never visited
            Edge e;
            if (make_edge(i, q, cloud, m.first, m.second, epsilon, &e)) row.push_back(e);
        }
    }
}

// 3x3 inverse (row-major), returns false if singular.
bool invert3(const double* m, double* inv) {
    const double c00 = m[4] * m[8] - m[5] * m[7];
    const double c01 = m[5] * m[6] - m[3] * m[8];
    const double c02 = m[3] * m[7] - m[4] * m[6];
    const double det = m[0] * c00 + m[1] * c01 + m[2] * c02;
    if (det == 0.0 || !std::isfinite(det)) return false;
    const double id = 1.0 / det;
    inv[0] = c00 * id;
    inv[1] = (m[2] * m[7] - m[1] * m[8]) * id;
    inv[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    inv[3] = c01 * id;
    inv[4] = (m[0] * m[8] - m[2] * m[6]) * id;
    inv[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    inv[6] = c02 * id;
    inv[7] = (m[1] * m[6] - m[0] * m[7]) * id;
    inv[8] = (m[0] * m[4] - m[1] * m[3]) * id;
    return true;
}

inline int64_t floordiv(int64_t a, int64_t b) {
    int64_t q = a / b;
    if ((a % b != 0) && ((a < 0) != (b < 0))) --q;
    return q;
}

// Cell list over atoms wrapped into the home cell.  Candidate generation only: every accepted pair is tested with
// the reference arithmetic on the stored (unwrapped) positions and the reference image cube |n_k| <= nimg is
// enforced on the reported image, so the result set equals the exhaustive scan.
void search_celllist(int64_t n, const double* x, const double* y, const double* z, const double* L,
                     double r_cutoff, double epsilon, int nimg, RowSink& sink) {
    double inv[9];
    if (!invert3(L, inv)) {  // degenerate lattice: fall back to the exhaustive scan
        search_brute(n, x, y, z, L, r_cutoff, epsilon, nimg, sink);
        return;
    }
    // frac_k = pos . col_k(inv);  perpendicular height h_k = 1 / |col_k(inv)|
    double h[3];
    for (int k = 0; k < 3; ++k) h[k] = 1.0 / std::sqrt(inv[k] * inv[k] + inv[3 + k] * inv[3 + k] + inv[6 + k] * inv[6 + k]);
    int64_t g[3], R[3];
    for (int k = 0; k < 3; ++k) {
        g[k] = std::max<int64_t>(1, static_cast<int64_t>(std::floor(h[k] / r_cutoff)));
        g[k] = std::min<int64_t>(g[k], 128);
        const double w = h[k] / static_cast<double>(g[k]);
        R[k] = static_cast<int64_t>(std::floor(r_cutoff / w + 1e-9)) + 1;
    }
    std::vector<int64_t> cell(n * 3), wrap(n * 3);
    for (int64_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) {
            const double f = x[i] * inv[k] + y[i] * inv[3 + k] + z[i] * inv[6 + k];
            const double w = std::floor(f);
            double fw = f - w;  // [0,1)
            int64_t c = static_cast<int64_t>(std::floor(fw * static_cast<double>(g[k])));
            if (c >= g[k]) c = g[k] - 1;
            if (c < 0) c = 0;
            cell[3 * i + k] = c;
            wrap[3 * i + k] = static_cast<int64_t>(w);
        }
    }
    const int64_t ncell = g[0] * g[1] * g[2];
    std::vector<std::vector<int32_t>> bins(ncell);
    for (int64_t i = 0; i < n; ++i)
        bins[(cell[3 * i] * g[1] + cell[3 * i + 1]) * g[2] + cell[3 * i + 2]].push_back(static_cast<int32_t>(i));

    const double r2 = r_cutoff * r_cutoff;
    for (int64_t i = 0; i < n; ++i) {
        const double q[3] = {x[i], y[i], z[i]};
        std::vector<Edge>& row = sink.rows[i];
        for (int64_t da = -R[0]; da <= R[0]; ++da)
            for (int64_t db = -R[1]; db <= R[1]; ++db)
                for (int64_t dc = -R[2]; dc <= R[2]; ++dc) {
                    const int64_t G[3] = {cell[3 * i] + da, cell[3 * i + 1] + db, cell[3 * i + 2] + dc};
                    int64_t m[3], c[3];
                    for (int k = 0; k < 3; ++k) {
                        m[k] = floordiv(G[k], g[k]);
                        c[k] = G[k] - m[k] * g[k];
                    }
                    const std::vector<int32_t>& bin = bins[(c[0] * g[1] + c[1]) * g[2] + c[2]];
                    for (int32_t j : bin) {
                        // image index relative to the stored (unwrapped) positions
                        int64_t nn[3];
                        bool in_cube = true;
                        for (int k = 0; k < 3; ++k) {
                            nn[k] = m[k] - wrap[3 * j + k] + wrap[3 * i + k];
                            if (nn[k] < -nimg || nn[k] > nimg) in_cube = false;
                        }
                        if (!in_cube) continue;
                        const Vec3 off = image_offset(L, static_cast<int>(nn[0]), static_cast<int>(nn[1]),
                                                      static_cast<int>(nn[2]));
                        const double p[3] = {x[j] + off.v[0], y[j] + off.v[1], z[j] + off.v[2]};
                        const double d2 = eval_metric(q, p);
                        if (near_cutoff(d2, r_cutoff)) ++sink.n_near;
                        if (d2 < r2) {
                            const double dist = std::sqrt(d2);
                            if (j == i && dist < epsilon) continue;
                            Edge e;
                            e.dist = dist;
                            e.dst = j;
                            for (int k = 0; k < 3; ++k) {
                                e.n[k] = static_cast<int8_t>(nn[k]);
                                e.disp[k] = p[k] - q[k];
                            }
                            row.push_back(e);
                        }
                    }
                }
    }
}

int run_search(int64_t n, const double* x, const double* y, const double* z, const double* L, double r_cutoff,
               double epsilon, int algo, int nimg, RowSink& sink) {
    sink.rows.assign(static_cast<size_t>(n), {});
    sink.n_near = 0;
    if (n == 0) return 0;
    switch (algo) {
        case ORC_ALGO_BRUTE: search_brute(n, x, y, z, L, r_cutoff, epsilon, nimg, sink); break;
        case ORC_ALGO_KDTREE: search_kdtree(n, x, y, z, L, r_cutoff, epsilon, nimg, sink); break;
        case ORC_ALGO_CELLLIST: search_celllist(n, x, y, z, L, r_cutoff, epsilon, nimg, sink); break;
        default: return -1;
    }
    return 0;
}

// Dynamic-schedule parallel for over [0,n) on `threads` std::threads (no OpenMP dependency).
template <class F>
void parallel_for(int64_t n, int threads, F&& body) {
    if (threads <= 1 || n <= 1) {
        for (int64_t i = 0; i < n; ++i) body(i, 0);
        return;
    }
    std::atomic<int64_t> next{0};
    std::vector<std::thread> pool;
    pool.reserve(static_cast<size_t>(threads));
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            for (;;) {
                const int64_t i = next.fetch_add(1);
                if (i >= n) break;
                body(i, t);
            }
        });
    for (auto& th : pool) th.join();
}

inline int resolve_threads(int threads) {
    if (threads > 0) return threads;
    const unsigned hc = std::thread::hardware_concurrency();
    return hc ? static_cast<int>(hc) : 1;
}

template <class T>
T* dup(const std::vector<T>& v) {
    T* p = static_cast<T*>(std::malloc(std::max<size_t>(1, v.size()) * sizeof(T)));
    if (!v.empty()) std::memcpy(p, v.data(), v.size() * sizeof(T));
    return p;
}

}  // namespace

extern "C" {

int orc_num_images(const double lattice[9], double r_cutoff) {
    return num_images(lattice, r_cutoff);
}

int orc_neighbor_list(int64_t n, const double* x, const double* y, const double* z, const double lattice[9],
                      double r_cutoff, uint64_t max_neighbors, double epsilon, int algo, orc_graph* out) {
    if (!out || n < 0) return -1;
    std::memset(out, 0, sizeof(*out));
    const int nimg = num_images(lattice, r_cutoff);
    if (nimg > 127) return -2;
    RowSink sink;
    if (run_search(n, x, y, z, lattice, r_cutoff, epsilon, algo, nimg, sink) != 0) return -1;

    std::vector<int64_t> row_ptr(static_cast<size_t>(n) + 1, 0);
    std::vector<int32_t> dst;
    std::vector<int8_t> image;
    std::vector<double> dist, disp;
    int64_t tie_rows = 0;
    for (int64_t i = 0; i < n; ++i) {
        std::vector<Edge>& row = sink.rows[i];
        std::sort(row.begin(), row.end(), canonical_less);  // neighbor_list.cpp:56-58 + canonical tie-break
        if (row.size() > max_neighbors) {                   // :60-62
            if (row[max_neighbors - 1].dist == row[max_neighbors].dist) ++tie_rows;
            row.resize(max_neighbors);
        }
        for (const Edge& e : row) {
            dst.push_back(e.dst);
            image.push_back(e.n[0]);
            image.push_back(e.n[1]);
            image.push_back(e.n[2]);
            image.push_back(0);
            dist.push_back(e.dist);
            disp.push_back(e.disp[0]);
            disp.push_back(e.disp[1]);
            disp.push_back(e.disp[2]);
        }
        row_ptr[i + 1] = static_cast<int64_t>(dst.size());
    }
    out->n_atoms = n;
    out->n_edges = static_cast<int64_t>(dst.size());
    out->row_ptr = dup(row_ptr);
    out->dst = dup(dst);
    out->image = dup(image);
    out->dist = dup(dist);
    out->disp = dup(disp);
    out->n_near_cutoff = sink.n_near;
    out->n_tie_rows = tie_rows;
    out->num_images = nimg;
    return 0;
}

void orc_free_graph(orc_graph* g) {
    if (!g) return;
    std::free(g->row_ptr);
    std::free(g->dst);
    std::free(g->image);
    std::free(g->dist);
    std::free(g->disp);
    std::memset(g, 0, sizeof(*g));
}

// edge_features.cpp:7-24
int orc_gaussian_rbf(double distance, double r_cutoff, double dr, double* out) {
    const int n = static_cast<int>(std::floor(r_cutoff / dr));
    if (!out) return n;
    const double sigma = r_cutoff / 3;
    const double inv_sigma_squared = 1 / std::pow(sigma, 2);
    const double norm = 1 / (sigma * std::sqrt(2 * M_PI));
    for (int k = 0; k < n; ++k) {
        const double center = k * dr;
        out[k] = norm * std::exp(-0.5 * std::pow(center - distance, 2) * inv_sigma_squared);
    }
    return n;
}

int orc_edge_attr_colmajor(int64_t n_edges, const double* dist, double r_cutoff, double dr, double* out) {
    const int K = orc_gaussian_rbf(0.0, r_cutoff, dr, nullptr);
    std::vector<double> g(static_cast<size_t>(std::max(K, 1)));
    for (int64_t e = 0; e < n_edges; ++e) {
        orc_gaussian_rbf(dist[e], r_cutoff, dr, g.data());
        for (int k = 0; k < K; ++k) out[static_cast<int64_t>(k) * n_edges + e] = g[k];  // crystal_graph.cpp:37
    }
    return K;
}

int orc_edge_attr_rowmajor(int64_t n_edges, const double* dist, double r_cutoff, double dr, double* out) {
    const int K = orc_gaussian_rbf(0.0, r_cutoff, dr, nullptr);
    for (int64_t e = 0; e < n_edges; ++e) orc_gaussian_rbf(dist[e], r_cutoff, dr, out + e * K);
    return K;
}

void orc_frac_to_cart(int64_t n, const double* frac, const double lattice[9], double* x, double* y, double* z) {
    double* outs[3] = {x, y, z};
    for (int64_t i = 0; i < n; ++i) {
        const double* f = frac + 3 * i;
        for (int k = 0; k < 3; ++k)
            outs[k][i] = (lattice[0 + k] * f[0] + lattice[3 + k] * f[1]) + lattice[6 + k] * f[2];
    }
}

// One structure the way the reference does it end to end (for timing): NeighborList then CrystalGraph edge part.
static void reference_style_structure(int64_t n, const double* x, const double* y, const double* z,
                                      const double* L, double r_cutoff, uint64_t max_neighbors, double epsilon,
                                      double rbf_r_cutoff, double dr, int algo, int64_t* n_edges,
                                      double* checksum) {
    const int nimg = num_images(L, r_cutoff);
    RowSink sink;
    run_search(n, x, y, z, L, r_cutoff, epsilon, algo, nimg, sink);
    std::vector<std::vector<Edge>> lists(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) {
        std::vector<Edge>& row = sink.rows[i];
        std::sort(row.begin(), row.end(), [](const Edge& a, const Edge& b) { return a.dist < b.dist; });
        if (row.size() > max_neighbors) row.resize(max_neighbors);
        lists[i] = row;  // neighbor_list.cpp:64 is a copy
    }
    // crystal_graph.cpp:23-40
    int edge_count = 0;
    for (int64_t i = 0; i < n; ++i) edge_count += static_cast<int>(lists[i].size());
    const int K = dr > 0 ? static_cast<int>(std::floor(rbf_r_cutoff / dr)) : 0;
    std::vector<int> edge_index(static_cast<size_t>(2) * edge_count);
    std::vector<double> edge_attr(static_cast<size_t>(edge_count) * std::max(K, 0));
    int edge = 0;
    double cs = 0;
    for (int64_t i = 0; i < n; ++i) {
        for (const Edge& nb : lists[i]) {
            edge_index[2 * edge] = static_cast<int>(i);
            edge_index[2 * edge + 1] = nb.dst;
            cs += nb.dist;
            if (K > 0) {
                std::vector<double> g(static_cast<size_t>(K));  // one heap VectorXd per edge (edge_features.cpp:15)
                orc_gaussian_rbf(nb.dist, rbf_r_cutoff, dr, g.data());
                for (int k = 0; k < K; ++k) edge_attr[static_cast<size_t>(k) * edge_count + edge] = g[k];
                cs += g[K / 2];
            }
            ++edge;
        }
    }
    *n_edges = edge_count;
    *checksum = cs + (edge_count ? edge_index[2 * (edge_count - 1) + 1] * 1e-9 : 0.0);
}

int orc_build_batch(int64_t n_structs, const int64_t* atom_offsets, const double* x, const double* y,
                    const double* z, const double* lattices, double r_cutoff, uint64_t max_neighbors,
                    double epsilon, double rbf_r_cutoff, double dr, int algo, int threads,
                    orc_batch_result* out) {
    if (!out) return -1;
    const int used = resolve_threads(threads);
    std::vector<int64_t> tot(static_cast<size_t>(used), 0);
    std::vector<double> css(static_cast<size_t>(used), 0.0);
    const auto t0 = std::chrono::steady_clock::now();
    parallel_for(n_structs, used, [&](int64_t s, int t) {
        const int64_t a0 = atom_offsets[s], a1 = atom_offsets[s + 1];
        int64_t ne = 0;
        double c = 0;
        reference_style_structure(a1 - a0, x + a0, y + a0, z + a0, lattices + 9 * s, r_cutoff, max_neighbors,
                                  epsilon, rbf_r_cutoff, dr, algo, &ne, &c);
        tot[static_cast<size_t>(t)] += ne;
        css[static_cast<size_t>(t)] += c;
    });
    const auto t1 = std::chrono::steady_clock::now();
    int64_t total = 0;
    double cs = 0;
    for (int t = 0; t < used; ++t) {
        total += tot[static_cast<size_t>(t)];
        cs += css[static_cast<size_t>(t)];
    }
    out->n_edges = total;
    out->checksum = cs;
    out->seconds = std::chrono::duration<double>(t1 - t0).count();
    out->threads = used;
    return 0;
}

int orc_batch_edge_counts(int64_t n_structs, const int64_t* atom_offsets, const double* x, const double* y,
                          const double* z, const double* lattices, double r_cutoff, uint64_t max_neighbors,
                          double epsilon, int algo, int threads, int64_t* counts) {
    parallel_for(n_structs, resolve_threads(threads), [&](int64_t s, int) {
        const int64_t a0 = atom_offsets[s], a1 = atom_offsets[s + 1];
        const double* L = lattices + 9 * s;
        RowSink sink;
        run_search(a1 - a0, x + a0, y + a0, z + a0, L, r_cutoff, epsilon, algo, num_images(L, r_cutoff), sink);
        int64_t e = 0;
        for (const auto& row : sink.rows) e += static_cast<int64_t>(std::min<uint64_t>(row.size(), max_neighbors));
        counts[s] = e;
    });
    return 0;
}

int orc_rows_brute(int64_t n, const double* x, const double* y, const double* z, const double lattice[9],
                   double r_cutoff, double epsilon, int64_t n_sources, const int64_t* sources, int64_t row_cap,
                   int64_t* deg, int32_t* dst, int8_t* image, double* dist) {
    const int nimg = num_images(lattice, r_cutoff);
    if (nimg > 127) return -2;
    const double r2 = r_cutoff * r_cutoff;
    const int W = 2 * nimg + 1;
    std::vector<Vec3> offs(static_cast<size_t>(W) * W * W);
    for (int na = -nimg; na <= nimg; ++na)
        for (int nb = -nimg; nb <= nimg; ++nb)
            for (int nc = -nimg; nc <= nimg; ++nc)
                offs[((na + nimg) * W + (nb + nimg)) * W + (nc + nimg)] = image_offset(lattice, na, nb, nc);
    parallel_for(n_sources, resolve_threads(0), [&](int64_t qi, int) {
        const int64_t i = sources[qi];
        const double q[3] = {x[i], y[i], z[i]};
        std::vector<Edge> row;
        for (int na = -nimg; na <= nimg; ++na)
            for (int nb = -nimg; nb <= nimg; ++nb)
                for (int nc = -nimg; nc <= nimg; ++nc) {
                    const Vec3& off = offs[((na + nimg) * W + (nb + nimg)) * W + (nc + nimg)];
                    for (int64_t j = 0; j < n; ++j) {
                        const double p[3] = {x[j] + off.v[0], y[j] + off.v[1], z[j] + off.v[2]};
                        const double d2 = eval_metric(q, p);
                        if (d2 < r2) {
                            const double d = std::sqrt(d2);
                            if (j == i && d < epsilon) continue;
                            Edge e;
                            e.dist = d;
                            e.dst = static_cast<int32_t>(j);
                            e.n[0] = static_cast<int8_t>(na);
                            e.n[1] = static_cast<int8_t>(nb);
                            e.n[2] = static_cast<int8_t>(nc);
                            e.disp[0] = e.disp[1] = e.disp[2] = 0;
                            row.push_back(e);
                        }
                    }
                }
        std::sort(row.begin(), row.end(), canonical_less);
        deg[qi] = static_cast<int64_t>(row.size());
        if (dst && image && dist) {
            const int64_t m = std::min<int64_t>(row_cap, static_cast<int64_t>(row.size()));
            for (int64_t r = 0; r < m; ++r) {
                dst[qi * row_cap + r] = row[r].dst;
                image[(qi * row_cap + r) * 4 + 0] = row[r].n[0];
                image[(qi * row_cap + r) * 4 + 1] = row[r].n[1];
                image[(qi * row_cap + r) * 4 + 2] = row[r].n[2];
                image[(qi * row_cap + r) * 4 + 3] = 0;
                dist[qi * row_cap + r] = row[r].dist;
            }
        }
    });
    return 0;
}

}  // extern "C"
