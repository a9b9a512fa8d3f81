// oracle/ref_shim/nanoflann.hpp -- TEST INFRASTRUCTURE ONLY (never on the product's include path).
//
// Interface shim for the five nanoflann names the reference touches (include/graph/neighbor_list.hpp:57-58,
// src/graph/neighbor_list.cpp:31-40): KDTreeSingleIndexAdaptor, L2_Simple_Adaptor, KDTreeSingleIndexAdaptorParams,
// ResultItem, radiusSearch.  nanoflann v1.5.5 is an un-vendored download of the reference (Makefile:177-186) and is
// absent from this image; its published algorithm is restated here as recalled (KDTreeBaseClass::divideTree /
// middleSplit_ / planeSplit / computeMinMax, KDTreeSingleIndexAdaptor::buildIndex / computeBoundingBox /
// computeInitialDistances / searchLevel / radiusSearch, RadiusResultSet, IndexDist_Sorter, L2_Simple_Adaptor), with
// its arithmetic kept operation for operation: evalMetric accumulates result += diff*diff over the dimensions from
// 0.0, the radius predicate is the strict `dist < radius` of RadiusResultSet::addPoint, a sub-tree is visited iff
// mindist * (1 + eps) <= radius (eps = 0 by default), and matches come back sorted by distance (SearchParameters
// default sorted = true).  With it the reference's own neighbor_list.cpp compiles verbatim (oracle/Makefile `ref`).
#pragma once

#include <algorithm>
#include <array>
#include <cstddef>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <utility>
#include <vector>

namespace nanoflann {

template <typename IndexType = size_t, typename DistanceType = double>
struct ResultItem {
    ResultItem() = default;
    ResultItem(const IndexType index, const DistanceType distance) : first(index), second(distance) {}
    IndexType first;
    DistanceType second;
};

struct IndexDist_Sorter {
    template <typename PairType>
    bool operator()(const PairType& p1, const PairType& p2) const {
        return p1.second < p2.second;
    }
};

template <typename DistanceType_, typename IndexType_ = size_t>
class RadiusResultSet {
public:
    using DistanceType = DistanceType_;
    using IndexType = IndexType_;
    const DistanceType radius;
    std::vector<ResultItem<IndexType, DistanceType>>& m_indices_dists;

    RadiusResultSet(DistanceType radius_, std::vector<ResultItem<IndexType, DistanceType>>& indices_dists)
        : radius(radius_), m_indices_dists(indices_dists) {
        init();
    }
    void init() { clear(); }
    void clear() { m_indices_dists.clear(); }
    size_t size() const { return m_indices_dists.size(); }
    bool empty() const { return m_indices_dists.empty(); }
    bool full() const { return true; }
    bool addPoint(DistanceType dist, IndexType index) {
        if (dist < radius) m_indices_dists.emplace_back(index, dist);
        return true;
    }
    DistanceType worstDist() const { return radius; }
};

template <class T, class DataSource, typename DistanceType_ = T, typename IndexType = uint32_t>
struct L2_Simple_Adaptor {
    using ElementType = T;
    using DistanceType = DistanceType_;
    const DataSource& data_source;
    explicit L2_Simple_Adaptor(const DataSource& data_source_) : data_source(data_source_) {}
    DistanceType evalMetric(const T* a, const IndexType b_idx, size_t size) const {
        DistanceType result = DistanceType();
        for (size_t i = 0; i < size; ++i) {
            const DistanceType diff = a[i] - data_source.kdtree_get_pt(b_idx, i);
            result += diff * diff;
        }
        return result;
    }
    template <typename U, typename V>
    DistanceType accum_dist(const U a, const V b, const size_t) const {
        return (a - b) * (a - b);
    }
};

enum class KDTreeSingleIndexAdaptorFlags { None = 0, SkipInitialBuildIndex = 1 };

struct KDTreeSingleIndexAdaptorParams {
    explicit KDTreeSingleIndexAdaptorParams(size_t leaf_max_size_ = 10,
                                            KDTreeSingleIndexAdaptorFlags flags_ = KDTreeSingleIndexAdaptorFlags::None,
                                            unsigned int n_thread_build_ = 1)
        : leaf_max_size(leaf_max_size_), flags(flags_), n_thread_build(n_thread_build_) {}
    size_t leaf_max_size;
    KDTreeSingleIndexAdaptorFlags flags;
    unsigned int n_thread_build;
};

struct SearchParameters {
    SearchParameters(float eps_ = 0, bool sorted_ = true) : eps(eps_), sorted(sorted_) {}
    float eps;
    bool sorted;
};

template <typename Distance, class DatasetAdaptor, int32_t DIM = -1, typename IndexType = uint32_t>
class KDTreeSingleIndexAdaptor {
public:
    using ElementType = typename Distance::ElementType;
    using DistanceType = typename Distance::DistanceType;
    using Offset = size_t;
    using Size = size_t;
    using Dimension = int32_t;

    struct Interval {
        ElementType low, high;
    };
    using BoundingBox = std::array<Interval, (DIM > 0 ? DIM : 3)>;
    using distance_vector_t = std::array<DistanceType, (DIM > 0 ? DIM : 3)>;

    struct Node {
        union {
            struct leaf {
                Offset left, right;
            } lr;
            struct nonleaf {
                Dimension divfeat;
                DistanceType divlow, divhigh;
            } sub;
        } node_type;
        Node *child1 = nullptr, *child2 = nullptr;
    };
    using NodePtr = Node*;

    KDTreeSingleIndexAdaptor(const KDTreeSingleIndexAdaptor&) = delete;

    explicit KDTreeSingleIndexAdaptor(const Dimension dimensionality, const DatasetAdaptor& inputData,
                                      const KDTreeSingleIndexAdaptorParams& params = KDTreeSingleIndexAdaptorParams())
        : dataset_(inputData), index_params_(params), distance_(inputData) {
        dim_ = (DIM > 0) ? DIM : dimensionality;
        leaf_max_size_ = params.leaf_max_size;
        if (!(static_cast<int>(params.flags) & static_cast<int>(KDTreeSingleIndexAdaptorFlags::SkipInitialBuildIndex))) {
            buildIndex();  // v1.5.x builds in the constructor unless told otherwise; the reference builds again
        }
    }

    void buildIndex() {
        size_ = dataset_.kdtree_get_point_count();
        size_at_index_build_ = size_;
        vAcc_.resize(size_);
        for (Size i = 0; i < size_; i++) vAcc_[i] = static_cast<IndexType>(i);
        pool_.clear();
        root_node_ = nullptr;
        if (size_ == 0) return;
        computeBoundingBox(root_bbox_);
        root_node_ = divideTree(0, size_, root_bbox_);
    }

    template <typename RESULTSET>
    bool findNeighbors(RESULTSET& result, const ElementType* vec, const SearchParameters& searchParams = {}) const {
        if (size_ == 0) return false;
        if (!root_node_) throw std::runtime_error("[nanoflann] findNeighbors() called before building the index.");
        float epsError = 1 + searchParams.eps;
        distance_vector_t dists;
        dists.fill(0);
        DistanceType dist = computeInitialDistances(vec, dists);
        searchLevel(result, vec, root_node_, dist, dists, epsError);
        return result.full();
    }

    Size radiusSearch(const ElementType* query_point, const DistanceType& radius,
                      std::vector<ResultItem<IndexType, DistanceType>>& IndicesDists,
                      const SearchParameters& searchParams = {}) const {
        RadiusResultSet<DistanceType, IndexType> resultSet(radius, IndicesDists);
        findNeighbors(resultSet, query_point, searchParams);
        const Size nFound = resultSet.size();
        if (searchParams.sorted) std::sort(IndicesDists.begin(), IndicesDists.end(), IndexDist_Sorter());
        return nFound;
    }

private:
    ElementType dataset_get(IndexType element, Dimension component) const {
        return dataset_.kdtree_get_pt(element, component);
    }

    void computeBoundingBox(BoundingBox& bbox) {
        if (dataset_.kdtree_get_bbox(bbox)) return;
        const Size N = dataset_.kdtree_get_point_count();
        if (!N) throw std::runtime_error("[nanoflann] computeBoundingBox() called but no data points found.");
        for (Dimension i = 0; i < dim_; ++i) bbox[i].low = bbox[i].high = dataset_get(vAcc_[0], i);
        for (Offset k = 1; k < N; ++k) {
            for (Dimension i = 0; i < dim_; ++i) {
                const auto val = dataset_get(vAcc_[k], i);
                if (val < bbox[i].low) bbox[i].low = val;
                if (val > bbox[i].high) bbox[i].high = val;
            }
        }
    }

    NodePtr divideTree(const Offset left, const Offset right, BoundingBox& bbox) {
        pool_.emplace_back(new Node());
        NodePtr node = pool_.back().get();
        if ((right - left) <= static_cast<Offset>(leaf_max_size_)) {
            node->child1 = node->child2 = nullptr;
            node->node_type.lr.left = left;
            node->node_type.lr.right = right;
            for (Dimension i = 0; i < dim_; ++i) {
                bbox[i].low = dataset_get(vAcc_[left], i);
                bbox[i].high = dataset_get(vAcc_[left], i);
            }
            for (Offset k = left + 1; k < right; ++k) {
                for (Dimension i = 0; i < dim_; ++i) {
                    const auto val = dataset_get(vAcc_[k], i);
                    if (bbox[i].low > val) bbox[i].low = val;
                    if (bbox[i].high < val) bbox[i].high = val;
                }
            }
        } else {
            Offset idx;
            Dimension cutfeat;
            DistanceType cutval;
            middleSplit_(left, right - left, idx, cutfeat, cutval, bbox);
            node->node_type.sub.divfeat = cutfeat;
            BoundingBox left_bbox(bbox);
            left_bbox[cutfeat].high = cutval;
            node->child1 = divideTree(left, left + idx, left_bbox);
            BoundingBox right_bbox(bbox);
            right_bbox[cutfeat].low = cutval;
            node->child2 = divideTree(left + idx, right, right_bbox);
            node->node_type.sub.divlow = left_bbox[cutfeat].high;
            node->node_type.sub.divhigh = right_bbox[cutfeat].low;
            for (Dimension i = 0; i < dim_; ++i) {
                bbox[i].low = std::min(left_bbox[i].low, right_bbox[i].low);
                bbox[i].high = std::max(left_bbox[i].high, right_bbox[i].high);
            }
        }
        return node;
    }

    void computeMinMax(const Offset ind, const Size count, const Dimension element, ElementType& min_elem,
                       ElementType& max_elem) const {
        min_elem = dataset_get(vAcc_[ind], element);
        max_elem = min_elem;
        for (Offset i = 1; i < count; ++i) {
            ElementType val = dataset_get(vAcc_[ind + i], element);
            if (val < min_elem) min_elem = val;
            if (val > max_elem) max_elem = val;
        }
    }

    void middleSplit_(const Offset ind, const Size count, Offset& index, Dimension& cutfeat, DistanceType& cutval,
                      const BoundingBox& bbox) {
        const auto EPS = static_cast<DistanceType>(0.00001);
        ElementType max_span = bbox[0].high - bbox[0].low;
        for (Dimension i = 1; i < dim_; ++i) {
            ElementType span = bbox[i].high - bbox[i].low;
            if (span > max_span) max_span = span;
        }
        ElementType max_spread = -1;
        cutfeat = 0;
        ElementType min_elem = 0, max_elem = 0;
        for (Dimension i = 0; i < dim_; ++i) {
            ElementType span = bbox[i].high - bbox[i].low;
            if (span > (1 - EPS) * max_span) {
                ElementType min_elem_, max_elem_;
                computeMinMax(ind, count, i, min_elem_, max_elem_);
                ElementType spread = max_elem_ - min_elem_;
                if (spread > max_spread) {
                    cutfeat = i;
                    max_spread = spread;
                    min_elem = min_elem_;
                    max_elem = max_elem_;
                }
            }
        }
        DistanceType split_val = (bbox[cutfeat].low + bbox[cutfeat].high) / 2;
        if (split_val < min_elem)
            cutval = min_elem;
        else if (split_val > max_elem)
            cutval = max_elem;
        else
            cutval = split_val;
        Offset lim1, lim2;
        planeSplit(ind, count, cutfeat, cutval, lim1, lim2);
        if (lim1 > count / 2)
            index = lim1;
        else if (lim2 < count / 2)
            index = lim2;
        else
            index = count / 2;
    }

    void planeSplit(const Offset ind, const Size count, const Dimension cutfeat, const DistanceType& cutval,
                    Offset& lim1, Offset& lim2) {
        Offset left = 0;
        Offset right = count - 1;
        for (;;) {
            while (left <= right && dataset_get(vAcc_[ind + left], cutfeat) < cutval) ++left;
            while (right && left <= right && dataset_get(vAcc_[ind + right], cutfeat) >= cutval) --right;
            if (left > right || !right) break;
            std::swap(vAcc_[ind + left], vAcc_[ind + right]);
            ++left;
            --right;
        }
        lim1 = left;
        right = count - 1;
        for (;;) {
            while (left <= right && dataset_get(vAcc_[ind + left], cutfeat) <= cutval) ++left;
            while (right && left <= right && dataset_get(vAcc_[ind + right], cutfeat) > cutval) --right;
            if (left > right || !right) break;
            std::swap(vAcc_[ind + left], vAcc_[ind + right]);
            ++left;
            --right;
        }
        lim2 = left;
    }

    DistanceType computeInitialDistances(const ElementType* vec, distance_vector_t& dists) const {
        DistanceType dist = DistanceType();
        for (Dimension i = 0; i < dim_; ++i) {
            if (vec[i] < root_bbox_[i].low) {
                dists[i] = distance_.accum_dist(vec[i], root_bbox_[i].low, i);
                dist += dists[i];
            }
            if (vec[i] > root_bbox_[i].high) {
                dists[i] = distance_.accum_dist(vec[i], root_bbox_[i].high, i);
                dist += dists[i];
            }
        }
        return dist;
    }

    template <class RESULTSET>
    bool searchLevel(RESULTSET& result_set, const ElementType* vec, const NodePtr node, DistanceType mindist,
                     distance_vector_t& dists, const float epsError) const {
        if ((node->child1 == nullptr) && (node->child2 == nullptr)) {
            DistanceType worst_dist = result_set.worstDist();
            for (Offset i = node->node_type.lr.left; i < node->node_type.lr.right; ++i) {
                const IndexType accessor = vAcc_[i];
                DistanceType dist = distance_.evalMetric(vec, accessor, dim_);
                if (dist < worst_dist) {
                    if (!result_set.addPoint(dist, vAcc_[i])) return false;
                }
            }
            return true;
        }
        Dimension idx = node->node_type.sub.divfeat;
        ElementType val = vec[idx];
        DistanceType diff1 = val - node->node_type.sub.divlow;
        DistanceType diff2 = val - node->node_type.sub.divhigh;
        NodePtr bestChild;
        NodePtr otherChild;
        DistanceType cut_dist;
        if ((diff1 + diff2) < 0) {
            bestChild = node->child1;
            otherChild = node->child2;
            cut_dist = distance_.accum_dist(val, node->node_type.sub.divhigh, idx);
        } else {
            bestChild = node->child2;
            otherChild = node->child1;
            cut_dist = distance_.accum_dist(val, node->node_type.sub.divlow, idx);
        }
        if (!searchLevel(result_set, vec, bestChild, mindist, dists, epsError)) return false;
        DistanceType dst = dists[idx];
        mindist = mindist + cut_dist - dst;
        dists[idx] = cut_dist;
        if (mindist * epsError <= result_set.worstDist()) {
            if (!searchLevel(result_set, vec, otherChild, mindist, dists, epsError)) return false;
        }
        dists[idx] = dst;
        return true;
    }

    const DatasetAdaptor& dataset_;
    KDTreeSingleIndexAdaptorParams index_params_;
    Distance distance_;
    std::vector<IndexType> vAcc_;
    std::vector<std::unique_ptr<Node>> pool_;
    NodePtr root_node_ = nullptr;
    Size leaf_max_size_ = 10;
    Size size_ = 0, size_at_index_build_ = 0;
    Dimension dim_ = 3;
    BoundingBox root_bbox_;
};

}  // namespace nanoflann
