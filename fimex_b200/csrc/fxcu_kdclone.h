// fxcu_kdclone.h -- host side of K9 (coord_kdtree): a kd-tree whose shape, leaf order and search order equal those of the
// nanoflann 1.1.7 index the reference builds in flannTranslatePointsToClosestInputCell (src/CDMInterpolator.cc:953-1028).
//
// Why the shape has to be equal and not merely "a kd-tree": the reference takes ret_matches.at(0) after std::sort on the
// distance alone (include/nanoflann/nanoflann.hpp:966-974, sorter :179-186), so among exactly equidistant source points
// (duplicated wrap columns, poles) the winner is decided by the order in which the tree walk meets them; and source points
// with NaN coordinates stay in the cloud (CDMInterpolator.cc:982-986), where every comparison against them is false, which
// bends the partitioning and makes the walk miss points that lie inside the radius.  Both effects are functions of
//   * the build:   nanoflann.hpp:1020-1039 (bounding box), :1053-1101 (node recursion), :1103-1112 (min/max),
//                  :1159-1198 (split selection, incl. the `cutfeat` slip at :1175), :1210-1237 (two-pass partition)
//   * the search:  :1239-1256 (initial distances), :1263-1310 (near child first, far child iff mindistsq <= r^2)
// restated here over flat arrays (the points are kept permuted in leaf order next to the index permutation, so every pass of
// the build is a linear sweep and the device reads a leaf as one contiguous run).  Every floating-point comparison is written
// in the sense the reference writes it (a < b is not !(a >= b) once NaN is around).
//
// Host-only C++: no CUDA in here.  fxcu_kdtree.cu uploads the result and walks it on the GPU; the host walk below is used for
// the few targets whose result depends on libstdc++'s introsort (more than 16 matches AND a tie at the minimum).
#pragma once

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <future>
#include <limits>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "fxcu_hostpar.h"

namespace fxcu {
namespace kdclone {

constexpr uint32_t LEAF_MAX = 12;    // KDTreeSingleIndexAdaptorParams(12), CDMInterpolator.cc:993
constexpr int MAX_DEPTH = 4096;      // far beyond any tree a finite cloud gives; a degenerate NaN cloud that makes the reference recurse without end stops here
constexpr uint32_t TASK_REF = 0x80000000u;

struct Node {            // 32 bytes, shared with the device
    double divlow, divhigh; // interior: tight upper bound of the low child / tight lower bound of the high child on `feat`
    uint32_t a, b;          // interior: child ids (low side, high side); leaf: [a, b) in leaf order
    int32_t feat;           // split dimension, -1 for a leaf
    int32_t pad_;
};

struct Box {
    double lo[3], hi[3];
};

struct Tree {
    std::vector<Node> nodes;   // nodes[root]
    uint32_t root = 0;
    std::vector<uint32_t> vind; // leaf order -> source index
    std::vector<double> pts;    // 3 doubles per point, in leaf order
    Box rootBox{};              // after the build (the recursion tightens it in place, :1093-1096)
    std::string error;          // non-empty: the reference's behaviour is undefined for this cloud
};

namespace detail {

struct Ctx {
    std::vector<Node> nodes;
};

struct Builder {
    uint32_t n;
    uint32_t* vind;
    double* p; // permuted with vind
    std::vector<Ctx*> tasks;
    std::mutex taskMutex;
    std::string error;
    std::mutex errMutex;
    int parDepth;
    std::atomic<bool> sawEmptyRange{false}; // an empty child reads its neighbour's first entry: only the sequential order defines what it sees

    inline double at(size_t i, int d) const { return p[3 * i + d]; }
    inline void exchange(size_t i, size_t j)
    {
        std::swap(vind[i], vind[j]);
        std::swap(p[3 * i], p[3 * j]);
        std::swap(p[3 * i + 1], p[3 * j + 1]);
        std::swap(p[3 * i + 2], p[3 * j + 2]);
    }
    void fail(const char* what)
    {
        std::lock_guard<std::mutex> g(errMutex);
        if (error.empty())
            error = what;
    }

    // :1103-1112 for all three dimensions in one sweep (each dimension's result is what the reference's separate sweep gives).
    // The sequential rule "start from entry 0, replace on v < min / v > max" means: NaN if entry 0 is NaN (no comparison ever
    // succeeds), else the extreme of the non-NaN entries -- which is what the chunked form computes for large ranges.
    void minmax3(size_t first, size_t count, double mn[3], double mx[3]) const
    {
        if (count >= (size_t(1) << 19) && parDepth > 0) {
            const size_t parts = 16, step = (count + parts - 1) / parts;
            double pmn[16][3], pmx[16][3];
            parallel_ranges(parts, 1, [&](size_t a, size_t b) {
                for (size_t c = a; c < b; ++c) {
                    const size_t lo = first + c * step, hi = std::min(first + count, lo + step);
                    for (int d = 0; d < 3; ++d) {
                        pmn[c][d] = std::numeric_limits<double>::infinity();
                        pmx[c][d] = -std::numeric_limits<double>::infinity();
                    }
                    for (size_t i = lo; i < hi; ++i)
                        for (int d = 0; d < 3; ++d) {
                            const double v = at(i, d);
                            if (v < pmn[c][d])
                                pmn[c][d] = v;
                            if (v > pmx[c][d])
                                pmx[c][d] = v;
                        }
                }
            });
            for (int d = 0; d < 3; ++d) {
                mn[d] = mx[d] = at(first, d);
                for (size_t c = 0; c < parts; ++c) {
                    if (pmn[c][d] < mn[d])
                        mn[d] = pmn[c][d];
                    if (pmx[c][d] > mx[d])
                        mx[d] = pmx[c][d];
                }
            }
            return;
        }
        for (int d = 0; d < 3; ++d)
            mn[d] = mx[d] = at(first, d);
        for (size_t i = first + 1; i < first + count; ++i)
            for (int d = 0; d < 3; ++d) {
                const double v = at(i, d);
                if (v < mn[d])
                    mn[d] = v;
                if (v > mx[d])
                    mx[d] = v;
            }
    }

    // :1210-1237: values below the cut to the front, then values equal to it; exchanges pair the next misplaced entry from
    // the left with the next one from the right, which fixes the order inside both parts
    void partition(size_t first, size_t count, int feat, double cut, size_t& lim1, size_t& lim2)
    {
        size_t l = 0, r = count - 1;
        for (;;) {
            while (l <= r && at(first + l, feat) < cut)
                ++l;
            while (r && l <= r && at(first + r, feat) >= cut)
                --r;
            if (l > r || !r)
                break;
            exchange(first + l, first + r);
            ++l;
            --r;
        }
        lim1 = l;
        r = count - 1;
        for (;;) {
            while (l <= r && at(first + l, feat) <= cut)
                ++l;
            while (r && l <= r && at(first + r, feat) > cut)
                --r;
            if (l > r || !r)
                break;
            exchange(first + l, first + r);
            ++l;
            --r;
        }
        lim2 = l;
    }

    // returns the node id inside ctx (or TASK_REF|task for a subtree built by another thread)
    uint32_t divide(Ctx& ctx, uint32_t left, uint32_t right, Box& box, int depth)
    {
        if (depth > MAX_DEPTH) {
            fail("coord_kdtree: the source cloud makes the reference's kd-tree build recurse without end (NaN coordinates); its result is undefined");
            return 0;
        }
        const uint32_t id = static_cast<uint32_t>(ctx.nodes.size());
        ctx.nodes.push_back(Node{0., 0., left, right, -1, 0});
        const uint32_t count = right - left;
        if (count <= LEAF_MAX) {
            // :1059-1073.  An empty range still reads the entry at `left`, as the reference does.
            if (count == 0)
                sawEmptyRange = true;
            if (left >= n) {
                fail("coord_kdtree: the reference's kd-tree build reads past its index array for this cloud (NaN coordinates); its result is undefined");
                return id;
            }
            for (int d = 0; d < 3; ++d)
                box.lo[d] = box.hi[d] = at(left, d);
            for (uint32_t k = left + 1; k < right; ++k)
                for (int d = 0; d < 3; ++d) {
                    const double v = at(k, d);
                    if (box.lo[d] > v)
                        box.lo[d] = v;
                    if (box.hi[d] < v)
                        box.hi[d] = v;
                }
            return id;
        }
        // ---- split selection, :1159-1198 ----
        const double EPS = 0.00001;
        double maxSpan = box.hi[0] - box.lo[0];
        for (int d = 1; d < 3; ++d) {
            const double span = box.hi[d] - box.lo[d];
            if (span > maxSpan)
                maxSpan = span;
        }
        double mn[3], mx[3];
        minmax3(left, count, mn, mx);
        double maxSpread = -1;
        int feat = 0;
        for (int d = 0; d < 3; ++d) {
            const double span = box.hi[d] - box.lo[d];
            if (span > (1 - EPS) * maxSpan) {
                const double spread = mx[feat] - mn[feat]; // the spread of the CURRENT candidate, not of d (:1175)
                if (spread > maxSpread) {
                    feat = d;
                    maxSpread = spread;
                }
            }
        }
        const double mid = (box.lo[feat] + box.hi[feat]) / 2;
        double cut;
        if (mid < mn[feat])
            cut = mn[feat];
        else if (mid > mx[feat])
            cut = mx[feat];
        else
            cut = mid;
        size_t lim1, lim2;
        partition(left, count, feat, cut, lim1, lim2);
        const size_t half = count / 2;
        const uint32_t idx = static_cast<uint32_t>(lim1 > half ? lim1 : (lim2 < half ? lim2 : half));

        // ---- children, :1081-1096 ----
        Box lowBox = box, highBox = box;
        lowBox.hi[feat] = cut;
        highBox.lo[feat] = cut;
        uint32_t c1, c2;
        if (depth < parDepth && count > 65536) {
            Ctx* sub = new Ctx;
            uint32_t task;
            {
                std::lock_guard<std::mutex> g(taskMutex);
                task = static_cast<uint32_t>(tasks.size());
                tasks.push_back(sub);
            }
            auto fut = std::async(std::launch::async, [&]() { return divide(*sub, left, left + idx, lowBox, depth + 1); });
            c2 = divide(ctx, left + idx, right, highBox, depth + 1);
            fut.get(); // root of the sub-context is its node 0
            c1 = TASK_REF | task;
        } else {
            c1 = divide(ctx, left, left + idx, lowBox, depth + 1);
            c2 = divide(ctx, left + idx, right, highBox, depth + 1);
        }
        Node& nd = ctx.nodes[id];
        nd.feat = feat;
        nd.a = c1;
        nd.b = c2;
        nd.divlow = lowBox.hi[feat];
        nd.divhigh = highBox.lo[feat];
        for (int d = 0; d < 3; ++d) {
            // std::min(a, b) is (b < a) ? b : a; std::max(a, b) is (a < b) ? b : a
            box.lo[d] = (highBox.lo[d] < lowBox.lo[d]) ? highBox.lo[d] : lowBox.lo[d];
            box.hi[d] = (lowBox.hi[d] < highBox.hi[d]) ? highBox.hi[d] : lowBox.hi[d];
        }
        return id;
    }
};

} // namespace detail

// Core: t.pts holds the cloud in source order (3 doubles per point, NaN triples for points without coordinates), t.vind has n
// entries.  Returns false when the build met an empty range while running threaded: the caller must restore t.pts to source order
// and call again with threads = 1 (only the sequential order defines what such a range reads).
inline bool build_in_place(Tree& t, unsigned threads)
{
    const size_t n = t.vind.size();
    t.error.clear();
    t.nodes.clear();
    if (n == 0) {
        t.error = "coord_kdtree: empty source cloud";
        return true;
    }
    parallel_ranges(n, 1 << 18, [&](size_t a, size_t b) {
        for (size_t i = a; i < b; ++i)
            t.vind[i] = static_cast<uint32_t>(i);
    });
    detail::Builder b;
    b.n = static_cast<uint32_t>(n);
    b.vind = t.vind.data();
    b.p = t.pts.data();
    b.parDepth = 0;
    for (unsigned c = 1; c < threads; c *= 2)
        ++b.parDepth;
    // :1020-1039 (the same rule as minmax3)
    b.minmax3(0, n, t.rootBox.lo, t.rootBox.hi);
    detail::Ctx main;
    b.divide(main, 0, static_cast<uint32_t>(n), t.rootBox, 0);
    if (b.sawEmptyRange && b.parDepth > 0) {
        for (detail::Ctx* c : b.tasks)
            delete c;
        return false; // degenerate (NaN-bent) tree: to be redone in the reference's own order
    }
    // flatten: the main context first, then the sub-contexts; references between contexts become global ids
    std::vector<size_t> offset(b.tasks.size());
    size_t total = main.nodes.size();
    for (size_t k = 0; k < b.tasks.size(); ++k) {
        offset[k] = total;
        total += b.tasks[k]->nodes.size();
    }
    t.nodes.reserve(total);
    auto append = [&](const std::vector<Node>& src, size_t base) {
        for (Node nd : src) {
            if (nd.feat >= 0) {
                nd.a = (nd.a & TASK_REF) ? static_cast<uint32_t>(offset[nd.a & ~TASK_REF]) : static_cast<uint32_t>(nd.a + base);
                nd.b = (nd.b & TASK_REF) ? static_cast<uint32_t>(offset[nd.b & ~TASK_REF]) : static_cast<uint32_t>(nd.b + base);
            }
            t.nodes.push_back(nd);
        }
    };
    append(main.nodes, 0);
    for (size_t k = 0; k < b.tasks.size(); ++k) {
        append(b.tasks[k]->nodes, offset[k]);
        delete b.tasks[k];
    }
    t.root = 0;
    t.error = b.error;
    return true;
}

// xyz: 3 doubles per source point in source order (NaN triples for points without coordinates).  threads <= 1: sequential.
inline void build(Tree& t, const double* xyz, size_t n, unsigned threads)
{
    t.vind.resize(n);
    t.pts.assign(xyz, xyz + 3 * n);
    if (!build_in_place(t, threads)) {
        t.pts.assign(xyz, xyz + 3 * n);
        build_in_place(t, 1);
    }
}

// :1239-1256
inline double initial_distances(const Tree& t, const double q[3], double dists[3])
{
    double distsq = 0.0;
    for (int d = 0; d < 3; ++d) {
        dists[d] = 0;
        if (q[d] < t.rootBox.lo[d]) {
            dists[d] = (q[d] - t.rootBox.lo[d]) * (q[d] - t.rootBox.lo[d]);
            distsq += dists[d];
        }
        if (q[d] > t.rootBox.hi[d]) {
            dists[d] = (q[d] - t.rootBox.hi[d]) * (q[d] - t.rootBox.hi[d]);
            distsq += dists[d];
        }
    }
    return distsq;
}

namespace detail {
inline void walk(const Tree& t, const double q[3], uint32_t node, double mindistsq, double dists[3], double r2,
                 std::vector<std::pair<size_t, double>>& out)
{
    const Node& nd = t.nodes[node];
    if (nd.feat < 0) {
        for (uint32_t i = nd.a; i < nd.b; ++i) {
            const double d0 = q[0] - t.pts[3 * size_t(i)], d1 = q[1] - t.pts[3 * size_t(i) + 1], d2 = q[2] - t.pts[3 * size_t(i) + 2];
            const double dist = d0 * d0 + d1 * d1 + d2 * d2; // PointCloud::kdtree_distance, CDMInterpolator.cc:924-930
            if (dist < r2)
                out.emplace_back(t.vind[i], dist);
        }
        return;
    }
    const int k = nd.feat;
    const double val = q[k];
    const double diff1 = val - nd.divlow, diff2 = val - nd.divhigh;
    uint32_t near, far;
    double cut;
    if ((diff1 + diff2) < 0) {
        near = nd.a;
        far = nd.b;
        cut = (val - nd.divhigh) * (val - nd.divhigh);
    } else {
        near = nd.b;
        far = nd.a;
        cut = (val - nd.divlow) * (val - nd.divlow);
    }
    walk(t, q, near, mindistsq, dists, r2, out);
    const double keep = dists[k];
    mindistsq = mindistsq + cut - keep;
    dists[k] = cut;
    if (mindistsq <= r2) // epsError is float(1 + 0): the product is exact
        walk(t, q, far, mindistsq, dists, r2, out);
    dists[k] = keep;
}
} // namespace detail

// radiusSearch(q, r2, matches, sorted=true) -> matches.at(0).first, or -1 (nanoflann.hpp:959-976).  The sort is the real
// std::sort of this libstdc++ on the same element type and comparator, fed in the same order.
inline long search_first(const Tree& t, const double q[3], double r2)
{
    std::vector<std::pair<size_t, double>> m;
    double dists[3];
    const double d0 = initial_distances(t, q, dists);
    detail::walk(t, q, t.root, d0, dists, r2, m);
    if (m.empty())
        return -1;
    std::sort(m.begin(), m.end(), [](const std::pair<size_t, double>& p1, const std::pair<size_t, double>& p2) { return p1.second < p2.second; });
    return static_cast<long>(m[0].first);
}

// lon/lat degrees -> unit-sphere xyz with the host's libm, CDMInterpolator.cc:969-988 (cloud) and :1002-1011 (targets)
inline void sphere_xyz(const double* lon, const double* lat, size_t first, size_t last, bool nan_to_nan, double* xyz)
{
    const double D2R = 0.017453292519943296; // deg_to_rad, include/fimex/MathUtils.h:158-162
    for (size_t i = first; i < last; ++i) {
        if (nan_to_nan && (std::isnan(lat[i]) || std::isnan(lon[i]))) {
            xyz[3 * i] = xyz[3 * i + 1] = xyz[3 * i + 2] = std::nan("");
            continue;
        }
        const double lat_rad = lat[i] * D2R, lon_rad = lon[i] * D2R;
        const double sinLat = std::sin(lat_rad), cosLat = std::cos(lat_rad), sinLon = std::sin(lon_rad), cosLon = std::cos(lon_rad);
        xyz[3 * i] = cosLat * cosLon;
        xyz[3 * i + 1] = cosLat * sinLon;
        xyz[3 * i + 2] = sinLat;
    }
}

} // namespace kdclone
} // namespace fxcu
