// fxcu_kdtree.cu -- K9, coord_kdtree with the reference's own tie and NaN behaviour.
//
// Replaces flannTranslatePointsToClosestInputCell (src/CDMInterpolator.cc:953-1028).  The reference's answer for a target is
// "element 0 of nanoflann's radius search after std::sort on the distance" -- a function of the tree's shape whenever two source
// points are exactly equidistant or the cloud holds NaN points (fxcu_kdclone.h explains why).  So this path
//   1. forms the unit-sphere coordinates of the cloud and of the targets with the HOST's libm (all host cores): sin/cos of the
//      device differ from glibc's in the last bit, and every later comparison must see the reference's numbers;
//   2. builds the tree on the host with the reference's split rules (fxcu_kdclone.h; ~0.25 s for a 2030x1354 swath);
//   3. walks it on the GPU, one thread per target, in the reference's order (near child first, far child iff
//      mindistsq <= r^2, leaf entries in index order), keeping the smallest distance, the FIRST entry met that has it, the
//      number of matches and the number of ties at the minimum.  All arithmetic is separately rounded f64 (the library is
//      compiled with --fmad=false), i.e. the numbers are the host's;
//   4. up to 16 matches libstdc++'s std::sort is an insertion sort, which keeps the first-met minimum in front; a unique minimum
//      is in front whatever the sort does.  Only targets with more than 16 matches AND a tie at the minimum (or a walk deeper
//      than the device stack) depend on introsort's partitioning: those few are redone on the host with the real std::sort.
// nanoflann_compat = 0 selects the tree-free search of fxcu_nn.cu instead (true nearest point, NaN points skipped, lower
// source index on ties).
#include "fxcu_common.h"
#include "fxcu_hostpar.h"
#include "fxcu_kdclone.h"

#include <thread>

namespace fxcu {

// defined in fxcu_nn.cu
int kdtree_translate_grid(double maxDist, double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals, const double* latVals,
                          size_t orgXDimSize, size_t orgYDimSize);

namespace {

constexpr int KD_STACK = 64;

struct __align__(16) DevNode {
    double divlow, divhigh;
    unsigned a, b;
    int feat;
    int pad_;
};
static_assert(sizeof(DevNode) == sizeof(kdclone::Node), "device and host node layouts must agree");

struct DevBox {
    double lo[3], hi[3];
};

__global__ void __launch_bounds__(128) kd_walk_kernel(const double* __restrict__ q, size_t nq, const DevNode* __restrict__ nodes,
                                                      const double* __restrict__ pts, const unsigned* __restrict__ vind, DevBox box, unsigned root,
                                                      double r2, unsigned nx, double* __restrict__ outx, double* __restrict__ outy,
                                                      unsigned* __restrict__ redo, unsigned* __restrict__ nredo)
{
    const size_t i = blockIdx.x * 128ull + threadIdx.x;
    if (i >= nq)
        return;
    const double v0 = q[3 * i], v1 = q[3 * i + 1], v2 = q[3 * i + 2];
    // computeInitialDistances, nanoflann.hpp:1239-1256
    double d0 = 0., d1 = 0., d2 = 0., md = 0.;
    if (v0 < box.lo[0]) { d0 = (v0 - box.lo[0]) * (v0 - box.lo[0]); md += d0; }
    if (v0 > box.hi[0]) { d0 = (v0 - box.hi[0]) * (v0 - box.hi[0]); md += d0; }
    if (v1 < box.lo[1]) { d1 = (v1 - box.lo[1]) * (v1 - box.lo[1]); md += d1; }
    if (v1 > box.hi[1]) { d1 = (v1 - box.hi[1]) * (v1 - box.hi[1]); md += d1; }
    if (v2 < box.lo[2]) { d2 = (v2 - box.lo[2]) * (v2 - box.lo[2]); md += d2; }
    if (v2 > box.hi[2]) { d2 = (v2 - box.hi[2]) * (v2 - box.hi[2]); md += d2; }

    unsigned sn[KD_STACK];
    double sm[KD_STACK], s0[KD_STACK], s1[KD_STACK], s2[KD_STACK];
    int sp = 0;
    bool overflow = false;

    double best = INFINITY;
    unsigned best_idx = 0xffffffffu, nmatch = 0, nties = 0;
    unsigned cur = root;
    for (;;) {
        // searchLevel, nanoflann.hpp:1263-1310: go down the near side, remember the far sides that can still hold a match
        for (;;) {
            const int4 meta = __ldg(reinterpret_cast<const int4*>(&nodes[cur]) + 1);
            if (meta.z < 0) {
                for (unsigned k = static_cast<unsigned>(meta.x); k < static_cast<unsigned>(meta.y); ++k) {
                    const double e0 = v0 - pts[3 * size_t(k)], e1 = v1 - pts[3 * size_t(k) + 1], e2 = v2 - pts[3 * size_t(k) + 2];
                    const double dist = e0 * e0 + e1 * e1 + e2 * e2; // kdtree_distance, CDMInterpolator.cc:924-930
                    if (dist < r2) { // RadiusResultSet::addPoint, :150-154
                        ++nmatch;
                        if (dist < best) {
                            best = dist;
                            best_idx = vind[k];
                            nties = 1;
                        } else if (dist == best) {
                            ++nties;
                        }
                    }
                }
                break;
            }
            const double2 div = __ldg(reinterpret_cast<const double2*>(&nodes[cur]));
            const int f = meta.z;
            const double val = f == 0 ? v0 : (f == 1 ? v1 : v2);
            const double diff1 = val - div.x, diff2 = val - div.y;
            unsigned near, far;
            double cut;
            if ((diff1 + diff2) < 0) {
                near = static_cast<unsigned>(meta.x);
                far = static_cast<unsigned>(meta.y);
                cut = (val - div.y) * (val - div.y);
            } else {
                near = static_cast<unsigned>(meta.y);
                far = static_cast<unsigned>(meta.x);
                cut = (val - div.x) * (val - div.x);
            }
            const double keep = f == 0 ? d0 : (f == 1 ? d1 : d2);
            const double fmd = md + cut - keep;
            if (fmd <= r2) {
                if (sp < KD_STACK) {
                    sn[sp] = far;
                    sm[sp] = fmd;
                    s0[sp] = f == 0 ? cut : d0;
                    s1[sp] = f == 1 ? cut : d1;
                    s2[sp] = f == 2 ? cut : d2;
                    ++sp;
                } else {
                    overflow = true;
                }
            }
            cur = near;
        }
        if (sp == 0)
            break;
        --sp;
        cur = sn[sp];
        md = sm[sp];
        d0 = s0[sp];
        d1 = s1[sp];
        d2 = s2[sp];
    }
    if (overflow || (nmatch > 16 && nties > 1)) {
        redo[atomicAdd(nredo, 1u)] = static_cast<unsigned>(i);
        return;
    }
    if (best_idx != 0xffffffffu) {
        outx[i] = static_cast<double>(best_idx % nx); // pos = ix + nx*iy, CDMInterpolator.cc:1015-1020
        outy[i] = static_cast<double>(best_idx / nx);
    } else {
        outx[i] = -1000.;
        outy[i] = -1000.;
    }
}

int kdtree_translate_clone(double maxDist, double* px, double* py, size_t nq, const double* lon, const double* lat, size_t nx, size_t ny)
{
    const size_t n = nx * ny;
    const double r = maxDist / 6371000.; // MIFI_EARTH_RADIUS_M; "all calculations on a sphere with unit 1" (:964)
    const double r2 = r * r;             // "using square since distance is not sqrt" (:998)
    const unsigned hw = host_threads();

    // the targets' coordinates are formed while the tree is being built
    std::vector<double> qxyz(3 * nq);
    kdclone::Tree tree;
    tree.vind.resize(n);
    tree.pts.resize(3 * n); // the cloud is written straight into the tree's point array and permuted there
    {
        struct Joiner {
            std::thread t;
            ~Joiner()
            {
                if (t.joinable())
                    t.join();
            }
        } targets{std::thread([&]() { parallel_ranges(nq, 1 << 16, [&](size_t a, size_t b) { kdclone::sphere_xyz(px, py, a, b, false, qxyz.data()); }); })};
        auto fill = [&]() { parallel_ranges(n, 1 << 16, [&](size_t a, size_t b) { kdclone::sphere_xyz(lon, lat, a, b, true, tree.pts.data()); }); };
        fill();
        if (!kdclone::build_in_place(tree, std::min(hw, 32u))) {
            fill();
            kdclone::build_in_place(tree, 1);
        }
    }
    if (!tree.error.empty())
        FXCU_FAIL(tree.error);
    if (nq == 0)
        return FIMEXCU_OK;
    if (nq >= 0xffffffffull)
        FXCU_FAIL("fimexcu_kdtree_translate: too many target points");

    DevBuf dNodes, dPts, dVind, dQ, dX, dY, dRedo, dCount;
    if (!dNodes.alloc(tree.nodes.size() * sizeof(DevNode), nullptr) || !dPts.alloc(3 * n * 8, nullptr) || !dVind.alloc(n * 4, nullptr) ||
        !dQ.alloc(3 * nq * 8, nullptr) || !dX.alloc(nq * 8, nullptr) || !dY.alloc(nq * 8, nullptr) || !dRedo.alloc(nq * 4, nullptr) ||
        !dCount.alloc(4, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(dNodes.p, tree.nodes.data(), tree.nodes.size() * sizeof(DevNode), cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dPts.p, tree.pts.data(), 3 * n * 8, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dVind.p, tree.vind.data(), n * 4, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dQ.p, qxyz.data(), 3 * nq * 8, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemsetAsync(dCount.p, 0, 4, nullptr));
    DevBox box;
    for (int d = 0; d < 3; ++d) {
        box.lo[d] = tree.rootBox.lo[d];
        box.hi[d] = tree.rootBox.hi[d];
    }
    kd_walk_kernel<<<blocks_for(nq, 128), 128>>>(dQ.as<double>(), nq, dNodes.as<DevNode>(), dPts.as<double>(), dVind.as<unsigned>(), box, tree.root, r2,
                                                 static_cast<unsigned>(nx), dX.as<double>(), dY.as<double>(), dRedo.as<unsigned>(),
                                                 dCount.as<unsigned>());
    FXCU_CUDA(cudaGetLastError());
    unsigned nredo = 0;
    FXCU_CUDA(cudaMemcpyAsync(&nredo, dCount.p, 4, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(px, dX.p, nq * 8, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(py, dY.p, nq * 8, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    if (nredo) {
        std::vector<unsigned> redo(nredo);
        FXCU_CUDA(cudaMemcpy(redo.data(), dRedo.p, size_t(nredo) * 4, cudaMemcpyDeviceToHost));
        parallel_ranges(nredo, 256, [&](size_t a, size_t b) {
            for (size_t k = a; k < b; ++k) {
                const size_t i = redo[k];
                const long pos = kdclone::search_first(tree, &qxyz[3 * i], r2);
                if (pos >= 0) {
                    px[i] = static_cast<double>(size_t(pos) % nx);
                    py[i] = static_cast<double>(size_t(pos) / nx);
                } else {
                    px[i] = py[i] = -1000.;
                }
            }
        });
    }
    return FIMEXCU_OK;
}

} // namespace
} // namespace fxcu

using namespace fxcu;

extern "C" {

int fimexcu_kdtree_translate_mode(double maxDist, double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals,
                                  const double* latVals, size_t orgXDimSize, size_t orgYDimSize, int nanoflann_compat)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t n = orgXDimSize * orgYDimSize;
    if (n == 0 || n >= 2147483647ull)
        FXCU_FAIL("fimexcu_kdtree_translate: source grid size out of range");
    if (!(maxDist > 0))
        FXCU_FAIL("fimexcu_kdtree_translate: maxDist must be positive"); // assert(maxDist != 0), CDMInterpolator.cc:961
    if (!nanoflann_compat)
        return kdtree_translate_grid(maxDist, pointsOnXAxis, pointsOnYAxis, pointsSize, lonVals, latVals, orgXDimSize, orgYDimSize);
    try {
        return kdtree_translate_clone(maxDist, pointsOnXAxis, pointsOnYAxis, pointsSize, lonVals, latVals, orgXDimSize, orgYDimSize);
    } catch (const std::exception& e) {
        FXCU_FAIL(std::string("fimexcu_kdtree_translate: ") + e.what());
    }
}

int fimexcu_kdtree_translate(double maxDist, double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals,
                             const double* latVals, size_t orgXDimSize, size_t orgYDimSize)
{
    return fimexcu_kdtree_translate_mode(maxDist, pointsOnXAxis, pointsOnYAxis, pointsSize, lonVals, latVals, orgXDimSize, orgYDimSize, 1);
}

} // extern "C"
