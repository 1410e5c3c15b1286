// host_selftest.cc -- exercises the C++ glue the way CDMInterpolator::getDataSlice does (src/CDMInterpolator.cc:219-263):
// createCachedInterpolation -> interpolateValues -> reprojectValues.  Reads a small binary case written by the pytest driver
// (tests/test_gpu_host_cpp.py), writes the results next to it; the driver compares them with the oracle.
//   file layout (little endian): int32 method, inX, inY, outX, outY, nz; f64 px[out], py[out]; f32 u[nz*in], v[nz*in]; f64 matrix[out*3]
#include "CudaCachedInterpolation.h"

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

using namespace MetNoFimex;

static bool read_all(FILE* f, void* p, size_t bytes) { return bytes == 0 || fread(p, 1, bytes, f) == bytes; }

int main(int argc, char** argv)
{
    if (argc != 3) {
        fprintf(stderr, "usage: host_selftest case.bin result.bin\n");
        return 2;
    }
    FILE* f = fopen(argv[1], "rb");
    if (!f)
        return 2;
    int32_t hdr[6];
    if (!read_all(f, hdr, sizeof(hdr)))
        return 2;
    const int method = hdr[0];
    const size_t inX = hdr[1], inY = hdr[2], outX = hdr[3], outY = hdr[4], nz = hdr[5];
    const size_t nout = outX * outY, nin = inX * inY;
    auto px = make_shared_array<double>(nout), py = make_shared_array<double>(nout);
    auto u = make_shared_array<float>(nz * nin), v = make_shared_array<float>(nz * nin);
    auto m = std::make_shared<reproject::Matrix>(static_cast<int>(outX), static_cast<int>(outY));
    if (!read_all(f, px.get(), nout * 8) || !read_all(f, py.get(), nout * 8) || !read_all(f, u.get(), nz * nin * 4) || !read_all(f, v.get(), nz * nin * 4) ||
        !read_all(f, m->mtx(), nout * 3 * 8))
        return 2;
    fclose(f);
    try {
        // an unknown method must throw like the reference's factory
        bool threw = false;
        try {
            createCudaCachedInterpolation("x", "y", 99, px, py, inX, inY, outX, outY);
        } catch (const CDMException&) {
            threw = true;
        }
        if (!threw)
            return 3;
        CachedInterpolationInterface_p ci = createCudaCachedInterpolation("x", "y", method, px, py, inX, inY, outX, outY);
        // the driver passes full-window tables (no reduction beyond offset 0), so the input is used as it is
        if (ci->getInX() != inX || ci->getInY() != inY)
            return 4;
        size_t newSize = 0, newSize2 = 0;
        shared_array<float> iu = ci->interpolateValues(u, nz * nin, newSize);
        shared_array<float> iv = ci->interpolateValues(v, nz * nin, newSize2);
        if (newSize != nz * nout || newSize2 != newSize)
            return 5;
        CudaCachedVectorReprojection cvr(m);
        cvr.reprojectValues(iu, iv, newSize);
        FILE* o = fopen(argv[2], "wb");
        fwrite(iu.get(), 4, newSize, o);
        fwrite(iv.get(), 4, newSize, o);
        fclose(o);
    } catch (const std::exception& e) {
        fprintf(stderr, "exception: %s\n", e.what());
        return 6;
    }
    printf("host_selftest ok\n");
    return 0;
}
