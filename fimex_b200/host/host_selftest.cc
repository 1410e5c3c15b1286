// host_selftest.cc -- exercises the C++ glue the way CDMInterpolator::getDataSlice does (src/CDMInterpolator.cc:219-263):
// createCachedInterpolation -> interpolateValues -> reprojectValues.  Reads a small binary case written by the pytest driver
// (tests/test_gpu_host_cpp.py), writes the results next to it; the driver compares them with the oracle.
//   file layout (little endian): int32 method, inX, inY, outX, outY, nz; f64 px[out], py[out]; f32 u[nz*in], v[nz*in]; f64 matrix[out*3]
#include "CudaCachedInterpolation.h"
#include "CudaInterpolatorProcess2d.h"
#include "CudaSlab.h"

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

using namespace MetNoFimex;

static bool read_all(FILE* f, void* p, size_t bytes) { return bytes == 0 || fread(p, 1, bytes, f) == bytes; }

// host_selftest fill case.bin result.bin -- the fill functors the way processArray_ drives them (src/CDMInterpolator.cc:163-186):
//   case: int32 kind (0 fill2d, 1 creepfill2d, 2 creepfillval2d), nx, ny, nz; f32 p0, p1, p2; f32 field[nz*ny*nx]
//         fill2d: p0 relaxCrit, p1 corrEff, p2 maxLoop; creepfill: p0 repeat, p1 setWeight, p2 defaultValue
//   result: field after operator() level by level, then field after processSlab()
static int fill_mode(const char* in_path, const char* out_path)
{
    FILE* f = fopen(in_path, "rb");
    if (!f)
        return 2;
    int32_t hdr[4];
    float p[3];
    if (!read_all(f, hdr, sizeof(hdr)) || !read_all(f, p, sizeof(p)))
        return 2;
    const size_t nx = hdr[1], ny = hdr[2], nz = hdr[3], n = nx * ny * nz;
    std::vector<float> a(n), b;
    if (!read_all(f, a.data(), n * 4))
        return 2;
    fclose(f);
    b = a;
    try {
        std::shared_ptr<CudaInterpolatorProcess2d> proc;
        if (hdr[0] == 0)
            proc = std::make_shared<CudaInterpolatorFill2d>(p[0], p[1], static_cast<size_t>(p[2]));
        else if (hdr[0] == 1)
            proc = std::make_shared<CudaInterpolatorCreepFill2d>(static_cast<unsigned short>(p[0]), static_cast<char>(p[1]));
        else
            proc = std::make_shared<CudaInterpolatorCreepFillVal2d>(static_cast<unsigned short>(p[0]), static_cast<char>(p[1]), p[2]);
        InterpolatorProcess2d& base = *proc; // the reference holds them as InterpolatorProcess2d_p
        for (size_t z = 0; z < nz; ++z)
            base(a.data() + z * nx * ny, nx, ny);
        proc->processSlab(b.data(), n, nx, ny);
    } catch (const std::exception& e) {
        fprintf(stderr, "exception: %s\n", e.what());
        return 6;
    }
    FILE* o = fopen(out_path, "wb");
    fwrite(a.data(), 4, n, o);
    fwrite(b.data(), 4, n, o);
    fclose(o);
    printf("host_selftest fill ok\n");
    return 0;
}

// host_selftest threads N case.bin result.bin -- the writer's loop in miniature (src/io/netcdf/NetCDF_CDMWriter.cc:599-618):
// ONE cache object, N host threads pulling different slabs through interpolateValues at the same time (each thread gets its GPU
// from cuda_glue::bind_thread(); with one GPU they share it).  Case file as below with nz = levels per slab and u holding
// nslabs = (file size) slabs back to back: int32 method,inX,inY,outX,outY,nz,nslabs; f64 px,py; f32 slabs[nslabs][nz*in].
// result: all regridded slabs in order, then int32 device ordinal used for each slab.
static int threads_mode(int nthreads, const char* in_path, const char* out_path)
{
    FILE* f = fopen(in_path, "rb");
    if (!f)
        return 2;
    int32_t hdr[7];
    if (!read_all(f, hdr, sizeof(hdr)))
        return 2;
    const int method = hdr[0];
    const size_t inX = hdr[1], inY = hdr[2], outX = hdr[3], outY = hdr[4], nz = hdr[5], nslabs = hdr[6];
    const size_t nout = outX * outY, nin = inX * inY;
    auto px = make_shared_array<double>(nout), py = make_shared_array<double>(nout);
    if (!read_all(f, px.get(), nout * 8) || !read_all(f, py.get(), nout * 8))
        return 2;
    std::vector<shared_array<float>> slabs(nslabs), results(nslabs);
    for (size_t k = 0; k < nslabs; ++k) {
        slabs[k] = make_shared_array<float>(nz * nin); // pageable, like the upstream reader's arrays
        if (!read_all(f, slabs[k].get(), nz * nin * 4))
            return 2;
    }
    fclose(f);
    std::vector<int32_t> device_of(nslabs, -1);
    std::vector<std::string> errors(nthreads);
    try {
        CachedInterpolationInterface_p ci = createCudaCachedInterpolation("x", "y", method, px, py, inX, inY, outX, outY);
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; ++t)
            pool.emplace_back([&, t]() {
                try {
                    for (size_t k = t; k < nslabs; k += nthreads) { // static schedule, like the reference's omp for
                        size_t newSize = 0;
                        results[k] = ci->interpolateValues(slabs[k], nz * nin, newSize);
                        device_of[k] = cuda_glue::bind_thread();
                        if (newSize != nz * nout)
                            errors[t] = "wrong newSize";
                    }
                } catch (const std::exception& e) {
                    errors[t] = e.what();
                }
            });
        for (auto& th : pool)
            th.join();
    } catch (const std::exception& e) {
        fprintf(stderr, "exception: %s\n", e.what());
        return 6;
    }
    for (const std::string& e : errors)
        if (!e.empty()) {
            fprintf(stderr, "thread error: %s\n", e.c_str());
            return 7;
        }
    FILE* o = fopen(out_path, "wb");
    for (size_t k = 0; k < nslabs; ++k)
        fwrite(results[k].get(), 4, nz * nout, o);
    fwrite(device_of.data(), 4, nslabs, o);
    fclose(o);
    printf("host_selftest threads ok (%zu devices)\n", cuda_glue::devices().size());
    return 0;
}

// host_selftest chain case.bin result.bin -- BASELINE configs[4] as a device-resident reader chain (CudaSlab.h): the ensemble field
// (hybrid, member, y, x) arrives as a PACKED short slab, is unpacked, regridded bilinearly (staged kernel, padded layout), the
// surface pressure is regridded the same way, the vertical interpolation to fixed pressure levels runs with the pressure formed from
// the REGRIDDED ps, the wind pair goes through the fused rotation -- and only the final results come back to the host.
//   case: int32 method(vertical), inX, inY, outX, outY, nLev, nMem, nOut; f64 px[out], py[out]; i16 field[nLev*nMem*in]; f64 fill;
//         f32 ps[nMem*in]; f64 ap[nLev], b[nLev], levels[nOut]; f64 matrix[out*3]
//   result: f32 vertical[nOut*nMem*out]; f32 u[nMem*out], v[nMem*out] (the rotated regridded ps / first level pair, as a K7 check)
static int chain_mode(const char* in_path, const char* out_path)
{
    FILE* f = fopen(in_path, "rb");
    if (!f)
        return 2;
    int32_t hdr[8];
    if (!read_all(f, hdr, sizeof(hdr)))
        return 2;
    const int vmethod = hdr[0];
    const size_t inX = hdr[1], inY = hdr[2], outX = hdr[3], outY = hdr[4], nLev = hdr[5], nMem = hdr[6], nOut = hdr[7];
    const size_t nout = outX * outY, nin = inX * inY;
    auto px = make_shared_array<double>(nout), py = make_shared_array<double>(nout);
    auto raw = make_shared_array<short>(nLev * nMem * nin);
    auto ps = make_shared_array<float>(nMem * nin);
    double fill = 0;
    std::vector<double> ap(nLev), b(nLev), levels(nOut);
    auto m = std::make_shared<reproject::Matrix>(static_cast<int>(outX), static_cast<int>(outY));
    if (!read_all(f, px.get(), nout * 8) || !read_all(f, py.get(), nout * 8) || !read_all(f, raw.get(), nLev * nMem * nin * 2) || !read_all(f, &fill, 8) ||
        !read_all(f, ps.get(), nMem * nin * 4) || !read_all(f, ap.data(), nLev * 8) || !read_all(f, b.data(), nLev * 8) ||
        !read_all(f, levels.data(), nOut * 8) || !read_all(f, m->mtx(), nout * 3 * 8))
        return 2;
    fclose(f);
    try {
        CudaCachedInterpolation ci("x", "y", FIMEXCU_INTERPOL_BILINEAR, px, py, inX, inY, outX, outY);
        if (ci.getInX() != inX || ci.getInY() != inY)
            return 4; // the driver passes full-window tables
        size_t ip, il, op, ol;
        fimexcu_cache_padded_layout(ci.handle(), &ip, &il, &op, &ol);
        // reader side: the packed field crosses PCIe as shorts, straight into the padded layout
        CudaSlab_p packed = CudaSlab::upload(raw, nLev * nMem, inY, inX, ip, il);
        CudaSlab_p field = unpackSlab(*packed, fill);
        packed.reset();
        CudaSlab_p psIn = uploadInterpolationInput(ci, ps, nMem * nin);
        // horizontal, then vertical with the pressure from the regridded ps (fimex.cc:862-863)
        CudaSlab_p hField = interpolateSlab(ci, *field);
        CudaSlab_p hPs = interpolateSlab(ci, *psIn);
        CudaSlab_p vert = cudaVerticalToFixedLevels(vmethod, *hField, nLev, ap, b, *hPs, 0.01, levels);
        // fused vector pass on two device slabs (ps and the first level of the field stand in for a wind pair)
        CudaCachedVectorReprojection cvr(m);
        CudaSlab_p firstLevel = CudaSlab::allocate(CudaSlab::FLOAT, nMem, inY, inX, ip, il);
        if (fimexcu_memcpy_d2d(firstLevel->devicePtr(), field->devicePtr(), nMem * il * sizeof(float)) != FIMEXCU_OK)
            return 5;
        auto uv = interpolateVectorSlabs(ci, cvr, *psIn, *firstLevel);
        // writer side
        shared_array<float> hv = vert->asFloat(), hu = uv.first->asFloat(), hvv = uv.second->asFloat();
        {
            // the same pair by the separate steps (interpolate each, rotate in place on the padded slabs): identical bits
            CudaSlab_p su = interpolateSlab(ci, *psIn), sv = interpolateSlab(ci, *firstLevel);
            rotateVectorSlabs(cvr, *su, *sv);
            shared_array<float> a = su->asFloat(), b2 = sv->asFloat();
            auto same = [](const float* x, const float* y, size_t n) { // NaN where the other is NaN, equal bits elsewhere
                for (size_t i = 0; i < n; ++i)
                    if ((x[i] != x[i]) != (y[i] != y[i]) || (x[i] == x[i] && std::memcmp(x + i, y + i, 4) != 0))
                        return false;
                return true;
            };
            if (!same(a.get(), hu.get(), uv.first->size()) || !same(b2.get(), hvv.get(), uv.second->size())) {
                fprintf(stderr, "rotateVectorSlabs differs from the fused vector pass\n");
                return 7;
            }
        }
        FILE* o = fopen(out_path, "wb");
        fwrite(hv.get(), 4, vert->size(), o);
        fwrite(hu.get(), 4, uv.first->size(), o);
        fwrite(hvv.get(), 4, uv.second->size(), o);
        fclose(o);
    } catch (const std::exception& e) {
        fprintf(stderr, "exception: %s\n", e.what());
        return 6;
    }
    printf("host_selftest chain ok\n");
    return 0;
}

int main(int argc, char** argv)
{
    if (argc == 4 && std::strcmp(argv[1], "chain") == 0)
        return chain_mode(argv[2], argv[3]);
    if (argc == 4 && std::strcmp(argv[1], "fill") == 0)
        return fill_mode(argv[2], argv[3]);
    if (argc == 5 && std::strcmp(argv[1], "threads") == 0)
        return threads_mode(std::atoi(argv[2]), argv[3], argv[4]);
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
