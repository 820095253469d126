// TEST INFRASTRUCTURE (fixture generator, run once in the build container where /root/reference exists).
// Builds a small Vec3f NanoVDB grid with the reference's own NanoVDB headers, writes it uncompressed with
// nanovdb::io::writeGrid, and samples it exactly like the reference's source stage does
// (integrator/uniform/integrator.cu:309-311: worldToIndex + SampleFromVoxels<Vec3fTree, 1, false>).
//   g++ -std=c++17 -O1 -DNANOVDB_USE_ZIP -I/root/reference/ext/openvdb/nanovdb tests/golden/make_nvdb_golden.cpp -o /tmp/make_nvdb_golden -lz
//   /tmp/make_nvdb_golden tests/golden/source_vec3f.nvdb tests/golden/source_vec3f_samples.txt tests/golden/source_vec3f_zip.nvdb
// (the optional third argument writes the same grid with io::Codec::ZIP, the reference's zlib codec, io/IO.h:221-236)
#include <nanovdb/NanoVDB.h>
#include <nanovdb/io/IO.h>
#include <nanovdb/math/SampleFromVoxels.h>
#include <nanovdb/tools/CreateNanoGrid.h>
#include <nanovdb/tools/GridBuilder.h>

#include <cmath>
#include <cstdio>
#include <random>

int main(int argc, char **argv)
{
    if (argc < 3)
        return 1;
    using namespace nanovdb;
    tools::build::Grid<Vec3f> grid(Vec3f(0.0f), "source", GridClass::Unknown);
    const double h = 0.25;
    grid.setTransform(h, Vec3d(-1.5, -1.0, -0.75));
    auto acc = grid.getAccessor();
    // a box of voxels that straddles leaf (8^3) and lower-node (128^3) boundaries inside one upper node, with a hole (inactive voxel)
    for (int i = 3; i < 17; ++i)
        for (int j = 122; j < 133; ++j)
            for (int k = 2; k < 11; ++k)
            {
                if (i == 9 && j == 127 && k == 5)
                    continue;
                acc.setValue(Coord(i, j, k), Vec3f(0.1f * i + 0.01f * j, std::sin(0.3f * j) + 0.2f * k, 0.05f * i * k - 0.3f));
            }
    auto handle = tools::createNanoGrid(grid);
    io::writeGrid(argv[1], handle, io::Codec::NONE);
    if (argc > 3)
        io::writeGrid(argv[3], handle, io::Codec::ZIP);
    const auto *g = handle.grid<Vec3f>();
    math::SampleFromVoxels<Vec3fTree, 1, false> sampler(g->tree());
    FILE *f = fopen(argv[2], "w");
    std::mt19937 rng(7);
    std::uniform_real_distribution<float> ux(-1.2f, 3.2f), uy(29.0f, 32.8f), uz(-0.7f, 2.4f);
    for (int n = 0; n < 400; ++n)
    {
        const Vec3f p(ux(rng), uy(rng), uz(rng));
        const Vec3f v = sampler(g->worldToIndex(p));
        fprintf(f, "%.9g %.9g %.9g %.9g %.9g %.9g\n", p[0], p[1], p[2], v[0], v[1], v[2]);
    }
    fclose(f);
    const auto bb = g->indexBBox();
    printf("grid %llu bytes, index bbox [%d %d %d]-[%d %d %d]\n", (unsigned long long)handle.size(), bb.min()[0], bb.min()[1], bb.min()[2], bb.max()[0], bb.max()[1],
           bb.max()[2]);
    return 0;
}
