// NanoVDB Vec3f source grids (row N3): host-side reader that turns a .nvdb file into the dense Vec3f grid the C ABI takes.
//
// Replaces nanovdb::io::readGrid + Vec3fGrid::worldToIndex + SampleFromVoxels<Vec3fTree, 1, false>
// (core/problem.cu:135-149, integrator/uniform/integrator.cu:296-313). NanoVDB itself is a third-party header library
// (ext/openvdb/nanovdb, 32.x); nothing of it is compiled in here. The byte layout below restates its published format as
// tabulated in PNanoVDB.h (grid 672 B, map at 296, tree 64 B, root / upper 32^3 / lower 16^3 / leaf 8^3 nodes, the
// Vec3f row of pnanovdb_grid_type_constants) and io/IO.h + NanoVDB.h:5817-5856 (FileHeader 16 B, FileMetaData 176 B + name).
// The tree is flattened over its index bounding box: getValue() semantics (leaf values incl. inactive voxels, tile values of
// childless table entries, background elsewhere), so that trilinear interpolation of the dense grid == the reference's
// order-1 sampler. Supported: files written with Codec::NONE or Codec::ZIP (io/IO.h:221-236,279-292: a 64-bit compressed
// size followed by one zlib stream that inflates to FileMetaData::gridSize bytes; inflated here with the system zlib) and raw
// grid buffers; Vec3f grids, axis-aligned uniform maps (the only kind a voxel size + origin can express), zero background
// (what lies outside the box reads as zero). Codec::BLOSC needs the blosc library, which this image does not have: such
// files are reported as unsupported.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <zlib.h>

namespace wnvdb
{
    struct DenseVec3f
    {
        int dims[3] = {0, 0, 0};       // nx, ny, nz
        int ijk_min[3] = {0, 0, 0};    // index-space coordinate of voxel (0,0,0) of the dense grid
        float origin[3] = {0, 0, 0};   // world position of that voxel
        float voxel = 1.0f;
        std::vector<float> data;       // [z][y][x][3]
    };

    namespace detail
    {
        constexpr uint64_t kMagicNumb = 0x304244566f6e614eULL, kMagicGrid = 0x314244566f6e614eULL, kMagicFile = 0x324244566f6e614eULL;
        // Vec3f row of PNanoVDB.h: pnanovdb_grid_type_constants[6]
        constexpr uint32_t kGridTypeVec3f = 6, kRootOffBackground = 28, kRootSize = 96, kRootTileSize = 32, kRootTileOffValue = 20, kTableStride = 16,
                           kUpperOffTable = 8256, kUpperOffChildMask = 4128, kLowerOffTable = 1088, kLowerOffChildMask = 544, kLeafOffTable = 128;
        struct Buf
        {
            const uint8_t *p;
            size_t n;
            bool ok(size_t off, size_t len) const { return off <= n && len <= n - off; }
            template <class T>
            T rd(size_t off) const
            {
                T v;
                memcpy(&v, p + off, sizeof(T));
                return v;
            }
        };
        struct Box
        {
            int lo[3], hi[3]; // inclusive
        };
        inline void fill(DenseVec3f &d, const Box &bb, const int lo[3], int size, const float v[3])
        {
            int a[3], b[3];
            for (int k = 0; k < 3; ++k)
            {
                a[k] = lo[k] > bb.lo[k] ? lo[k] : bb.lo[k];
                const long long hi = (long long)lo[k] + size - 1;
                b[k] = hi < bb.hi[k] ? (int)hi : bb.hi[k];
                if (a[k] > b[k])
                    return;
            }
            for (int z = a[2]; z <= b[2]; ++z)
                for (int y = a[1]; y <= b[1]; ++y)
                    for (int x = a[0]; x <= b[0]; ++x)
                    {
                        float *o = &d.data[((((size_t)(z - bb.lo[2]) * d.dims[1]) + (y - bb.lo[1])) * d.dims[0] + (x - bb.lo[0])) * 3];
                        o[0] = v[0], o[1] = v[1], o[2] = v[2];
                    }
        }
        inline bool mask_bit(const Buf &b, size_t mask_off, uint32_t n) { return (b.rd<uint32_t>(mask_off + 4u * (n >> 5)) >> (n & 31u)) & 1u; }
    } // namespace detail

    // -> "" on success, else the reason
    inline std::string read_vec3f(const std::string &path, DenseVec3f &out, size_t max_voxels = (size_t)1 << 28)
    {
        using namespace detail;
        std::vector<uint8_t> bytes;
        {
            FILE *f = fopen(path.c_str(), "rb");
            if (!f)
                return "cannot open " + path;
            fseek(f, 0, SEEK_END);
            const long sz = ftell(f);
            fseek(f, 0, SEEK_SET);
            bytes.resize(sz > 0 ? (size_t)sz : 0);
            const size_t got = bytes.empty() ? 0 : fread(bytes.data(), 1, bytes.size(), f);
            fclose(f);
            if (got != bytes.size() || bytes.size() < 16)
                return "cannot read " + path;
        }
        Buf b{bytes.data(), bytes.size()};
        size_t grid = 0;
        const uint64_t magic = b.rd<uint64_t>(0);
        if (magic == kMagicNumb || magic == kMagicFile)
        {
            // could still be a raw grid written with the old magic: a file header has gridCount / codec at 12 / 14
            const uint16_t grid_count = b.rd<uint16_t>(12), codec = b.rd<uint16_t>(14);
            if (grid_count == 0)
                return "nvdb: file holds no grid";
            if (codec > 1)
                return "nvdb: compressed file with the BLOSC codec is not supported (no blosc library), write with Codec::NONE or Codec::ZIP";
            size_t pos = 16;
            uint16_t first_codec = 0;
            uint64_t first_grid_size = 0;
            for (uint16_t g = 0; g < grid_count; ++g)
            {
                if (!b.ok(pos, 176))
                    return "nvdb: truncated metadata";
                const uint32_t name_size = b.rd<uint32_t>(pos + 136); // FileMetaData::nameSize
                const uint16_t gcodec = b.rd<uint16_t>(pos + 168);
                if (gcodec > 1)
                    return "nvdb: compressed grid with the BLOSC codec is not supported (no blosc library)";
                if (g == 0)
                    first_codec = gcodec, first_grid_size = b.rd<uint64_t>(pos); // FileMetaData::gridSize
                pos += 176 + name_size;
            }
            grid = pos; // first grid of the first segment
            if (first_codec == 1)
            {
                // Codec::ZIP (io/IO.h:279-292): uint64 size of the zlib stream, then the stream; inflates to gridSize bytes
                if (!b.ok(grid, 8))
                    return "nvdb: truncated compressed grid";
                const uint64_t csize = b.rd<uint64_t>(grid);
                if (!b.ok(grid + 8, (size_t)csize) || first_grid_size < 672 || first_grid_size > ((uint64_t)1 << 34))
                    return "nvdb: truncated compressed grid";
                std::vector<uint8_t> raw((size_t)first_grid_size);
                uLongf n = (uLongf)raw.size();
                if (uncompress(raw.data(), &n, bytes.data() + grid + 8, (uLong)csize) != Z_OK || (uint64_t)n != first_grid_size)
                    return "nvdb: compressed grid does not inflate to its recorded size";
                bytes.swap(raw);
                b = Buf{bytes.data(), bytes.size()};
                grid = 0;
            }
        }
        else if (magic != kMagicGrid)
            return "nvdb: not a NanoVDB file (bad magic)";
        if (!b.ok(grid, 672 + 64))
            return "nvdb: truncated grid";
        const uint64_t gmagic = b.rd<uint64_t>(grid);
        if (gmagic != kMagicNumb && gmagic != kMagicGrid)
            return "nvdb: bad grid magic";
        if (b.rd<uint32_t>(grid + 636) != kGridTypeVec3f)
            return "nvdb: the source grid must be of type Vec3f";
        const uint64_t grid_size = b.rd<uint64_t>(grid + 32);
        if (!b.ok(grid, (size_t)grid_size))
            return "nvdb: truncated grid data";
        // Map (float part): mat[9] at +0, invmat[9] at +36, vec[3] at +72 of the map at grid + 296
        float mat[9], vec[3];
        for (int i = 0; i < 9; ++i)
            mat[i] = b.rd<float>(grid + 296 + 4u * i);
        for (int i = 0; i < 3; ++i)
            vec[i] = b.rd<float>(grid + 296 + 72 + 4u * i);
        for (int i = 0; i < 9; ++i)
            if ((i % 4 != 0 && mat[i] != 0.0f) || (i % 4 == 0 && mat[i] != mat[0]))
                return "nvdb: only axis-aligned maps with one voxel size are supported";
        if (!(mat[0] > 0.0f))
            return "nvdb: bad voxel size";
        const size_t tree = grid + 672;
        const size_t root = tree + (size_t)b.rd<uint64_t>(tree + 24);
        if (!b.ok(root, kRootSize))
            return "nvdb: bad root offset";
        Box bb;
        for (int k = 0; k < 3; ++k)
        {
            bb.lo[k] = b.rd<int32_t>(root + 4u * k);
            bb.hi[k] = b.rd<int32_t>(root + 12 + 4u * k);
        }
        const uint32_t tiles = b.rd<uint32_t>(root + 24);
        float bg[3];
        for (int k = 0; k < 3; ++k)
            bg[k] = b.rd<float>(root + kRootOffBackground + 4u * k);
        if (bg[0] != 0.0f || bg[1] != 0.0f || bg[2] != 0.0f)
            return "nvdb: non-zero background values are not supported (the dense grid reads zero outside its box)";
        if (tiles == 0 || bb.lo[0] > bb.hi[0] || bb.lo[1] > bb.hi[1] || bb.lo[2] > bb.hi[2])
            return "nvdb: empty grid";
        size_t nvox = 1;
        for (int k = 0; k < 3; ++k)
        {
            out.dims[k] = bb.hi[k] - bb.lo[k] + 1;
            out.ijk_min[k] = bb.lo[k];
            nvox *= (size_t)out.dims[k];
            if (nvox > max_voxels)
                return "nvdb: index bounding box too large to flatten";
        }
        out.voxel = mat[0];
        for (int k = 0; k < 3; ++k)
            out.origin[k] = vec[k] + (float)bb.lo[k] * mat[0];
        out.data.assign(nvox * 3, 0.0f);
        for (uint32_t t = 0; t < tiles; ++t)
        {
            const size_t tile = root + kRootSize + (size_t)t * kRootTileSize;
            if (!b.ok(tile, kRootTileSize))
                return "nvdb: truncated root table";
            const uint64_t key = b.rd<uint64_t>(tile);
            const int64_t child = b.rd<int64_t>(tile + 8);
            // pnanovdb_coord_to_key: key = (z >> 12) | (y >> 12) << 21 | (x >> 12) << 42 on the unsigned coordinates
            int uo[3];
            uo[0] = (int32_t)((uint32_t)((key >> 42) & 0x1fffffu) << 12);
            uo[1] = (int32_t)((uint32_t)((key >> 21) & 0x1fffffu) << 12);
            uo[2] = (int32_t)((uint32_t)(key & 0x1fffffu) << 12);
            if (child == 0)
            {
                float v[3];
                for (int k = 0; k < 3; ++k)
                    v[k] = b.rd<float>(tile + kRootTileOffValue + 4u * k);
                fill(out, bb, uo, 4096, v);
                continue;
            }
            const size_t upper = root + (size_t)child;
            if (!b.ok(upper, kUpperOffTable + 32768u * kTableStride))
                return "nvdb: bad upper node offset";
            for (uint32_t n = 0; n < 32768u; ++n)
            {
                int lo[3] = {uo[0] + (int)(n >> 10) * 128, uo[1] + (int)((n >> 5) & 31u) * 128, uo[2] + (int)(n & 31u) * 128};
                if (lo[0] > bb.hi[0] || lo[1] > bb.hi[1] || lo[2] > bb.hi[2] || lo[0] + 127 < bb.lo[0] || lo[1] + 127 < bb.lo[1] || lo[2] + 127 < bb.lo[2])
                    continue;
                const size_t ue = upper + kUpperOffTable + (size_t)n * kTableStride;
                if (!mask_bit(b, upper + kUpperOffChildMask, n))
                {
                    float v[3];
                    for (int k = 0; k < 3; ++k)
                        v[k] = b.rd<float>(ue + 4u * k);
                    fill(out, bb, lo, 128, v);
                    continue;
                }
                const size_t lower = upper + (size_t)b.rd<int64_t>(ue);
                if (!b.ok(lower, kLowerOffTable + 4096u * kTableStride))
                    return "nvdb: bad lower node offset";
                for (uint32_t m = 0; m < 4096u; ++m)
                {
                    int ll[3] = {lo[0] + (int)(m >> 8) * 8, lo[1] + (int)((m >> 4) & 15u) * 8, lo[2] + (int)(m & 15u) * 8};
                    if (ll[0] > bb.hi[0] || ll[1] > bb.hi[1] || ll[2] > bb.hi[2] || ll[0] + 7 < bb.lo[0] || ll[1] + 7 < bb.lo[1] || ll[2] + 7 < bb.lo[2])
                        continue;
                    const size_t le = lower + kLowerOffTable + (size_t)m * kTableStride;
                    if (!mask_bit(b, lower + kLowerOffChildMask, m))
                    {
                        float v[3];
                        for (int k = 0; k < 3; ++k)
                            v[k] = b.rd<float>(le + 4u * k);
                        fill(out, bb, ll, 8, v);
                        continue;
                    }
                    const size_t leaf = lower + (size_t)b.rd<int64_t>(le);
                    if (!b.ok(leaf, kLeafOffTable + 512u * 12u))
                        return "nvdb: bad leaf offset";
                    for (uint32_t q = 0; q < 512u; ++q) // pnanovdb_leaf_coord_to_offset: x << 6 | y << 3 | z
                    {
                        const int c[3] = {ll[0] + (int)(q >> 6), ll[1] + (int)((q >> 3) & 7u), ll[2] + (int)(q & 7u)};
                        float v[3];
                        for (int k = 0; k < 3; ++k)
                            v[k] = b.rd<float>(leaf + kLeafOffTable + (size_t)q * 12u + 4u * k);
                        fill(out, bb, c, 1, v);
                    }
                }
            }
        }
        return "";
    }
} // namespace wnvdb
