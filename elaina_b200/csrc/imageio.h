// Minimal image writers/readers for the film export surface.
//
// The reference saves every exported film twice, as <name>.exr and <name>.png, both flipped vertically
// (integrator/common.h:165-240 -> util/film.h:94-144 -> core/texture.cu:82-115 -> util/image.cpp:30-104):
//   EXR: channels A,B,G,R stored as HALF, scan lines in increasing y;   PNG: 8-bit RGBA, v -> clamp(int(v * 255), 0, 255).
// It uses tinyexr and stb for that; these are independent single-purpose encoders: an uncompressed scan-line
// OpenEXR 2.0 file and a PNG whose zlib stream consists of stored (uncompressed) deflate blocks. Any EXR/PNG reader
// accepts both. Also a binary PGM/PPM (P5/P6) reader for mask images (the reference reads masks through stb).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace wimg
{
    inline uint16_t float_to_half(float f)
    {
        uint32_t x;
        memcpy(&x, &f, 4);
        const uint32_t sign = (x >> 16) & 0x8000u;
        const int32_t exp = (int32_t)((x >> 23) & 0xff) - 127 + 15;
        uint32_t man = x & 0x7fffffu;
        if (((x >> 23) & 0xff) == 0xff)
            return (uint16_t)(sign | 0x7c00u | (man ? 0x200u : 0u)); // inf / nan
        if (exp >= 31)
            return (uint16_t)(sign | 0x7c00u); // overflow -> inf
        if (exp <= 0)
        {
            if (exp < -10)
                return (uint16_t)sign; // underflow -> 0
            man |= 0x800000u;
            const int shift = 14 - exp;
            uint32_t h = man >> shift;
            const uint32_t rem = man & ((1u << shift) - 1u), half = 1u << (shift - 1);
            if (rem > half || (rem == half && (h & 1u)))
                ++h;
            return (uint16_t)(sign | h);
        }
        uint32_t h = ((uint32_t)exp << 10) | (man >> 13);
        const uint32_t rem = man & 0x1fffu;
        if (rem > 0x1000u || (rem == 0x1000u && (h & 1u)))
            ++h; // may carry into the exponent: still correct (rounds up to the next binade / inf)
        return (uint16_t)(sign | h);
    }

    // ---- OpenEXR, single part, scan line, NO_COMPRESSION, 4 HALF channels. rgba: w*h*4 floats, row 0 first.
    inline bool write_exr(const std::string &path, int w, int h, const float *rgba, bool flip)
    {
        std::vector<uint8_t> out;
        auto put = [&](const void *p, size_t n)
        { out.insert(out.end(), (const uint8_t *)p, (const uint8_t *)p + n); };
        auto put_i32 = [&](int32_t v)
        { put(&v, 4); };
        auto put_str = [&](const char *s)
        { put(s, strlen(s) + 1); };
        auto attr = [&](const char *name, const char *type, const void *data, int32_t size)
        {
            put_str(name);
            put_str(type);
            put_i32(size);
            put(data, (size_t)size);
        };
        const uint32_t magic = 20000630u, version = 2u;
        put(&magic, 4);
        put(&version, 4);
        {
            std::vector<uint8_t> ch;
            const char *names[4] = {"A", "B", "G", "R"}; // alphabetical, as the format requires
            for (int c = 0; c < 4; ++c)
            {
                ch.insert(ch.end(), names[c], names[c] + 2);
                const int32_t pixel_type = 1; // HALF
                const uint8_t plinear[4] = {0, 0, 0, 0};
                const int32_t xs = 1, ys = 1;
                ch.insert(ch.end(), (const uint8_t *)&pixel_type, (const uint8_t *)&pixel_type + 4);
                ch.insert(ch.end(), plinear, plinear + 4);
                ch.insert(ch.end(), (const uint8_t *)&xs, (const uint8_t *)&xs + 4);
                ch.insert(ch.end(), (const uint8_t *)&ys, (const uint8_t *)&ys + 4);
            }
            ch.push_back(0);
            attr("channels", "chlist", ch.data(), (int32_t)ch.size());
        }
        const uint8_t compression = 0;
        attr("compression", "compression", &compression, 1);
        const int32_t window[4] = {0, 0, w - 1, h - 1};
        attr("dataWindow", "box2i", window, 16);
        attr("displayWindow", "box2i", window, 16);
        const uint8_t line_order = 0;
        attr("lineOrder", "lineOrder", &line_order, 1);
        const float aspect = 1.0f, center[2] = {0.0f, 0.0f}, width = 1.0f;
        attr("pixelAspectRatio", "float", &aspect, 4);
        attr("screenWindowCenter", "v2f", center, 8);
        attr("screenWindowWidth", "float", &width, 4);
        out.push_back(0); // end of header
        const size_t table = out.size();
        const size_t line_bytes = (size_t)w * 4 * 2, chunk = 8 + line_bytes;
        out.resize(table + (size_t)h * 8);
        for (int y = 0; y < h; ++y)
        {
            const uint64_t off = (uint64_t)(table + (size_t)h * 8 + (size_t)y * chunk);
            memcpy(&out[table + (size_t)y * 8], &off, 8);
        }
        std::vector<uint16_t> line((size_t)w * 4);
        const int order[4] = {3, 2, 1, 0}; // A B G R from r g b a
        for (int y = 0; y < h; ++y)
        {
            const float *src = rgba + (size_t)(flip ? h - 1 - y : y) * w * 4;
            for (int c = 0; c < 4; ++c)
                for (int x = 0; x < w; ++x)
                    line[(size_t)c * w + x] = float_to_half(src[(size_t)x * 4 + order[c]]);
            put_i32(y);
            put_i32((int32_t)line_bytes);
            put(line.data(), line_bytes);
        }
        FILE *f = fopen(path.c_str(), "wb");
        if (!f)
            return false;
        const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
        fclose(f);
        return ok;
    }

    // ---- PNG, 8-bit RGBA, zlib stream of stored blocks
    inline uint32_t crc32_update(uint32_t crc, const uint8_t *p, size_t n)
    {
        static uint32_t table[256];
        static bool init = false;
        if (!init)
        {
            for (uint32_t i = 0; i < 256; ++i)
            {
                uint32_t c = i;
                for (int k = 0; k < 8; ++k)
                    c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
                table[i] = c;
            }
            init = true;
        }
        crc = ~crc;
        for (size_t i = 0; i < n; ++i)
            crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
        return ~crc;
    }
    inline bool write_png(const std::string &path, int w, int h, const uint8_t *rgba, bool flip)
    {
        std::vector<uint8_t> raw;
        raw.reserve((size_t)h * ((size_t)w * 4 + 1));
        for (int y = 0; y < h; ++y)
        {
            raw.push_back(0); // filter: none
            const uint8_t *src = rgba + (size_t)(flip ? h - 1 - y : y) * w * 4;
            raw.insert(raw.end(), src, src + (size_t)w * 4);
        }
        std::vector<uint8_t> z;
        z.push_back(0x78);
        z.push_back(0x01);
        uint32_t a = 1, b = 0;
        for (size_t pos = 0; pos < raw.size() || pos == 0;)
        {
            const size_t n = std::min<size_t>(65535, raw.size() - pos);
            z.push_back(pos + n >= raw.size() ? 1 : 0);
            z.push_back((uint8_t)(n & 0xff));
            z.push_back((uint8_t)(n >> 8));
            z.push_back((uint8_t)(~n & 0xff));
            z.push_back((uint8_t)((~n >> 8) & 0xff));
            z.insert(z.end(), raw.begin() + (long)pos, raw.begin() + (long)(pos + n));
            for (size_t i = pos; i < pos + n; ++i)
            {
                a = (a + raw[i]) % 65521u;
                b = (b + a) % 65521u;
            }
            pos += n;
            if (n == 0)
                break;
        }
        const uint32_t adler = (b << 16) | a;
        for (int k = 3; k >= 0; --k)
            z.push_back((uint8_t)(adler >> (8 * k)));
        std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
        auto be32 = [](std::vector<uint8_t> &v, uint32_t x)
        {
            for (int k = 3; k >= 0; --k)
                v.push_back((uint8_t)(x >> (8 * k)));
        };
        auto chunk = [&](const char *type, const std::vector<uint8_t> &data)
        {
            be32(out, (uint32_t)data.size());
            std::vector<uint8_t> body(type, type + 4);
            body.insert(body.end(), data.begin(), data.end());
            out.insert(out.end(), body.begin(), body.end());
            be32(out, crc32_update(0, body.data(), body.size()));
        };
        std::vector<uint8_t> ihdr;
        be32(ihdr, (uint32_t)w);
        be32(ihdr, (uint32_t)h);
        const uint8_t tail[5] = {8, 6, 0, 0, 0}; // 8 bit, RGBA, deflate, adaptive filter set, no interlace
        ihdr.insert(ihdr.end(), tail, tail + 5);
        chunk("IHDR", ihdr);
        chunk("IDAT", z);
        chunk("IEND", {});
        FILE *f = fopen(path.c_str(), "wb");
        if (!f)
            return false;
        const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
        fclose(f);
        return ok;
    }

    // ---- binary PGM (P5) / PPM (P6), maxval <= 255: mask images. out: w*h bytes, nonzero = any channel nonzero
    inline bool read_pnm_mask(const std::string &path, int &w, int &h, std::vector<uint8_t> &out)
    {
        FILE *f = fopen(path.c_str(), "rb");
        if (!f)
            return false;
        auto token = [&](char *buf, int cap) -> bool
        {
            int c = fgetc(f);
            for (;;)
            {
                while (c == ' ' || c == '\n' || c == '\r' || c == '\t')
                    c = fgetc(f);
                if (c == '#')
                {
                    while (c != '\n' && c != EOF)
                        c = fgetc(f);
                    continue;
                }
                break;
            }
            int n = 0;
            while (c != EOF && c != ' ' && c != '\n' && c != '\r' && c != '\t' && n < cap - 1)
            {
                buf[n++] = (char)c;
                c = fgetc(f);
            }
            buf[n] = 0;
            return n > 0;
        };
        char t[32];
        bool ok = token(t, 32);
        const int ch = (ok && !strcmp(t, "P5")) ? 1 : ((ok && !strcmp(t, "P6")) ? 3 : 0);
        int maxv = 0;
        ok = ch && token(t, 32) && (w = atoi(t)) > 0 && token(t, 32) && (h = atoi(t)) > 0 && token(t, 32) && (maxv = atoi(t)) > 0 && maxv <= 255;
        if (ok)
        {
            std::vector<uint8_t> px((size_t)w * h * ch);
            ok = fread(px.data(), 1, px.size(), f) == px.size();
            out.assign((size_t)w * h, 0);
            for (size_t i = 0; ok && i < (size_t)w * h; ++i)
                for (int c = 0; c < ch; ++c)
                    if (px[i * ch + c])
                        out[i] = 1;
        }
        fclose(f);
        return ok;
    }
} // namespace wimg
