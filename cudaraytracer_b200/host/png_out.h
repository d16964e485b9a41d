// png_out.h — PNG output of the render path.
//
// The reference writes frames with stb_image_write v1.16
// (CudaTest/stb-master/stb_image_write.h:176,1215, called from
// core/render.h:135-157). stb is third-party code that does not travel with
// this repo, so the encoder here is written from the PNG (RFC 2083) and zlib /
// DEFLATE (RFC 1950 / 1951) specifications: filter type 0 on every row, one
// zlib stream of fixed-Huffman blocks with a greedy hash-chain-free LZ77
// (one candidate per 3-byte hash), CRC-32 per chunk, Adler-32 trailer.
// stbi_write_png keeps stb's signature so reference-style code compiles
// unchanged; the decoded pixels are what counts (any PNG reader gives back
// exactly the bytes passed in), the compressed bytes differ from stb's.
//
// WritePng reproduces core/render.h:135-157: vertical flip (row 0 of the
// colorBuffer is the bottom row), char(255.99 * c) per channel, alpha 255,
// 4 bytes per pixel, file "<prefix><frameIndex>.png".
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "vec3.h"

namespace crt_png {

inline uint32_t crc32(const uint8_t* p, size_t n, uint32_t crc = 0)
{
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        ready = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return ~crc;
}

inline uint32_t adler32(const uint8_t* p, size_t n)
{
    uint32_t a = 1, b = 0;
    while (n > 0) {
        const size_t run = n < 5552 ? n : 5552;                       // largest run that cannot overflow 32 bits
        for (size_t i = 0; i < run; i++) { a += p[i]; b += a; }
        a %= 65521u; b %= 65521u;
        p += run; n -= run;
    }
    return (b << 16) | a;
}

// LSB-first bit writer of a DEFLATE stream.
struct BitSink {
    std::vector<uint8_t> out;
    uint32_t acc = 0;
    int nbits = 0;
    void put(uint32_t value, int count)                               // value's low `count` bits, LSB first
    {
        acc |= value << nbits;
        nbits += count;
        while (nbits >= 8) { out.push_back((uint8_t)acc); acc >>= 8; nbits -= 8; }
    }
    void put_huffman(uint32_t code, int len)                          // Huffman codes go MSB first
    {
        uint32_t r = 0;
        for (int i = 0; i < len; i++) r |= ((code >> i) & 1u) << (len - 1 - i);
        put(r, len);
    }
    void flush() { if (nbits > 0) { out.push_back((uint8_t)acc); acc = 0; nbits = 0; } }
};

// Fixed Huffman code of RFC 1951 section 3.2.6 for literal/length symbol s.
inline void put_litlen(BitSink& b, int s)
{
    if (s < 144) b.put_huffman(0x30 + s, 8);
    else if (s < 256) b.put_huffman(0x190 + (s - 144), 9);
    else if (s < 280) b.put_huffman(s - 256, 7);
    else b.put_huffman(0xC0 + (s - 280), 8);
}

inline void put_length(BitSink& b, int len)                           // 3..258
{
    static const int base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const int extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    int i = 28;
    while (base[i] > len) i--;
    put_litlen(b, 257 + i);
    if (extra[i]) b.put((uint32_t)(len - base[i]), extra[i]);
}

inline void put_distance(BitSink& b, int dist)                        // 1..32768
{
    static const int base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static const int extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    int i = 29;
    while (base[i] > dist) i--;
    b.put_huffman((uint32_t)i, 5);
    if (extra[i]) b.put((uint32_t)(dist - base[i]), extra[i]);
}

// zlib stream (RFC 1950) of `n` bytes: one final fixed-Huffman block.
inline std::vector<uint8_t> zlib_compress(const uint8_t* data, size_t n)
{
    BitSink b;
    b.out.reserve(n / 2 + 64);
    b.out.push_back(0x78);                                            // CM = 8 (deflate), CINFO = 7 (32 KiB window)
    b.out.push_back(0x5E);                                            // FLEVEL 1, no dictionary; (0x78 << 8 | 0x5E) % 31 == 0
    b.put(1, 1);                                                      // BFINAL
    b.put(1, 2);                                                      // BTYPE = 01, fixed Huffman
    const int HASH_BITS = 15;
    std::vector<int64_t> head((size_t)1 << HASH_BITS, -1);
    size_t i = 0;
    while (i < n) {
        int best_len = 0;
        size_t best_pos = 0;
        if (i + 3 <= n) {
            const uint32_t h = ((uint32_t)data[i] | ((uint32_t)data[i + 1] << 8) | ((uint32_t)data[i + 2] << 16)) * 2654435761u >> (32 - HASH_BITS);
            const int64_t cand = head[h];
            head[h] = (int64_t)i;
            if (cand >= 0 && i - (size_t)cand <= 32768) {
                const size_t limit = n - i < 258 ? n - i : 258;
                size_t l = 0;
                while (l < limit && data[(size_t)cand + l] == data[i + l]) l++;
                if (l >= 3) { best_len = (int)l; best_pos = (size_t)cand; }
            }
        }
        if (best_len >= 3) {
            put_length(b, best_len);
            put_distance(b, (int)(i - best_pos));
            i += (size_t)best_len;
        } else {
            put_litlen(b, data[i]);
            i++;
        }
    }
    put_litlen(b, 256);                                               // end of block
    b.flush();
    const uint32_t ad = adler32(data, n);
    for (int s = 24; s >= 0; s -= 8) b.out.push_back((uint8_t)(ad >> s));
    return b.out;
}

inline void put_be32(std::vector<uint8_t>& v, uint32_t x)
{
    for (int s = 24; s >= 0; s -= 8) v.push_back((uint8_t)(x >> s));
}

inline void put_chunk(std::vector<uint8_t>& png, const char type[4], const uint8_t* body, size_t n)
{
    put_be32(png, (uint32_t)n);
    const size_t start = png.size();
    png.insert(png.end(), type, type + 4);
    if (n) png.insert(png.end(), body, body + n);
    put_be32(png, crc32(png.data() + start, n + 4));
}

// Encodes w x h pixels of `comp` 8-bit channels (1 = grey, 2 = grey + alpha,
// 3 = RGB, 4 = RGBA), rows `stride` bytes apart (0 = tightly packed).
inline std::vector<uint8_t> encode(int w, int h, int comp, const void* data, int stride)
{
    static const uint8_t colour_type[5] = {0, 0, 4, 2, 6};
    std::vector<uint8_t> png;
    if (w <= 0 || h <= 0 || comp < 1 || comp > 4 || !data) return png;
    const size_t row = (size_t)w * (size_t)comp;
    if (stride == 0) stride = (int)row;
    std::vector<uint8_t> raw((row + 1) * (size_t)h);
    for (int y = 0; y < h; y++) {
        raw[(row + 1) * (size_t)y] = 0;                                // filter type 0 (None)
        std::memcpy(&raw[(row + 1) * (size_t)y + 1], (const uint8_t*)data + (size_t)stride * (size_t)y, row);
    }
    const std::vector<uint8_t> z = zlib_compress(raw.data(), raw.size());
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    png.insert(png.end(), sig, sig + 8);
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, (uint32_t)w);
    put_be32(ihdr, (uint32_t)h);
    ihdr.push_back(8);                                                // bit depth
    ihdr.push_back(colour_type[comp]);
    ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);          // deflate, adaptive filtering, no interlace
    put_chunk(png, "IHDR", ihdr.data(), ihdr.size());
    put_chunk(png, "IDAT", z.data(), z.size());
    put_chunk(png, "IEND", nullptr, 0);
    return png;
}

}  // namespace crt_png

// Same signature and return convention (non-zero on success) as
// stb_image_write.h:176.
inline int stbi_write_png(char const* filename, int w, int h, int comp, const void* data, int stride_in_bytes)
{
    const std::vector<uint8_t> png = crt_png::encode(w, h, comp, data, stride_in_bytes);
    if (png.empty()) return 0;
    FILE* f = std::fopen(filename, "wb");
    if (!f) return 0;
    const size_t put = std::fwrite(png.data(), 1, png.size(), f);
    const int closed = std::fclose(f);
    return put == png.size() && closed == 0;
}

struct RGBColor {                                                     // core/render.h:33-39
    unsigned char r, g, b, a;
    RGBColor() {}
    RGBColor(unsigned char r_, unsigned char g_, unsigned char b_, unsigned char a_) : r(r_), g(g_), b(b_), a(a_) {}
};

// core/render.h:137-146 on the host (crt_pack_rgba8_device is the same
// conversion on the GPU).
inline void colorBufferToRGBA(int nx, int ny, const vec3* colorBuffer, RGBColor* rgb)
{
    for (int i = 0; i < ny; i++)
        for (int j = 0; j < nx; j++) {
            const size_t pixel_index = (size_t)(ny - 1 - i) * nx + j;
            RGBColor& o = rgb[(size_t)i * nx + j];
            o.r = (unsigned char)(int)(255.99 * colorBuffer[pixel_index].r());
            o.g = (unsigned char)(int)(255.99 * colorBuffer[pixel_index].g());
            o.b = (unsigned char)(int)(255.99 * colorBuffer[pixel_index].b());
            o.a = 255;
        }
}

// core/render.h:135-157. The reference hard-codes the prefix
// "images/moveTest/picture_"; it is the default here.
inline bool WritePng(int nx, int ny, int frameIndex, const vec3* colorBuffer,
                     const char* prefix = "images/moveTest/picture_")
{
    std::vector<RGBColor> rgb((size_t)nx * ny);
    colorBufferToRGBA(nx, ny, colorBuffer, rgb.data());
    const std::string path = std::string(prefix) + std::to_string(frameIndex) + ".png";
    return stbi_write_png(path.c_str(), nx, ny, (int)sizeof(RGBColor), rgb.data(), 0) != 0;
}
