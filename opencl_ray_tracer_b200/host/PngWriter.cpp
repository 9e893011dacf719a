#include "PngWriter.h"

#include <cstdint>
#include <cstdio>
#include <cstring>

namespace rtbh {

namespace {

uint32_t g_crc_table[256];
bool g_crc_ready = false;

void crc_init() {
    for (uint32_t n = 0; n < 256; ++n) {
        uint32_t c = n;
        for (int k = 0; k < 8; ++k) c = (c & 1u) ? (0xEDB88320u ^ (c >> 1)) : (c >> 1);
        g_crc_table[n] = c;
    }
    g_crc_ready = true;
}

uint32_t crc_update(uint32_t crc, const unsigned char* p, size_t n) {
    for (size_t i = 0; i < n; ++i) crc = g_crc_table[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return crc;
}

void put32(unsigned char* p, uint32_t v) {
    p[0] = (unsigned char)(v >> 24); p[1] = (unsigned char)(v >> 16);
    p[2] = (unsigned char)(v >> 8);  p[3] = (unsigned char)v;
}

bool write_chunk(FILE* f, const char type[4], const unsigned char* data, size_t n) {
    unsigned char hdr[8];
    put32(hdr, (uint32_t)n);
    std::memcpy(hdr + 4, type, 4);
    uint32_t crc = crc_update(0xFFFFFFFFu, hdr + 4, 4);
    if (n) crc = crc_update(crc, data, n);
    unsigned char tail[4];
    put32(tail, crc ^ 0xFFFFFFFFu);
    return std::fwrite(hdr, 1, 8, f) == 8 && (n == 0 || std::fwrite(data, 1, n, f) == n) &&
           std::fwrite(tail, 1, 4, f) == 4;
}

}  // namespace

unsigned encodePNG(const char* filename, const unsigned char* rgba, unsigned width, unsigned height) {
    if (!filename || !rgba || width == 0 || height == 0) return 1;
    if (!g_crc_ready) crc_init();
    // raw scanlines: filter byte 0 + 4*width bytes
    const size_t stride = 1 + 4 * (size_t)width;
    const size_t raw_n = stride * height;
    std::vector<unsigned char> z;
    z.reserve(raw_n + raw_n / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);  // zlib header, no preset dictionary
    uint32_t a = 1, b = 0;                  // adler32 of the raw stream
    std::vector<unsigned char> raw(raw_n);
    for (unsigned y = 0; y < height; ++y) {
        raw[y * stride] = 0;
        std::memcpy(&raw[y * stride + 1], rgba + 4 * (size_t)width * y, 4 * (size_t)width);
    }
    for (size_t i = 0; i < raw_n;) {
        const size_t blk = (raw_n - i < 5552) ? raw_n - i : 5552;  // keeps a, b below 2^32
        for (size_t k = 0; k < blk; ++k) { a += raw[i + k]; b += a; }
        a %= 65521u; b %= 65521u;
        i += blk;
    }
    for (size_t i = 0; i < raw_n;) {
        const size_t blk = (raw_n - i < 65535) ? raw_n - i : 65535;
        const bool last = (i + blk == raw_n);
        z.push_back(last ? 1 : 0);  // BFINAL + BTYPE=00 (stored)
        z.push_back((unsigned char)(blk & 0xFF)); z.push_back((unsigned char)(blk >> 8));
        z.push_back((unsigned char)(~blk & 0xFF)); z.push_back((unsigned char)((~blk >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + i, raw.begin() + i + blk);
        i += blk;
    }
    unsigned char ad[4];
    put32(ad, (b << 16) | a);
    z.insert(z.end(), ad, ad + 4);

    FILE* f = std::fopen(filename, "wb");
    if (!f) return 2;
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    unsigned char ihdr[13];
    put32(ihdr, width); put32(ihdr + 4, height);
    ihdr[8] = 8; ihdr[9] = 6; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;  // 8-bit RGBA
    bool ok = std::fwrite(sig, 1, 8, f) == 8 && write_chunk(f, "IHDR", ihdr, 13) &&
              write_chunk(f, "IDAT", z.data(), z.size()) && write_chunk(f, "IEND", nullptr, 0);
    ok = (std::fclose(f) == 0) && ok;
    return ok ? 0 : 3;
}

}  // namespace rtbh
