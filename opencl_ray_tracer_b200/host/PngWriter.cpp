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

// ---- deflate (RFC 1951): one block with the fixed Huffman codes, LZ77 matches from hash chains ----------------
struct BitWriter {
    std::vector<unsigned char>& out;
    uint64_t acc = 0;
    int n = 0;
    explicit BitWriter(std::vector<unsigned char>& o) : out(o) {}
    void put(uint32_t bits, int count) {  // LSB first
        acc |= (uint64_t)bits << n;
        n += count;
        while (n >= 8) { out.push_back((unsigned char)(acc & 0xFF)); acc >>= 8; n -= 8; }
    }
    void put_code(uint32_t code, int len) {  // Huffman codes go in MSB first
        uint32_t r = 0;
        for (int i = 0; i < len; ++i) r |= ((code >> i) & 1u) << (len - 1 - i);
        put(r, len);
    }
    void flush() { if (n > 0) { out.push_back((unsigned char)(acc & 0xFF)); acc = 0; n = 0; } }
};

void put_symbol(BitWriter& bw, int sym) {  // fixed literal/length code, RFC 1951 3.2.6
    if (sym < 144) bw.put_code(0x30 + sym, 8);
    else if (sym < 256) bw.put_code(0x190 + (sym - 144), 9);
    else if (sym < 280) bw.put_code(sym - 256, 7);
    else bw.put_code(0xC0 + (sym - 280), 8);
}

const unsigned short kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const unsigned char kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const unsigned short kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const unsigned char kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

void put_match(BitWriter& bw, int len, int dist) {
    int lc = 28;
    while (kLenBase[lc] > len) --lc;
    put_symbol(bw, 257 + lc);
    if (kLenExtra[lc]) bw.put((uint32_t)(len - kLenBase[lc]), kLenExtra[lc]);
    int dc = 29;
    while (kDistBase[dc] > dist) --dc;
    bw.put_code((uint32_t)dc, 5);
    if (kDistExtra[dc]) bw.put((uint32_t)(dist - kDistBase[dc]), kDistExtra[dc]);
}

void deflate_fixed(const unsigned char* src, size_t n, std::vector<unsigned char>& out) {
    constexpr int kHashBits = 15, kWindow = 32768, kMaxLen = 258, kMinLen = 3, kMaxChain = 32, kGoodLen = 64;
    std::vector<int32_t> head((size_t)1 << kHashBits, -1), prev(kWindow, -1);
    BitWriter bw(out);
    bw.put(1, 1);  // BFINAL
    bw.put(1, 2);  // BTYPE = 01: fixed Huffman codes
    auto hash = [&](size_t i) {
        const uint32_t v = (uint32_t)src[i] | ((uint32_t)src[i + 1] << 8) | ((uint32_t)src[i + 2] << 16);
        return (v * 2654435761u) >> (32 - kHashBits);
    };
    auto insert = [&](size_t i) {
        const uint32_t hh = hash(i);
        prev[i & (kWindow - 1)] = head[hh];
        head[hh] = (int32_t)i;
    };
    size_t i = 0;
    while (i < n) {
        int best_len = 0, best_dist = 0;
        if (i + kMinLen <= n) {
            const int max_len = (n - i < (size_t)kMaxLen) ? (int)(n - i) : kMaxLen;
            int32_t c = head[hash(i)];
            for (int chain = 0; c >= 0 && (size_t)c + kWindow > i && chain < kMaxChain; ++chain) {
                const unsigned char* a = src + c;
                const unsigned char* b = src + i;
                if (a[best_len] == b[best_len]) {
                    int l = 0;
                    while (l < max_len && a[l] == b[l]) ++l;
                    if (l > best_len) { best_len = l; best_dist = (int)(i - (size_t)c); if (l >= kGoodLen || l == max_len) break; }
                }
                const int32_t nx = prev[(size_t)c & (kWindow - 1)];
                if (nx >= c) break;  // the slot was overwritten by a newer position
                c = nx;
            }
        }
        if (best_len >= kMinLen) {
            put_match(bw, best_len, best_dist);
            const size_t stop = i + (size_t)best_len;
            for (; i < stop; ++i)
                if (i + kMinLen <= n) insert(i);
        } else {
            put_symbol(bw, src[i]);
            if (i + kMinLen <= n) insert(i);
            ++i;
        }
    }
    put_symbol(bw, 256);  // end of block
    bw.flush();
}

int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

}  // namespace

unsigned encodePNG(const char* filename, const unsigned char* rgba, unsigned width, unsigned height) {
    if (!filename || !rgba || width == 0 || height == 0) return 1;
    if (!g_crc_ready) crc_init();
    // scanlines: filter byte + 4*width bytes; per line the filter (None / Sub / Up / Average / Paeth) with the smallest
    // sum of absolute filtered values (the PNG specification's heuristic, lodepng's default strategy as well)
    const size_t row = 4 * (size_t)width;
    const size_t stride = 1 + row;
    const size_t raw_n = stride * height;
    std::vector<unsigned char> raw(raw_n);
    std::vector<unsigned char> cand(row), zero(row, 0);
    for (unsigned y = 0; y < height; ++y) {
        const unsigned char* cur = rgba + row * y;
        const unsigned char* up = y ? rgba + row * (y - 1) : zero.data();
        unsigned long best = ~0ul;
        for (int f = 0; f < 5; ++f) {
            unsigned long sum = 0;
            for (size_t x = 0; x < row; ++x) {
                const int a = x >= 4 ? cur[x - 4] : 0, b = up[x], c = x >= 4 ? up[x - 4] : 0;
                const int pred = f == 0 ? 0 : f == 1 ? a : f == 2 ? b : f == 3 ? ((a + b) >> 1) : paeth(a, b, c);
                const unsigned char v = (unsigned char)(cur[x] - pred);
                cand[x] = v;
                sum += v < 128 ? v : 256 - v;
            }
            if (sum < best) {
                best = sum;
                raw[y * stride] = (unsigned char)f;
                std::memcpy(&raw[y * stride + 1], cand.data(), row);
            }
        }
    }
    uint32_t a = 1, b = 0;                  // adler32 of the raw stream
    for (size_t i = 0; i < raw_n;) {
        const size_t blk = (raw_n - i < 5552) ? raw_n - i : 5552;  // keeps a, b below 2^32
        for (size_t k = 0; k < blk; ++k) { a += raw[i + k]; b += a; }
        a %= 65521u; b %= 65521u;
        i += blk;
    }
    std::vector<unsigned char> z;
    z.reserve(raw_n / 8 + 64);
    z.push_back(0x78); z.push_back(0x01);  // zlib header: deflate, 32K window, no preset dictionary
    deflate_fixed(raw.data(), raw_n, z);
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
