// PngWriter.h -- RGBA8 PNG output for the headless render state.  Stands in for the reference's
// lodepng wrapper MainState::encodePNG (RayTrace/states/MainState.cpp:410-417, lodepng.h:236-241):
// same argument meaning (filename, RGBA8 bytes, width, height), returns 0 on success.
#ifndef RTB_HOST_PNGWRITER_H_
#define RTB_HOST_PNGWRITER_H_

#include <vector>

namespace rtbh {

// 8-bit RGBA PNG: per-scanline adaptive filter (minimum sum of absolute values), zlib stream = one deflate block with
// the fixed Huffman codes and LZ77 matches from hash chains (32K window).
unsigned encodePNG(const char* filename, const unsigned char* rgba, unsigned width, unsigned height);
inline unsigned encodePNG(const char* filename, const std::vector<unsigned char>& imageData, unsigned width,
                          unsigned height) {
    if (imageData.size() < (size_t)width * height * 4) return 1;
    return encodePNG(filename, imageData.data(), width, height);
}

}  // namespace rtbh

#endif
