/* png.h — PNG decoder (the role of ext/lodepng in src/textures/PNGTexture.cpp:76-77): any
 * non-interlaced or Adam7 PNG -> 8-bit RGBA, the conversion lodepng::decode applies by default. */
#pragma once
#include "nori.h"

namespace nori {
/* rgba: width*height*4 bytes, row 0 first */
void decodePng(const std::string &path, std::vector<uint8_t> &rgba, unsigned &width, unsigned &height);
/* Bitmap::saveToLDR (src/utils/bitmap.cpp:109-140): linear RGB floats (h x w x 3) -> 8-bit sRGB PNG */
void writePng(const std::string &path, const float *data, size_t h, size_t w);
}
