/* codecs.h — the three decompressors the on-disk formats either side of the path need, written
 * from their published specifications (no zlib / lz4 / blosc development files in this image):
 *   inflate     RFC 1950/1951 (zlib stream, DEFLATE)   — PNG textures (lodepng in the reference,
 *               ext/lodepng) and OpenVDB's COMPRESS_ZIP
 *   lz4 block   LZ4 block format                        — inside Blosc
 *   blosc 1     c-blosc 1.x frame (16-byte header, block starts, optional split + byte shuffle)
 *               — OpenVDB's COMPRESS_BLOSC, what Blender writes (scenes/project/volume/*.vdb)
 * All throw NoriException on malformed input and never read or write out of bounds. */
#pragma once
#include "nori.h"

namespace nori {

/* zlib stream -> bytes. `expected` (if non-zero) reserves the output. */
std::vector<uint8_t> zlibInflate(const uint8_t *src, size_t n, size_t expected = 0);

/* one LZ4 block of `n` bytes into exactly `dst_len` bytes */
void lz4DecodeBlock(const uint8_t *src, size_t n, uint8_t *dst, size_t dst_len);

/* a Blosc-1 frame; returns the decompressed payload (header.nbytes bytes) */
std::vector<uint8_t> bloscDecode(const uint8_t *src, size_t n);

}
