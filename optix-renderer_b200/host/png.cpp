/* png.cpp — PNG (ISO/IEC 15948) decoder written from the specification: chunk walk, zlib
 * inflate (codecs.cpp), the five scanline filters, Adam7, all colour types and bit depths,
 * PLTE/tRNS, converted to RGBA8 the way lodepng's default decode does (16-bit samples keep their
 * high byte, grey samples below 8 bits are scaled to 0..255). CRCs are not verified. */
#include "png.h"
#include "codecs.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>

namespace nori {
namespace {

uint32_t be32(const uint8_t *p) { return ((uint32_t) p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }

int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

/* undo the filters of one (sub)image in place; `data` holds h rows of 1 + stride bytes */
void unfilter(uint8_t *data, size_t h, size_t stride, size_t bpp, std::vector<uint8_t> &out) {
    out.assign(h * stride, 0);
    for (size_t y = 0; y < h; ++y) {
        const uint8_t type = data[y * (stride + 1)];
        const uint8_t *in = data + y * (stride + 1) + 1;
        uint8_t *cur = out.data() + y * stride;
        const uint8_t *up = y ? cur - stride : nullptr;
        for (size_t x = 0; x < stride; ++x) {
            const int a = x >= bpp ? cur[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0;
            int v = in[x];
            switch (type) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: throw NoriException("png: unknown filter type %d", (int) type);
            }
            cur[x] = (uint8_t) v;
        }
    }
}

struct Header { unsigned w, h; int depth, color, interlace; };

int channels(int color) { return color == 0 ? 1 : color == 2 ? 3 : color == 3 ? 1 : color == 4 ? 2 : 4; }

/* one unfiltered pass (pw x ph samples) -> RGBA8 into the full image at (x0 + x*dx, y0 + y*dy) */
void emit(const Header &hd, const std::vector<uint8_t> &raw, size_t pw, size_t ph, size_t stride, const std::vector<uint8_t> &plte,
          const std::vector<uint8_t> &trns, unsigned x0, unsigned y0, unsigned dx, unsigned dy, std::vector<uint8_t> &rgba) {
    const int ch = channels(hd.color), depth = hd.depth;
    for (size_t y = 0; y < ph; ++y) {
        const uint8_t *row = raw.data() + y * stride;
        for (size_t x = 0; x < pw; ++x) {
            unsigned s[4] = { 0, 0, 0, 0 };                   /* raw samples at the file's bit depth */
            for (int c = 0; c < ch; ++c) {
                const size_t i = x * ch + c;
                if (depth == 8) s[c] = row[i];
                else if (depth == 16) s[c] = ((unsigned) row[2 * i] << 8) | row[2 * i + 1];
                else { const size_t bit = i * depth; s[c] = (row[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1u); }
            }
            uint8_t *o = rgba.data() + (((size_t) y0 + y * dy) * hd.w + x0 + x * dx) * 4;
            auto to8 = [&](unsigned v) -> uint8_t { return depth == 16 ? (uint8_t) (v >> 8) : depth == 8 ? (uint8_t) v : (uint8_t) ((v * 255u) / ((1u << depth) - 1u)); };
            if (hd.color == 3) {
                if (3 * s[0] + 2 >= plte.size()) { o[0] = o[1] = o[2] = 0; o[3] = 255; }
                else { o[0] = plte[3 * s[0]]; o[1] = plte[3 * s[0] + 1]; o[2] = plte[3 * s[0] + 2]; o[3] = s[0] < trns.size() ? trns[s[0]] : 255; }
            } else if (hd.color == 0 || hd.color == 4) {
                o[0] = o[1] = o[2] = to8(s[0]);
                if (hd.color == 4) o[3] = to8(s[1]);
                else o[3] = (trns.size() >= 2 && s[0] == (((unsigned) trns[0] << 8) | trns[1])) ? 0 : 255;
            } else {
                o[0] = to8(s[0]); o[1] = to8(s[1]); o[2] = to8(s[2]);
                if (hd.color == 6) o[3] = to8(s[3]);
                else o[3] = (trns.size() >= 6 && s[0] == (((unsigned) trns[0] << 8) | trns[1]) && s[1] == (((unsigned) trns[2] << 8) | trns[3]) &&
                             s[2] == (((unsigned) trns[4] << 8) | trns[5])) ? 0 : 255;
            }
        }
    }
}

}  // namespace

void decodePng(const std::string &path, std::vector<uint8_t> &rgba, unsigned &width, unsigned &height) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw NoriException("PNGTexture: image file not found %s", path);
    std::vector<uint8_t> d((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    static const uint8_t sig[8] = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
    if (d.size() < 8 || memcmp(d.data(), sig, 8) != 0) throw NoriException("%s: not a PNG file", path);
    Header hd{ 0, 0, 0, 0, 0 };
    std::vector<uint8_t> idat, plte, trns;
    bool seenHeader = false;
    for (size_t pos = 8; pos + 12 <= d.size();) {
        const uint32_t len = be32(d.data() + pos);
        const char *type = (const char *) d.data() + pos + 4;
        const uint8_t *body = d.data() + pos + 8;
        if (pos + 12 + (size_t) len > d.size()) throw NoriException("%s: truncated PNG chunk", path);
        if (!memcmp(type, "IHDR", 4)) {
            if (len < 13) throw NoriException("%s: bad IHDR", path);
            hd.w = be32(body); hd.h = be32(body + 4); hd.depth = body[8]; hd.color = body[9]; hd.interlace = body[12];
            if (body[10] != 0 || body[11] != 0) throw NoriException("%s: unknown PNG compression/filter method", path);
            seenHeader = true;
        } else if (!memcmp(type, "PLTE", 4)) plte.assign(body, body + len);
        else if (!memcmp(type, "tRNS", 4)) trns.assign(body, body + len);
        else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!memcmp(type, "IEND", 4)) break;
        pos += 12 + (size_t) len;
    }
    const bool depthOk = hd.color == 0 ? (hd.depth == 1 || hd.depth == 2 || hd.depth == 4 || hd.depth == 8 || hd.depth == 16)
                       : hd.color == 3 ? (hd.depth == 1 || hd.depth == 2 || hd.depth == 4 || hd.depth == 8)
                       : (hd.color == 2 || hd.color == 4 || hd.color == 6) ? (hd.depth == 8 || hd.depth == 16) : false;
    if (!seenHeader || hd.w == 0 || hd.h == 0 || !depthOk || hd.interlace > 1) throw NoriException("%s: unsupported or corrupt PNG header", path);
    if ((uint64_t) hd.w * hd.h > (1ull << 28)) throw NoriException("%s: image too large", path);
    const size_t bits = (size_t) channels(hd.color) * hd.depth, bpp = std::max<size_t>(1, bits / 8);
    std::vector<uint8_t> data = zlibInflate(idat.data(), idat.size());
    width = hd.w; height = hd.h;
    rgba.assign((size_t) hd.w * hd.h * 4, 0);
    std::vector<uint8_t> raw;
    if (hd.interlace == 0) {
        const size_t stride = (hd.w * bits + 7) / 8;
        if (data.size() < (stride + 1) * hd.h) throw NoriException("%s: PNG pixel data too short", path);
        unfilter(data.data(), hd.h, stride, bpp, raw);
        emit(hd, raw, hd.w, hd.h, stride, plte, trns, 0, 0, 1, 1, rgba);
    } else {
        static const unsigned X0[7] = { 0, 4, 0, 2, 0, 1, 0 }, Y0[7] = { 0, 0, 4, 0, 2, 0, 1 }, DX[7] = { 8, 8, 4, 4, 2, 2, 1 }, DY[7] = { 8, 8, 8, 4, 4, 2, 2 };
        size_t at = 0;
        for (int p = 0; p < 7; ++p) {
            const size_t pw = (hd.w + DX[p] - X0[p] - 1) / DX[p], ph = (hd.h + DY[p] - Y0[p] - 1) / DY[p];
            if (pw == 0 || ph == 0) continue;
            const size_t stride = (pw * bits + 7) / 8;
            if (data.size() < at + (stride + 1) * ph) throw NoriException("%s: PNG pixel data too short", path);
            unfilter(data.data() + at, ph, stride, bpp, raw);
            emit(hd, raw, pw, ph, stride, plte, trns, X0[p], Y0[p], DX[p], DY[p], rgba);
            at += (stride + 1) * ph;
        }
    }
}


/* Bitmap::saveToLDR (src/utils/bitmap.cpp:109-140): sRGB transfer, x 255 + 0.5, clamped, 8-bit RGB — written as a PNG whose
 * zlib stream uses stored blocks (no compressor on this side; every PNG reader takes them) */
static uint32_t crc32Of(const uint8_t *p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) { for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; } ready = true; }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xffu] ^ (crc >> 8);
    return ~crc;
}
void writePng(const std::string &path, const float *data, size_t h, size_t w) {
    std::vector<uint8_t> raw;
    raw.reserve(h * (1 + 3 * w));
    for (size_t y = 0; y < h; ++y) {
        raw.push_back(0);                                                   /* filter type: none */
        for (size_t i = 0; i < 3 * w; ++i) {
            const float v = data[y * 3 * w + i];
            const float g = v <= 0.0031308f ? 12.92f * v : 1.055f * std::pow(v, 1.f / 2.4f) - 0.055f;
            const float b = 255.f * g + 0.5f;
            raw.push_back((uint8_t) (b < 0.f ? 0.f : b > 255.f ? 255.f : b));
        }
    }
    std::vector<uint8_t> z = { 0x78, 0x01 };                                /* zlib header, then stored deflate blocks of <= 65535 bytes */
    uint32_t a = 1, b = 0;
    for (size_t at = 0; at < raw.size() || at == 0; at += 65535) {
        const size_t n = std::min<size_t>(65535, raw.size() - at);
        z.push_back(at + n >= raw.size() ? 1 : 0);
        z.push_back((uint8_t) (n & 0xff)); z.push_back((uint8_t) (n >> 8)); z.push_back((uint8_t) (~n & 0xff)); z.push_back((uint8_t) ((~n >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + at, raw.begin() + at + n);
        if (at + n >= raw.size()) break;
    }
    for (uint8_t c : raw) { a = (a + c) % 65521u; b = (b + a) % 65521u; }    /* Adler-32 */
    const uint32_t adler = (b << 16) | a;
    for (int k = 3; k >= 0; --k) z.push_back((uint8_t) (adler >> (8 * k)));
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw NoriException("cannot write %s", path);
    auto chunk = [&](const char *type, const std::vector<uint8_t> &body) {
        std::vector<uint8_t> c(type, type + 4);
        c.insert(c.end(), body.begin(), body.end());
        const uint32_t n = (uint32_t) body.size(), crc = crc32Of(c.data(), c.size());
        const uint8_t len[4] = { (uint8_t) (n >> 24), (uint8_t) (n >> 16), (uint8_t) (n >> 8), (uint8_t) n };
        const uint8_t tail[4] = { (uint8_t) (crc >> 24), (uint8_t) (crc >> 16), (uint8_t) (crc >> 8), (uint8_t) crc };
        fwrite(len, 1, 4, f); fwrite(c.data(), 1, c.size(), f); fwrite(tail, 1, 4, f);
    };
    const uint8_t sig[8] = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
    fwrite(sig, 1, 8, f);
    std::vector<uint8_t> ihdr = { (uint8_t) (w >> 24), (uint8_t) (w >> 16), (uint8_t) (w >> 8), (uint8_t) w,
                                  (uint8_t) (h >> 24), (uint8_t) (h >> 16), (uint8_t) (h >> 8), (uint8_t) h, 8, 2, 0, 0, 0 };
    chunk("IHDR", ihdr); chunk("IDAT", z); chunk("IEND", {});
    if (fclose(f) != 0) throw NoriException("cannot write %s", path);
}


} // namespace nori
