/* codecs.cpp — inflate (RFC 1950/1951), LZ4 block, Blosc-1 frame. See codecs.h. */
#include "codecs.h"
#include <cstring>

namespace nori {

/* ---- inflate ----------------------------------------------------------------------------------- */
namespace {

struct BitReader {
    const uint8_t *p; size_t n, pos = 0; uint32_t buf = 0; int cnt = 0;
    uint32_t bits(int need) {
        while (cnt < need) {
            if (pos >= n) throw NoriException("inflate: unexpected end of stream");
            buf |= (uint32_t) p[pos++] << cnt; cnt += 8;
        }
        const uint32_t v = need ? (buf & ((1u << need) - 1u)) : 0u;
        buf >>= need; cnt -= need;
        return v;
    }
    void alignByte() { buf = 0; cnt = 0; }
};

/* canonical Huffman code: count[len] codes of each length, symbols ordered by (length, value) */
struct Huffman {
    uint16_t count[16]; uint16_t symbol[288];
    void build(const uint8_t *lengths, int n) {
        memset(count, 0, sizeof(count));
        for (int i = 0; i < n; ++i) count[lengths[i]]++;
        count[0] = 0;
        int left = 1;
        for (int len = 1; len < 16; ++len) {
            left = (left << 1) - count[len];
            if (left < 0) throw NoriException("inflate: over-subscribed Huffman code");
        }
        uint16_t offs[16]; offs[1] = 0;
        for (int len = 1; len < 15; ++len) offs[len + 1] = offs[len] + count[len];
        for (int i = 0; i < n; ++i) if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t) i;
    }
    int decode(BitReader &br) const {
        int code = 0, first = 0, index = 0;
        for (int len = 1; len < 16; ++len) {
            code |= (int) br.bits(1);
            const int c = count[len];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        throw NoriException("inflate: invalid Huffman code");
    }
};

const uint16_t kLenBase[29] = { 3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258 };
const uint8_t kLenExtra[29] = { 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0 };
const uint16_t kDistBase[30] = { 1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577 };
const uint8_t kDistExtra[30] = { 0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13 };

void inflateBlockData(BitReader &br, std::vector<uint8_t> &out, const Huffman &lit, const Huffman &dist) {
    for (;;) {
        const int sym = lit.decode(br);
        if (sym < 256) out.push_back((uint8_t) sym);
        else if (sym == 256) return;
        else {
            if (sym > 285) throw NoriException("inflate: bad length symbol");
            const size_t len = kLenBase[sym - 257] + br.bits(kLenExtra[sym - 257]);
            const int ds = dist.decode(br);
            if (ds > 29) throw NoriException("inflate: bad distance symbol");
            const size_t d = kDistBase[ds] + br.bits(kDistExtra[ds]);
            if (d > out.size()) throw NoriException("inflate: distance too far back");
            size_t from = out.size() - d;
            for (size_t i = 0; i < len; ++i) out.push_back(out[from + i]);     /* may overlap: byte by byte */
        }
    }
}

}  // namespace

std::vector<uint8_t> zlibInflate(const uint8_t *src, size_t n, size_t expected) {
    if (n < 2 || (src[0] & 0x0f) != 8 || ((src[0] << 8) | src[1]) % 31 != 0 || (src[1] & 0x20))
        throw NoriException("inflate: not a zlib stream");
    BitReader br{ src + 2, n - 2 };
    std::vector<uint8_t> out;
    if (expected) out.reserve(expected);
    Huffman fixedLit, fixedDist;
    {
        uint8_t l[288];
        for (int i = 0; i < 144; ++i) l[i] = 8;
        for (int i = 144; i < 256; ++i) l[i] = 9;
        for (int i = 256; i < 280; ++i) l[i] = 7;
        for (int i = 280; i < 288; ++i) l[i] = 8;
        fixedLit.build(l, 288);
        for (int i = 0; i < 30; ++i) l[i] = 5;
        fixedDist.build(l, 30);
    }
    for (bool last = false; !last;) {
        last = br.bits(1) != 0;
        const uint32_t type = br.bits(2);
        if (type == 0) {
            br.alignByte();
            if (br.pos + 4 > br.n) throw NoriException("inflate: truncated stored block");
            const uint32_t len = br.p[br.pos] | (br.p[br.pos + 1] << 8), nlen = br.p[br.pos + 2] | (br.p[br.pos + 3] << 8);
            if ((len ^ 0xffffu) != nlen) throw NoriException("inflate: stored block length mismatch");
            br.pos += 4;
            if (br.pos + len > br.n) throw NoriException("inflate: truncated stored block");
            out.insert(out.end(), br.p + br.pos, br.p + br.pos + len);
            br.pos += len;
        } else if (type == 1) {
            inflateBlockData(br, out, fixedLit, fixedDist);
        } else if (type == 2) {
            const int nlen = (int) br.bits(5) + 257, ndist = (int) br.bits(5) + 1, ncode = (int) br.bits(4) + 4;
            if (nlen > 286 || ndist > 30) throw NoriException("inflate: bad code counts");
            static const uint8_t order[19] = { 16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15 };
            uint8_t lengths[320];
            memset(lengths, 0, sizeof(lengths));
            for (int i = 0; i < ncode; ++i) lengths[order[i]] = (uint8_t) br.bits(3);
            Huffman lencode; lencode.build(lengths, 19);
            int idx = 0;
            memset(lengths, 0, sizeof(lengths));
            while (idx < nlen + ndist) {
                const int sym = lencode.decode(br);
                if (sym < 16) lengths[idx++] = (uint8_t) sym;
                else {
                    uint8_t val = 0; int rep;
                    if (sym == 16) { if (idx == 0) throw NoriException("inflate: repeat without a previous length"); val = lengths[idx - 1]; rep = 3 + (int) br.bits(2); }
                    else if (sym == 17) rep = 3 + (int) br.bits(3);
                    else rep = 11 + (int) br.bits(7);
                    if (idx + rep > nlen + ndist) throw NoriException("inflate: too many code lengths");
                    while (rep--) lengths[idx++] = val;
                }
            }
            if (lengths[256] == 0) throw NoriException("inflate: no end-of-block code");
            Huffman lit, dist;
            lit.build(lengths, nlen);
            dist.build(lengths + nlen, ndist);
            inflateBlockData(br, out, lit, dist);
        } else
            throw NoriException("inflate: reserved block type");
    }
    return out;
}

/* ---- LZ4 block --------------------------------------------------------------------------------- */
void lz4DecodeBlock(const uint8_t *src, size_t n, uint8_t *dst, size_t dst_len) {
    size_t ip = 0, op = 0;
    while (ip < n) {
        const uint8_t token = src[ip++];
        size_t lit = token >> 4;
        if (lit == 15) { uint8_t b; do { if (ip >= n) throw NoriException("lz4: truncated literal length"); b = src[ip++]; lit += b; } while (b == 255); }
        if (ip + lit > n || op + lit > dst_len) throw NoriException("lz4: literal run out of bounds");
        memcpy(dst + op, src + ip, lit); ip += lit; op += lit;
        if (ip >= n) break;                                   /* the last sequence has no match */
        if (ip + 2 > n) throw NoriException("lz4: truncated offset");
        const size_t off = src[ip] | ((size_t) src[ip + 1] << 8); ip += 2;
        if (off == 0 || off > op) throw NoriException("lz4: bad match offset");
        size_t len = token & 15;
        if (len == 15) { uint8_t b; do { if (ip >= n) throw NoriException("lz4: truncated match length"); b = src[ip++]; len += b; } while (b == 255); }
        len += 4;
        if (op + len > dst_len) throw NoriException("lz4: match out of bounds");
        for (size_t i = 0; i < len; ++i, ++op) dst[op] = dst[op - off];
    }
    if (op != dst_len) throw NoriException("lz4: block decoded to %d bytes, expected %d", (int) op, (int) dst_len);
}

/* ---- Blosc 1 ----------------------------------------------------------------------------------- */
static uint32_t rd32(const uint8_t *p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t) p[3] << 24); }

std::vector<uint8_t> bloscDecode(const uint8_t *src, size_t n) {
    if (n < 16) throw NoriException("blosc: frame shorter than its header");
    const uint8_t flags = src[2];
    const size_t typesize = src[3], nbytes = rd32(src + 4), blocksize = rd32(src + 8), cbytes = rd32(src + 12);
    if (cbytes > n) throw NoriException("blosc: frame says %d bytes, %d available", (int) cbytes, (int) n);
    std::vector<uint8_t> out(nbytes);
    if (nbytes == 0) return out;
    if (flags & 0x2) {                                        /* stored */
        if (16 + nbytes > n) throw NoriException("blosc: truncated stored frame");
        memcpy(out.data(), src + 16, nbytes);
        return out;
    }
    if (flags & 0x4) throw NoriException("blosc: bit-shuffled frames are not supported");
    const int codec = flags >> 5;
    if (codec != 1 && codec != 3) throw NoriException("blosc: inner codec %d is not supported (LZ4 and zlib are)", codec);
    if (blocksize == 0 || typesize == 0) throw NoriException("blosc: bad header");
    const size_t nblocks = (nbytes + blocksize - 1) / blocksize;
    if (16 + 4 * nblocks > n) throw NoriException("blosc: truncated block table");
    const bool shuffle = (flags & 0x1) && typesize > 1;
    std::vector<uint8_t> tmp(blocksize);
    for (size_t b = 0; b < nblocks; ++b) {
        const size_t bsize = (b == nblocks - 1 && nbytes % blocksize) ? nbytes % blocksize : blocksize;
        const bool leftover = bsize != blocksize;
        const size_t nsplits = (!(flags & 0x10) && typesize <= 16 && blocksize / typesize >= 128 && !leftover) ? typesize : 1;
        const size_t neblock = bsize / nsplits;
        size_t ip = rd32(src + 16 + 4 * b);
        uint8_t *dst = shuffle ? tmp.data() : out.data() + b * blocksize;
        for (size_t s = 0; s < nsplits; ++s) {
            if (ip + 4 > n) throw NoriException("blosc: truncated block");
            const size_t c = rd32(src + ip); ip += 4;
            if (ip + c > n) throw NoriException("blosc: truncated block");
            if (c == neblock) memcpy(dst + s * neblock, src + ip, neblock);
            else if (codec == 1) lz4DecodeBlock(src + ip, c, dst + s * neblock, neblock);
            else {
                std::vector<uint8_t> z = zlibInflate(src + ip, c, neblock);
                if (z.size() != neblock) throw NoriException("blosc: zlib block size mismatch");
                memcpy(dst + s * neblock, z.data(), neblock);
            }
            ip += c;
        }
        if (shuffle) {                                        /* byte j of element i was stored at j*nelem + i */
            uint8_t *o = out.data() + b * blocksize;
            const size_t nelem = bsize / typesize;
            for (size_t j = 0; j < typesize; ++j)
                for (size_t i = 0; i < nelem; ++i) o[i * typesize + j] = tmp[j * nelem + i];
            memcpy(o + nelem * typesize, tmp.data() + nelem * typesize, bsize - nelem * typesize);
        }
    }
    return out;
}

}
