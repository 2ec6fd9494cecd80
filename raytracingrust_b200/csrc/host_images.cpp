// Host image IO of the render path's inputs and outputs (no CUDA here):
//   rtr_image_load   SkyMap::new (lib/backgrounds.rs:58-80): the first RGBA layer of an OpenEXR file, texel
//                    image[x + y*width] with row 0 the first row of the data window.  The reference goes through the
//                    `exr` crate (Cargo.toml: exr ^1.4.2), which is not under /root/reference; this is a reader of
//                    the published OpenEXR 2 file layout written from the format description: single-part
//                    scan-line files, channels R,G,B(,A) of type half / float / uint, compressions NONE, RLE,
//                    ZIPS, ZIP, PIZ and PXR24.  Also Radiance RGBE (.hdr), which the reference cannot read at all.
//   rtr_write_exr    the fp32 accumulation buffer as a float ZIP scan-line EXR
// Checked in tests/test_images.py against files written and read back by OpenCV's OpenEXR codec.
#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rtr.h"

namespace rtr { void set_global_error(const std::string& msg); }

struct rtr_host_image {
    uint32_t width = 0, height = 0;
    std::vector<float> rgb;  // 3 floats per texel, row-major, row 0 = top
};

namespace {

struct ImageError { std::string msg; };
[[noreturn]] void fail(const std::string& m) { throw ImageError{m}; }

bool read_file(const char* path, std::vector<uint8_t>& out) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); return false; }
    out.resize((size_t)n);
    bool ok = n == 0 || std::fread(out.data(), 1, (size_t)n, f) == (size_t)n;
    std::fclose(f);
    return ok;
}

float half_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1Fu, man = h & 0x3FFu, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {  // subnormal half -> normal float
            int e = -1;
            do { ++e; man <<= 1; } while ((man & 0x400u) == 0);
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3FFu) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7F800000u | (man << 13);
    else bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

// ------------------------------------------------------------------ little-endian cursor
struct Cursor {
    const uint8_t* p;
    const uint8_t* end;
    void need(size_t n) const { if (p > end || (size_t)(end - p) < n) fail("unexpected end of file"); }
    uint8_t u8() { need(1); return *p++; }
    uint32_t u32() { need(4); uint32_t v = p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); p += 4; return v; }
    int32_t i32() { return (int32_t)u32(); }
    uint64_t u64() { uint64_t lo = u32(), hi = u32(); return lo | (hi << 32); }
    std::string cstr() {
        std::string s;
        for (;;) { uint8_t c = u8(); if (!c) break; s.push_back((char)c); if (s.size() > 255) fail("attribute name too long"); }
        return s;
    }
};

void inflate_into(const uint8_t* src, size_t n, std::vector<uint8_t>& dst, size_t cap) {
    dst.resize(cap);
    uLongf len = (uLongf)cap;
    int rc = uncompress(dst.data(), &len, src, (uLong)n);
    if (rc != Z_OK) fail("zlib inflate failed");
    dst.resize(len);
}

// predictor + byte de-interleave shared by RLE / ZIPS / ZIP blocks
void unpredict_deinterleave(std::vector<uint8_t>& t, std::vector<uint8_t>& out) {
    const size_t n = t.size();
    for (size_t i = 1; i < n; ++i) t[i] = (uint8_t)(t[i - 1] + t[i] - 128);
    out.resize(n);
    const size_t half = (n + 1) / 2;
    for (size_t i = 0; i < n; ++i) out[i] = (i & 1) ? t[half + i / 2] : t[i / 2];
}

void rle_decode(const uint8_t* src, size_t n, std::vector<uint8_t>& dst, size_t cap) {
    dst.clear();
    size_t i = 0;
    while (i < n) {
        const int8_t c = (int8_t)src[i++];
        if (c < 0) {
            const size_t k = (size_t)(-(int)c);
            if (i + k > n || dst.size() + k > cap) fail("corrupt RLE block");
            dst.insert(dst.end(), src + i, src + i + k);
            i += k;
        } else {
            const size_t k = (size_t)c + 1;
            if (i >= n || dst.size() + k > cap) fail("corrupt RLE block");
            dst.insert(dst.end(), k, src[i++]);
        }
    }
}

// ------------------------------------------------------------------ PIZ: Huffman + Haar wavelet + value LUT
// Layout of a PIZ block (OpenEXR file layout documentation, "PIZ compression"): u16 minNonZero, u16 maxNonZero,
// the bytes [minNonZero, maxNonZero] of an 8192-byte "value occurs" bitmap, i32 length, Huffman stream.
constexpr int HUF_ENCBITS = 16;
constexpr int HUF_ENCSIZE = (1 << HUF_ENCBITS) + 1;  // 65536 values + the run-length symbol
constexpr int HUF_LUTBITS = 14;
constexpr int HUF_MAXLEN = 58;

struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint64_t c = 0;
    int lc = 0;
    void fill() { while (lc <= 56) { c = (c << 8) | (p < end ? *p++ : 0u); lc += 8; } }
    uint32_t peek(int n) const { return (uint32_t)((c >> (lc - n)) & ((1ull << n) - 1ull)); }
    uint64_t peek64(int n) const { return (c >> (lc - n)) & ((1ull << n) - 1ull); }
    void drop(int n) { lc -= n; }
};

void huf_uncompress(const uint8_t* src, size_t n_src, uint16_t* raw, size_t n_raw) {
    if (n_src == 0) { if (n_raw) fail("PIZ: empty Huffman stream"); return; }
    if (n_src < 20) fail("PIZ: truncated Huffman header");
    auto rd = [&](size_t o) { return (uint32_t)src[o] | (src[o + 1] << 8) | (src[o + 2] << 16) | ((uint32_t)src[o + 3] << 24); };
    const uint32_t im = rd(0), iM = rd(4), n_bits = rd(12);
    if (im >= (uint32_t)HUF_ENCSIZE || iM >= (uint32_t)HUF_ENCSIZE || im > iM) fail("PIZ: bad Huffman symbol range");
    // code lengths, 6 bits each, with zero-run escapes (59..62: short run of 2..5 zeros, 63: 8-bit run + 6)
    std::vector<uint8_t> len(HUF_ENCSIZE, 0);
    const uint8_t* p = src + 20;
    const uint8_t* end = src + n_src;
    {
        uint64_t c = 0; int lc = 0;
        auto get = [&](int nb) -> uint32_t {
            while (lc < nb) { if (p >= end) fail("PIZ: truncated code table"); c = (c << 8) | *p++; lc += 8; }
            lc -= nb;
            return (uint32_t)((c >> lc) & ((1u << nb) - 1u));
        };
        for (uint32_t i = im; i <= iM; ++i) {
            const uint32_t l = get(6);
            if (l == 63u) {
                uint32_t run = get(8) + 6u;
                if (i + run > iM + 1u) fail("PIZ: code table overruns");
                i += run - 1u;
            } else if (l >= 59u) {
                uint32_t run = l - 59u + 2u;
                if (i + run > iM + 1u) fail("PIZ: code table overruns");
                i += run - 1u;
            } else len[i] = (uint8_t)l;
        }
    }
    if ((uint64_t)(end - p) * 8u < n_bits) fail("PIZ: truncated Huffman data");
    // canonical codes: lengths longest-first, code values ascending with the symbol index within one length
    uint64_t count[HUF_MAXLEN + 1] = {0}, first[HUF_MAXLEN + 1] = {0};
    for (int i = 0; i < HUF_ENCSIZE; ++i) count[len[i]]++;
    {
        uint64_t c = 0;
        for (int l = HUF_MAXLEN; l > 0; --l) { const uint64_t nc = (c + count[l]) >> 1; first[l] = c; c = nc; }
    }
    uint32_t offset[HUF_MAXLEN + 2];
    offset[1] = 0;
    for (int l = 1; l <= HUF_MAXLEN; ++l) offset[l + 1] = offset[l] + (uint32_t)count[l];
    std::vector<uint32_t> sorted(offset[HUF_MAXLEN + 1]);
    {
        uint32_t fill[HUF_MAXLEN + 1];
        for (int l = 1; l <= HUF_MAXLEN; ++l) fill[l] = offset[l];
        for (int i = 0; i < HUF_ENCSIZE; ++i) if (len[i]) sorted[fill[len[i]]++] = (uint32_t)i;
    }
    struct Fast { uint32_t sym; uint32_t len; };
    std::vector<Fast> lut((size_t)1 << HUF_LUTBITS, Fast{0, 0});
    for (int l = 1; l <= HUF_LUTBITS; ++l)
        for (uint64_t k = 0; k < count[l]; ++k) {
            const uint64_t code = first[l] + k;
            if (code >> l) fail("PIZ: invalid code table");
            const size_t lo = (size_t)(code << (HUF_LUTBITS - l)), span = (size_t)1 << (HUF_LUTBITS - l);
            for (size_t j = 0; j < span; ++j) lut[lo + j] = Fast{sorted[offset[l] + k], (uint32_t)l};
        }
    BitReader br{p, end};
    const uint32_t rlc = iM;
    size_t out = 0;
    int64_t bits_left = (int64_t)n_bits;
    while (out < n_raw) {
        if (bits_left <= 0) fail("PIZ: Huffman data ends early");
        br.fill();
        uint32_t sym, l;
        const Fast f = lut[br.peek(HUF_LUTBITS)];
        if (f.len) { sym = f.sym; l = f.len; }
        else {
            l = 0; sym = 0;
            for (int k = HUF_LUTBITS + 1; k <= HUF_MAXLEN; ++k) {
                if (!count[k]) continue;
                const uint64_t v = br.peek64(k);
                if (v >= first[k] && v - first[k] < count[k]) { sym = sorted[offset[k] + (uint32_t)(v - first[k])]; l = (uint32_t)k; break; }
            }
            if (!l) fail("PIZ: invalid Huffman code");
        }
        br.drop((int)l);
        bits_left -= l;
        if (sym == rlc) {
            br.fill();
            const uint32_t run = br.peek(8);
            br.drop(8);
            bits_left -= 8;
            if (out == 0 || out + run > n_raw) fail("PIZ: bad run");
            const uint16_t v = raw[out - 1];
            for (uint32_t k = 0; k < run; ++k) raw[out++] = v;
        } else raw[out++] = (uint16_t)sym;
    }
}

inline void wdec14(uint16_t l, uint16_t h, uint16_t& a, uint16_t& b) {
    const int ls = (int16_t)l, hs = (int16_t)h;
    const int ai = ls + (hs & 1) + (hs >> 1);
    a = (uint16_t)(int16_t)ai;
    b = (uint16_t)(int16_t)(ai - hs);
}
inline void wdec16(uint16_t l, uint16_t h, uint16_t& a, uint16_t& b) {
    const int m = l, d = h;
    const int bb = (m - (d >> 1)) & 0xFFFF;
    const int aa = (d + bb - 0x8000) & 0xFFFF;
    b = (uint16_t)bb;
    a = (uint16_t)aa;
}
// inverse 2-D Haar transform in place over nx x ny values with strides ox, oy
void wav2_decode(uint16_t* in, int nx, int ox, int ny, int oy, uint16_t mx) {
    const bool w14 = mx < (1 << 14);
    const int n = nx > ny ? ny : nx;
    int p = 1;
    while (p <= n) p <<= 1;
    p >>= 1;
    int p2 = p;
    p >>= 1;
    while (p >= 1) {
        uint16_t* py = in;
        uint16_t* ey = in + (ptrdiff_t)oy * (ny - p2);
        const ptrdiff_t oy1 = (ptrdiff_t)oy * p, oy2 = (ptrdiff_t)oy * p2, ox1 = (ptrdiff_t)ox * p, ox2 = (ptrdiff_t)ox * p2;
        uint16_t i00, i01, i10, i11;
        for (; py <= ey; py += oy2) {
            uint16_t* px = py;
            uint16_t* ex = py + (ptrdiff_t)ox * (nx - p2);
            for (; px <= ex; px += ox2) {
                uint16_t* p01 = px + ox1;
                uint16_t* p10 = px + oy1;
                uint16_t* p11 = p10 + ox1;
                if (w14) {
                    wdec14(*px, *p10, i00, i10); wdec14(*p01, *p11, i01, i11);
                    wdec14(i00, i01, *px, *p01); wdec14(i10, i11, *p10, *p11);
                } else {
                    wdec16(*px, *p10, i00, i10); wdec16(*p01, *p11, i01, i11);
                    wdec16(i00, i01, *px, *p01); wdec16(i10, i11, *p10, *p11);
                }
            }
            if (nx & p) {
                uint16_t* p10 = px + oy1;
                if (w14) wdec14(*px, *p10, i00, *p10); else wdec16(*px, *p10, i00, *p10);
                *px = i00;
            }
        }
        if (ny & p) {
            uint16_t* px = py;
            uint16_t* ex = py + (ptrdiff_t)ox * (nx - p2);
            for (; px <= ex; px += ox2) {
                uint16_t* p01 = px + ox1;
                if (w14) wdec14(*px, *p01, i00, *p01); else wdec16(*px, *p01, i00, *p01);
                *px = i00;
            }
        }
        p2 = p;
        p >>= 1;
    }
}

// ------------------------------------------------------------------ OpenEXR scan-line reader
struct Channel { std::string name; int type; int bytes; };  // type 0 uint, 1 half, 2 float
enum { EXR_NONE = 0, EXR_RLE = 1, EXR_ZIPS = 2, EXR_ZIP = 3, EXR_PIZ = 4, EXR_PXR24 = 5 };

void piz_block(const uint8_t* src, size_t n_src, const std::vector<Channel>& ch, int width, int lines,
               std::vector<uint8_t>& out, size_t raw_size) {
    Cursor c{src, src + n_src};
    c.need(4);
    const uint32_t min_nz = c.p[0] | (c.p[1] << 8), max_nz = c.p[2] | (c.p[3] << 8);
    c.p += 4;
    std::vector<uint8_t> bitmap(8192, 0);
    if (max_nz >= 8192) fail("PIZ: bad bitmap range");
    if (min_nz <= max_nz) { c.need(max_nz - min_nz + 1); std::memcpy(bitmap.data() + min_nz, c.p, max_nz - min_nz + 1); c.p += max_nz - min_nz + 1; }
    std::vector<uint16_t> lut(65536, 0);
    uint32_t k = 0;
    for (uint32_t i = 0; i < 65536; ++i) if (i == 0 || (bitmap[i >> 3] & (1 << (i & 7)))) lut[k++] = (uint16_t)i;
    const uint16_t max_value = (uint16_t)(k - 1);
    const int32_t length = c.i32();
    if (length < 0 || (size_t)length > (size_t)(c.end - c.p)) fail("PIZ: bad Huffman length");
    const size_t n_shorts = raw_size / 2;
    std::vector<uint16_t> tmp(n_shorts);
    huf_uncompress(c.p, (size_t)length, tmp.data(), n_shorts);
    std::vector<size_t> start(ch.size());
    size_t at = 0;
    for (size_t i = 0; i < ch.size(); ++i) {
        const int size = ch[i].bytes / 2;
        start[i] = at;
        for (int j = 0; j < size; ++j) wav2_decode(tmp.data() + at + j, width, size, lines, width * size, max_value);
        at += (size_t)width * lines * size;
    }
    for (size_t i = 0; i < n_shorts; ++i) tmp[i] = lut[tmp[i]];
    out.resize(raw_size);
    uint8_t* o = out.data();
    std::vector<size_t> cur = start;
    for (int y = 0; y < lines; ++y)
        for (size_t i = 0; i < ch.size(); ++i) {
            const size_t n = (size_t)width * (ch[i].bytes / 2);
            for (size_t j = 0; j < n; ++j) { const uint16_t v = tmp[cur[i] + j]; *o++ = (uint8_t)(v & 0xFF); *o++ = (uint8_t)(v >> 8); }
            cur[i] += n;
        }
}

void pxr24_block(const uint8_t* src, size_t n_src, const std::vector<Channel>& ch, int width, int lines,
                 std::vector<uint8_t>& out, size_t raw_size) {
    std::vector<uint8_t> tmp;
    inflate_into(src, n_src, tmp, raw_size);
    out.resize(raw_size);
    uint8_t* o = out.data();
    size_t at = 0;
    const size_t n = (size_t)width;
    for (int y = 0; y < lines; ++y)
        for (const Channel& c : ch) {
            const int planes = c.type == 1 ? 2 : (c.type == 2 ? 3 : 4);
            if (at + n * planes > tmp.size()) fail("PXR24: block too short");
            uint32_t pixel = 0;
            for (size_t j = 0; j < n; ++j) {
                uint32_t diff = 0;
                if (c.type == 1) diff = ((uint32_t)tmp[at + j] << 8) | tmp[at + n + j];
                else if (c.type == 2) diff = ((uint32_t)tmp[at + j] << 24) | ((uint32_t)tmp[at + n + j] << 16) | ((uint32_t)tmp[at + 2 * n + j] << 8);
                else diff = ((uint32_t)tmp[at + j] << 24) | ((uint32_t)tmp[at + n + j] << 16) | ((uint32_t)tmp[at + 2 * n + j] << 8) | tmp[at + 3 * n + j];
                pixel += diff;
                if (c.type == 1) { *o++ = (uint8_t)(pixel & 0xFF); *o++ = (uint8_t)((pixel >> 8) & 0xFF); }
                else { *o++ = (uint8_t)pixel; *o++ = (uint8_t)(pixel >> 8); *o++ = (uint8_t)(pixel >> 16); *o++ = (uint8_t)(pixel >> 24); }
            }
            at += n * planes;
        }
}

void load_exr(const std::vector<uint8_t>& file, rtr_host_image& img) {
    Cursor c{file.data(), file.data() + file.size()};
    if (c.u32() != 20000630u) fail("not an OpenEXR file");
    const uint32_t version = c.u32();
    if ((version & 0xFFu) != 2u) fail("unsupported OpenEXR version");
    if (version & 0x200u) fail("tiled OpenEXR files are not supported");
    if (version & 0x800u) fail("deep OpenEXR files are not supported");
    if (version & 0x1000u) fail("multi-part OpenEXR files are not supported");
    std::vector<Channel> ch;
    int compression = -1, xmin = 0, ymin = 0, xmax = -1, ymax = -1;
    for (;;) {
        const std::string name = c.cstr();
        if (name.empty()) break;
        const std::string type = c.cstr();
        const int32_t size = c.i32();
        if (size < 0) fail("bad attribute size");
        c.need((size_t)size);
        Cursor a{c.p, c.p + size};
        c.p += size;
        if (name == "channels" && type == "chlist") {
            for (;;) {
                const std::string cn = a.cstr();
                if (cn.empty()) break;
                Channel k;
                k.name = cn;
                k.type = a.i32();
                a.u8(); a.u8(); a.u8(); a.u8();  // pLinear + reserved
                const int32_t xs = a.i32(), ys = a.i32();
                if (xs != 1 || ys != 1) fail("sub-sampled channels are not supported");
                if (k.type < 0 || k.type > 2) fail("bad channel type");
                k.bytes = k.type == 1 ? 2 : 4;
                ch.push_back(k);
            }
        } else if (name == "compression") compression = a.u8();
        else if (name == "dataWindow") { xmin = a.i32(); ymin = a.i32(); xmax = a.i32(); ymax = a.i32(); }
    }
    if (ch.empty() || compression < 0 || xmax < xmin || ymax < ymin) fail("incomplete OpenEXR header");
    if (compression > EXR_PXR24) fail("OpenEXR compression " + std::to_string(compression) + " (B44/DWA) is not supported");
    const int64_t width = (int64_t)xmax - xmin + 1, height = (int64_t)ymax - ymin + 1;
    if (width > 65536 || height > 65536) fail("OpenEXR image too large");
    // first RGB(A) layer: plain R,G,B or the first "<layer>.R" with its G and B; a lone Y is grey
    int ir = -1, ig = -1, ib = -1;
    auto find = [&](const std::string& n) { for (size_t i = 0; i < ch.size(); ++i) if (ch[i].name == n) return (int)i; return -1; };
    ir = find("R"); ig = find("G"); ib = find("B");
    if (ir < 0 || ig < 0 || ib < 0) {
        ir = ig = ib = -1;
        for (size_t i = 0; i < ch.size() && ir < 0; ++i) {
            const std::string& n = ch[i].name;
            if (n.size() > 2 && n.compare(n.size() - 2, 2, ".R") == 0) {
                const std::string base = n.substr(0, n.size() - 1);
                const int g = find(base + "G"), b = find(base + "B");
                if (g >= 0 && b >= 0) { ir = (int)i; ig = g; ib = b; }
            }
        }
        if (ir < 0) { const int y = find("Y"); if (y >= 0) ir = ig = ib = y; }
        if (ir < 0) fail("no RGB layer in the OpenEXR file");
    }
    size_t line_bytes = 0;
    std::vector<size_t> ch_off(ch.size());
    for (size_t i = 0; i < ch.size(); ++i) { ch_off[i] = line_bytes; line_bytes += (size_t)ch[i].bytes * (size_t)width; }
    const int lines_per_block = compression == EXR_ZIP || compression == EXR_PXR24 ? 16 : (compression == EXR_PIZ ? 32 : 1);
    const size_t n_blocks = (size_t)((height + lines_per_block - 1) / lines_per_block);
    std::vector<uint64_t> offsets(n_blocks);
    for (auto& o : offsets) o = c.u64();
    img.width = (uint32_t)width;
    img.height = (uint32_t)height;
    img.rgb.assign((size_t)width * height * 3, 0.0f);
    std::vector<uint8_t> tmp, raw;
    for (size_t b = 0; b < n_blocks; ++b) {
        if (offsets[b] > file.size() || file.size() - offsets[b] < 8) fail("bad chunk offset");  // no u64 wrap-around
        Cursor k{file.data() + offsets[b], file.data() + file.size()};
        const int64_t y0 = (int64_t)k.i32() - ymin;
        const int32_t size = k.i32();
        if (y0 < 0 || y0 >= height || size < 0) fail("bad chunk header");
        k.need((size_t)size);
        const int lines = (int)std::min<int64_t>(lines_per_block, height - y0);
        const size_t raw_size = line_bytes * (size_t)lines;
        const uint8_t* data = k.p;
        if ((size_t)size >= raw_size || compression == EXR_NONE) {  // stored
            if ((size_t)size < raw_size) fail("short chunk");
        } else if (compression == EXR_ZIPS || compression == EXR_ZIP) {
            inflate_into(k.p, (size_t)size, tmp, raw_size);
            unpredict_deinterleave(tmp, raw);
            data = raw.data();
        } else if (compression == EXR_RLE) {
            rle_decode(k.p, (size_t)size, tmp, raw_size);
            unpredict_deinterleave(tmp, raw);
            data = raw.data();
        } else if (compression == EXR_PIZ) {
            piz_block(k.p, (size_t)size, ch, (int)width, lines, raw, raw_size);
            data = raw.data();
        } else {
            pxr24_block(k.p, (size_t)size, ch, (int)width, lines, raw, raw_size);
            data = raw.data();
        }
        if (data == raw.data() && raw.size() != raw_size) fail("chunk decodes to the wrong size");
        const int src_ch[3] = {ir, ig, ib};
        for (int y = 0; y < lines; ++y)
            for (int comp = 0; comp < 3; ++comp) {
                const Channel& cc = ch[src_ch[comp]];
                const uint8_t* s = data + (size_t)y * line_bytes + ch_off[src_ch[comp]];
                float* d = img.rgb.data() + ((size_t)(y0 + y) * width) * 3 + comp;
                for (int64_t x = 0; x < width; ++x, d += 3) {
                    if (cc.type == 1) *d = half_to_float((uint16_t)(s[2 * x] | (s[2 * x + 1] << 8)));
                    else {
                        const uint32_t v = s[4 * x] | (s[4 * x + 1] << 8) | (s[4 * x + 2] << 16) | ((uint32_t)s[4 * x + 3] << 24);
                        if (cc.type == 2) std::memcpy(d, &v, 4); else *d = (float)v;
                    }
                }
            }
    }
}

// ------------------------------------------------------------------ Radiance RGBE (.hdr / .pic)
void load_hdr(const std::vector<uint8_t>& file, rtr_host_image& img) {
    const uint8_t* p = file.data();
    const uint8_t* end = p + file.size();
    auto line = [&]() { std::string s; while (p < end && *p != '\n') s.push_back((char)*p++); if (p < end) ++p; return s; };
    std::string l = line();
    if (l.compare(0, 2, "#?") != 0) fail("not a Radiance HDR file");
    bool rgbe = false;
    for (;;) {
        if (p >= end) fail("truncated HDR header");
        l = line();
        if (l.empty()) break;
        if (l.compare(0, 7, "FORMAT=") == 0) rgbe = l == "FORMAT=32-bit_rle_rgbe";
    }
    if (!rgbe) fail("HDR file is not 32-bit_rle_rgbe");
    l = line();
    char ys = 0, xs = 0, yc = 0, xc = 0;
    long H = 0, W = 0;
    if (std::sscanf(l.c_str(), "%c%c %ld %c%c %ld", &ys, &yc, &H, &xs, &xc, &W) != 6 || yc != 'Y' || xc != 'X' || xs != '+' ||
        (ys != '-' && ys != '+') || W <= 0 || H <= 0 || W > 65536 || H > 65536)
        fail("unsupported HDR resolution line '" + l + "'");
    img.width = (uint32_t)W;
    img.height = (uint32_t)H;
    img.rgb.assign((size_t)W * H * 3, 0.0f);
    std::vector<uint8_t> scan((size_t)W * 4);
    for (long row = 0; row < H; ++row) {
        if (end - p < 4) fail("truncated HDR data");
        if (W >= 8 && W < 32768 && p[0] == 2 && p[1] == 2 && ((p[2] << 8) | p[3]) == W) {  // new run-length form: 4 planes
            p += 4;
            for (int comp = 0; comp < 4; ++comp) {
                long x = 0;
                while (x < W) {
                    if (p >= end) fail("truncated HDR scanline");
                    int c = *p++;
                    if (c > 128) {
                        c -= 128;
                        if (p >= end || x + c > W) fail("corrupt HDR run");
                        const uint8_t v = *p++;
                        while (c--) scan[(size_t)(x++) * 4 + comp] = v;
                    } else {
                        if (c == 0 || end - p < c || x + c > W) fail("corrupt HDR run");
                        while (c--) scan[(size_t)(x++) * 4 + comp] = *p++;
                    }
                }
            }
        } else {  // flat RGBE texels
            if ((size_t)(end - p) < (size_t)W * 4) fail("truncated HDR data");
            std::memcpy(scan.data(), p, (size_t)W * 4);
            p += (size_t)W * 4;
        }
        const long y = ys == '-' ? row : H - 1 - row;
        float* d = img.rgb.data() + (size_t)y * W * 3;
        for (long x = 0; x < W; ++x, d += 3) {
            const uint8_t* q = &scan[(size_t)x * 4];
            if (q[3] == 0) { d[0] = d[1] = d[2] = 0.0f; continue; }
            const float f = std::ldexp(1.0f, (int)q[3] - (128 + 8));
            d[0] = q[0] * f; d[1] = q[1] * f; d[2] = q[2] * f;
        }
    }
}

// ------------------------------------------------------------------ OpenEXR writer: float channels, ZIP blocks
void put32(std::vector<uint8_t>& o, uint32_t v) { for (int i = 0; i < 4; ++i) o.push_back((uint8_t)(v >> (8 * i))); }
void put64(std::vector<uint8_t>& o, uint64_t v) { for (int i = 0; i < 8; ++i) o.push_back((uint8_t)(v >> (8 * i))); }
void puts0(std::vector<uint8_t>& o, const char* s) { while (*s) o.push_back((uint8_t)*s++); o.push_back(0); }
void putf(std::vector<uint8_t>& o, float f) { uint32_t v; std::memcpy(&v, &f, 4); put32(o, v); }
void attr(std::vector<uint8_t>& o, const char* name, const char* type, const std::vector<uint8_t>& data) {
    puts0(o, name); puts0(o, type); put32(o, (uint32_t)data.size());
    o.insert(o.end(), data.begin(), data.end());
}

bool write_exr(const char* path, const float* px, uint32_t W, uint32_t H, uint32_t nch) {
    std::vector<uint8_t> o;
    put32(o, 20000630u);
    put32(o, 2u);
    const char* names4[4] = {"A", "B", "G", "R"};
    const int src4[4] = {3, 2, 1, 0};
    const char* const* names = nch == 4 ? names4 : names4 + 1;
    const int* src = nch == 4 ? src4 : src4 + 1;
    {
        std::vector<uint8_t> d;
        for (uint32_t i = 0; i < nch; ++i) { puts0(d, names[i]); put32(d, 2u); put32(d, 0u); put32(d, 1u); put32(d, 1u); }
        d.push_back(0);
        attr(o, "channels", "chlist", d);
    }
    attr(o, "compression", "compression", {(uint8_t)EXR_ZIP});
    {
        std::vector<uint8_t> d;
        put32(d, 0); put32(d, 0); put32(d, W - 1); put32(d, H - 1);
        attr(o, "dataWindow", "box2i", d);
        attr(o, "displayWindow", "box2i", d);
    }
    attr(o, "lineOrder", "lineOrder", {0});
    { std::vector<uint8_t> d; putf(d, 1.0f); attr(o, "pixelAspectRatio", "float", d); }
    { std::vector<uint8_t> d; putf(d, 0.0f); putf(d, 0.0f); attr(o, "screenWindowCenter", "v2f", d); }
    { std::vector<uint8_t> d; putf(d, 1.0f); attr(o, "screenWindowWidth", "float", d); }
    o.push_back(0);
    const uint32_t n_blocks = (H + 15u) / 16u;
    const size_t table_at = o.size();
    o.resize(o.size() + (size_t)n_blocks * 8, 0);
    const size_t line_bytes = (size_t)W * 4 * nch;
    std::vector<uint8_t> raw, t, z;
    for (uint32_t b = 0; b < n_blocks; ++b) {
        const uint32_t y0 = b * 16u, lines = std::min(16u, H - y0);
        raw.resize(line_bytes * lines);
        uint8_t* r = raw.data();
        for (uint32_t y = 0; y < lines; ++y)
            for (uint32_t c = 0; c < nch; ++c)
                for (uint32_t x = 0; x < W; ++x, r += 4) std::memcpy(r, px + ((size_t)(y0 + y) * W + x) * nch + src[c], 4);
        const size_t n = raw.size(), half = (n + 1) / 2;
        t.resize(n);
        for (size_t i = 0; i < n; ++i) t[(i & 1) ? half + i / 2 : i / 2] = raw[i];
        uint8_t prev = t.empty() ? 0 : t[0];
        for (size_t i = 1; i < n; ++i) { const uint8_t cur = t[i]; t[i] = (uint8_t)(cur - prev + 128); prev = cur; }
        uLongf zl = compressBound((uLong)n);
        z.resize(zl);
        if (compress2(z.data(), &zl, t.data(), (uLong)n, 6) != Z_OK) return false;
        const bool stored = zl >= n;
        uint64_t at = o.size();
        for (int i = 0; i < 8; ++i) o[table_at + (size_t)b * 8 + i] = (uint8_t)(at >> (8 * i));
        put32(o, y0);
        put32(o, (uint32_t)(stored ? n : zl));
        if (stored) o.insert(o.end(), raw.begin(), raw.end());
        else o.insert(o.end(), z.begin(), z.begin() + zl);
    }
    (void)put64;
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    const bool ok = std::fwrite(o.data(), 1, o.size(), f) == o.size();
    return std::fclose(f) == 0 && ok;
}

}  // namespace

extern "C" {

rtr_status rtr_image_load(const char* path, rtr_host_image** out) {
    if (!path || !out) { rtr::set_global_error("rtr_image_load: null argument"); return RTR_INVALID_ARG; }
    *out = nullptr;
    std::vector<uint8_t> file;
    if (!read_file(path, file)) { rtr::set_global_error(std::string("rtr_image_load: cannot read '") + path + "'"); return RTR_INVALID_ARG; }
    rtr_host_image* img = new rtr_host_image();
    try {
        if (file.size() >= 4 && file[0] == 0x76 && file[1] == 0x2f && file[2] == 0x31 && file[3] == 0x01) load_exr(file, *img);
        else if (file.size() >= 2 && file[0] == '#' && file[1] == '?') load_hdr(file, *img);
        else fail("neither OpenEXR nor Radiance HDR");
    } catch (const ImageError& e) {
        delete img;
        rtr::set_global_error(std::string("rtr_image_load: '") + path + "': " + e.msg);
        return RTR_UNSUPPORTED;
    } catch (const std::exception& e) {
        delete img;
        rtr::set_global_error(std::string("rtr_image_load: '") + path + "': " + e.what());
        return RTR_OOM;
    }
    *out = img;
    return RTR_OK;
}
uint32_t rtr_host_image_width(const rtr_host_image* img) { return img ? img->width : 0; }
uint32_t rtr_host_image_height(const rtr_host_image* img) { return img ? img->height : 0; }
const float* rtr_host_image_rgb(const rtr_host_image* img) { return img ? img->rgb.data() : nullptr; }
void rtr_host_image_destroy(rtr_host_image* img) { delete img; }

rtr_status rtr_write_exr(const char* path, const float* pixels, uint32_t width, uint32_t height, uint32_t channels) {
    if (!path || !pixels || width == 0 || height == 0 || (channels != 3 && channels != 4)) {
        rtr::set_global_error("rtr_write_exr: bad argument (channels must be 3 or 4)");
        return RTR_INVALID_ARG;
    }
    if (!write_exr(path, pixels, width, height, channels)) {
        rtr::set_global_error(std::string("rtr_write_exr: cannot write '") + path + "'");
        return RTR_INVALID_ARG;
    }
    return RTR_OK;
}

}  // extern "C"
