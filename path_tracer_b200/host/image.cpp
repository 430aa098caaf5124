// Host scene layer, part 5: image decoding for LoadTexture.
//
// The reference loads every texture and skybox through `stbi_loadf(Path, &W, &H, &N, 4)`
// (/root/reference/src/scene/scene.cpp:294-297; model import :718), i.e. whatever stb_image
// decodes, delivered as linear RGBA32F.  This file restates the parts of that contract the
// reference's assets use, from the published formats, with stb_image's conversions:
//   * PNG  (ISO/IEC 15948): all colour types and bit depths, tRNS, Adam7 interlace; inflate by zlib.
//          16-bit samples are reduced to their high byte, sub-byte grey is scaled to 0..255,
//          a tRNS colour key gives alpha 0 where the pixel equals it (stb_image.h: "convert_16_to_8",
//          "depth_scale_table", "compute_transparency").
//   * Radiance .hdr (RGBE, "#?RADIANCE" / "#?RGBE", FORMAT=32-bit_rle_rgbe, "-Y h +X w"):
//          new-style per-channel run-length scanlines and flat pixels; rgb = mantissa * 2^(e-136), alpha 1.
//   * JPEG (ITU T.81) baseline and progressive Huffman, 8-bit, 1, 3 or 4 (CMYK / YCCK) components, any sampling factors the
//          decoder below can upsample (1x1, 2x1, 1x2, 2x2 and generic integer factors).  stb_image's integer
//          IDCT, its triangle-filter chroma upsampling and its 20-bit fixed-point YCbCr conversion are
//          followed so that 8-bit output is identical with the reference's.
//   * TGA (types 1/2/3 and run-length 9/10/11; 8/15/16/24/32 bits) and BMP (1/4/8-bit paletted, 16/24/32-bit, BITFIELDS; no RLE),
//          with stb_image's conversions (5-5-5 scaling, bit-field replication, the "all alpha zero means opaque" rule).
//   * GIF (first frame, as stb_image composes it), PSD (flattened RGB image, raw or PackBits), Softimage PIC, PGM / PPM (binary P5 / P6, 8 and 16 bit) as stb_image reads them, quirks included (see decode_pnm_stb); PFM (PF, linear float,
//          rows bottom to top) is this repository's own addition.
// 8-bit sources become float the way stbi_loadf does it (stb_image.h "ldr_to_hdr"): colour channels
// pow(v / 255, 2.2), alpha v / 255.
// Pinned against the reference's own vendored stb_image.h compiled into oracle/_ref
// (tests/test_image_decode.py; golden files written by tests/golden/make_image_golden.py).
#include "scene.hpp"

#include <zlib.h>

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace ptb { namespace host {

namespace {

typedef unsigned char u8;
typedef unsigned short u16;

bool read_file(const char* path, std::vector<u8>& out) {
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

// 8-bit RGBA -> linear float RGBA as stbi_loadf delivers it.
void ldr_to_float(const std::vector<u8>& rgba, std::vector<vec4>& px) {
    float lut[256];
    for (int v = 0; v < 256; v++) lut[v] = (float)(std::pow((double)(v / 255.0f), (double)2.2f) * (double)1.0f);
    px.resize(rgba.size() / 4);
    for (size_t i = 0; i < px.size(); i++)
        px[i] = {lut[rgba[4 * i]], lut[rgba[4 * i + 1]], lut[rgba[4 * i + 2]], rgba[4 * i + 3] / 255.0f};
}

// ------------------------------------------------------------------ PNG
uint32_t be32(const u8* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

int paeth(int a, int b, int c) {
    int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// Unfilters one (sub)image of w x h pixels from `raw` and writes native-depth samples (channels per pixel) to `out`.
bool png_unfilter(const u8*& raw, const u8* end, uint32_t w, uint32_t h, int channels, int depth, std::vector<u16>& out) {
    const size_t bits = (size_t)channels * depth;
    const size_t stride = (w * bits + 7) / 8, bpp = bits >= 8 ? bits / 8 : 1;
    std::vector<u8> prev(stride, 0), cur(stride);
    out.resize((size_t)w * h * channels);
    for (uint32_t y = 0; y < h; y++) {
        if (raw + 1 + stride > end) return false;
        const int filter = *raw++;
        if (filter > 4) return false;
        for (size_t i = 0; i < stride; i++) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int x = raw[i];
            switch (filter) {
                case 1: x += a; break;
                case 2: x += b; break;
                case 3: x += (a + b) >> 1; break;
                case 4: x += paeth(a, b, c); break;
                default: break;
            }
            cur[i] = (u8)x;
        }
        raw += stride;
        u16* o = &out[(size_t)y * w * channels];
        const size_t n = (size_t)w * channels;
        if (depth == 8) for (size_t i = 0; i < n; i++) o[i] = cur[i];
        else if (depth == 16) for (size_t i = 0; i < n; i++) o[i] = (u16)((cur[2 * i] << 8) | cur[2 * i + 1]);
        else for (size_t i = 0; i < n; i++) {                      // 1, 2, 4 bits: most significant bits first
            const size_t bit = i * depth;
            o[i] = (u16)((cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1));
        }
        prev.swap(cur);
    }
    return true;
}

bool decode_png(const std::vector<u8>& file, int& W, int& H, std::vector<u8>& rgba) {
    static const u8 sig[8] = {137, 80, 78, 71, 13, 10, 26, 10};
    if (file.size() < 8 || std::memcmp(file.data(), sig, 8)) return false;
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = 0, interlace = 0;
    u8 palette[256][4];
    for (auto& p : palette) { p[0] = p[1] = p[2] = 0; p[3] = 255; }
    bool has_key = false, seen_ihdr = false;
    u16 key[3] = {0, 0, 0};
    std::vector<u8> idat;
    size_t pos = 8;
    while (pos + 8 <= file.size()) {
        const uint32_t len = be32(&file[pos]);
        const u8* type = &file[pos + 4];
        const u8* data = &file[pos + 8];
        if (pos + 12 + (size_t)len > file.size()) return false;
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len != 13) return false;
            w = be32(data); h = be32(data + 4);
            depth = data[8]; ctype = data[9]; interlace = data[12];
            if (data[10] != 0 || data[11] != 0 || interlace > 1) return false;
            if (w == 0 || h == 0 || w > (1u << 24) || h > (1u << 24)) return false;
            if (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16) return false;
            if (ctype != 0 && ctype != 2 && ctype != 3 && ctype != 4 && ctype != 6) return false;
            if (ctype == 3 && depth == 16) return false;
            if ((ctype == 2 || ctype == 4 || ctype == 6) && depth < 8) return false;
            seen_ihdr = true;
        } else if (!seen_ihdr) {
            return false;
        } else if (!std::memcmp(type, "PLTE", 4)) {
            if (len > 768 || len % 3) return false;
            for (uint32_t i = 0; i < len / 3; i++) { palette[i][0] = data[3 * i]; palette[i][1] = data[3 * i + 1]; palette[i][2] = data[3 * i + 2]; }
        } else if (!std::memcmp(type, "tRNS", 4)) {
            if (ctype == 3) { for (uint32_t i = 0; i < len && i < 256; i++) palette[i][3] = data[i]; }
            else if (ctype == 0 && len >= 2) { key[0] = (u16)((data[0] << 8) | data[1]); has_key = true; }
            else if (ctype == 2 && len >= 6) { for (int k = 0; k < 3; k++) key[k] = (u16)((data[2 * k] << 8) | data[2 * k + 1]); has_key = true; }
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), data, data + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)len;
    }
    if (!seen_ihdr || idat.empty()) return false;
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;

    // inflate: the exact decompressed size follows from the header
    static const int xorig[7] = {0, 4, 0, 2, 0, 1, 0}, yorig[7] = {0, 0, 4, 0, 2, 0, 1};
    static const int xspc[7] = {8, 8, 4, 4, 2, 2, 1}, yspc[7] = {8, 8, 8, 4, 4, 2, 2};
    size_t need = 0;
    if (!interlace) need = (size_t)h * (1 + ((size_t)w * channels * depth + 7) / 8);
    else for (int p = 0; p < 7; p++) {
        const size_t pw = (w - xorig[p] + xspc[p] - 1) / xspc[p], ph = (h - yorig[p] + yspc[p] - 1) / yspc[p];
        if (pw && ph) need += ph * (1 + (pw * channels * depth + 7) / 8);
    }
    std::vector<u8> raw(need);
    {
        z_stream zs;
        std::memset(&zs, 0, sizeof zs);
        if (inflateInit(&zs) != Z_OK) return false;
        zs.next_in = idat.data(); zs.avail_in = (uInt)idat.size();
        zs.next_out = raw.data(); zs.avail_out = (uInt)raw.size();
        const int rc = inflate(&zs, Z_FINISH);
        const size_t got = zs.total_out;
        inflateEnd(&zs);
        if ((rc != Z_STREAM_END && rc != Z_BUF_ERROR && rc != Z_OK) || got < need) return false;
    }

    std::vector<u16> samples((size_t)w * h * channels);
    const u8* rp = raw.data();
    const u8* rend = raw.data() + raw.size();
    if (!interlace) {
        if (!png_unfilter(rp, rend, w, h, channels, depth, samples)) return false;
    } else {
        std::vector<u16> sub;
        for (int p = 0; p < 7; p++) {
            const uint32_t pw = (w - xorig[p] + xspc[p] - 1) / xspc[p], ph = (h - yorig[p] + yspc[p] - 1) / yspc[p];
            if (!pw || !ph) continue;
            if (!png_unfilter(rp, rend, pw, ph, channels, depth, sub)) return false;
            for (uint32_t j = 0; j < ph; j++)
                for (uint32_t i = 0; i < pw; i++)
                    std::memcpy(&samples[(((size_t)j * yspc[p] + yorig[p]) * w + (size_t)i * xspc[p] + xorig[p]) * channels],
                                &sub[((size_t)j * pw + i) * channels], sizeof(u16) * channels);
        }
    }

    // to 8-bit RGBA
    static const int depth_scale[9] = {0, 0xff, 0x55, 0, 0x11, 0, 0, 0, 0x01};
    W = (int)w; H = (int)h;
    rgba.resize((size_t)w * h * 4);
    const size_t n = (size_t)w * h;
    auto to8 = [&](u16 v) -> u8 { return depth == 16 ? (u8)(v >> 8) : depth == 8 ? (u8)v : (u8)(v * depth_scale[depth]); };
    // the colour key is compared at 16 bits for 16-bit files and on the scaled 8-bit values otherwise
    u16 k8[3];
    for (int k = 0; k < 3; k++) k8[k] = depth == 16 ? key[k] : (u16)((key[k] & 255) * depth_scale[depth]);
    for (size_t i = 0; i < n; i++) {
        const u16* s = &samples[i * channels];
        u8* o = &rgba[4 * i];
        switch (ctype) {
            case 0: {
                o[0] = o[1] = o[2] = to8(s[0]);
                const u16 cmp = depth == 16 ? s[0] : (u16)to8(s[0]);
                o[3] = (has_key && cmp == k8[0]) ? 0 : 255;
                break;
            }
            case 2: {
                for (int k = 0; k < 3; k++) o[k] = to8(s[k]);
                bool eq = has_key;
                for (int k = 0; k < 3 && eq; k++) eq = (depth == 16 ? s[k] : (u16)to8(s[k])) == k8[k];
                o[3] = eq ? 0 : 255;
                break;
            }
            case 3: std::memcpy(o, palette[s[0] & 255], 4); break;
            case 4: o[0] = o[1] = o[2] = to8(s[0]); o[3] = to8(s[1]); break;
            default: for (int k = 0; k < 4; k++) o[k] = to8(s[k]); break;
        }
    }
    return true;
}

// ------------------------------------------------------------------ Radiance HDR
bool hdr_line(const std::vector<u8>& f, size_t& pos, std::string& line) {
    line.clear();
    if (pos >= f.size()) return false;
    while (pos < f.size() && f[pos] != '\n') { if (line.size() < 1023) line.push_back((char)f[pos]); pos++; }
    pos++;
    return true;
}

void rgbe_to_float(const u8* in, vec4& out) {
    if (in[3] != 0) {
        const float f1 = (float)std::ldexp(1.0f, (int)in[3] - (128 + 8));
        out = {in[0] * f1, in[1] * f1, in[2] * f1, 1.0f};
    } else {
        out = {0.0f, 0.0f, 0.0f, 1.0f};
    }
}

bool decode_hdr(const std::vector<u8>& file, int& W, int& H, std::vector<vec4>& px) {
    size_t pos = 0;
    std::string line;
    if (!hdr_line(file, pos, line) || (line != "#?RADIANCE" && line != "#?RGBE")) return false;
    bool format_ok = false;
    for (;;) {
        if (!hdr_line(file, pos, line)) return false;
        if (line.empty()) break;
        if (line == "FORMAT=32-bit_rle_rgbe") format_ok = true;
    }
    if (!format_ok || !hdr_line(file, pos, line)) return false;
    long h = 0, w = 0;
    {
        if (line.compare(0, 3, "-Y ")) return false;
        char* end = nullptr;
        h = std::strtol(line.c_str() + 3, &end, 10);
        while (*end == ' ') end++;
        if (std::strncmp(end, "+X ", 3)) return false;
        w = std::strtol(end + 3, nullptr, 10);
    }
    if (w <= 0 || h <= 0 || w > (1 << 24) || h > (1 << 24)) return false;
    W = (int)w; H = (int)h;
    px.resize((size_t)w * h);
    auto flat_from = [&](size_t first_pixel) -> bool {
        for (size_t i = first_pixel; i < (size_t)w * h; i++) {
            if (pos + 4 > file.size()) return false;
            rgbe_to_float(&file[pos], px[i]);
            pos += 4;
        }
        return true;
    };
    if (w < 8 || w >= 32768) return flat_from(0);
    std::vector<u8> scan((size_t)w * 4);
    for (long j = 0; j < h; j++) {
        if (pos + 4 > file.size()) return false;
        const u8 c1 = file[pos], c2 = file[pos + 1], len = file[pos + 2];
        if (c1 != 2 || c2 != 2 || (len & 0x80)) {
            // not run-length encoded: these four bytes are a pixel, and the rest of the file is flat
            if (j != 0) return false;
            return flat_from(0);
        }
        if ((((long)len << 8) | file[pos + 3]) != w) return false;
        pos += 4;
        for (int k = 0; k < 4; k++) {
            long i = 0;
            while (i < w) {
                if (pos >= file.size()) return false;
                int count = file[pos++];
                if (count > 128) {
                    if (pos >= file.size()) return false;
                    const u8 value = file[pos++];
                    count -= 128;
                    if (count == 0 || count > w - i) return false;
                    for (int z = 0; z < count; z++) scan[(size_t)(i++) * 4 + k] = value;
                } else {
                    if (count == 0 || count > w - i || pos + (size_t)count > file.size()) return false;
                    for (int z = 0; z < count; z++) scan[(size_t)(i++) * 4 + k] = file[pos++];
                }
            }
        }
        for (long i = 0; i < w; i++) rgbe_to_float(&scan[(size_t)i * 4], px[(size_t)j * w + i]);
    }
    return true;
}

// ------------------------------------------------------------------ JPEG (baseline + progressive Huffman)
// Follows ITU T.81 for the entropy decoding and stb_image's integer arithmetic for everything after it
// (dequantise, IDCT with 12-bit constants, triangle-filter upsampling, fixed-point YCbCr -> RGB).
struct jpeg_huff {
    u8 size[257];
    u16 code[256];
    u8 values[256];
    int maxcode[18];
    int delta[17];
    int count = 0;
    jpeg_huff() { std::memset(size, 0, sizeof size); std::memset(code, 0, sizeof code); std::memset(values, 0, sizeof values);
                  std::memset(maxcode, 0, sizeof maxcode); std::memset(delta, 0, sizeof delta); maxcode[17] = 0x7fffffff; }
    bool build(const int* counts) {
        int k = 0;
        for (int i = 0; i < 16; i++)
            for (int j = 0; j < counts[i]; j++) { if (k >= 256) return false; size[k++] = (u8)(i + 1); }
        size[k] = 0;
        count = k;
        int c = 0;
        k = 0;
        for (int j = 1; j <= 16; j++) {
            delta[j] = k - c;
            if (size[k] == j) {
                while (size[k] == j) code[k++] = (u16)(c++);
                if (c - 1 >= (1 << j)) return false;
            }
            maxcode[j] = c << (16 - j);
            c <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        return true;
    }
};

struct jpeg_component {
    int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0, dc_pred = 0;
    int w2 = 0, h2 = 0;                    // padded plane size in pixels
    int coeff_w = 0, coeff_h = 0;          // blocks
    std::vector<u8> plane;
    std::vector<short> coeff;              // progressive only
};

struct jpeg_decoder {
    const u8* p; const u8* end;
    jpeg_huff huff_dc[4], huff_ac[4];
    u16 dequant[4][64];
    jpeg_component comp[4];
    int ncomp = 0, width = 0, height = 0, hmax = 1, vmax = 1, mcu_x = 0, mcu_y = 0, mcu_w = 0, mcu_h = 0;
    bool progressive = false;
    int restart_interval = 0, todo = 0;
    int spec_start = 0, spec_end = 0, succ_high = 0, succ_low = 0, eob_run = 0;
    int scan_n = 0, order[4];
    int app14_transform = -1;
    bool jfif = false;
    // bit reader
    uint32_t code_buffer = 0; int code_bits = 0; int marker = 0xff; bool nomore = false;

    int get8() { return p < end ? *p++ : 0; }
    int get16() { int a = get8(); return (a << 8) | get8(); }

    void grow() {
        do {
            unsigned b = nomore ? 0 : (unsigned)get8();
            if (b == 0xff) {
                int c = get8();
                while (c == 0xff) c = get8();
                if (c != 0) { marker = c; nomore = true; return; }
            }
            code_buffer |= b << (24 - code_bits);
            code_bits += 8;
        } while (code_bits <= 24);
    }
    int huff_decode(const jpeg_huff& h) {
        if (code_bits < 16) grow();
        const unsigned temp = code_buffer >> 16;
        int k;
        for (k = 1; k < 17; k++) if ((int)temp < h.maxcode[k]) break;       // (a table the file never defined has no codes: k ends at 17)
        if (k == 17) { code_bits -= 16; return -1; }
        if (k > code_bits) return -1;
        const int c = (int)((code_buffer >> (32 - k)) & ((1u << k) - 1)) + h.delta[k];
        if (c < 0 || c >= 256) return -1;
        code_bits -= k;
        code_buffer <<= k;
        return h.values[c];
    }
    int extend_receive(int n) {              // n bits, sign-extended per T.81 F.2.2.1
        if (n == 0) return 0;
        if (code_bits < n) grow();
        if (code_bits < n) return 0;
        const int sgn = (int)(code_buffer >> 31);
        const unsigned k = (code_buffer << n) | (code_buffer >> (32 - n));      // rotate left
        const unsigned mask = (1u << n) - 1;
        code_buffer = k & ~mask;
        code_bits -= n;
        return (int)(k & mask) + ((sgn - 1) & (int)((~0u << n) + 1));
    }
    int get_bits(int n) {
        if (n == 0) return 0;
        if (code_bits < n) grow();
        if (code_bits < n) return 0;
        const unsigned k = (code_buffer << n) | (code_buffer >> (32 - n));
        const unsigned mask = (1u << n) - 1;
        code_buffer = k & ~mask;
        code_bits -= n;
        return (int)(k & mask);
    }
    int get_bit() {
        if (code_bits < 1) grow();
        if (code_bits < 1) return 0;
        const unsigned k = code_buffer;
        code_buffer <<= 1;
        --code_bits;
        return (int)(k & 0x80000000u);
    }
    void reset() {
        code_bits = 0; code_buffer = 0; nomore = false;
        for (int i = 0; i < 4; i++) comp[i].dc_pred = 0;
        marker = 0xff;
        todo = restart_interval ? restart_interval : 0x7fffffff;
        eob_run = 0;
    }
};

const u8 jpeg_dezigzag[64 + 15] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
    63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

bool jpeg_block_baseline(jpeg_decoder& j, short data[64], const jpeg_huff& hdc, const jpeg_huff& hac, int b, const u16* dq) {
    if (j.code_bits < 16) j.grow();
    const int t = j.huff_decode(hdc);
    if (t < 0 || t > 15) return false;
    std::memset(data, 0, 64 * sizeof(short));
    const int diff = t ? j.extend_receive(t) : 0;
    const int dc = j.comp[b].dc_pred + diff;
    j.comp[b].dc_pred = dc;
    data[0] = (short)(dc * dq[0]);
    int k = 1;
    do {
        const int rs = j.huff_decode(hac);
        if (rs < 0) return false;
        const int s = rs & 15, r = rs >> 4;
        if (s == 0) {
            if (rs != 0xf0) break;
            k += 16;
        } else {
            k += r;
            const unsigned zig = jpeg_dezigzag[k++];
            data[zig] = (short)(j.extend_receive(s) * dq[zig]);
        }
    } while (k < 64);
    return true;
}

bool jpeg_block_prog_dc(jpeg_decoder& j, short data[64], const jpeg_huff& hdc, int b) {
    if (j.spec_end != 0) return false;
    if (j.code_bits < 16) j.grow();
    if (j.succ_high == 0) {
        std::memset(data, 0, 64 * sizeof(short));
        const int t = j.huff_decode(hdc);
        if (t < 0 || t > 15) return false;
        const int diff = t ? j.extend_receive(t) : 0;
        const int dc = j.comp[b].dc_pred + diff;
        j.comp[b].dc_pred = dc;
        data[0] = (short)(dc * (1 << j.succ_low));
    } else {
        if (j.get_bit()) data[0] += (short)(1 << j.succ_low);
    }
    return true;
}

bool jpeg_block_prog_ac(jpeg_decoder& j, short data[64], const jpeg_huff& hac) {
    if (j.spec_start == 0) return false;
    if (j.succ_high == 0) {
        const int shift = j.succ_low;
        if (j.eob_run) { --j.eob_run; return true; }
        int k = j.spec_start;
        do {
            const int rs = j.huff_decode(hac);
            if (rs < 0) return false;
            const int s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (r < 15) {
                    j.eob_run = 1 << r;
                    if (r) j.eob_run += j.get_bits(r);
                    --j.eob_run;
                    break;
                }
                k += 16;
            } else {
                k += r;
                const unsigned zig = jpeg_dezigzag[k++];
                data[zig] = (short)(j.extend_receive(s) * (1 << shift));
            }
        } while (k <= j.spec_end);
    } else {
        const short bit = (short)(1 << j.succ_low);
        if (j.eob_run) {
            --j.eob_run;
            for (int k = j.spec_start; k <= j.spec_end; k++) {
                short* p = &data[jpeg_dezigzag[k]];
                if (*p != 0 && j.get_bit() && (*p & bit) == 0) { if (*p > 0) *p += bit; else *p -= bit; }
            }
        } else {
            int k = j.spec_start;
            do {
                const int rs = j.huff_decode(hac);
                if (rs < 0) return false;
                int s = rs & 15, r = rs >> 4;
                if (s == 0) {
                    if (r < 15) {
                        j.eob_run = (1 << r) - 1;
                        if (r) j.eob_run += j.get_bits(r);
                        r = 64;            // force the end of the block
                    }
                } else {
                    if (s != 1) return false;
                    s = j.get_bit() ? bit : -bit;
                }
                while (k <= j.spec_end) {
                    short* p = &data[jpeg_dezigzag[k++]];
                    if (*p != 0) {
                        if (j.get_bit() && (*p & bit) == 0) { if (*p > 0) *p += bit; else *p -= bit; }
                    } else {
                        if (r == 0) { *p = (short)s; break; }
                        --r;
                    }
                }
            } while (k <= j.spec_end);
        }
    }
    return true;
}

// (64-bit intermediates: identical results for every valid coefficient range, and no signed overflow on a corrupt file)
inline u8 jpeg_clamp(long long x) { return x < 0 ? 0 : x > 255 ? 255 : (u8)x; }

#define PTB_F2F(x) ((int)(((x) * 4096 + 0.5)))
#define PTB_FSH(x) ((x) * 4096)
#define PTB_IDCT_1D(s0, s1, s2, s3, s4, s5, s6, s7)            \
    long long t0, t1, t2, t3, p1, p2, p3, p4, p5, x0, x1, x2, x3; \
    p2 = s2; p3 = s6;                                           \
    p1 = (p2 + p3) * PTB_F2F(0.5411961f);                       \
    t2 = p1 + p3 * PTB_F2F(-1.847759065f);                      \
    t3 = p1 + p2 * PTB_F2F(0.765366865f);                       \
    p2 = s0; p3 = s4;                                           \
    t0 = PTB_FSH(p2 + p3);                                      \
    t1 = PTB_FSH(p2 - p3);                                      \
    x0 = t0 + t3; x3 = t0 - t3; x1 = t1 + t2; x2 = t1 - t2;     \
    t0 = s7; t1 = s5; t2 = s3; t3 = s1;                         \
    p3 = t0 + t2; p4 = t1 + t3; p1 = t0 + t3; p2 = t1 + t2;     \
    p5 = (p3 + p4) * PTB_F2F(1.175875602f);                     \
    t0 = t0 * PTB_F2F(0.298631336f);                            \
    t1 = t1 * PTB_F2F(2.053119869f);                            \
    t2 = t2 * PTB_F2F(3.072711026f);                            \
    t3 = t3 * PTB_F2F(1.501321110f);                            \
    p1 = p5 + p1 * PTB_F2F(-0.899976223f);                      \
    p2 = p5 + p2 * PTB_F2F(-2.562915447f);                      \
    p3 = p3 * PTB_F2F(-1.961570560f);                           \
    p4 = p4 * PTB_F2F(-0.390180644f);                           \
    t3 += p1 + p4; t2 += p2 + p3; t1 += p2 + p4; t0 += p1 + p3;

void jpeg_idct(u8* out, int out_stride, short data[64]) {
    long long val[64], *v = val;
    short* d = data;
    for (int i = 0; i < 8; ++i, ++d, ++v) {
        if (d[8] == 0 && d[16] == 0 && d[24] == 0 && d[32] == 0 && d[40] == 0 && d[48] == 0 && d[56] == 0) {
            const long long dcterm = d[0] * 4;
            v[0] = v[8] = v[16] = v[24] = v[32] = v[40] = v[48] = v[56] = dcterm;
        } else {
            PTB_IDCT_1D(d[0], d[8], d[16], d[24], d[32], d[40], d[48], d[56])
            x0 += 512; x1 += 512; x2 += 512; x3 += 512;
            v[0] = (x0 + t3) >> 10; v[56] = (x0 - t3) >> 10;
            v[8] = (x1 + t2) >> 10; v[48] = (x1 - t2) >> 10;
            v[16] = (x2 + t1) >> 10; v[40] = (x2 - t1) >> 10;
            v[24] = (x3 + t0) >> 10; v[32] = (x3 - t0) >> 10;
        }
    }
    v = val;
    u8* o = out;
    for (int i = 0; i < 8; ++i, v += 8, o += out_stride) {
        PTB_IDCT_1D(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7])
        x0 += 65536 + (128 << 17); x1 += 65536 + (128 << 17); x2 += 65536 + (128 << 17); x3 += 65536 + (128 << 17);
        o[0] = jpeg_clamp((x0 + t3) >> 17); o[7] = jpeg_clamp((x0 - t3) >> 17);
        o[1] = jpeg_clamp((x1 + t2) >> 17); o[6] = jpeg_clamp((x1 - t2) >> 17);
        o[2] = jpeg_clamp((x2 + t1) >> 17); o[5] = jpeg_clamp((x2 - t1) >> 17);
        o[3] = jpeg_clamp((x3 + t0) >> 17); o[4] = jpeg_clamp((x3 - t0) >> 17);
    }
}

// markers between scans; returns false on a malformed segment
bool jpeg_segment(jpeg_decoder& j, int m) {
    switch (m) {
        case 0xDD:
            if (j.get16() != 4) return false;
            j.restart_interval = j.get16();
            return true;
        case 0xDB: {
            int L = j.get16() - 2;
            while (L > 0) {
                const int q = j.get8(), prec = q >> 4, t = q & 15;
                if ((prec != 0 && prec != 1) || t > 3) return false;
                for (int i = 0; i < 64; i++) j.dequant[t][jpeg_dezigzag[i]] = (u16)(prec ? j.get16() : j.get8());
                L -= prec ? 129 : 65;
            }
            return L == 0;
        }
        case 0xC4: {
            int L = j.get16() - 2;
            while (L > 0) {
                int sizes[16], n = 0;
                const int q = j.get8(), tc = q >> 4, th = q & 15;
                if (tc > 1 || th > 3) return false;
                for (int i = 0; i < 16; i++) { sizes[i] = j.get8(); n += sizes[i]; }
                if (n > 256) return false;
                L -= 17;
                jpeg_huff& h = tc == 0 ? j.huff_dc[th] : j.huff_ac[th];
                if (!h.build(sizes)) return false;
                for (int i = 0; i < n; i++) h.values[i] = (u8)j.get8();
                L -= n;
            }
            return L == 0;
        }
        default: break;
    }
    if ((m >= 0xE0 && m <= 0xEF) || m == 0xFE) {
        int L = j.get16();
        if (L < 2) return false;
        L -= 2;
        if (m == 0xE0 && L >= 5) {
            static const u8 tag[5] = {'J', 'F', 'I', 'F', 0};
            bool ok = true;
            for (int i = 0; i < 5; i++) if (j.get8() != tag[i]) ok = false;
            L -= 5;
            if (ok) j.jfif = true;
        } else if (m == 0xEE && L >= 12) {
            static const u8 tag[6] = {'A', 'd', 'o', 'b', 'e', 0};
            bool ok = true;
            for (int i = 0; i < 6; i++) if (j.get8() != tag[i]) ok = false;
            L -= 6;
            if (ok) { j.get8(); j.get16(); j.get16(); j.app14_transform = j.get8(); L -= 6; }
        }
        j.p += L;
        if (j.p > j.end) j.p = j.end;
        return true;
    }
    return false;
}

bool jpeg_frame_header(jpeg_decoder& j) {
    const int Lf = j.get16();
    if (Lf < 11) return false;
    if (j.get8() != 8) return false;
    j.height = j.get16(); j.width = j.get16();
    if (j.height == 0 || j.width == 0) return false;
    const int c = j.get8();
    if (c != 1 && c != 3 && c != 4) return false;   // 4: CMYK / YCCK with the Adobe marker, combined as stb_image combines them
    j.ncomp = c;
    if (Lf != 8 + 3 * c) return false;
    for (int i = 0; i < c; i++) {
        j.comp[i].id = j.get8();
        const int q = j.get8();
        j.comp[i].h = q >> 4; j.comp[i].v = q & 15;
        if (j.comp[i].h < 1 || j.comp[i].h > 4 || j.comp[i].v < 1 || j.comp[i].v > 4) return false;
        j.comp[i].tq = j.get8();
        if (j.comp[i].tq > 3) return false;
    }
    j.hmax = j.vmax = 1;
    for (int i = 0; i < c; i++) { j.hmax = std::max(j.hmax, j.comp[i].h); j.vmax = std::max(j.vmax, j.comp[i].v); }
    for (int i = 0; i < c; i++) if (j.hmax % j.comp[i].h || j.vmax % j.comp[i].v) return false;
    j.mcu_w = j.hmax * 8; j.mcu_h = j.vmax * 8;
    j.mcu_x = (j.width + j.mcu_w - 1) / j.mcu_w;
    j.mcu_y = (j.height + j.mcu_h - 1) / j.mcu_h;
    for (int i = 0; i < c; i++) {
        jpeg_component& k = j.comp[i];
        k.w2 = j.mcu_x * k.h * 8; k.h2 = j.mcu_y * k.v * 8;
        k.plane.assign((size_t)k.w2 * k.h2, 0);
        if (j.progressive) {
            k.coeff_w = k.w2 / 8; k.coeff_h = k.h2 / 8;
            k.coeff.assign((size_t)k.w2 * k.h2, 0);
        }
    }
    return true;
}

bool jpeg_scan_header(jpeg_decoder& j) {
    const int Ls = j.get16();
    j.scan_n = j.get8();
    if (j.scan_n < 1 || j.scan_n > 4 || j.scan_n > j.ncomp) return false;
    if (Ls != 6 + 2 * j.scan_n) return false;
    for (int i = 0; i < j.scan_n; i++) {
        const int id = j.get8(), q = j.get8();
        int which = 0;
        for (; which < j.ncomp; which++) if (j.comp[which].id == id) break;
        if (which == j.ncomp) return false;
        j.comp[which].hd = q >> 4; j.comp[which].ha = q & 15;
        if (j.comp[which].hd > 3 || j.comp[which].ha > 3) return false;
        j.order[i] = which;
    }
    j.spec_start = j.get8(); j.spec_end = j.get8();
    const int aa = j.get8();
    j.succ_high = aa >> 4; j.succ_low = aa & 15;
    if (j.progressive) {
        if (j.spec_start > 63 || j.spec_end > 63 || j.spec_start > j.spec_end || j.succ_high > 13 || j.succ_low > 13) return false;
    } else {
        if (j.spec_start != 0 || j.succ_high != 0 || j.succ_low != 0) return false;
        j.spec_end = 63;
    }
    return true;
}

bool jpeg_restart(jpeg_decoder& j) {          // at the end of a restart interval: expect RSTn
    if (j.code_bits < 24) j.grow();
    if (!(j.marker >= 0xD0 && j.marker <= 0xD7)) return false;       // the scan ends here
    j.reset();
    return true;
}

bool jpeg_entropy_scan(jpeg_decoder& j) {
    j.reset();
    short data[64];
    if (!j.progressive) {
        if (j.scan_n == 1) {
            jpeg_component& k = j.comp[j.order[0]];
            const int n = j.order[0], w = (((j.width * k.h + j.hmax - 1) / j.hmax) + 7) >> 3,
                      h = (((j.height * k.v + j.vmax - 1) / j.vmax) + 7) >> 3;
            for (int y = 0; y < h; y++)
                for (int x = 0; x < w; x++) {
                    if (!jpeg_block_baseline(j, data, j.huff_dc[k.hd], j.huff_ac[k.ha], n, j.dequant[k.tq])) return false;
                    jpeg_idct(&k.plane[(size_t)k.w2 * y * 8 + x * 8], k.w2, data);
                    if (--j.todo <= 0) { if (!jpeg_restart(j)) return true; }
                }
            return true;
        }
        for (int my = 0; my < j.mcu_y; my++)
            for (int mx = 0; mx < j.mcu_x; mx++) {
                for (int c = 0; c < j.scan_n; c++) {
                    const int n = j.order[c];
                    jpeg_component& k = j.comp[n];
                    for (int y = 0; y < k.v; y++)
                        for (int x = 0; x < k.h; x++) {
                            const int x2 = (mx * k.h + x) * 8, y2 = (my * k.v + y) * 8;
                            if (!jpeg_block_baseline(j, data, j.huff_dc[k.hd], j.huff_ac[k.ha], n, j.dequant[k.tq])) return false;
                            jpeg_idct(&k.plane[(size_t)k.w2 * y2 + x2], k.w2, data);
                        }
                }
                if (--j.todo <= 0) { if (!jpeg_restart(j)) return true; }
            }
        return true;
    }
    if (j.scan_n == 1) {
        const int n = j.order[0];
        jpeg_component& k = j.comp[n];
        const int w = (((j.width * k.h + j.hmax - 1) / j.hmax) + 7) >> 3, h = (((j.height * k.v + j.vmax - 1) / j.vmax) + 7) >> 3;
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                short* blk = &k.coeff[64 * ((size_t)x + (size_t)y * k.coeff_w)];
                if (j.spec_start == 0) { if (!jpeg_block_prog_dc(j, blk, j.huff_dc[k.hd], n)) return false; }
                else if (!jpeg_block_prog_ac(j, blk, j.huff_ac[k.ha])) return false;
                if (--j.todo <= 0) { if (!jpeg_restart(j)) return true; }
            }
        return true;
    }
    for (int my = 0; my < j.mcu_y; my++)
        for (int mx = 0; mx < j.mcu_x; mx++) {
            for (int c = 0; c < j.scan_n; c++) {
                const int n = j.order[c];
                jpeg_component& k = j.comp[n];
                for (int y = 0; y < k.v; y++)
                    for (int x = 0; x < k.h; x++) {
                        const int x2 = mx * k.h + x, y2 = my * k.v + y;
                        short* blk = &k.coeff[64 * ((size_t)x2 + (size_t)y2 * k.coeff_w)];
                        if (!jpeg_block_prog_dc(j, blk, j.huff_dc[k.hd], n)) return false;
                    }
            }
            if (--j.todo <= 0) { if (!jpeg_restart(j)) return true; }
        }
    return true;
}

void jpeg_finish_progressive(jpeg_decoder& j) {
    for (int n = 0; n < j.ncomp; n++) {
        jpeg_component& k = j.comp[n];
        const int w = (((j.width * k.h + j.hmax - 1) / j.hmax) + 7) >> 3, h = (((j.height * k.v + j.vmax - 1) / j.vmax) + 7) >> 3;
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                short* blk = &k.coeff[64 * ((size_t)x + (size_t)y * k.coeff_w)];
                const u16* dq = j.dequant[k.tq];
                for (int i = 0; i < 64; i++) blk[i] = (short)(blk[i] * dq[i]);
                jpeg_idct(&k.plane[(size_t)k.w2 * y * 8 + x * 8], k.w2, blk);
            }
    }
}

// Row resamplers: `near` is the row closer to the output row, `far` the other one; w = input samples.
typedef u8* (*jpeg_resample_fn)(u8* out, u8* near, u8* far, int w, int hs);
u8* jpeg_resample_1(u8*, u8* near, u8*, int, int) { return near; }
u8* jpeg_resample_v2(u8* out, u8* near, u8* far, int w, int) {
    for (int i = 0; i < w; i++) out[i] = (u8)((3 * near[i] + far[i] + 2) >> 2);
    return out;
}
u8* jpeg_resample_h2(u8* out, u8* in, u8*, int w, int) {
    if (w == 1) { out[0] = out[1] = in[0]; return out; }
    out[0] = in[0];
    out[1] = (u8)((in[0] * 3 + in[1] + 2) >> 2);
    int i;
    for (i = 1; i < w - 1; i++) {
        const int n = 3 * in[i] + 2;
        out[i * 2 + 0] = (u8)((n + in[i - 1]) >> 2);
        out[i * 2 + 1] = (u8)((n + in[i + 1]) >> 2);
    }
    out[i * 2 + 0] = (u8)((in[w - 2] * 3 + in[w - 1] + 2) >> 2);
    out[i * 2 + 1] = in[w - 1];
    return out;
}
u8* jpeg_resample_hv2(u8* out, u8* near, u8* far, int w, int) {
    if (w == 1) { out[0] = out[1] = (u8)((3 * near[0] + far[0] + 2) >> 2); return out; }
    int t1 = 3 * near[0] + far[0], t0;
    out[0] = (u8)((t1 + 2) >> 2);
    for (int i = 1; i < w; i++) {
        t0 = t1;
        t1 = 3 * near[i] + far[i];
        out[i * 2 - 1] = (u8)((3 * t0 + t1 + 8) >> 4);
        out[i * 2] = (u8)((3 * t1 + t0 + 8) >> 4);
    }
    out[w * 2 - 1] = (u8)((t1 + 2) >> 2);
    return out;
}
u8* jpeg_resample_generic(u8* out, u8* near, u8*, int w, int hs) {
    for (int i = 0; i < w; i++) for (int k = 0; k < hs; k++) out[i * hs + k] = near[i];
    return out;
}

bool decode_jpeg(const std::vector<u8>& file, int& W, int& H, std::vector<u8>& rgba) {
    if (file.size() < 4 || file[0] != 0xFF || file[1] != 0xD8) return false;
    jpeg_decoder* jp = new jpeg_decoder();
    struct guard { jpeg_decoder* p; ~guard() { delete p; } } g{jp};
    jpeg_decoder& j = *jp;
    j.p = file.data() + 2; j.end = file.data() + file.size();
    std::memset(j.dequant, 0, sizeof j.dequant);
    auto next_marker = [&]() -> int {
        if (j.marker != 0xff) { const int m = j.marker; j.marker = 0xff; return m; }
        int x = j.get8();
        if (x != 0xff) return 0xff;
        while (x == 0xff) x = j.get8();
        return x;
    };
    int m = next_marker();
    bool have_frame = false;
    while (!(m == 0xC0 || m == 0xC1 || m == 0xC2)) {
        if (!jpeg_segment(j, m)) return false;
        m = next_marker();
        while (m == 0xff) { if (j.p >= j.end) return false; m = next_marker(); }
    }
    j.progressive = m == 0xC2;
    if (!jpeg_frame_header(j)) return false;
    have_frame = true;
    m = next_marker();
    while (m != 0xD9) {
        if (m == 0xDA) {
            if (!jpeg_scan_header(j)) return false;
            if (!jpeg_entropy_scan(j)) return false;
            if (j.marker == 0xff) {
                // skip padding / garbage up to the next marker
                while (j.p < j.end) {
                    const int x = j.get8();
                    if (x == 0xff) { j.marker = j.get8(); break; }
                }
            }
        } else if (m == 0xDC) {
            const int Ld = j.get16(), NL = j.get16();
            if (Ld != 4 || NL != j.height) return false;
        } else if (m == 0xff) {
            if (j.p >= j.end) break;
        } else if (!jpeg_segment(j, m)) {
            return false;
        }
        m = next_marker();
    }
    if (!have_frame) return false;
    if (j.progressive) jpeg_finish_progressive(j);

    // upsample to full resolution and convert to RGB
    W = j.width; H = j.height;
    rgba.resize((size_t)W * H * 4);
    struct resample { jpeg_resample_fn fn; u8 *line0, *line1; int hs, vs, w_lores, ystep, ypos; std::vector<u8> buf; } rs[4];
    for (int k = 0; k < j.ncomp; k++) {
        resample& r = rs[k];
        r.buf.assign((size_t)W + 3, 0);
        r.hs = j.hmax / j.comp[k].h; r.vs = j.vmax / j.comp[k].v;
        r.ystep = r.vs >> 1;
        r.w_lores = (W + r.hs - 1) / r.hs;
        r.ypos = 0;
        r.line0 = r.line1 = j.comp[k].plane.data();
        if (r.hs == 1 && r.vs == 1) r.fn = jpeg_resample_1;
        else if (r.hs == 1 && r.vs == 2) r.fn = jpeg_resample_v2;
        else if (r.hs == 2 && r.vs == 1) r.fn = jpeg_resample_h2;
        else if (r.hs == 2 && r.vs == 2) r.fn = jpeg_resample_hv2;
        else r.fn = jpeg_resample_generic;
    }
    // stb_image: three components are RGB as they are if their ids spell "RGB", or if an Adobe marker says "no transform" and
    // there is no JFIF marker; four are CMYK (transform 0), YCCK (transform 2) or YCbCr plus a channel that is ignored.
    const bool ids_rgb = j.ncomp == 3 && j.comp[0].id == 'R' && j.comp[1].id == 'G' && j.comp[2].id == 'B';
    const bool is_rgb = j.ncomp == 3 && (ids_rgb || (j.app14_transform == 0 && !j.jfif));
    auto blinn = [](int x, int m) { const unsigned t = (unsigned)x * (unsigned)m + 128u; return (u8)((t + (t >> 8)) >> 8); };
    auto ycbcr_row = [](u8* o, const u8* yr, const u8* cbr, const u8* crr, int count) {
        for (int x = 0; x < count; x++) {
            const int yf = (yr[x] << 20) + (1 << 19);
            const int cr = crr[x] - 128, cb = cbr[x] - 128;
            int r = yf + cr * (int)(1.40200f * 4096.0f + 0.5f) * 256;
            int gch = yf + cr * -((int)(0.71414f * 4096.0f + 0.5f) * 256) + ((cb * -((int)(0.34414f * 4096.0f + 0.5f) * 256)) & 0xffff0000);
            int b = yf + cb * (int)(1.77200f * 4096.0f + 0.5f) * 256;
            r >>= 20; gch >>= 20; b >>= 20;
            o[4 * x] = jpeg_clamp(r); o[4 * x + 1] = jpeg_clamp(gch); o[4 * x + 2] = jpeg_clamp(b); o[4 * x + 3] = 255;
        }
    };
    for (int y = 0; y < H; y++) {
        u8* rows[4] = {nullptr, nullptr, nullptr, nullptr};
        for (int k = 0; k < j.ncomp; k++) {
            resample& r = rs[k];
            const bool y_bot = r.ystep >= (r.vs >> 1);
            rows[k] = r.fn(r.buf.data(), y_bot ? r.line1 : r.line0, y_bot ? r.line0 : r.line1, r.w_lores, r.hs);
            if (++r.ystep >= r.vs) {
                r.ystep = 0;
                r.line0 = r.line1;
                if (++r.ypos < ((j.height * j.comp[k].v + j.vmax - 1) / j.vmax)) r.line1 += j.comp[k].w2;
            }
        }
        u8* o = &rgba[(size_t)y * W * 4];
        if (j.ncomp == 1) {
            for (int x = 0; x < W; x++) { o[4 * x] = o[4 * x + 1] = o[4 * x + 2] = rows[0][x]; o[4 * x + 3] = 255; }
        } else if (is_rgb) {
            for (int x = 0; x < W; x++) { o[4 * x] = rows[0][x]; o[4 * x + 1] = rows[1][x]; o[4 * x + 2] = rows[2][x]; o[4 * x + 3] = 255; }
        } else if (j.ncomp == 4 && j.app14_transform == 0) {             // CMYK: each ink times black
            for (int x = 0; x < W; x++) {
                const int m = rows[3][x];
                o[4 * x] = blinn(rows[0][x], m); o[4 * x + 1] = blinn(rows[1][x], m); o[4 * x + 2] = blinn(rows[2][x], m); o[4 * x + 3] = 255;
            }
        } else {
            ycbcr_row(o, rows[0], rows[1], rows[2], W);
            if (j.ncomp == 4 && j.app14_transform == 2)                 // YCCK: the complement of the converted colour times black
                for (int x = 0; x < W; x++) {
                    const int m = rows[3][x];
                    o[4 * x] = blinn(255 - o[4 * x], m); o[4 * x + 1] = blinn(255 - o[4 * x + 1], m); o[4 * x + 2] = blinn(255 - o[4 * x + 2], m);
                }
        }
    }
    return true;
}

// ------------------------------------------------------------------ TGA (Truevision, as stb_image reads it)
// Types 1 / 2 / 3 (colour-mapped, true colour, grey) and their run-length forms 9 / 10 / 11; 8, 15, 16, 24, 32 bits per pixel or
// palette entry; 15/16-bit colour is 5-5-5 scaled by 255/31; 16-bit grey is grey + alpha; rows bottom-up unless bit 5 of the
// descriptor is set.  One stb quirk is kept because files depend on it: the palette's "first entry" field is skipped as BYTES.
uint32_t le16(const u8* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }

bool tga_header_ok(const std::vector<u8>& f) {
    if (f.size() < 18) return false;
    const int cmap = f[1], type = f[2], pal_bits = f[7], bpp = f[16];
    if (cmap > 1) return false;
    if (cmap == 1) {
        if (type != 1 && type != 9) return false;
        if (pal_bits != 8 && pal_bits != 15 && pal_bits != 16 && pal_bits != 24 && pal_bits != 32) return false;
    } else if (type != 2 && type != 3 && type != 10 && type != 11) return false;
    if (le16(&f[12]) < 1 || le16(&f[14]) < 1) return false;
    if (cmap == 1 && bpp != 8 && bpp != 16) return false;
    return bpp == 8 || bpp == 15 || bpp == 16 || bpp == 24 || bpp == 32;
}

int tga_components(int bits, bool grey, bool& rgb16) {
    rgb16 = false;
    switch (bits) {
        case 8: return 1;
        case 16: if (grey) return 2;   // fall through: 5-5-5 colour
        case 15: rgb16 = true; return 3;
        case 24: case 32: return bits / 8;
        default: return 0;
    }
}

bool decode_tga(const std::vector<u8>& f, int& W, int& H, std::vector<u8>& rgba) {
    if (!tga_header_ok(f)) return false;
    const int id_len = f[0], cmap = f[1];
    int type = f[2];
    const uint32_t pal_start = le16(&f[3]), pal_len = le16(&f[5]);
    const int pal_bits = f[7], bpp = f[16];
    const uint32_t w = le16(&f[12]), h = le16(&f[14]);
    const bool bottom_up = !((f[17] >> 5) & 1);
    const bool rle = type >= 8;
    if (rle) type -= 8;
    const bool indexed = cmap == 1;
    bool rgb16 = false;
    const int comp = indexed ? tga_components(pal_bits, false, rgb16) : tga_components(bpp, type == 3, rgb16);
    if (!comp) return false;
    size_t pos = 18 + (size_t)id_len;
    auto need = [&](size_t n) { return pos + n <= f.size(); };
    auto read_555 = [&](u8* out) {
        const uint32_t px = le16(&f[pos]); pos += 2;
        out[0] = (u8)((((px >> 10) & 31) * 255) / 31); out[1] = (u8)((((px >> 5) & 31) * 255) / 31); out[2] = (u8)(((px & 31) * 255) / 31);
    };
    std::vector<u8> palette;
    if (indexed) {
        if (pal_len == 0) return false;
        pos += pal_start;
        palette.resize((size_t)pal_len * comp);
        if (rgb16) { if (!need((size_t)pal_len * 2)) return false; for (uint32_t i = 0; i < pal_len; i++) read_555(&palette[(size_t)i * comp]); }
        else { if (!need(palette.size())) return false; std::memcpy(palette.data(), &f[pos], palette.size()); pos += palette.size(); }
    }
    std::vector<u8> px((size_t)w * h * comp);
    u8 cur[4] = {0, 0, 0, 0};
    int count = 0;
    bool run = false, fetch = true;
    for (size_t i = 0; i < (size_t)w * h; i++) {
        if (rle) {
            if (count == 0) { if (!need(1)) return false; const int cmd = f[pos++]; count = 1 + (cmd & 127); run = (cmd >> 7) != 0; fetch = true; }
            else if (!run) fetch = true;
        } else fetch = true;
        if (fetch) {
            if (indexed) {
                if (!need(bpp == 8 ? 1 : 2)) return false;
                uint32_t idx = bpp == 8 ? f[pos] : le16(&f[pos]);
                pos += bpp == 8 ? 1 : 2;
                if (idx >= pal_len) idx = 0;
                for (int k = 0; k < comp; k++) cur[k] = palette[(size_t)idx * comp + k];
            } else if (rgb16) { if (!need(2)) return false; read_555(cur); }
            else { if (!need((size_t)comp)) return false; for (int k = 0; k < comp; k++) cur[k] = f[pos++]; }
            fetch = false;
        }
        for (int k = 0; k < comp; k++) px[i * comp + k] = cur[k];
        --count;
    }
    W = (int)w; H = (int)h;
    rgba.resize((size_t)w * h * 4);
    for (uint32_t y = 0; y < h; y++) {
        const u8* src = &px[(size_t)(bottom_up ? h - 1 - y : y) * w * comp];
        u8* o = &rgba[(size_t)y * w * 4];
        for (uint32_t x = 0; x < w; x++, src += comp, o += 4) {
            if (comp == 1) { o[0] = o[1] = o[2] = src[0]; o[3] = 255; }
            else if (comp == 2) { o[0] = o[1] = o[2] = src[0]; o[3] = src[1]; }
            else {
                const bool swap = !rgb16;                         // 24 / 32-bit pixels are stored blue first
                o[0] = src[swap ? 2 : 0]; o[1] = src[1]; o[2] = src[swap ? 0 : 2]; o[3] = comp == 4 ? src[3] : 255;
            }
        }
    }
    return true;
}

// ------------------------------------------------------------------ BMP (Windows / OS2 bitmaps, uncompressed or BITFIELDS)
int bmp_high_bit(uint32_t z) { int n = -1; while (z) { n++; z >>= 1; } return n; }
int bmp_bit_count(uint32_t a) { int n = 0; while (a) { n += a & 1; a >>= 1; } return n; }
// a masked field of `bits` bits scaled to 0..255 by bit replication
int bmp_field(uint32_t v, int shift, int bits) {
    static const uint32_t mul[9] = {0, 0xff, 0x55, 0x49, 0x11, 0x21, 0x41, 0x81, 0x01};
    static const uint32_t shr[9] = {0, 0, 0, 1, 0, 2, 4, 6, 0};
    if (shift < 0) v <<= -shift; else v >>= shift;
    if (bits < 1 || bits > 8) return 0;
    v >>= (8 - bits);
    return (int)((v * mul[bits]) >> shr[bits]);
}

bool decode_bmp(const std::vector<u8>& f, int& W, int& H, std::vector<u8>& rgba) {
    if (f.size() < 26 || f[0] != 'B' || f[1] != 'M') return false;
    auto r16 = [&](size_t p) -> uint32_t { return p + 2 <= f.size() ? le16(&f[p]) : 0u; };
    auto r32 = [&](size_t p) -> uint32_t { return p + 4 <= f.size() ? (le16(&f[p]) | (le16(&f[p + 2]) << 16)) : 0u; };
    const uint32_t offset = r32(10), hsz = r32(14);
    if (hsz != 12 && hsz != 40 && hsz != 56 && hsz != 108 && hsz != 124) return false;
    int width, height;
    uint32_t bpp, compress = 0;
    uint32_t mr = 0, mg = 0, mb = 0, ma = 0;
    bool all_alpha_unknown = false;                    // 32 bits without masks: alpha counts only if some pixel has any
    if (hsz == 12) { width = (int)r16(18); height = (int)r16(20); if (r16(22) != 1) return false; bpp = r16(24); }
    else {
        width = (int)r32(18); height = (int)r32(22);
        if (r16(26) != 1) return false;
        bpp = r16(28); compress = r32(30);
        if (compress == 1 || compress == 2) return false;                       // run-length forms
        if (compress >= 4) return false;
        if (compress == 3 && bpp != 16 && bpp != 32) return false;
        if (hsz == 40 || hsz == 56) {
            if (bpp == 16 || bpp == 32) {
                if (compress == 0) {
                    if (bpp == 32) { mr = 0xffu << 16; mg = 0xffu << 8; mb = 0xffu; ma = 0xffu << 24; all_alpha_unknown = true; }
                    else { mr = 31u << 10; mg = 31u << 5; mb = 31u; }
                } else {
                    const size_t m = hsz == 56 ? 54 : 54;
                    mr = r32(m); mg = r32(m + 4); mb = r32(m + 8);
                    if (mr == mg && mg == mb) return false;
                }
            }
        } else {
            mr = r32(54); mg = r32(58); mb = r32(62); ma = r32(66);
            if (compress != 3 && (bpp == 16 || bpp == 32)) {
                if (bpp == 32) { mr = 0xffu << 16; mg = 0xffu << 8; mb = 0xffu; ma = 0xffu << 24; all_alpha_unknown = true; }
                else { mr = 31u << 10; mg = 31u << 5; mb = 31u; ma = 0; }
            }
        }
    }
    if (width <= 0 || height == 0 || width > (1 << 24)) return false;
    const bool flip = height > 0;
    const uint32_t h = (uint32_t)(height < 0 ? -(long long)height : height), w = (uint32_t)width;
    if (h > (1u << 24)) return false;
    // header bytes consumed before the palette / pixel data: 14 + hsz (+ 12 for the masks that follow a 40-byte BITFIELDS header)
    size_t header_end = 14 + (size_t)hsz + ((hsz == 40 && compress == 3) ? 12 : 0);
    int psize = 0;
    if (hsz == 12) { if (bpp < 24) psize = (int)((offset - header_end) / 3); }
    else if (bpp < 16) psize = (int)((offset - header_end) >> 2);
    if (offset < header_end || offset > f.size()) return false;
    W = (int)w; H = (int)h;
    rgba.assign((size_t)w * h * 4, 255);
    bool any_alpha = false;
    if (bpp < 16) {
        if (psize == 0 || psize > 256) return false;
        u8 pal[256][4];
        size_t pos = header_end;
        for (int i = 0; i < psize; i++) {
            if (pos + 3 > f.size()) return false;
            pal[i][2] = f[pos]; pal[i][1] = f[pos + 1]; pal[i][0] = f[pos + 2]; pal[i][3] = 255;
            pos += hsz == 12 ? 3 : 4;
        }
        if (bpp != 1 && bpp != 4 && bpp != 8) return false;
        const size_t row = (((size_t)w * bpp + 7) / 8 + 3) & ~(size_t)3;
        if (offset + row * h > f.size()) return false;
        for (uint32_t y = 0; y < h; y++) {
            const u8* src = &f[offset + row * y];
            u8* o = &rgba[(size_t)(flip ? h - 1 - y : y) * w * 4];
            for (uint32_t x = 0; x < w; x++) {
                int v;
                if (bpp == 8) v = src[x];
                else if (bpp == 4) v = (src[x >> 1] >> ((x & 1) ? 0 : 4)) & 15;
                else v = (src[x >> 3] >> (7 - (x & 7))) & 1;
                if (v >= psize) v = 0;
                o[4 * x] = pal[v][0]; o[4 * x + 1] = pal[v][1]; o[4 * x + 2] = pal[v][2]; o[4 * x + 3] = 255;
            }
        }
        return true;
    }
    if (bpp != 16 && bpp != 24 && bpp != 32) return false;
    const size_t bytes = bpp / 8, row = ((size_t)w * bytes + 3) & ~(size_t)3;
    if (offset + (row * (h - 1) + (size_t)w * bytes) > f.size()) return false;
    const bool bgr24 = bpp == 24, bgra32 = bpp == 32 && mb == 0xffu && mg == 0xff00u && mr == 0x00ff0000u && ma == 0xff000000u;
    int rs = 0, gs = 0, bs = 0, as = 0, rc = 0, gc = 0, bc = 0, ac = 0;
    if (!bgr24 && !bgra32) {
        if (!mr || !mg || !mb) return false;
        rs = bmp_high_bit(mr) - 7; rc = bmp_bit_count(mr); gs = bmp_high_bit(mg) - 7; gc = bmp_bit_count(mg);
        bs = bmp_high_bit(mb) - 7; bc = bmp_bit_count(mb); as = bmp_high_bit(ma) - 7; ac = bmp_bit_count(ma);
        if (rc > 8 || gc > 8 || bc > 8 || ac > 8) return false;
    }
    const bool has_alpha = ma != 0;
    for (uint32_t y = 0; y < h; y++) {
        const u8* src = &f[offset + row * y];
        u8* o = &rgba[(size_t)(flip ? h - 1 - y : y) * w * 4];
        for (uint32_t x = 0; x < w; x++, src += bytes, o += 4) {
            if (bgr24 || bgra32) {
                o[0] = src[2]; o[1] = src[1]; o[2] = src[0];
                const u8 a = bgra32 ? src[3] : 255;
                any_alpha |= a != 0;
                o[3] = has_alpha ? a : 255;
            } else {
                const uint32_t v = bpp == 16 ? le16(src) : (le16(src) | (le16(src + 2) << 16));
                o[0] = (u8)bmp_field(v & mr, rs, rc); o[1] = (u8)bmp_field(v & mg, gs, gc); o[2] = (u8)bmp_field(v & mb, bs, bc);
                const int a = ma ? bmp_field(v & ma, as, ac) : 255;
                any_alpha |= a != 0;
                o[3] = has_alpha ? (u8)a : 255;
            }
        }
    }
    if (has_alpha && all_alpha_unknown && !any_alpha) for (size_t i = 3; i < rgba.size(); i += 4) rgba[i] = 255;
    return true;
}

// ------------------------------------------------------------------ GIF, the first frame, as stb_image returns it
// stb_image.h stbi__gif_load / stbi__gif_load_next / stbi__process_gif_raster.  Always four components.  What it does beyond the
// format: pixels the frame does not draw - outside its rectangle, or drawn with the transparent index - stay (0, 0, 0, 0), unless
// the header names a background index > 0, in which case every untouched pixel gets that palette entry COPIED AS STORED, and the
// palette is stored blue-first (so red and blue are swapped in a background fill - a stb_image quirk files render with in the
// reference).  The transparent index of a graphic-control extension clears the alpha of the GLOBAL palette entry; a local palette
// gets it when it is parsed.  A stream that ends (0x3B) before any image, a raster without a leading clear code, more than 8192
// codes, or a frame rectangle outside the screen is refused.
bool decode_gif(const std::vector<u8>& f, int& W, int& H, std::vector<u8>& rgba) {
    size_t pos = 0;
    auto get8 = [&]() -> int { return pos < f.size() ? f[pos++] : 0; };
    auto get16 = [&]() -> int { int a = get8(); return a | (get8() << 8); };
    auto skip = [&](int n) { pos = std::min(f.size(), pos + (size_t)std::max(0, n)); };
    if (get8() != 'G' || get8() != 'I' || get8() != 'F' || get8() != '8') return false;
    const int version = get8();
    if (version != '7' && version != '9') return false;
    if (get8() != 'a') return false;
    const int w = get16(), h = get16(), flags = get8(), bgindex = get8();
    get8();                                                    // aspect ratio
    if (w > (1 << 24) || h > (1 << 24)) return false;
    u8 pal[256][4] = {}, lpal[256][4] = {};
    auto colortable = [&](u8 (*table)[4], int entries, int transparent) {
        for (int i = 0; i < entries; i++) {
            table[i][2] = (u8)get8(); table[i][1] = (u8)get8(); table[i][0] = (u8)get8();     // stored blue, green, red
            table[i][3] = transparent == i ? 0 : 255;
        }
    };
    if (flags & 0x80) colortable(pal, 2 << (flags & 7), -1);
    if ((uint64_t)w * (uint64_t)h > ((uint64_t)1 << 29)) return false;
    const size_t pcount = (size_t)w * h;
    std::vector<u8> out(pcount * 4, 0), history(pcount, 0);
    int transparent = -1, eflags = 0;
    for (;;) {
        const int tag = get8();
        if (tag == 0x2C) {
            const int x = get16(), y = get16(), fw = get16(), fh = get16();
            if (x + fw > w || y + fh > h) return false;
            const long long line = (long long)w * 4;
            const long long start_x = (long long)x * 4, start_y = (long long)y * line, max_x = start_x + (long long)fw * 4, max_y = start_y + (long long)fh * line;
            long long cur_x = start_x, cur_y = fw == 0 ? max_y : start_y, step;
            int parse;
            const int lflags = get8();
            if (lflags & 0x40) { step = 8 * line; parse = 3; } else { step = line; parse = 0; }
            u8 (*table)[4];
            if (lflags & 0x80) { colortable(lpal, 2 << (lflags & 7), (eflags & 1) ? transparent : -1); table = lpal; }
            else if (flags & 0x80) table = pal;
            else return false;
            // LZW raster
            const int lzw_cs = get8();
            if (lzw_cs > 12) return false;
            struct code_t { int16_t prefix; u8 first, suffix; };
            std::vector<code_t> codes(8192);
            const int clear = 1 << lzw_cs;
            bool first = true;
            int codesize = lzw_cs + 1, codemask = (1 << codesize) - 1, avail = clear + 2, oldcode = -1, len = 0, valid_bits = 0;
            int32_t bits = 0;
            for (int i = 0; i < clear; i++) { codes[i].prefix = -1; codes[i].first = (u8)i; codes[i].suffix = (u8)i; }
            std::vector<u8> chain;
            auto emit = [&](int code) {                       // the string of `code`, first symbol first
                chain.clear();
                for (int c = code; c >= 0; c = codes[c].prefix) chain.push_back(codes[c].suffix);
                for (size_t k = chain.size(); k-- > 0;) {
                    if (cur_y >= max_y) continue;
                    const long long idx = cur_x + cur_y;
                    u8* p = &out[(size_t)idx];
                    history[(size_t)(idx / 4)] = 1;
                    const u8* c = table[chain[k]];
                    if (c[3] > 128) { p[0] = c[2]; p[1] = c[1]; p[2] = c[0]; p[3] = c[3]; }
                    cur_x += 4;
                    if (cur_x >= max_x) {
                        cur_x = start_x;
                        cur_y += step;
                        while (cur_y >= max_y && parse > 0) {
                            step = ((long long)1 << parse) * line;
                            cur_y = start_y + (step >> 1);
                            --parse;
                        }
                    }
                }
            };
            bool done = false;
            while (!done) {
                if (valid_bits < codesize) {
                    if (len == 0) {
                        len = get8();
                        if (len == 0) break;                   // raster ends with the blocks
                    }
                    --len;
                    bits |= (int32_t)((uint32_t)get8() << valid_bits);
                    valid_bits += 8;
                } else {
                    const int code = bits & codemask;
                    bits >>= codesize;
                    valid_bits -= codesize;
                    if (code == clear) {
                        codesize = lzw_cs + 1; codemask = (1 << codesize) - 1; avail = clear + 2; oldcode = -1; first = false;
                    } else if (code == clear + 1) {
                        skip(len);
                        while ((len = get8()) > 0) skip(len);
                        done = true;
                    } else if (code <= avail) {
                        if (first) return false;
                        if (oldcode >= 0) {
                            if (avail + 1 > 8192) return false;
                            code_t& n = codes[avail++];
                            n.prefix = (int16_t)oldcode;
                            n.first = codes[oldcode].first;
                            n.suffix = (code == avail) ? n.first : codes[code].first;
                        } else if (code == avail) return false;
                        emit(code);
                        if ((avail & codemask) == 0 && avail <= 0x0FFF) { codesize++; codemask = (1 << codesize) - 1; }
                        oldcode = code;
                    } else return false;
                }
            }
            if (bgindex > 0)
                for (size_t pi = 0; pi < pcount; pi++)
                    if (!history[pi]) { pal[bgindex][3] = 255; std::memcpy(&out[pi * 4], pal[bgindex], 4); }
            W = w; H = h;
            rgba.swap(out);
            return w > 0 && h > 0;
        } else if (tag == 0x21) {
            const int ext = get8();
            int len;
            if (ext == 0xF9) {
                len = get8();
                if (len == 4) {
                    eflags = get8();
                    get16();                                   // delay
                    if (transparent >= 0) pal[transparent][3] = 255;
                    if (eflags & 1) { transparent = get8(); pal[transparent][3] = 0; }
                    else { skip(1); transparent = -1; }
                } else { skip(len); continue; }
            }
            while ((len = get8()) != 0) skip(len);
        } else {
            return false;                                      // 0x3B before any image, or an unknown block
        }
    }
}

// ------------------------------------------------------------------ PSD, the flattened image, as stb_image reads it
// stb_image.h stbi__psd_load (8-bit output).  "8BPS" version 1, RGB mode only, 8 or 16 bits per channel, raw or PackBits rows
// (the run-length reader takes one BYTE per pixel whatever the depth, as stb_image's does), channels stored one after the other; the
// first four are used, missing ones are 0 (colour) or 255 (alpha), 16-bit samples keep their high byte.  With an alpha channel the
// colour is "un-matted" from white, in float: c' = c / a + 255 (1 - 1 / a) for 0 < a < 255, converted to a byte the way the
// compiler converts it (truncation to a 32-bit integer, low byte).  Reads past the end of the file yield zeros, as in stb_image.
bool decode_psd(const std::vector<u8>& f, int& W, int& H, std::vector<u8>& rgba) {
    size_t pos = 0;
    auto get8 = [&]() -> uint32_t { return pos < f.size() ? f[pos++] : 0u; };
    auto get16 = [&]() -> uint32_t { uint32_t a = get8(); return (a << 8) | get8(); };
    auto get32 = [&]() -> uint32_t { uint32_t a = get16(); return (a << 16) | get16(); };
    auto skip = [&](uint32_t n) { pos = (size_t)std::min<uint64_t>(f.size(), (uint64_t)pos + n); };
    if (get32() != 0x38425053u) return false;
    if (get16() != 1) return false;
    skip(6);
    const int channels = (int)get16();
    if (channels < 0 || channels > 16) return false;
    const uint32_t h = get32(), w = get32();
    if (h > (1u << 24) || w > (1u << 24)) return false;
    const int depth = (int)get16();
    if (depth != 8 && depth != 16) return false;
    if (get16() != 3) return false;
    skip(get32());
    skip(get32());
    skip(get32());
    const int compression = (int)get16();
    if (compression > 1) return false;
    if ((uint64_t)w * h > ((uint64_t)1 << 29) || w == 0 || h == 0) return false;
    const size_t count = (size_t)w * h;
    rgba.assign(count * 4, 0);
    if (compression) {
        skip((uint32_t)std::min<uint64_t>(0xFFFFFFFFu, (uint64_t)h * channels * 2));
        for (int c = 0; c < 4; c++) {
            u8* p = rgba.data() + c;
            if (c >= channels) { for (size_t i = 0; i < count; i++, p += 4) *p = c == 3 ? 255 : 0; continue; }
            size_t done = 0;
            while (done < count) {
                const size_t left = count - done;
                uint32_t len = get8();
                if (len == 128) continue;
                if (len < 128) {
                    len++;
                    if (len > left) return false;
                    done += len;
                    while (len--) { *p = (u8)get8(); p += 4; }
                } else {
                    len = 257 - len;
                    if (len > left) return false;
                    const u8 v = (u8)get8();
                    done += len;
                    while (len--) { *p = v; p += 4; }
                }
            }
        }
    } else {
        for (int c = 0; c < 4; c++) {
            u8* p = rgba.data() + c;
            if (c >= channels) { for (size_t i = 0; i < count; i++, p += 4) *p = c == 3 ? 255 : 0; }
            else if (depth == 16) { for (size_t i = 0; i < count; i++, p += 4) *p = (u8)(get16() >> 8); }
            else { for (size_t i = 0; i < count; i++, p += 4) *p = (u8)get8(); }
        }
    }
    if (channels >= 4) {
        auto to_byte = [](float v) -> u8 {                    // (unsigned char) of a float: cvttss2si, low byte
            const int32_t t = (v >= -2147483648.0f && v < 2147483648.0f) ? (int32_t)v : INT32_MIN;
            return (u8)((uint32_t)t & 0xFFu);
        };
        for (size_t i = 0; i < count; i++) {
            u8* px = &rgba[4 * i];
            if (px[3] != 0 && px[3] != 255) {
                const float a = px[3] / 255.0f, ra = 1.0f / a, inv_a = 255.0f * (1 - ra);
                px[0] = to_byte(px[0] * ra + inv_a); px[1] = to_byte(px[1] * ra + inv_a); px[2] = to_byte(px[2] * ra + inv_a);
            }
        }
    }
    W = (int)w; H = (int)h;
    return true;
}

// ------------------------------------------------------------------ Softimage PIC, as stb_image reads it
// stb_image.h stbi__pic_load / stbi__pic_load_core.  Magic 53 80 F6 34 and "PICT" at byte 88; width and height (big-endian 16 bit) at
// byte 92; from byte 104 a chain of up to ten packets {chained, bits (8), type, channel mask: 0x80 R, 0x40 G, 0x20 B, 0x10 A}, then for
// every row the data of every packet in turn: type 0 raw, 1 runs of {count, value}, 2 mixed (count >= 128: a run of count - 127, or of a
// 16-bit count after 128; else count + 1 raw values).  Channels no packet carries stay 255.
bool decode_pic(const std::vector<u8>& f, int& W, int& H, std::vector<u8>& rgba) {
    if (f.size() < 104 || f[0] != 0x53 || f[1] != 0x80 || f[2] != 0xF6 || f[3] != 0x34 || std::memcmp(&f[88], "PICT", 4) != 0) return false;
    size_t pos = 92;
    auto at_end = [&]() { return pos >= f.size(); };
    auto get8 = [&]() -> int { return pos < f.size() ? f[pos++] : 0; };
    auto get16 = [&]() -> int { int a = get8(); return (a << 8) | get8(); };
    const int w = get16(), h = get16();
    if (at_end()) return false;
    pos += 8;                                                  // ratio, fields, pad
    struct packet { int size, type, channel; } packets[10];
    int count = 0, chained;
    do {
        if (count == 10) return false;
        packet& p = packets[count++];
        chained = get8(); p.size = get8(); p.type = get8(); p.channel = get8();
        if (at_end()) return false;
        if (p.size != 8) return false;
    } while (chained);
    if (w == 0 || h == 0) return false;
    rgba.assign((size_t)w * h * 4, 0xFF);
    auto read_value = [&](int channel, u8* dest) {
        for (int i = 0, mask = 0x80; i < 4; i++, mask >>= 1)
            if (channel & mask) { if (at_end()) return false; dest[i] = (u8)get8(); }
        return true;
    };
    auto copy_value = [](int channel, u8* dest, const u8* src) {
        for (int i = 0, mask = 0x80; i < 4; i++, mask >>= 1) if (channel & mask) dest[i] = src[i];
    };
    for (int y = 0; y < h; y++)
        for (int k = 0; k < count; k++) {
            const packet& p = packets[k];
            u8* dest = &rgba[(size_t)y * w * 4];
            if (p.type == 0) {
                for (int x = 0; x < w; x++, dest += 4) if (!read_value(p.channel, dest)) return false;
            } else if (p.type == 1) {
                int left = w;
                while (left > 0) {
                    int n = get8();
                    if (at_end()) return false;
                    if (n > left) n = (u8)left;
                    u8 value[4] = {0, 0, 0, 0};
                    if (!read_value(p.channel, value)) return false;
                    for (int i = 0; i < n; i++, dest += 4) copy_value(p.channel, dest, value);
                    left -= n;
                }
            } else if (p.type == 2) {
                int left = w;
                while (left > 0) {
                    int n = get8();
                    if (at_end()) return false;
                    if (n >= 128) {
                        n = n == 128 ? get16() : n - 127;
                        if (n > left) return false;
                        u8 value[4] = {0, 0, 0, 0};
                        if (!read_value(p.channel, value)) return false;
                        for (int i = 0; i < n; i++, dest += 4) copy_value(p.channel, dest, value);
                    } else {
                        n++;
                        if (n > left) return false;
                        for (int i = 0; i < n; i++, dest += 4) if (!read_value(p.channel, dest)) return false;
                    }
                    left -= n;
                }
            } else {
                return false;
            }
        }
    W = w; H = h;
    return true;
}

// ------------------------------------------------------------------ PGM / PPM as stb_image reads them, and PFM
// stb_image.h:7491-7618.  "P5" (grey) and "P6" (RGB), binary.  Header: width, height, maximum value, separated by blanks
// (space, \t, \n, \v, \f, \r) and comments (# to the end of the line); ONE byte after the maximum value, then the samples.
// The maximum value only chooses the sample size (<= 255: one byte, <= 65535: two) - samples are not rescaled by it - and
// two-byte samples are read in HOST byte order and reduced to their high byte, which on a little-endian host is the SECOND
// byte of the big-endian pair the format prescribes.  Both quirks are stb_image's; files decode as they do in the reference.
// PFM ("PF", little-endian, rows bottom to top, linear float) is this repository's own addition: stb_image does not read it.
bool pnm_blank(u8 c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }

bool decode_pnm_stb(const std::vector<u8>& f, int& W, int& H, std::vector<u8>& rgba) {
    if (f.size() < 3 || f[0] != 'P' || (f[1] != '5' && f[1] != '6')) return false;
    const int comp = f[1] == '6' ? 3 : 1;
    size_t pos = 2;
    auto get = [&]() -> u8 { return pos < f.size() ? f[pos++] : 0; };
    auto at_end = [&]() { return pos >= f.size(); };
    u8 c = get();
    auto skip = [&]() {
        for (;;) {
            while (!at_end() && pnm_blank(c)) c = get();
            if (at_end() || c != '#') break;
            while (!at_end() && c != '\n' && c != '\r') c = get();
        }
    };
    bool overflow = false;
    auto integer = [&]() {
        long long v = 0;
        while (!at_end() && c >= '0' && c <= '9') {
            v = v * 10 + (c - '0');
            c = get();
            if (v > 214748364 || (v == 214748364 && c > '7')) { overflow = true; return 0; }
        }
        return (int)v;
    };
    skip();
    const int w = integer();
    if (overflow || w == 0) return false;
    skip();
    const int h = integer();
    if (overflow || h == 0) return false;
    skip();
    const int maxv = integer();
    if (overflow || maxv > 65535) return false;
    const int bytes = maxv > 255 ? 2 : 1;
    if (w > (1 << 24) || h > (1 << 24)) return false;
    const size_t count = (size_t)w * h * comp;
    if (count > ((size_t)1 << 31) / bytes || pos + count * bytes > f.size()) return false;
    rgba.resize((size_t)w * h * 4);
    for (size_t i = 0; i < (size_t)w * h; i++) {
        u8 v[3];
        for (int k = 0; k < comp; k++) {
            const size_t at = pos + (i * comp + k) * bytes;
            if (bytes == 1) v[k] = f[at];
            else { uint16_t native; std::memcpy(&native, &f[at], 2); v[k] = (u8)(native >> 8); }
        }
        if (comp == 1) v[1] = v[2] = v[0];
        rgba[4 * i] = v[0]; rgba[4 * i + 1] = v[1]; rgba[4 * i + 2] = v[2]; rgba[4 * i + 3] = 255;
    }
    W = w; H = h;
    return true;
}

bool pnm_token(const std::vector<u8>& f, size_t& pos, std::string& tok) {
    tok.clear();
    while (pos < f.size()) {
        if (f[pos] == '#') { while (pos < f.size() && f[pos] != '\n') pos++; }
        else if (std::isspace(f[pos])) pos++;
        else break;
    }
    while (pos < f.size() && !std::isspace(f[pos])) tok.push_back((char)f[pos++]);
    return !tok.empty();
}

bool decode_pfm(const std::vector<u8>& file, int& W, int& H, std::vector<vec4>& px) {
    size_t pos = 0;
    std::string magic, a, b, c;
    if (!pnm_token(file, pos, magic) || !pnm_token(file, pos, a) || !pnm_token(file, pos, b) || !pnm_token(file, pos, c)) return false;
    const int w = std::atoi(a.c_str()), h = std::atoi(b.c_str());
    if (magic != "PF" || w <= 0 || h <= 0) return false;
    pos++;                                   // the single whitespace byte that ends the header
    const float scale = (float)std::atof(c.c_str());
    if (!(scale < 0) || pos + (size_t)w * h * 12 > file.size()) return false;      // little-endian only
    px.resize((size_t)w * h);
    for (int y = 0; y < h; y++)                                           // PFM rows run bottom to top
        for (int x = 0; x < w; x++) {
            float s[3];
            std::memcpy(s, &file[pos + ((size_t)(h - 1 - y) * w + x) * 12], 12);
            px[(size_t)y * w + x] = {s[0], s[1], s[2], 1.0f};
        }
    W = w; H = h;
    return true;
}

}  // namespace

// What stbi_loadf(Path, &W, &H, &N, 4) returns for the formats above: linear RGBA32F, row 0 at the top.
bool DecodeImageFile(char const* Path, int& Width, int& Height, std::vector<vec4>& Pixels) {
    std::vector<u8> file;
    if (!read_file(Path, file) || file.size() < 4) return false;
    std::vector<u8> rgba;
    if (file[0] == 137 && file[1] == 'P') {
        if (!decode_png(file, Width, Height, rgba)) return false;
        ldr_to_float(rgba, Pixels);
        return true;
    }
    if (file[0] == 0xFF && file[1] == 0xD8) {
        if (!decode_jpeg(file, Width, Height, rgba)) return false;
        ldr_to_float(rgba, Pixels);
        return true;
    }
    if (file[0] == '#' && file[1] == '?') return decode_hdr(file, Width, Height, Pixels);
    if (file[0] == 'P' && (file[1] == '5' || file[1] == '6')) {
        if (!decode_pnm_stb(file, Width, Height, rgba)) return false;
        ldr_to_float(rgba, Pixels);
        return true;
    }
    if (file[0] == 'P' && file[1] == 'F') return decode_pfm(file, Width, Height, Pixels);
    if (file[0] == 'G' && file[1] == 'I' && file[2] == 'F' && file[3] == '8') {
        if (!decode_gif(file, Width, Height, rgba)) return false;
        ldr_to_float(rgba, Pixels);
        return true;
    }
    if (file[0] == '8' && file[1] == 'B' && file[2] == 'P' && file[3] == 'S') {
        if (!decode_psd(file, Width, Height, rgba)) return false;
        ldr_to_float(rgba, Pixels);
        return true;
    }
    if (file.size() >= 92 && file[0] == 0x53 && file[1] == 0x80 && file[2] == 0xF6 && file[3] == 0x34 && std::memcmp(&file[88], "PICT", 4) == 0) {
        if (!decode_pic(file, Width, Height, rgba)) return false;
        ldr_to_float(rgba, Pixels);
        return true;
    }
    if (file[0] == 'B' && file[1] == 'M') {
        if (!decode_bmp(file, Width, Height, rgba)) return false;
        ldr_to_float(rgba, Pixels);
        return true;
    }
    if (decode_tga(file, Width, Height, rgba)) {               // no magic number: tried last, as stb_image does
        ldr_to_float(rgba, Pixels);
        return true;
    }
    return false;
}

}}  // namespace ptb::host
