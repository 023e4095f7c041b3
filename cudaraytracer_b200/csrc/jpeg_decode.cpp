// jpeg_decode.cpp — JPEG (ITU-T T.81) -> tightly packed RGB8, the layout Image::value reads (Texture.cuh:76-105).
//
// Replaces the decode step of the reference's texture path: LoadImage -> stbi_load(filename, &w, &h, &n, 0)
// (Utils/RawStbImage.h:11-22, called from CudaLayer::ImageAllocation, CudaLayer.cpp:874-916). stb_image is third-party code
// vendored by the reference (vendor/stb); it is not carried here. This is a separate implementation of the standard:
//   * baseline / extended sequential (SOF0, SOF1) and progressive (SOF2) Huffman JPEG, 8-bit, 1 or 3 components,
//     any sampling factors, restart intervals, interleaved and non-interleaved scans;
//   * the reconstruction arithmetic follows what stb_image documents for its decoder, so that a texture decoded here has the
//     texels the reference renders with: the 12-bit fixed-point "slow integer" inverse DCT of the IJG code (two passes, 10 and
//     17 bit descale), triangle-filter chroma upsampling for 2x1 / 1x2 / 2x2 subsampling (replication otherwise), and the
//     20-bit fixed-point YCbCr -> RGB matrix. tests/test_host_camera_io.py pins it bit-exactly against the vendored decoder
//     (compiled from /root/reference into oracle/_ref when that tree is present) and against committed fixtures;
//   * entropy decoding of a baseline scan is split at its restart markers and the pieces are decoded by several threads
//     (the 8k textures of the reference carry restart intervals); the IDCT + colour conversion run threaded over MCU rows.
// Not supported (CRT_ERR_UNSUPPORTED): arithmetic coding, 12-bit samples, lossless and hierarchical processes, 4 components.
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "crt_internal.h"

namespace {

const uint8_t kZigzag[64 + 15] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13,
                                  6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31,
                                  39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
                                  // a corrupt run may step past 63: keep those writes inside the block
                                  63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

struct Huffman
{
    // canonical code: for each length l (1..16) the first code value, the index of its first symbol and the count
    int32_t mincode[17], maxcode[18], valptr[17];
    uint8_t values[256];
    uint8_t look_len[512]; // 9-bit lookahead: code length (0 = longer than 9 bits)
    uint8_t look_sym[512];
    bool present = false;

    bool build(const uint8_t counts[16], const uint8_t* syms, int n_syms)
    {
        if (n_syms > 256) return false;
        memcpy(values, syms, (size_t)n_syms);
        memset(look_len, 0, sizeof(look_len));
        int code = 0, k = 0;
        for (int l = 1; l <= 16; l++) {
            valptr[l] = k;
            mincode[l] = code;
            const int cnt = counts[l - 1];
            if (cnt && code + cnt > (1 << l)) return false;
            for (int i = 0; i < cnt; i++, k++, code++) {
                if (l <= 9) {
                    const int first = code << (9 - l);
                    for (int j = 0; j < (1 << (9 - l)); j++) {
                        look_len[first + j] = (uint8_t)l;
                        look_sym[first + j] = values[k];
                    }
                }
            }
            maxcode[l] = cnt ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        present = true;
        return k == n_syms;
    }
};

// Entropy-coded segment reader: removes 0xFF00 stuffing, stops at any marker (and remembers it), then feeds zero bits.
struct BitReader
{
    const uint8_t* p = nullptr;
    const uint8_t* end = nullptr;
    uint32_t buf = 0;
    int bits = 0;
    int marker = 0; // marker byte met in the stream (0 = none yet)

    void init(const uint8_t* b, const uint8_t* e) { p = b; end = e; buf = 0; bits = 0; marker = 0; }

    void fill()
    {
        while (bits <= 24) {
            uint32_t byte = 0;
            if (!marker && p < end) {
                byte = *p++;
                if (byte == 0xff) {
                    uint32_t c = p < end ? *p : 0xd9u;
                    while (c == 0xff && p + 1 < end) { p++; c = *p; } // fill bytes
                    if (c == 0) p++;                                  // stuffed zero: a literal 0xFF
                    else { marker = (int)c; p++; byte = 0; }
                }
            }
            buf |= byte << (24 - bits);
            bits += 8;
        }
    }
    int get_bits(int n)
    {
        if (n == 0) return 0;
        if (bits < n) fill();
        const int v = (int)(buf >> (32 - n));
        buf <<= n;
        bits -= n;
        return v;
    }
    int get_bit() { return get_bits(1); }
    // T.81 F.2.2.1 EXTEND: n magnitude bits -> signed value
    int receive_extend(int n)
    {
        if (n == 0) return 0;
        const int v = get_bits(n);
        return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
    }
    int decode(const Huffman& h)
    {
        if (bits < 16) fill();
        const int look = (int)(buf >> 23);
        int l = h.look_len[look];
        if (l) {
            buf <<= l;
            bits -= l;
            return h.look_sym[look];
        }
        for (l = 10; l <= 16; l++) { // T.81 F.2.2.3 DECODE: the first length whose largest code is not below the prefix
            const int code = (int)(buf >> (32 - l));
            if (code <= h.maxcode[l]) {
                if (code < h.mincode[l]) return -1;
                buf <<= l;
                bits -= l;
                return h.values[h.valptr[l] + code - h.mincode[l]];
            }
        }
        return -1; // not a code of this table
    }
};

struct Component
{
    int id = 0, h = 1, v = 1, tq = 0;
    int hd = 0, ha = 0;          // Huffman table selectors of the current scan
    int dc_pred = 0;
    int x = 0, y = 0;            // size in samples
    int w2 = 0, h2 = 0;          // padded to whole MCUs
    int bw = 0;                  // blocks per padded row (w2 / 8)
    uint8_t* data = nullptr;     // w2 x h2 reconstructed samples (calloc: large zero pages cost nothing until touched)
    int16_t* coeff = nullptr;    // progressive: bw x (h2 / 8) x 64 coefficients, natural order
    Component() = default;
    Component(const Component&) = delete;
    Component& operator=(const Component&) = delete;
    ~Component() { free(data); free(coeff); }
};

struct Decoder
{
    const uint8_t* base = nullptr;
    const uint8_t* end = nullptr;
    int width = 0, height = 0, n_comp = 0;
    bool progressive = false, jfif = false;
    int adobe_transform = -1;
    int h_max = 1, v_max = 1, mcu_w = 8, mcu_h = 8, mcus_x = 0, mcus_y = 0;
    int restart_interval = 0;
    uint16_t dequant[4][64] = {};
    Huffman dc[4], ac[4];
    Component comp[3];
    // current scan
    int scan_n = 0, order[3] = {0, 0, 0};
    int spec_start = 0, spec_end = 63, succ_high = 0, succ_low = 0;
    std::string err;
    bool unsupported = false;

    bool fail(const char* m) { err = m; return false; }
    bool nosupport(const char* m) { err = m; unsupported = true; return false; }
};

// ---- block decoders (T.81 F.2.2 sequential, G.1.2 progressive) -------------------------------------------------------------
// Sequential: one block, dequantised while decoding, natural order. Products are stored as 16-bit like the IDCT input.
bool decode_block_sequential(BitReader& br, const Huffman& hdc, const Huffman& hac, const uint16_t* dq, int& dc_pred, int16_t* blk)
{
    memset(blk, 0, 64 * sizeof(int16_t));
    const int t = br.decode(hdc);
    if (t < 0 || t > 15) return false;
    dc_pred += t ? br.receive_extend(t) : 0;
    blk[0] = (int16_t)(dc_pred * dq[0]);
    for (int k = 1; k < 64;) {
        const int rs = br.decode(hac);
        if (rs < 0) return false;
        const int s = rs & 15, r = rs >> 4;
        if (s == 0) {
            if (rs != 0xf0) break; // EOB
            k += 16;
        }
        else {
            k += r;
            const int z = kZigzag[k++];
            blk[z] = (int16_t)(br.receive_extend(s) * dq[z]);
        }
    }
    return true;
}

bool decode_block_prog_dc(BitReader& br, const Huffman& hdc, int& dc_pred, int succ_high, int succ_low, int16_t* blk)
{
    if (succ_high == 0) { // first scan of the DC coefficient
        memset(blk, 0, 64 * sizeof(int16_t));
        const int t = br.decode(hdc);
        if (t < 0 || t > 15) return false;
        dc_pred += t ? br.receive_extend(t) : 0;
        blk[0] = (int16_t)(dc_pred * (1 << succ_low));
    }
    else if (br.get_bit()) // refinement: one more bit
        blk[0] = (int16_t)(blk[0] + (1 << succ_low));
    return true;
}

bool decode_block_prog_ac(BitReader& br, const Huffman& hac, int ss, int se, int succ_high, int succ_low, int& eob_run, int16_t* blk)
{
    if (succ_high == 0) { // first scan of this band
        if (eob_run) { eob_run--; return true; }
        int k = ss;
        do {
            const int rs = br.decode(hac);
            if (rs < 0) return false;
            const int s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (r < 15) { // EOBn: this block and eob_run more end here
                    eob_run = (1 << r);
                    if (r) eob_run += br.get_bits(r);
                    eob_run--;
                    break;
                }
                k += 16;
            }
            else {
                k += r;
                const int z = kZigzag[k++];
                blk[z] = (int16_t)(br.receive_extend(s) * (1 << succ_low));
            }
        } while (k <= se);
        return true;
    }
    // refinement scan: correction bits for the coefficients that are already non-zero, new +-1 coefficients in between
    const int16_t bit = (int16_t)(1 << succ_low);
    auto refine = [&](int16_t* p) {
        if (br.get_bit() && (*p & bit) == 0) *p = (int16_t)(*p > 0 ? *p + bit : *p - bit);
    };
    if (eob_run) {
        eob_run--;
        for (int k = ss; k <= se; k++) {
            int16_t* p = &blk[kZigzag[k]];
            if (*p != 0) refine(p);
        }
        return true;
    }
    int k = ss;
    do {
        const int rs = br.decode(hac);
        if (rs < 0) return false;
        int s = rs & 15, r = rs >> 4;
        if (s == 0) {
            if (r < 15) {
                eob_run = (1 << r) - 1;
                if (r) eob_run += br.get_bits(r);
                r = 64; // run to the end of the band, refining on the way
            }
            // r == 15: skip 16 zero coefficients (non-zero ones in between are refined and not counted)
        }
        else {
            if (s != 1) return false;
            s = br.get_bit() ? bit : -bit;
        }
        while (k <= se) {
            int16_t* p = &blk[kZigzag[k++]];
            if (*p != 0) refine(p);
            else {
                if (r == 0) { *p = (int16_t)s; break; }
                r--;
            }
        }
    } while (k <= se);
    return true;
}

// ---- reconstruction ---------------------------------------------------------------------------------------------------------
// The IJG "slow integer" inverse DCT (Loeffler-Ligtenberg-Moshovitz factorisation, 12-bit constants), columns then rows,
// descale by 10 and 17 bits with the +128 level shift folded into the second rounding constant, output clamped to 0..255.
inline int fx(double x) { return (int)(x * 4096.0 + 0.5); }
inline uint8_t clamp8(int x) { return (uint8_t)(x < 0 ? 0 : x > 255 ? 255 : x); }

struct Idct1D { int x0, x1, x2, x3, t0, t1, t2, t3; };

inline Idct1D idct_1d(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7)
{
    static const int c0541 = fx(0.5411961), c1847 = fx(-1.847759065), c0765 = fx(0.765366865), c1175 = fx(1.175875602),
                     c0298 = fx(0.298631336), c2053 = fx(2.053119869), c3072 = fx(3.072711026), c1501 = fx(1.501321110),
                     c0899 = fx(-0.899976223), c2562 = fx(-2.562915447), c1961 = fx(-1.961570560), c0390 = fx(-0.390180644);
    Idct1D o;
    int p2 = s2, p3 = s6;
    int p1 = (p2 + p3) * c0541;
    int t2 = p1 + p3 * c1847;
    int t3 = p1 + p2 * c0765;
    p2 = s0; p3 = s4;
    int t0 = (p2 + p3) * 4096;
    int t1 = (p2 - p3) * 4096;
    o.x0 = t0 + t3; o.x3 = t0 - t3; o.x1 = t1 + t2; o.x2 = t1 - t2;
    t0 = s7; t1 = s5; t2 = s3; t3 = s1;
    p3 = t0 + t2;
    int p4 = t1 + t3;
    p1 = t0 + t3;
    p2 = t1 + t2;
    const int p5 = (p3 + p4) * c1175;
    t0 *= c0298; t1 *= c2053; t2 *= c3072; t3 *= c1501;
    p1 = p5 + p1 * c0899;
    p2 = p5 + p2 * c2562;
    p3 *= c1961;
    p4 *= c0390;
    o.t3 = t3 + p1 + p4;
    o.t2 = t2 + p2 + p3;
    o.t1 = t1 + p2 + p4;
    o.t0 = t0 + p1 + p3;
    return o;
}

void idct_block(uint8_t* out, int stride, const int16_t* d)
{
    int val[64];
    for (int i = 0; i < 8; i++) { // columns
        const int16_t* c = d + i;
        int* v = val + i;
        if (c[8] == 0 && c[16] == 0 && c[24] == 0 && c[32] == 0 && c[40] == 0 && c[48] == 0 && c[56] == 0) {
            const int dcterm = c[0] * 4; // a column with only a DC term: constant, scaled like the full pass
            v[0] = v[8] = v[16] = v[24] = v[32] = v[40] = v[48] = v[56] = dcterm;
        }
        else {
            Idct1D o = idct_1d(c[0], c[8], c[16], c[24], c[32], c[40], c[48], c[56]);
            o.x0 += 512; o.x1 += 512; o.x2 += 512; o.x3 += 512;
            v[0] = (o.x0 + o.t3) >> 10; v[56] = (o.x0 - o.t3) >> 10;
            v[8] = (o.x1 + o.t2) >> 10; v[48] = (o.x1 - o.t2) >> 10;
            v[16] = (o.x2 + o.t1) >> 10; v[40] = (o.x2 - o.t1) >> 10;
            v[24] = (o.x3 + o.t0) >> 10; v[32] = (o.x3 - o.t0) >> 10;
        }
    }
    for (int i = 0; i < 8; i++) { // rows
        const int* v = val + 8 * i;
        uint8_t* o8 = out + (size_t)i * stride;
        Idct1D o = idct_1d(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
        const int r = 65536 + (128 << 17);
        o.x0 += r; o.x1 += r; o.x2 += r; o.x3 += r;
        o8[0] = clamp8((o.x0 + o.t3) >> 17); o8[7] = clamp8((o.x0 - o.t3) >> 17);
        o8[1] = clamp8((o.x1 + o.t2) >> 17); o8[6] = clamp8((o.x1 - o.t2) >> 17);
        o8[2] = clamp8((o.x2 + o.t1) >> 17); o8[5] = clamp8((o.x2 - o.t1) >> 17);
        o8[3] = clamp8((o.x3 + o.t0) >> 17); o8[4] = clamp8((o.x3 - o.t0) >> 17);
    }
}

// ---- header segments --------------------------------------------------------------------------------------------------------
inline int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

bool parse_dqt(Decoder& D, const uint8_t* p, int len)
{
    while (len > 0) {
        const int pq = p[0] >> 4, tq = p[0] & 15;
        if (pq > 1 || tq > 3) return D.fail("jpeg: bad quantisation table header");
        const int need = 1 + 64 * (pq + 1);
        if (len < need) return D.fail("jpeg: truncated quantisation table");
        for (int i = 0; i < 64; i++) D.dequant[tq][kZigzag[i]] = (uint16_t)(pq ? be16(p + 1 + 2 * i) : p[1 + i]);
        p += need;
        len -= need;
    }
    return true;
}

bool parse_dht(Decoder& D, const uint8_t* p, int len)
{
    while (len > 0) {
        if (len < 17) return D.fail("jpeg: truncated Huffman table");
        const int tc = p[0] >> 4, th = p[0] & 15;
        if (tc > 1 || th > 3) return D.fail("jpeg: bad Huffman table header");
        int n = 0;
        for (int i = 0; i < 16; i++) n += p[1 + i];
        if (n > 256 || len < 17 + n) return D.fail("jpeg: truncated Huffman table");
        Huffman& h = tc ? D.ac[th] : D.dc[th];
        if (!h.build(p + 1, p + 17, n)) return D.fail("jpeg: inconsistent Huffman code lengths");
        p += 17 + n;
        len -= 17 + n;
    }
    return true;
}

bool parse_sof(Decoder& D, const uint8_t* p, int len, bool progressive)
{
    if (len < 6) return D.fail("jpeg: truncated frame header");
    if (p[0] != 8) return D.nosupport("jpeg: only 8-bit samples are supported");
    D.height = be16(p + 1);
    D.width = be16(p + 3);
    D.n_comp = p[5];
    D.progressive = progressive;
    if (D.width <= 0 || D.height <= 0) return D.fail("jpeg: empty image (or a DNL-defined height)");
    if (D.n_comp != 1 && D.n_comp != 3) return D.nosupport("jpeg: only 1 or 3 components are supported");
    if (len < 6 + 3 * D.n_comp) return D.fail("jpeg: truncated frame header");
    if ((uint64_t)D.width * D.height > (1ull << 28)) return D.nosupport("jpeg: image too large (more than 2^28 pixels)");
    D.h_max = D.v_max = 1;
    for (int i = 0; i < D.n_comp; i++) {
        Component& c = D.comp[i];
        c.id = p[6 + 3 * i];
        c.h = p[7 + 3 * i] >> 4;
        c.v = p[7 + 3 * i] & 15;
        c.tq = p[8 + 3 * i];
        if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) return D.fail("jpeg: bad component header");
        D.h_max = std::max(D.h_max, c.h);
        D.v_max = std::max(D.v_max, c.v);
    }
    for (int i = 0; i < D.n_comp; i++)
        if (D.h_max % D.comp[i].h || D.v_max % D.comp[i].v) return D.nosupport("jpeg: fractional sampling ratios are not supported");
    D.mcu_w = 8 * D.h_max;
    D.mcu_h = 8 * D.v_max;
    D.mcus_x = (D.width + D.mcu_w - 1) / D.mcu_w;
    D.mcus_y = (D.height + D.mcu_h - 1) / D.mcu_h;
    for (int i = 0; i < D.n_comp; i++) {
        Component& c = D.comp[i];
        c.x = (D.width * c.h + D.h_max - 1) / D.h_max;
        c.y = (D.height * c.v + D.v_max - 1) / D.v_max;
        c.w2 = D.mcus_x * c.h * 8;
        c.h2 = D.mcus_y * c.v * 8;
        c.bw = c.w2 / 8;
        c.data = (uint8_t*)calloc((size_t)c.w2 * c.h2, 1);
        if (progressive) c.coeff = (int16_t*)calloc((size_t)c.w2 * c.h2, sizeof(int16_t));
        if (!c.data || (progressive && !c.coeff)) return D.fail("jpeg: out of memory");
    }
    return true;
}

bool parse_sos(Decoder& D, const uint8_t* p, int len)
{
    if (len < 1) return D.fail("jpeg: truncated scan header");
    D.scan_n = p[0];
    if (D.scan_n < 1 || D.scan_n > D.n_comp || len != 4 + 2 * D.scan_n) return D.fail("jpeg: bad scan header");
    for (int i = 0; i < D.scan_n; i++) {
        const int id = p[1 + 2 * i], t = p[2 + 2 * i];
        int which = -1;
        for (int k = 0; k < D.n_comp; k++)
            if (D.comp[k].id == id) which = k;
        if (which < 0) return D.fail("jpeg: scan names an unknown component");
        D.comp[which].hd = t >> 4;
        D.comp[which].ha = t & 15;
        if (D.comp[which].hd > 3 || D.comp[which].ha > 3) return D.fail("jpeg: bad Huffman table selector");
        D.order[i] = which;
    }
    const uint8_t* q = p + 1 + 2 * D.scan_n;
    D.spec_start = q[0];
    D.spec_end = q[1];
    D.succ_high = q[2] >> 4;
    D.succ_low = q[2] & 15;
    if (D.progressive) {
        if (D.spec_start > 63 || D.spec_end > 63 || D.spec_start > D.spec_end || D.succ_high > 13 || D.succ_low > 13)
            return D.fail("jpeg: bad progressive scan parameters");
        if (D.spec_start == 0 && D.spec_end != 0) return D.fail("jpeg: a progressive DC scan must not carry AC coefficients");
        if (D.spec_start != 0 && D.scan_n != 1) return D.fail("jpeg: a progressive AC scan must have one component");
    }
    else {
        if (D.spec_start != 0 || D.succ_high != 0 || D.succ_low != 0) return D.fail("jpeg: bad sequential scan parameters");
        D.spec_end = 63;
    }
    return true;
}

// ---- one scan ---------------------------------------------------------------------------------------------------------------
struct Segment { const uint8_t* b; const uint8_t* e; };

// Splits the entropy-coded data that starts at p into the pieces between restart markers; returns where the scan ends
// (the first marker that is not RSTn, or the end of the file).
const uint8_t* split_scan(const uint8_t* p, const uint8_t* end, std::vector<Segment>& segs)
{
    const uint8_t* b = p;
    while (p < end) {
        const uint8_t* f = (const uint8_t*)memchr(p, 0xff, (size_t)(end - p));
        if (!f || f + 1 >= end) { p = end; break; }
        const uint8_t m = f[1];
        if (m == 0x00 || m == 0xff) { p = f + 1; continue; } // stuffing / fill byte
        if (m >= 0xd0 && m <= 0xd7) {
            segs.push_back({b, f});
            b = p = f + 2;
            continue;
        }
        segs.push_back({b, f});
        return f;
    }
    segs.push_back({b, end});
    return end;
}

// Decodes the units [u0, u1) of the current scan from one restart segment. A unit is an MCU of an interleaved scan or one
// block of a single-component scan. Sequential blocks are reconstructed at once; progressive ones go to the coefficient store.
bool decode_units(Decoder& D, Segment sg, int u0, int u1)
{
    BitReader br;
    br.init(sg.b, sg.e);
    int pred[3] = {0, 0, 0};
    int eob_run = 0;
    int16_t blk[64];
    auto one_block = [&](int ci, int bx, int by) -> bool {
        Component& c = D.comp[ci];
        if (!D.progressive) {
            if (!D.dc[c.hd].present || !D.ac[c.ha].present) return false;
            if (!decode_block_sequential(br, D.dc[c.hd], D.ac[c.ha], D.dequant[c.tq], pred[ci], blk)) return false;
            idct_block(c.data + (size_t)by * 8 * c.w2 + (size_t)bx * 8, c.w2, blk);
            return true;
        }
        int16_t* co = c.coeff + 64 * ((size_t)by * c.bw + bx);
        if (D.spec_start == 0) {
            if (D.succ_high == 0 && !D.dc[c.hd].present) return false;
            return decode_block_prog_dc(br, D.dc[c.hd], pred[ci], D.succ_high, D.succ_low, co);
        }
        if (!D.ac[c.ha].present) return false;
        return decode_block_prog_ac(br, D.ac[c.ha], D.spec_start, D.spec_end, D.succ_high, D.succ_low, eob_run, co);
    };
    if (D.scan_n == 1) {
        const int ci = D.order[0];
        const int bw = (D.comp[ci].x + 7) >> 3;
        for (int u = u0; u < u1; u++)
            if (!one_block(ci, u % bw, u / bw)) return false;
        return true;
    }
    for (int u = u0; u < u1; u++) {
        const int mx = u % D.mcus_x, my = u / D.mcus_x;
        for (int k = 0; k < D.scan_n; k++) {
            const int ci = D.order[k];
            const Component& c = D.comp[ci];
            for (int y = 0; y < c.v; y++)
                for (int x = 0; x < c.h; x++)
                    if (!one_block(ci, mx * c.h + x, my * c.v + y)) return false;
        }
    }
    return true;
}

int worker_count(size_t jobs)
{
    unsigned hw = std::thread::hardware_concurrency();
    if (const char* e = getenv("CRT_JPEG_THREADS")) hw = (unsigned)std::max(1, atoi(e));
    return (int)std::max<size_t>(1, std::min<size_t>(std::min<unsigned>(hw ? hw : 1, 32), jobs));
}

template <class F> void parallel_for(size_t n, F&& f)
{
    const int nt = worker_count(n);
    if (nt <= 1) {
        for (size_t i = 0; i < n; i++) f(i);
        return;
    }
    std::atomic<size_t> next{0};
    std::vector<std::thread> th;
    auto body = [&] { for (size_t i; (i = next.fetch_add(1)) < n;) f(i); };
    for (int t = 1; t < nt; t++) th.emplace_back(body);
    body();
    for (auto& t : th) t.join();
}

bool decode_scan(Decoder& D, const uint8_t*& p)
{
    std::vector<Segment> segs;
    p = split_scan(p, D.end, segs);
    int units;
    if (D.scan_n == 1) {
        const Component& c = D.comp[D.order[0]];
        units = ((c.x + 7) >> 3) * ((c.y + 7) >> 3);
    }
    else units = D.mcus_x * D.mcus_y;
    const int ri = D.restart_interval > 0 ? D.restart_interval : units;
    const size_t need = (size_t)(units + ri - 1) / ri;
    // A stream that ends early (or lacks an expected restart marker) keeps what was decoded, like the vendored decoder.
    const size_t n = std::min(need, segs.size());
    std::atomic<bool> ok{true};
    parallel_for(n, [&](size_t s) {
        const int u0 = (int)s * ri, u1 = std::min(units, u0 + ri);
        if (!decode_units(D, segs[s], u0, u1)) ok = false;
    });
    if (!ok) return D.fail("jpeg: corrupt entropy-coded data");
    return true;
}

// Progressive: every scan has been merged into the coefficient store; dequantise and reconstruct, threaded over block rows.
void finish_progressive(Decoder& D)
{
    for (int ci = 0; ci < D.n_comp; ci++) {
        Component& c = D.comp[ci];
        const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
        const uint16_t* dq = D.dequant[c.tq];
        parallel_for((size_t)bh, [&](size_t by) {
            int16_t blk[64];
            for (int bx = 0; bx < bw; bx++) {
                const int16_t* co = c.coeff + 64 * (by * c.bw + bx);
                for (int k = 0; k < 64; k++) blk[k] = (int16_t)(co[k] * dq[k]);
                idct_block(c.data + by * 8 * c.w2 + (size_t)bx * 8, c.w2, blk);
            }
        });
    }
}

// ---- chroma upsampling: one output row from the nearer and the farther low-resolution row -------------------------------------
// Triangle filter (3/4 near + 1/4 far per axis) for the 2x cases, replication otherwise.
void up_none(uint8_t* out, const uint8_t* near, const uint8_t*, int w, int) { memcpy(out, near, (size_t)w); }

void up_v2(uint8_t* out, const uint8_t* near, const uint8_t* far, int w, int)
{
    for (int i = 0; i < w; i++) out[i] = (uint8_t)((3 * near[i] + far[i] + 2) >> 2);
}

void up_h2(uint8_t* out, const uint8_t* in, const uint8_t*, int w, int)
{
    if (w == 1) { out[0] = out[1] = in[0]; return; }
    out[0] = in[0];
    out[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
    int i;
    for (i = 1; i < w - 1; i++) {
        const int n = 3 * in[i] + 2;
        out[i * 2 + 0] = (uint8_t)((n + in[i - 1]) >> 2);
        out[i * 2 + 1] = (uint8_t)((n + in[i + 1]) >> 2);
    }
    out[i * 2 + 0] = (uint8_t)((in[w - 2] * 3 + in[w - 1] + 2) >> 2);
    out[i * 2 + 1] = in[w - 1];
}

void up_hv2(uint8_t* out, const uint8_t* near, const uint8_t* far, int w, int)
{
    if (w == 1) { out[0] = out[1] = (uint8_t)((3 * near[0] + far[0] + 2) >> 2); return; }
    int t1 = 3 * near[0] + far[0];
    out[0] = (uint8_t)((t1 + 2) >> 2);
    for (int i = 1; i < w; i++) {
        const int t0 = t1;
        t1 = 3 * near[i] + far[i];
        out[i * 2 - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
        out[i * 2] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
    }
    out[w * 2 - 1] = (uint8_t)((t1 + 2) >> 2);
}

void up_generic(uint8_t* out, const uint8_t* near, const uint8_t*, int w, int hs)
{
    for (int i = 0; i < w; i++)
        for (int j = 0; j < hs; j++) out[i * hs + j] = near[i];
}

typedef void (*UpFn)(uint8_t*, const uint8_t*, const uint8_t*, int, int);

// YCbCr (full range, JFIF) -> RGB with a 20-bit fixed-point matrix; the Cb term of green is truncated to 16 fractional bits.
inline int fixed12_8(float x) { return ((int)(x * 4096.0f + 0.5f)) << 8; }

void ycc_to_rgb_row(uint8_t* out, const uint8_t* y, const uint8_t* pcb, const uint8_t* pcr, int n)
{
    static const int k_r_cr = fixed12_8(1.40200f), k_g_cr = -fixed12_8(0.71414f), k_g_cb = -fixed12_8(0.34414f),
                     k_b_cb = fixed12_8(1.77200f);
    for (int i = 0; i < n; i++) {
        const int yf = (y[i] << 20) + (1 << 19);
        const int cr = pcr[i] - 128, cb = pcb[i] - 128;
        int r = yf + cr * k_r_cr;
        int g = yf + cr * k_g_cr + (int)((unsigned)(cb * k_g_cb) & 0xffff0000u);
        int b = yf + cb * k_b_cb;
        r >>= 20; g >>= 20; b >>= 20;
        out[3 * i + 0] = clamp8(r);
        out[3 * i + 1] = clamp8(g);
        out[3 * i + 2] = clamp8(b);
    }
}

// Builds the RGB8 image from the reconstructed component planes.
bool emit_rgb(Decoder& D, uint8_t* rgb)
{
    struct Plan { UpFn fn; int hs, vs, w_lores; std::vector<int> near, far; } plan[3];
    for (int k = 0; k < D.n_comp; k++) {
        const Component& c = D.comp[k];
        Plan& P = plan[k];
        P.hs = D.h_max / c.h;
        P.vs = D.v_max / c.v;
        P.w_lores = (D.width + P.hs - 1) / P.hs;
        P.fn = P.hs == 1 && P.vs == 1 ? up_none : P.hs == 1 && P.vs == 2 ? up_v2 : P.hs == 2 && P.vs == 1 ? up_h2
             : P.hs == 2 && P.vs == 2 ? up_hv2 : up_generic;
        // which low-resolution rows feed output row j: a two-row window that steps down every vs output rows, the nearer row
        // being the lower one in the second half of a step
        P.near.resize((size_t)D.height);
        P.far.resize((size_t)D.height);
        int ystep = P.vs >> 1, ypos = 0, line0 = 0, line1 = 0;
        for (int j = 0; j < D.height; j++) {
            const bool bot = ystep >= (P.vs >> 1);
            P.near[j] = bot ? line1 : line0;
            P.far[j] = bot ? line0 : line1;
            if (++ystep >= P.vs) {
                ystep = 0;
                line0 = line1;
                if (++ypos < c.y) line1++;
            }
        }
    }
    bool is_rgb = false;
    if (D.n_comp == 3) {
        const bool ids_rgb = D.comp[0].id == 'R' && D.comp[1].id == 'G' && D.comp[2].id == 'B';
        is_rgb = ids_rgb || (D.adobe_transform == 0 && !D.jfif);
    }
    const size_t chunk = 16;
    parallel_for(((size_t)D.height + chunk - 1) / chunk, [&](size_t ch) {
        std::vector<uint8_t> line[3];
        for (int k = 0; k < D.n_comp; k++) line[k].resize((size_t)plan[k].w_lores * plan[k].hs + 8);
        const int j1 = (int)std::min<size_t>((size_t)D.height, (ch + 1) * chunk);
        for (int j = (int)(ch * chunk); j < j1; j++) {
            const uint8_t* row[3] = {nullptr, nullptr, nullptr};
            for (int k = 0; k < D.n_comp; k++) {
                const Component& c = D.comp[k];
                const Plan& P = plan[k];
                const uint8_t* near = c.data + (size_t)P.near[j] * c.w2;
                const uint8_t* far = c.data + (size_t)P.far[j] * c.w2;
                if (P.fn == up_none) row[k] = near;
                else {
                    P.fn(line[k].data(), near, far, P.w_lores, P.hs);
                    row[k] = line[k].data();
                }
            }
            uint8_t* out = rgb + (size_t)j * D.width * 3;
            if (D.n_comp == 1) {
                for (int i = 0; i < D.width; i++) out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = row[0][i];
            }
            else if (is_rgb) {
                for (int i = 0; i < D.width; i++) { out[3 * i] = row[0][i]; out[3 * i + 1] = row[1][i]; out[3 * i + 2] = row[2][i]; }
            }
            else ycc_to_rgb_row(out, row[0], row[1], row[2], D.width);
        }
    });
    return true;
}

bool decode_jpeg(Decoder& D, const uint8_t* data, size_t n, uint8_t** rgb)
{
    D.base = data;
    D.end = data + n;
    if (n < 4 || data[0] != 0xff || data[1] != 0xd8) return D.fail("jpeg: no SOI marker");
    const uint8_t* p = data + 2;
    bool have_frame = false, have_scan = false;
    while (p + 1 < D.end) {
        if (p[0] != 0xff) { p++; continue; } // resynchronise on garbage between segments
        const int m = p[1];
        if (m == 0xff) { p++; continue; }
        p += 2;
        if (m == 0xd9) break;                                  // EOI
        if (m == 0x01 || (m >= 0xd0 && m <= 0xd7)) continue;   // TEM, stray RSTn: no payload
        if (p + 2 > D.end) return D.fail("jpeg: truncated segment");
        const int len = be16(p);
        if (len < 2 || p + len > D.end) return D.fail("jpeg: bad segment length");
        const uint8_t* q = p + 2;
        const int pl = len - 2;
        p += len;
        switch (m) {
        case 0xdb: if (!parse_dqt(D, q, pl)) return false; break;
        case 0xc4: if (!parse_dht(D, q, pl)) return false; break;
        case 0xc0: case 0xc1: case 0xc2:
            if (have_frame) return D.fail("jpeg: more than one frame header");
            if (!parse_sof(D, q, pl, m == 0xc2)) return false;
            have_frame = true;
            break;
        case 0xc3: case 0xc5: case 0xc6: case 0xc7: case 0xc9: case 0xca: case 0xcb: case 0xcd: case 0xce: case 0xcf:
            return D.nosupport("jpeg: lossless, hierarchical and arithmetic-coded processes are not supported");
        case 0xdd:
            if (pl != 2) return D.fail("jpeg: bad restart interval segment");
            D.restart_interval = be16(q);
            break;
        case 0xe0: if (pl >= 5 && !memcmp(q, "JFIF\0", 5)) D.jfif = true; break;
        case 0xee: if (pl >= 12 && !memcmp(q, "Adobe\0", 6)) D.adobe_transform = q[11]; break;
        case 0xda:
            if (!have_frame) return D.fail("jpeg: scan before the frame header");
            if (!parse_sos(D, q, pl)) return false;
            if (!decode_scan(D, p)) return false;
            have_scan = true;
            break;
        default: break; // APPn, COM, DNL, ...: skipped
        }
    }
    if (!have_frame || !have_scan) return D.fail("jpeg: no image data");
    if (D.progressive) finish_progressive(D);
    *rgb = (uint8_t*)malloc((size_t)D.width * D.height * 3);
    if (!*rgb) return D.fail("jpeg: out of memory");
    return emit_rgb(D, *rgb);
}

} // namespace

extern "C" {

crt_status crt_decode_jpeg(const uint8_t* data, uint64_t n_bytes, uint8_t** out_pixels, int32_t* out_width, int32_t* out_height,
                           int32_t* out_file_channels)
{
    if (!data || !out_pixels || !out_width || !out_height) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_decode_jpeg: NULL argument");
    *out_pixels = nullptr;
    Decoder D;
    uint8_t* px = nullptr;
    if (!decode_jpeg(D, data, (size_t)n_bytes, &px)) {
        free(px);
        return crt_set_error(nullptr, D.unsupported ? CRT_ERR_UNSUPPORTED : CRT_ERR_INVALID, D.err.c_str());
    }
    *out_pixels = px;
    *out_width = D.width;
    *out_height = D.height;
    if (out_file_channels) *out_file_channels = D.n_comp;
    return CRT_OK;
}

crt_status crt_load_image(const char* path, uint8_t** out_pixels, int32_t* out_width, int32_t* out_height, int32_t* out_file_channels)
{
    if (!path || !out_pixels || !out_width || !out_height) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_image: NULL argument");
    *out_pixels = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_image: cannot open the file");
    unsigned char magic[2] = {0, 0};
    const size_t got = fread(magic, 1, 2, f);
    if (got == 2 && magic[0] == 'P' && magic[1] == '6') {
        fclose(f);
        if (out_file_channels) *out_file_channels = 3;
        return crt_load_ppm(path, out_pixels, out_width, out_height);
    }
    if (got != 2 || magic[0] != 0xff || magic[1] != 0xd8) {
        fclose(f);
        return crt_set_error(nullptr, CRT_ERR_UNSUPPORTED, "crt_load_image: neither a JPEG nor a binary PPM file");
    }
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (sz <= 0) { fclose(f); return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_image: empty file"); }
    std::vector<uint8_t> buf((size_t)sz);
    const bool ok = fread(buf.data(), 1, (size_t)sz, f) == (size_t)sz;
    fclose(f);
    if (!ok) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_image: read error");
    return crt_decode_jpeg(buf.data(), (uint64_t)sz, out_pixels, out_width, out_height, out_file_channels);
}

} // extern "C"
