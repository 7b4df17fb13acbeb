// jpeg_decode.h -- host-side JPEG decoder behind rtow::api::load_image().
//
// The reference's scene builders read ./input_images/*.jpg through its vendored stb_image v2.22
// (scenes.h:47,140 -> stbi_load(path,&nx,&ny,&nn,0)); the texels then feed ImageTexture::value
// (texture.h:65-79).  For the flattened scene to be byte-identical to the reference's, the decoded
// texels must be byte-identical too, so this decoder follows the arithmetic stb_image uses, which is
// fully determined by the JPEG standard (ITU T.81: Huffman / spectral-selection / successive-
// approximation coefficient decoding) plus three integer kernels:
//   * the inverse DCT: the jidctint "ISLOW" factorisation with 12-bit constants, a column pass that
//     keeps 2 extra bits (+512 >> 10) and a row pass that removes 17 (+65536 + (128<<17) >> 17)
//     (stb_image.h:2242-2330);
//   * chroma upsampling: the 3/4-1/4 triangle filters (stb_image.h:3251-3345), row pairs chosen by
//     the per-component ystep/ypos walk (stb_image.h:3700-3718);
//   * YCbCr -> RGB in 20-bit fixed point with the green term's low 16 bits masked
//     (stb_image.h:3457-3480).
// stb's SSE2 paths are written to give the same bytes as these scalar definitions.
// tests/test_jpeg_decode.py checks this decoder against the reference's own stbi_load (oracle/_ref)
// on every JPEG under the reference's input_images/ (baseline 4:2:0 and progressive files).
//
// Scope: 8-bit baseline (SOF0/SOF1) and progressive (SOF2) Huffman JPEGs, 1 or 3 components, any
// sampling factors <= 4, restart intervals.  Arithmetic coding, 12-bit, lossless, CMYK: rejected
// (nullptr), as stb rejects or the scene API cannot use them.  Host-side set-up code: never on the
// timed path, no intersection or shading arithmetic here.
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

namespace rtow {
namespace jpeg {

struct Huff {
    // canonical Huffman code (T.81 annex C): codes of length L are maxcode-bounded, values in symbol order
    uint8_t bits[17];
    uint8_t vals[256];
    int mincode[17], maxcode[18], valptr[17];
    bool present = false;
    void build() {
        int code = 0, k = 0;
        for (int L = 1; L <= 16; ++L) {
            valptr[L] = k;
            mincode[L] = code;
            code += bits[L];
            k += bits[L];
            maxcode[L] = bits[L] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        present = true;
    }
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0;
    int td = 0, ta = 0;           // Huffman table selectors of the current scan
    int x = 0, y = 0;             // sample dimensions of this component
    int bw = 0, bh = 0;           // blocks per row / column in the padded (MCU-aligned) plane
    int w2 = 0, h2 = 0;           // padded plane size in samples
    int dc_pred = 0;
    std::vector<short> coeff;     // progressive: bw*bh*64 coefficients
    std::vector<uint8_t> plane;   // w2*h2 samples after the IDCT
};

class Decoder {
  public:
    bool decode(const uint8_t *bytes, size_t len, std::vector<uint8_t> &rgb, int &nx, int &ny, int &channels) {
        d_ = bytes; n_ = len; pos_ = 0;
        if (len < 4 || bytes[0] != 0xFF || bytes[1] != 0xD8) return false;
        pos_ = 2;
        bool have_frame = false;
        for (;;) {
            int m = next_marker();
            if (m < 0) return false;
            if (m == 0xD9) break;                                   // EOI
            if (m == 0xDA) {                                        // SOS
                if (!have_frame || !read_sos()) return false;
                if (!decode_scan()) return false;
                continue;
            }
            if (m == 0xC0 || m == 0xC1 || m == 0xC2) {
                if (have_frame || !read_sof(m == 0xC2)) return false;
                have_frame = true;
                continue;
            }
            if ((m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) return false;   // lossless / arithmetic / hierarchical
            if (!read_segment(m)) return false;
        }
        if (!have_frame) return false;
        if (progressive_) finish_progressive();
        return output(rgb, nx, ny, channels);
    }

  private:
    const uint8_t *d_ = nullptr;
    size_t n_ = 0, pos_ = 0;
    // frame
    int width_ = 0, height_ = 0, ncomp_ = 0, hmax_ = 1, vmax_ = 1, mcux_ = 0, mcuy_ = 0;
    bool progressive_ = false;
    Component comp_[4];
    uint16_t dequant_[4][64];
    Huff hdc_[4], hac_[4];
    int restart_interval_ = 0;
    bool jfif_ = false;
    int adobe_transform_ = -1;
    int rgb_ids_ = 0;
    // scan
    int scan_n_ = 0, order_[4] = {0, 0, 0, 0};
    int ss_ = 0, se_ = 63, ah_ = 0, al_ = 0;
    int eobrun_ = 0;
    // bit reader (MSB first; FF00 -> FF, any other marker ends the entropy-coded segment and reads as zeros)
    uint32_t acc_ = 0;
    int nbits_ = 0;
    bool hit_marker_ = false;

    static const uint8_t *zigzag() {
        static const uint8_t z[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                      41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                      30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
        return z;
    }

    int u8() { return pos_ < n_ ? d_[pos_++] : -1; }
    int u16() { int a = u8(), b = u8(); return (a < 0 || b < 0) ? -1 : (a << 8) | b; }

    int next_marker() {
        for (;;) {
            int c = u8();
            while (c >= 0 && c != 0xFF) c = u8();   // tolerate stray bytes between segments (unread tail of a scan)
            while (c == 0xFF) c = u8();             // fill bytes
            if (c != 0x00) return c;                // FF00 is a stuffed data byte, not a marker
        }
    }

    bool read_segment(int m) {
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) return true;     // standalone markers
        int L = u16();
        if (L < 2 || pos_ + (size_t)(L - 2) > n_) return false;
        const size_t end = pos_ + (size_t)(L - 2);
        if (m == 0xDB) {                                            // DQT
            while (pos_ < end) {
                int q = u8();
                int p = q >> 4, t = q & 15;
                if ((p != 0 && p != 1) || t > 3) return false;
                for (int i = 0; i < 64; ++i) {
                    int v = p ? u16() : u8();
                    if (v < 0) return false;
                    dequant_[t][zigzag()[i]] = (uint16_t)v;
                }
            }
        } else if (m == 0xC4) {                                     // DHT
            while (pos_ < end) {
                int q = u8();
                int tc = q >> 4, th = q & 15;
                if (tc > 1 || th > 3) return false;
                Huff &h = tc ? hac_[th] : hdc_[th];
                int total = 0;
                h.bits[0] = 0;
                for (int i = 1; i <= 16; ++i) { int b = u8(); if (b < 0) return false; h.bits[i] = (uint8_t)b; total += b; }
                if (total > 256) return false;
                for (int i = 0; i < total; ++i) { int v = u8(); if (v < 0) return false; h.vals[i] = (uint8_t)v; }
                h.build();
            }
        } else if (m == 0xDD) {                                     // DRI
            if (L != 4) return false;
            restart_interval_ = u16();
        } else if (m == 0xE0) {                                     // APP0: "JFIF\0"
            if (L >= 7 && std::memcmp(d_ + pos_, "JFIF\0", 5) == 0) jfif_ = true;
        } else if (m == 0xEE) {                                     // APP14: "Adobe\0" + version(1) flags(2,2) transform(1)
            if (L >= 14 && std::memcmp(d_ + pos_, "Adobe\0", 6) == 0) adobe_transform_ = d_[pos_ + 11];
        }
        pos_ = end;
        return true;
    }

    bool read_sof(bool progressive) {
        int L = u16();
        if (L < 11) return false;
        int p = u8();
        height_ = u16(); width_ = u16(); ncomp_ = u8();
        if (p != 8 || height_ <= 0 || width_ <= 0) return false;
        if (ncomp_ != 1 && ncomp_ != 3) return false;               // CMYK / YCCK: not usable as an RGB8 texture
        if (L != 8 + 3 * ncomp_) return false;
        progressive_ = progressive;
        rgb_ids_ = 0;
        static const char rgbname[3] = {'R', 'G', 'B'};
        for (int i = 0; i < ncomp_; ++i) {
            Component &c = comp_[i];
            c.id = u8();
            if (ncomp_ == 3 && c.id == rgbname[i]) ++rgb_ids_;
            int q = u8();
            c.h = q >> 4; c.v = q & 15; c.tq = u8();
            if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) return false;
            hmax_ = c.h > hmax_ ? c.h : hmax_;
            vmax_ = c.v > vmax_ ? c.v : vmax_;
        }
        const int mw = hmax_ * 8, mh = vmax_ * 8;
        mcux_ = (width_ + mw - 1) / mw;
        mcuy_ = (height_ + mh - 1) / mh;
        for (int i = 0; i < ncomp_; ++i) {
            Component &c = comp_[i];
            c.x = (width_ * c.h + hmax_ - 1) / hmax_;
            c.y = (height_ * c.v + vmax_ - 1) / vmax_;
            c.bw = mcux_ * c.h; c.bh = mcuy_ * c.v;
            c.w2 = c.bw * 8; c.h2 = c.bh * 8;
            c.plane.assign((size_t)c.w2 * c.h2, 0);
            if (progressive_) c.coeff.assign((size_t)c.bw * c.bh * 64, 0);
        }
        return true;
    }

    bool read_sos() {
        int L = u16();
        scan_n_ = u8();
        if (scan_n_ < 1 || scan_n_ > ncomp_ || L != 6 + 2 * scan_n_) return false;
        for (int i = 0; i < scan_n_; ++i) {
            int id = u8(), q = u8();
            int which = -1;
            for (int k = 0; k < ncomp_; ++k) if (comp_[k].id == id) which = k;
            if (which < 0) return false;
            comp_[which].td = q >> 4; comp_[which].ta = q & 15;
            if (comp_[which].td > 3 || comp_[which].ta > 3) return false;
            order_[i] = which;
        }
        ss_ = u8(); se_ = u8();
        int a = u8();
        ah_ = a >> 4; al_ = a & 15;
        if (progressive_) {
            if (ss_ > 63 || se_ > 63 || ss_ > se_ || ah_ > 13 || al_ > 13) return false;
        } else {
            if (ss_ != 0 || ah_ != 0 || al_ != 0) return false;
            se_ = 63;
        }
        return true;
    }

    // ---- entropy-coded segment ----
    void reset_bits() { acc_ = 0; nbits_ = 0; hit_marker_ = false; }
    void fill() {
        while (nbits_ <= 24) {
            int b = 0;
            if (!hit_marker_ && pos_ < n_) {
                b = d_[pos_];
                if (b == 0xFF) {
                    int c = pos_ + 1 < n_ ? d_[pos_ + 1] : 0xD9;
                    if (c == 0) pos_ += 2;                       // stuffed zero
                    else { hit_marker_ = true; b = 0; }           // leave the marker for the segment loop
                } else {
                    ++pos_;
                }
            } else {
                hit_marker_ = true;
            }
            acc_ |= (uint32_t)b << (24 - nbits_);
            nbits_ += 8;
        }
    }
    int get_bits(int n) {
        if (n == 0) return 0;
        if (nbits_ < n) fill();
        int v = (int)(acc_ >> (32 - n));
        acc_ <<= n; nbits_ -= n;
        return v;
    }
    int get_bit() { return get_bits(1); }
    int decode_symbol(const Huff &h) {
        if (!h.present) return -1;
        if (nbits_ < 16) fill();
        int code = 0;
        for (int L = 1; L <= 16; ++L) {
            code = (int)(acc_ >> (32 - L));
            if (h.maxcode[L] >= 0 && code <= h.maxcode[L] && code >= h.mincode[L]) {
                acc_ <<= L; nbits_ -= L;
                return h.vals[h.valptr[L] + code - h.mincode[L]];
            }
        }
        return -1;
    }
    // T.81 F.2.2.1 EXTEND: n magnitude bits, the leading bit tells the sign
    int receive_extend(int n) {
        if (n == 0) return 0;
        int v = get_bits(n);
        return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
    }

    bool baseline_block(Component &c, short blk[64]) {
        std::memset(blk, 0, 64 * sizeof(short));
        const uint16_t *dq = dequant_[c.tq];
        int t = decode_symbol(hdc_[c.td]);
        if (t < 0 || t > 15) return false;
        int diff = receive_extend(t);
        c.dc_pred += diff;
        blk[0] = (short)(c.dc_pred * dq[0]);
        const Huff &ac = hac_[c.ta];
        for (int k = 1; k < 64;) {
            int rs = decode_symbol(ac);
            if (rs < 0) return false;
            int r = rs >> 4, s = rs & 15;
            if (s == 0) {
                if (rs != 0xF0) break;                   // EOB
                k += 16;
            } else {
                k += r;
                if (k > 63) return false;
                const int z = zigzag()[k++];
                blk[z] = (short)(receive_extend(s) * dq[z]);
            }
        }
        return true;
    }

    bool prog_dc(Component &c, short *blk) {
        if (ah_ == 0) {
            int t = decode_symbol(hdc_[c.td]);
            if (t < 0 || t > 15) return false;
            c.dc_pred += receive_extend(t);
            blk[0] = (short)(c.dc_pred * (1 << al_));
        } else if (get_bit()) {
            blk[0] = (short)(blk[0] + (1 << al_));
        }
        return true;
    }

    bool prog_ac(Component &c, short *blk) {
        const Huff &ac = hac_[c.ta];
        if (ah_ == 0) {                                    // first pass over this band
            if (eobrun_) { --eobrun_; return true; }
            for (int k = ss_; k <= se_;) {
                int rs = decode_symbol(ac);
                if (rs < 0) return false;
                int r = rs >> 4, s = rs & 15;
                if (s == 0) {
                    if (r < 15) {
                        eobrun_ = 1 << r;
                        if (r) eobrun_ += get_bits(r);
                        --eobrun_;
                        break;
                    }
                    k += 16;
                } else {
                    k += r;
                    if (k > 63) return false;
                    blk[zigzag()[k++]] = (short)(receive_extend(s) * (1 << al_));
                }
            }
            return true;
        }
        // refinement pass (T.81 G.1.2.3): new non-zero coefficients are +-1 << al; already non-zero ones get a correction bit
        const short bit = (short)(1 << al_);
        auto refine = [&](short &p) {
            if (get_bit() && (p & bit) == 0) p = (short)(p > 0 ? p + bit : p - bit);
        };
        if (eobrun_) {
            --eobrun_;
            for (int k = ss_; k <= se_; ++k) { short &p = blk[zigzag()[k]]; if (p != 0) refine(p); }
            return true;
        }
        int k = ss_;
        do {
            int rs = decode_symbol(ac);
            if (rs < 0) return false;
            int r = rs >> 4, s = rs & 15;
            if (s == 0) {
                if (r < 15) {
                    eobrun_ = (1 << r) - 1;
                    if (r) eobrun_ += get_bits(r);
                    r = 64;                                // run to the end of the band
                }
                // r == 15: skip 16 zero-history coefficients
            } else {
                if (s != 1) return false;
                s = get_bit() ? bit : -bit;
            }
            while (k <= se_) {
                short &p = blk[zigzag()[k++]];
                if (p != 0) {
                    refine(p);
                } else {
                    if (r == 0) { p = (short)s; break; }
                    --r;
                }
            }
        } while (k <= se_);
        return true;
    }

    // inverse DCT of one dequantised block into out[8 rows x stride] (stb_image.h:2242-2330)
    static inline int f2f(double x) { return (int)(x * 4096 + 0.5); }
    struct Idct1D { int x0, x1, x2, x3, t0, t1, t2, t3; };
    static inline Idct1D idct_1d(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7) {
        // even part
        int p2 = s2, p3 = s6;
        int p1 = (p2 + p3) * f2f(0.5411961f);
        int t2 = p1 + p3 * f2f(-1.847759065f);
        int t3 = p1 + p2 * f2f(0.765366865f);
        p2 = s0; p3 = s4;
        int t0 = (p2 + p3) * 4096, t1 = (p2 - p3) * 4096;
        Idct1D r;
        r.x0 = t0 + t3; r.x3 = t0 - t3; r.x1 = t1 + t2; r.x2 = t1 - t2;
        // odd part
        t0 = s7; t1 = s5; t2 = s3; t3 = s1;
        p3 = t0 + t2; int p4 = t1 + t3;
        p1 = t0 + t3; p2 = t1 + t2;
        int p5 = (p3 + p4) * f2f(1.175875602f);
        t0 = t0 * f2f(0.298631336f);
        t1 = t1 * f2f(2.053119869f);
        t2 = t2 * f2f(3.072711026f);
        t3 = t3 * f2f(1.501321110f);
        p1 = p5 + p1 * f2f(-0.899976223f);
        p2 = p5 + p2 * f2f(-2.562915447f);
        p3 = p3 * f2f(-1.961570560f);
        p4 = p4 * f2f(-0.390180644f);
        r.t3 = t3 + p1 + p4; r.t2 = t2 + p2 + p3; r.t1 = t1 + p2 + p4; r.t0 = t0 + p1 + p3;
        return r;
    }
    static inline uint8_t clamp8(int x) { return (uint8_t)(x < 0 ? 0 : x > 255 ? 255 : x); }
    static void idct_block(uint8_t *out, int stride, const short d[64]) {
        int val[64];
        for (int i = 0; i < 8; ++i) {                       // columns; keep 2 extra bits
            const short *c = d + i;
            int *v = val + i;
            if (c[8] == 0 && c[16] == 0 && c[24] == 0 && c[32] == 0 && c[40] == 0 && c[48] == 0 && c[56] == 0) {
                int dc = c[0] * 4;
                for (int k = 0; k < 8; ++k) v[8 * k] = dc;
                continue;
            }
            Idct1D r = idct_1d(c[0], c[8], c[16], c[24], c[32], c[40], c[48], c[56]);
            r.x0 += 512; r.x1 += 512; r.x2 += 512; r.x3 += 512;
            v[0] = (r.x0 + r.t3) >> 10; v[56] = (r.x0 - r.t3) >> 10;
            v[8] = (r.x1 + r.t2) >> 10; v[48] = (r.x1 - r.t2) >> 10;
            v[16] = (r.x2 + r.t1) >> 10; v[40] = (r.x2 - r.t1) >> 10;
            v[24] = (r.x3 + r.t0) >> 10; v[32] = (r.x3 - r.t0) >> 10;
        }
        for (int i = 0; i < 8; ++i) {                       // rows; remove 17 bits, round, re-centre on 128
            const int *v = val + 8 * i;
            uint8_t *o = out + (size_t)i * stride;
            Idct1D r = idct_1d(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
            const int bias = 65536 + (128 << 17);
            r.x0 += bias; r.x1 += bias; r.x2 += bias; r.x3 += bias;
            o[0] = clamp8((r.x0 + r.t3) >> 17); o[7] = clamp8((r.x0 - r.t3) >> 17);
            o[1] = clamp8((r.x1 + r.t2) >> 17); o[6] = clamp8((r.x1 - r.t2) >> 17);
            o[2] = clamp8((r.x2 + r.t1) >> 17); o[5] = clamp8((r.x2 - r.t1) >> 17);
            o[3] = clamp8((r.x3 + r.t0) >> 17); o[4] = clamp8((r.x3 - r.t0) >> 17);
        }
    }

    bool restart_if_due(int &todo) {
        if (restart_interval_ == 0) return true;
        if (--todo > 0) return true;
        // byte-align, expect RSTn; reset predictors
        if (nbits_ < 24) fill();
        if (!hit_marker_) {
            // skip whatever is left of the segment up to the marker
            while (pos_ + 1 < n_ && !(d_[pos_] == 0xFF && d_[pos_ + 1] != 0 && d_[pos_ + 1] != 0xFF)) ++pos_;
        }
        if (pos_ + 1 < n_ && d_[pos_] == 0xFF && d_[pos_ + 1] >= 0xD0 && d_[pos_ + 1] <= 0xD7) pos_ += 2;
        else return true;                                   // no RST where one was due: let the data run out (stb stops here too)
        reset_bits();
        for (int i = 0; i < ncomp_; ++i) comp_[i].dc_pred = 0;
        eobrun_ = 0;
        todo = restart_interval_;
        return true;
    }

    bool decode_scan() {
        reset_bits();
        for (int i = 0; i < ncomp_; ++i) comp_[i].dc_pred = 0;
        eobrun_ = 0;
        int todo = restart_interval_ ? restart_interval_ : 0x7fffffff;
        short blk[64];
        if (scan_n_ == 1) {
            // non-interleaved: the component's own block grid, ceil(x/8) x ceil(y/8)
            Component &c = comp_[order_[0]];
            const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
            for (int j = 0; j < h; ++j)
                for (int i = 0; i < w; ++i) {
                    if (progressive_) {
                        short *b = c.coeff.data() + 64 * ((size_t)i + (size_t)j * c.bw);
                        if (!(ss_ == 0 ? prog_dc(c, b) : prog_ac(c, b))) return false;
                    } else {
                        if (!baseline_block(c, blk)) return false;
                        idct_block(c.plane.data() + (size_t)c.w2 * j * 8 + (size_t)i * 8, c.w2, blk);
                    }
                    restart_if_due(todo);
                }
        } else {
            for (int j = 0; j < mcuy_; ++j)
                for (int i = 0; i < mcux_; ++i) {
                    for (int k = 0; k < scan_n_; ++k) {
                        Component &c = comp_[order_[k]];
                        for (int y = 0; y < c.v; ++y)
                            for (int x = 0; x < c.h; ++x) {
                                const int bx = i * c.h + x, by = j * c.v + y;
                                if (progressive_) {
                                    if (ss_ != 0) return false;     // AC scans are never interleaved
                                    if (!prog_dc(c, c.coeff.data() + 64 * ((size_t)bx + (size_t)by * c.bw))) return false;
                                } else {
                                    if (!baseline_block(c, blk)) return false;
                                    idct_block(c.plane.data() + (size_t)c.w2 * by * 8 + (size_t)bx * 8, c.w2, blk);
                                }
                            }
                    }
                    restart_if_due(todo);
                }
        }
        return true;
    }

    void finish_progressive() {
        short blk[64];
        for (int n = 0; n < ncomp_; ++n) {
            Component &c = comp_[n];
            const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
            const uint16_t *dq = dequant_[c.tq];
            for (int j = 0; j < h; ++j)
                for (int i = 0; i < w; ++i) {
                    const short *src = c.coeff.data() + 64 * ((size_t)i + (size_t)j * c.bw);
                    for (int k = 0; k < 64; ++k) blk[k] = (short)(src[k] * dq[k]);     // 16-bit product, as stored
                    idct_block(c.plane.data() + (size_t)c.w2 * j * 8 + (size_t)i * 8, c.w2, blk);
                }
        }
    }

    // ---- upsampling (stb_image.h:3251-3345) ----
    static const uint8_t *up_row(uint8_t *out, const uint8_t *nearr, const uint8_t *farr, int w, int hs, int vs) {
        if (hs == 1 && vs == 1) return nearr;
        if (hs == 1 && vs == 2) {
            for (int i = 0; i < w; ++i) out[i] = (uint8_t)((3 * nearr[i] + farr[i] + 2) >> 2);
            return out;
        }
        if (hs == 2 && vs == 1) {
            if (w == 1) { out[0] = out[1] = nearr[0]; return out; }
            out[0] = nearr[0];
            out[1] = (uint8_t)((nearr[0] * 3 + nearr[1] + 2) >> 2);
            int i;
            for (i = 1; i < w - 1; ++i) {
                const int n = 3 * nearr[i] + 2;
                out[i * 2 + 0] = (uint8_t)((n + nearr[i - 1]) >> 2);
                out[i * 2 + 1] = (uint8_t)((n + nearr[i + 1]) >> 2);
            }
            out[i * 2 + 0] = (uint8_t)((nearr[w - 2] * 3 + nearr[w - 1] + 2) >> 2);
            out[i * 2 + 1] = nearr[w - 1];
            return out;
        }
        if (hs == 2 && vs == 2) {
            if (w == 1) { out[0] = out[1] = (uint8_t)((3 * nearr[0] + farr[0] + 2) >> 2); return out; }
            int t1 = 3 * nearr[0] + farr[0];
            out[0] = (uint8_t)((t1 + 2) >> 2);
            for (int i = 1; i < w; ++i) {
                const int t0 = t1;
                t1 = 3 * nearr[i] + farr[i];
                out[i * 2 - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
                out[i * 2] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
            }
            out[w * 2 - 1] = (uint8_t)((t1 + 2) >> 2);
            return out;
        }
        for (int i = 0; i < w; ++i)                          // other factors: nearest
            for (int j = 0; j < hs; ++j) out[i * hs + j] = nearr[i];
        return out;
    }

    bool output(std::vector<uint8_t> &rgb, int &nx, int &ny, int &channels) {
        nx = width_; ny = height_;
        channels = ncomp_ >= 3 ? 3 : 1;
        const bool is_rgb = ncomp_ == 3 && (rgb_ids_ == 3 || (adobe_transform_ == 0 && !jfif_));
        rgb.assign((size_t)width_ * height_ * channels, 0);
        struct Up { int hs, vs, wl, ystep, ypos; const uint8_t *line0, *line1; std::vector<uint8_t> buf; } up[3];
        for (int k = 0; k < ncomp_; ++k) {
            Component &c = comp_[k];
            up[k].hs = hmax_ / c.h; up[k].vs = vmax_ / c.v;
            up[k].ystep = up[k].vs >> 1;
            up[k].wl = (width_ + up[k].hs - 1) / up[k].hs;
            up[k].ypos = 0;
            up[k].line0 = up[k].line1 = c.plane.data();
            up[k].buf.assign((size_t)up[k].wl * up[k].hs + 16, 0);
        }
        const int cr20 = ((int)(1.40200f * 4096.0f + 0.5f)) << 8, cg_cr = ((int)(0.71414f * 4096.0f + 0.5f)) << 8;
        const int cg_cb = ((int)(0.34414f * 4096.0f + 0.5f)) << 8, cb20 = ((int)(1.77200f * 4096.0f + 0.5f)) << 8;
        for (int j = 0; j < height_; ++j) {
            const uint8_t *row[3] = {nullptr, nullptr, nullptr};
            for (int k = 0; k < ncomp_; ++k) {
                Up &r = up[k];
                const bool bot = r.ystep >= (r.vs >> 1);
                row[k] = up_row(r.buf.data(), bot ? r.line1 : r.line0, bot ? r.line0 : r.line1, r.wl, r.hs, r.vs);
                if (++r.ystep >= r.vs) {
                    r.ystep = 0;
                    r.line0 = r.line1;
                    if (++r.ypos < comp_[k].y) r.line1 += comp_[k].w2;
                }
            }
            uint8_t *out = rgb.data() + (size_t)j * width_ * channels;
            if (channels == 1) {
                std::memcpy(out, row[0], (size_t)width_);
            } else if (is_rgb) {
                for (int i = 0; i < width_; ++i) { out[3 * i] = row[0][i]; out[3 * i + 1] = row[1][i]; out[3 * i + 2] = row[2][i]; }
            } else {
                for (int i = 0; i < width_; ++i) {           // stb_image.h:3457-3480
                    const int yf = (row[0][i] << 20) + (1 << 19);
                    const int cr = row[2][i] - 128, cb = row[1][i] - 128;
                    int r = yf + cr * cr20;
                    int g = yf + (cr * -cg_cr) + (int)((unsigned)(cb * -cg_cb) & 0xffff0000u);
                    int b = yf + cb * cb20;
                    r >>= 20; g >>= 20; b >>= 20;
                    out[3 * i] = clamp8(r); out[3 * i + 1] = clamp8(g); out[3 * i + 2] = clamp8(b);
                }
            }
        }
        return true;
    }
};

// Decodes a JPEG file to RGB8 (or grey8).  Returns a new[]-allocated buffer like the scene API's other loaders, or nullptr.
inline unsigned char *load_file(const char *path, int *nx, int *ny, int *channels) {
    FILE *f = std::fopen(path, "rb");
    if (!f) return nullptr;
    std::vector<uint8_t> bytes;
    uint8_t chunk[65536];
    size_t got;
    while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) bytes.insert(bytes.end(), chunk, chunk + got);
    std::fclose(f);
    std::vector<uint8_t> px;
    Decoder dec;
    int w = 0, h = 0, c = 0;
    if (!dec.decode(bytes.data(), bytes.size(), px, w, h, c)) return nullptr;
    unsigned char *out = new unsigned char[px.size()];
    std::memcpy(out, px.data(), px.size());
    *nx = w; *ny = h; *channels = c;
    return out;
}

}  // namespace jpeg
}  // namespace rtow
