// wire.h — the one numerically relevant piece of the proto surface: `Update`
// (src/protos/serverless_learn.proto:81-83), `repeated double delta = 1` in proto3 = packed:
// tag 0x0A, varint byte length, little-endian doubles.  An empty field serialises to nothing.
#pragma once
#include <stdint.h>
#include <string.h>

namespace slw {

inline long update_encode(const double* delta, int n, void* out_, long cap) {
    if (n <= 0) return 0;
    unsigned char* out = static_cast<unsigned char*>(out_);
    const uint64_t len = 8ull * (uint64_t)n;
    unsigned char hdr[12];
    int h = 0;
    hdr[h++] = 0x0A;
    for (uint64_t v = len;;) {
        if (v >= 0x80) { hdr[h++] = (unsigned char)(v | 0x80); v >>= 7; }
        else { hdr[h++] = (unsigned char)v; break; }
    }
    if ((long)(h + len) > cap) return -1;
    memcpy(out, hdr, (size_t)h);
    memcpy(out + h, delta, len);
    return (long)(h + len);
}

inline int update_decode(const void* buf_, long len, double* out, int cap) {
    const unsigned char* buf = static_cast<const unsigned char*>(buf_);
    long pos = 0;
    int n = 0;
    while (pos < len) {
        const unsigned char tag = buf[pos++];
        if (tag == 0x0A) {
            uint64_t l = 0;
            for (int shift = 0;; shift += 7) {
                if (pos >= len || shift > 63) return -1;
                const unsigned char b = buf[pos++];
                l |= (uint64_t)(b & 0x7F) << shift;
                if (!(b & 0x80)) break;
            }
            // lengths are compared unsigned and bounded by what is left of the buffer and of
            // `out` BEFORE any narrowing: a varint of 2^63 would otherwise turn negative
            if (l % 8 != 0 || l > (uint64_t)(len - pos)) return -1;
            if (l / 8 > (uint64_t)(cap - n)) return -1;
            const int k = (int)(l / 8);
            memcpy(out + n, buf + pos, l);
            n += k;
            pos += (long)l;
        } else if (tag == 0x09) {  // unpacked element: legal input for a proto3 parser
            if (pos + 8 > len || n + 1 > cap) return -1;
            memcpy(out + n, buf + pos, 8);
            n += 1;
            pos += 8;
        } else {
            return -1;
        }
    }
    return n;
}

}  // namespace slw
