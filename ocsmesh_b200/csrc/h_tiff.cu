// Host side of the raster read (on-disk format): the
// byte decoders a windowed GeoTIFF read needs for the files GDAL usually
// writes.  The reference reads through rasterio / GDAL / libtiff
// (ocsmesh/raster.py:502-531); none of them is part of the reference tree, so
// the two schemes are restated from the TIFF 6.0 specification (section 9
// PackBits, section 13 LZW).  DEFLATE goes through Python's zlib, predictors
// are undone in numpy (ocsmesh_b200/geotiff.py).  Pure host code, no ctx.
#include "ocsb_common.cuh"

namespace {

// TIFF LZW: codes are packed MSB first, start at 9 bits, 256 = Clear, 257 =
// EndOfInformation, and the width grows one code EARLY (when the next free
// entry is 511 / 1023 / 2047), as every post-1988 writer (libtiff) does.
int64_t lzw_decode(const uint8_t *src, int64_t n, uint8_t *dst, int64_t cap) {
    constexpr int CLEAR = 256, EOI = 257, FIRST = 258, MAXC = 4096;
    std::vector<int32_t> prefix(MAXC), length(MAXC);
    std::vector<uint8_t> suffix(MAXC), first(MAXC);
    for (int i = 0; i < 256; ++i) {
        prefix[i] = -1; length[i] = 1; suffix[i] = (uint8_t)i; first[i] = (uint8_t)i;
    }
    int next = FIRST, bits = 9, old = -1;
    uint64_t acc = 0;
    int have = 0;
    int64_t ip = 0, op = 0;
    auto emit = [&](int code) -> bool {
        const int len = length[code];
        if (op + len > cap) return false;
        uint8_t *p = dst + op + len;
        for (int c = code; c >= 0; c = prefix[c]) *--p = suffix[c];
        op += len;
        return true;
    };
    for (;;) {
        while (have < bits && ip < n) { acc = (acc << 8) | src[ip++]; have += 8; }
        if (have < bits) break;                       // data ended without EOI
        const int code = (int)((acc >> (have - bits)) & ((1u << bits) - 1));
        have -= bits;
        if (code == EOI) break;
        if (code == CLEAR) { next = FIRST; bits = 9; old = -1; continue; }
        if (old < 0) {                                // first code after a Clear
            if (code >= 256) return -1;
            if (!emit(code)) return -2;
            old = code;
            continue;
        }
        if (code < next) {
            if (!emit(code)) return -2;
            if (next < MAXC) {
                prefix[next] = old; suffix[next] = first[code];
                first[next] = first[old]; length[next] = length[old] + 1; ++next;
            }
        } else if (code == next && next < MAXC) {     // the KwKwK case
            prefix[next] = old; suffix[next] = first[old];
            first[next] = first[old]; length[next] = length[old] + 1; ++next;
            if (!emit(code)) return -2;
        } else {
            return -1;                                // corrupt stream
        }
        old = code;
        if (next == (1 << bits) - 1 && bits < 12) ++bits;
    }
    return op;
}

int64_t packbits_decode(const uint8_t *src, int64_t n, uint8_t *dst, int64_t cap) {
    int64_t ip = 0, op = 0;
    while (ip < n) {
        const int8_t c = (int8_t)src[ip++];
        if (c >= 0) {
            const int64_t len = (int64_t)c + 1;
            if (ip + len > n) return -1;
            if (op + len > cap) return -2;
            memcpy(dst + op, src + ip, (size_t)len);
            ip += len; op += len;
        } else if (c != -128) {
            const int64_t len = 1 - (int64_t)c;
            if (ip >= n) return -1;
            if (op + len > cap) return -2;
            memset(dst + op, src[ip++], (size_t)len);
            op += len;
        }
    }
    return op;
}

}  // namespace

extern "C" int32_t ocsb_tiff_decode(int32_t compression, const uint8_t *src, int64_t n,
                                    uint8_t *dst, int64_t cap, int64_t *n_out) {
    OCSB_REQUIRE(src && dst && n_out && n >= 0 && cap >= 0,
                 "ocsb_tiff_decode: bad arguments");
    int64_t got;
    if (compression == 5)
        got = lzw_decode(src, n, dst, cap);
    else if (compression == 32773)
        got = packbits_decode(src, n, dst, cap);
    else
        OCSB_REQUIRE(false, "ocsb_tiff_decode: unsupported compression %d", compression);
    OCSB_REQUIRE(got != -1, "ocsb_tiff_decode: corrupt %s stream",
                 compression == 5 ? "LZW" : "PackBits");
    OCSB_REQUIRE(got != -2, "ocsb_tiff_decode: the segment decodes to more than %lld bytes",
                 (long long)cap);
    *n_out = got;
    return OCSB_OK;
}
