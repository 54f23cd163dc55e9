/* png_io.c -- see png_io.h. */
#ifndef _GNU_SOURCE
#define _GNU_SOURCE /* realpath, st_mtim */
#endif
#include "png_io.h"

#include <limits.h>
#include <stdio.h>
#include <sys/stat.h>
#include <unistd.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

static const unsigned char kSig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};

static uint32_t be32(const unsigned char* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
}

typedef struct {
    uint32_t w, h;
    int depth, ctype, interlace;
    unsigned char* idat; /* concatenated IDAT payloads (only when wanted) */
    size_t idat_len;
} png_file;

/* Reads `len` payload bytes + the 4-byte CRC of a chunk whose type is `type` and checks the CRC (PNG spec 5.3). */
static int read_checked(FILE* fp, const unsigned char* type, unsigned char* dst, uint32_t len) {
    unsigned char c[4];
    if (len && fread(dst, 1, len, fp) != len) return 0;
    if (fread(c, 1, 4, fp) != 4) return 0;
    uLong crc = crc32(0L, (const Bytef*)type, 4);
    if (len) crc = crc32(crc, dst, len);
    return (uint32_t)crc == be32(c);
}

static int parse(const char* path, png_file* pf, int want_idat) {
    memset(pf, 0, sizeof *pf);
    FILE* fp = fopen(path, "rb");
    if (!fp) return RTPNG_EOPEN;
    unsigned char hdr[8];
    if (fread(hdr, 1, 8, fp) != 8 || memcmp(hdr, kSig, 8) != 0) { fclose(fp); return RTPNG_ENOTPNG; }
    int have_ihdr = 0, rc = RTPNG_ECORRUPT;
    size_t cap = 0;
    for (;;) {
        unsigned char ch[8];
        if (fread(ch, 1, 8, fp) != 8) break;
        uint32_t len = be32(ch);
        if (len > 0x7fffffffu) break;
        if (!memcmp(ch + 4, "IHDR", 4)) {
            unsigned char ih[13];
            if (len != 13 || !read_checked(fp, ch + 4, ih, 13)) break;
            pf->w = be32(ih); pf->h = be32(ih + 4);
            pf->depth = ih[8]; pf->ctype = ih[9]; pf->interlace = ih[12];
            have_ihdr = 1;
            if (!want_idat) { rc = RTPNG_OK; break; }
        } else if (!memcmp(ch + 4, "IDAT", 4) && want_idat) {
            if (pf->idat_len + len > cap) {
                cap = (pf->idat_len + len) * 2 + 4096;
                unsigned char* n = (unsigned char*)realloc(pf->idat, cap);
                if (!n) { rc = RTPNG_ENOMEM; break; }
                pf->idat = n;
            }
            if (!read_checked(fp, ch + 4, pf->idat + pf->idat_len, len)) break; /* short read or bad chunk CRC */
            pf->idat_len += len;
        } else if (!memcmp(ch + 4, "IEND", 4)) {
            rc = have_ihdr ? RTPNG_OK : RTPNG_ECORRUPT;
            break;
        } else {
            if (fseek(fp, (long)len + 4, SEEK_CUR)) break;
        }
    }
    fclose(fp);
    if (rc != RTPNG_OK) { free(pf->idat); pf->idat = NULL; }
    return rc;
}

int rtpng_info(const char* path, uint32_t* w, uint32_t* h, int* bit_depth, int* colour_type, int* interlace) {
    png_file pf;
    int rc = parse(path, &pf, 0);
    if (rc) return rc;
    if (w) *w = pf.w;
    if (h) *h = pf.h;
    if (bit_depth) *bit_depth = pf.depth;
    if (colour_type) *colour_type = pf.ctype;
    if (interlace) *interlace = pf.interlace;
    return RTPNG_OK;
}

static inline int paeth(int a, int b, int c) {
    int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

/* Undo the five PNG filters of one (sub-)image in place: h rows of {filter byte, w RGB8 pixels} (bpp = 3).
 * One loop per filter type (the type is per row); the first pixel has no left neighbour. */
static int unfilter_rgb8(unsigned char* raw, uint32_t w, uint32_t h) {
    const size_t stride = (size_t)w * 3;
    for (uint32_t y = 0; y < h; y++) {
        unsigned char* cur = raw + (stride + 1) * y + 1;
        const unsigned char* up = y ? raw + (stride + 1) * (y - 1) + 1 : NULL;
        int ft = cur[-1];
        if (ft > 4) return RTPNG_ECORRUPT;
        const size_t head = stride < 3 ? stride : 3;
        switch (ft) {
            case 1:
                for (size_t i = 3; i < stride; i++) cur[i] = (unsigned char)(cur[i] + cur[i - 3]);
                break;
            case 2:
                if (up) for (size_t i = 0; i < stride; i++) cur[i] = (unsigned char)(cur[i] + up[i]);
                break;
            case 3:
                for (size_t i = 0; i < head; i++) cur[i] = (unsigned char)(cur[i] + ((up ? up[i] : 0) >> 1));
                if (up) for (size_t i = 3; i < stride; i++) cur[i] = (unsigned char)(cur[i] + ((cur[i - 3] + up[i]) >> 1));
                else for (size_t i = 3; i < stride; i++) cur[i] = (unsigned char)(cur[i] + (cur[i - 3] >> 1));
                break;
            case 4:
                for (size_t i = 0; i < head; i++) cur[i] = (unsigned char)(cur[i] + paeth(0, up ? up[i] : 0, 0));
                if (up) for (size_t i = 3; i < stride; i++) cur[i] = (unsigned char)(cur[i] + paeth(cur[i - 3], up[i], up[i - 3]));
                else for (size_t i = 3; i < stride; i++) cur[i] = (unsigned char)(cur[i] + cur[i - 3]); /* paeth(a,0,0) = a */
                break;
            default: break;
        }
    }
    return RTPNG_OK;
}

/* Adam7 (PNG spec 8.2): pass p holds the pixels (x0 + i*dx, y0 + j*dy). */
static const uint32_t kA7x0[7] = {0, 4, 0, 2, 0, 1, 0}, kA7y0[7] = {0, 0, 4, 0, 2, 0, 1};
static const uint32_t kA7dx[7] = {8, 8, 4, 4, 2, 2, 1}, kA7dy[7] = {8, 8, 8, 4, 4, 2, 2};

int rtpng_read_rgb8_to_rgba(const char* path, uint8_t* dst, uint32_t w, uint32_t h) {
    png_file pf;
    int rc = parse(path, &pf, 1);
    if (rc) return rc;
    if (pf.w != w || pf.h != h || pf.depth != 8 || pf.ctype != 2 || pf.interlace > 1) { free(pf.idat); return RTPNG_EFORMAT; }
    /* the inflated stream: one image (interlace 0) or the seven Adam7 sub-images back to back, empty ones omitted */
    uint32_t pw[7], ph[7];
    int npass = 1;
    size_t raw_len = 0;
    if (pf.interlace == 0) { pw[0] = w; ph[0] = h; }
    else {
        npass = 7;
        for (int p = 0; p < 7; p++) {
            pw[p] = w > kA7x0[p] ? (w - kA7x0[p] + kA7dx[p] - 1) / kA7dx[p] : 0;
            ph[p] = h > kA7y0[p] ? (h - kA7y0[p] + kA7dy[p] - 1) / kA7dy[p] : 0;
        }
    }
    for (int p = 0; p < npass; p++) if (pw[p] && ph[p]) raw_len += ((size_t)pw[p] * 3 + 1) * ph[p];
    unsigned char* raw = (unsigned char*)malloc(raw_len + 16); /* the slack catches a stream that holds too much */
    if (!raw) { free(pf.idat); return RTPNG_ENOMEM; }
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    if (inflateInit(&zs) != Z_OK) { free(raw); free(pf.idat); return RTPNG_ECORRUPT; }
    size_t in_off = 0, out_off = 0;
    int zr = Z_OK;
    while (zr != Z_STREAM_END && out_off <= raw_len) { /* feed in < 4 GiB pieces: avail_in/out are 32-bit */
        size_t in_n = pf.idat_len - in_off, out_n = raw_len + 16 - out_off;
        if (in_n > 0x40000000u) in_n = 0x40000000u;
        if (out_n > 0x40000000u) out_n = 0x40000000u;
        zs.next_in = pf.idat + in_off; zs.avail_in = (uInt)in_n;
        zs.next_out = raw + out_off; zs.avail_out = (uInt)out_n;
        zr = inflate(&zs, Z_NO_FLUSH);
        in_off += in_n - zs.avail_in;
        out_off += out_n - zs.avail_out;
        if (zr != Z_OK && zr != Z_STREAM_END) break;
        if (zr == Z_OK && in_n == 0) break; /* input exhausted before the end of the stream */
    }
    inflateEnd(&zs);
    free(pf.idat);
    /* exactly the expected number of bytes AND the end of the zlib stream reached (inflate has then verified the
     * Adler-32 trailer): a truncated or padded stream is a corrupt file, not a texture */
    if (zr != Z_STREAM_END || out_off != raw_len) { free(raw); return RTPNG_ECORRUPT; }
    unsigned char* sub = raw;
    for (int p = 0; p < npass; p++) {
        if (!pw[p] || !ph[p]) continue;
        if (unfilter_rgb8(sub, pw[p], ph[p]) != RTPNG_OK) { free(raw); return RTPNG_ECORRUPT; }
        const size_t stride = (size_t)pw[p] * 3 + 1;
        const uint32_t x0 = pf.interlace ? kA7x0[p] : 0, y0 = pf.interlace ? kA7y0[p] : 0;
        const uint32_t dx = pf.interlace ? kA7dx[p] : 1, dy = pf.interlace ? kA7dy[p] : 1;
        for (uint32_t j = 0; j < ph[p]; j++) {
            const unsigned char* cur = sub + stride * j + 1;
            uint8_t* o = dst + ((size_t)(y0 + j * dy) * w + x0) * 4;
            for (uint32_t i = 0; i < pw[p]; i++, o += 4 * (size_t)dx) {
                o[0] = cur[3 * i]; o[1] = cur[3 * i + 1]; o[2] = cur[3 * i + 2]; o[3] = 255;
            }
        }
        sub += stride * ph[p];
    }
    free(raw);
    return RTPNG_OK;
}

/* ---- N3: decoded-texel cache ------------------------------------------------------------
 * One file per source image under cache_dir: a 40-byte header {magic, w, h, source size, source mtime (ns)}
 * followed by the w*h RGBA8 texels exactly as rtpng_read_rgb8_to_rgba produces them.  A hit is a plain read;
 * any mismatch (other size/mtime/dimensions, short file) is a miss and the entry is rewritten.  Failing to
 * write the cache is not an error. */
typedef struct {
    char magic[8];
    uint32_t w, h;
    uint64_t src_size, src_mtime_ns;
    uint64_t reserved;
} tex_cache_header;

static uint64_t fnv1a64(const char* s) {
    uint64_t hsh = 1469598103934665603ull;
    for (; *s; s++) { hsh ^= (unsigned char)*s; hsh *= 1099511628211ull; }
    return hsh;
}

int rtpng_read_rgb8_to_rgba_cached(const char* path, uint8_t* dst, uint32_t w, uint32_t h, const char* cache_dir, int* hit) {
    if (hit) *hit = 0;
    struct stat sb;
    if (!cache_dir || !*cache_dir || stat(path, &sb) != 0) return rtpng_read_rgb8_to_rgba(path, dst, w, h);
    char full[PATH_MAX];
    if (!realpath(path, full)) return rtpng_read_rgb8_to_rgba(path, dst, w, h);
    tex_cache_header want;
    memset(&want, 0, sizeof want);
    memcpy(want.magic, "RTBTEX1", 8);
    want.w = w; want.h = h;
    want.src_size = (uint64_t)sb.st_size;
    want.src_mtime_ns = (uint64_t)sb.st_mtim.tv_sec * 1000000000ull + (uint64_t)sb.st_mtim.tv_nsec;
    const size_t nbytes = (size_t)w * h * 4;
    char entry[PATH_MAX + 64];
    snprintf(entry, sizeof entry, "%s/%016llx.rgba", cache_dir, (unsigned long long)fnv1a64(full));
    FILE* fp = fopen(entry, "rb");
    if (fp) {
        tex_cache_header got;
        int ok = fread(&got, sizeof got, 1, fp) == 1 && memcmp(&got, &want, sizeof got) == 0 &&
                 fread(dst, 1, nbytes, fp) == nbytes && fgetc(fp) == EOF;
        fclose(fp);
        if (ok) { if (hit) *hit = 1; return RTPNG_OK; }
    }
    int rc = rtpng_read_rgb8_to_rgba(path, dst, w, h);
    if (rc != RTPNG_OK) return rc;
    char tmp[PATH_MAX + 96];
    snprintf(tmp, sizeof tmp, "%s.%ld.tmp", entry, (long)getpid());
    fp = fopen(tmp, "wb");
    if (fp) { /* write aside, then rename: readers never see a half-written entry */
        int ok = fwrite(&want, sizeof want, 1, fp) == 1 && fwrite(dst, 1, nbytes, fp) == nbytes;
        ok = (fclose(fp) == 0) && ok;
        if (!ok || rename(tmp, entry) != 0) remove(tmp);
    }
    return RTPNG_OK;
}

static void put_chunk(FILE* fp, const char* type, const unsigned char* data, uint32_t len) {
    unsigned char b[4] = {(unsigned char)(len >> 24), (unsigned char)(len >> 16), (unsigned char)(len >> 8), (unsigned char)len};
    fwrite(b, 1, 4, fp);
    fwrite(type, 1, 4, fp);
    if (len) fwrite(data, 1, len, fp);
    uLong crc = crc32(0L, (const Bytef*)type, 4);
    if (len) crc = crc32(crc, data, len);
    unsigned char c[4] = {(unsigned char)(crc >> 24), (unsigned char)(crc >> 16), (unsigned char)(crc >> 8), (unsigned char)crc};
    fwrite(c, 1, 4, fp);
}

int rtpng_write_rgb8_fp(FILE* fp, const uint8_t* rgb, uint32_t w, uint32_t h) {
    const size_t stride = (size_t)w * 3, raw_len = (stride + 1) * h;
    if (!fp) return RTPNG_EOPEN;
    if (raw_len > 0x7ff00000u) return RTPNG_EFORMAT;
    unsigned char* raw = (unsigned char*)malloc(raw_len ? raw_len : 1);
    if (!raw) return RTPNG_ENOMEM;
    for (uint32_t y = 0; y < h; y++) {
        raw[(stride + 1) * y] = 0;
        memcpy(raw + (stride + 1) * y + 1, rgb + stride * y, stride);
    }
    uLongf zcap = compressBound((uLong)raw_len);
    unsigned char* z = (unsigned char*)malloc(zcap);
    if (!z) { free(raw); return RTPNG_ENOMEM; }
    if (compress2(z, &zcap, raw, (uLong)raw_len, 6) != Z_OK) { free(raw); free(z); return RTPNG_ECORRUPT; }
    free(raw);
    fwrite(kSig, 1, 8, fp);
    unsigned char ih[13] = {(unsigned char)(w >> 24), (unsigned char)(w >> 16), (unsigned char)(w >> 8), (unsigned char)w,
                            (unsigned char)(h >> 24), (unsigned char)(h >> 16), (unsigned char)(h >> 8), (unsigned char)h,
                            8, 2, 0, 0, 0};
    put_chunk(fp, "IHDR", ih, 13);
    put_chunk(fp, "IDAT", z, (uint32_t)zcap);
    put_chunk(fp, "IEND", NULL, 0);
    free(z);
    return ferror(fp) ? RTPNG_EOPEN : RTPNG_OK;
}

int rtpng_write_rgb8(const char* path, const uint8_t* rgb, uint32_t w, uint32_t h) {
    FILE* fp = fopen(path, "wb");
    if (!fp) return RTPNG_EOPEN;
    int rc = rtpng_write_rgb8_fp(fp, rgb, w, h);
    if (fclose(fp) != 0 && rc == RTPNG_OK) rc = RTPNG_EOPEN;
    return rc;
}
