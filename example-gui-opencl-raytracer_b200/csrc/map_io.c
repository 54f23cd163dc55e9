/* map_io.c -- scene archive reader/writer.
 *
 * v1 is the reference's format, byte for byte (/root/reference/src/cpu_obj.c:51-101): three blocks of
 * {1-byte count, raw struct array} for rsphere (96 B), rplane (96 B), rlight (48 B) -- at most 255 per type.
 * v2 lifts that limit and can carry the triangle extension: "RTM2", four little-endian uint32 counts
 * (spheres, planes, lights, triangles), then the four raw arrays (96/96/48/112-byte records).
 * The loader accepts both; a file is v1 exactly when it parses as v1 with no byte left over. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/rtb200.h"

static const size_t kRec[4] = {96, 96, 48, 112};

static int parse_v1(const unsigned char* d, size_t n, uint32_t cnt[4], size_t off[4]) {
    size_t o = 0;
    for (int k = 0; k < 3; k++) {
        if (o >= n) return 0;
        cnt[k] = d[o++];
        off[k] = o;
        o += (size_t)cnt[k] * kRec[k];
        if (o > n) return 0;
    }
    cnt[3] = 0; off[3] = o;
    return o == n;
}
static int parse_v2(const unsigned char* d, size_t n, uint32_t cnt[4], size_t off[4]) {
    if (n < 20 || memcmp(d, "RTM2", 4) != 0) return 0;
    size_t o = 20;
    for (int k = 0; k < 4; k++) {
        memcpy(&cnt[k], d + 4 + 4 * k, 4);
        off[k] = o;
        o += (size_t)cnt[k] * kRec[k];
        if (o > n) return 0;
    }
    return o == n;
}

int rtb_map_load(const char* path, rtb_map* out) {
    if (!path || !out) return -1;
    memset(out, 0, sizeof *out);
    FILE* fp = fopen(path, "rb");
    if (!fp) return -2;
    fseek(fp, 0, SEEK_END);
    long len = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    if (len < 3) { fclose(fp); return -3; }
    unsigned char* d = (unsigned char*)malloc((size_t)len);
    if (!d) { fclose(fp); return -4; }
    if (fread(d, 1, (size_t)len, fp) != (size_t)len) { free(d); fclose(fp); return -2; }
    fclose(fp);
    uint32_t cnt[4];
    size_t off[4];
    int ver = parse_v1(d, (size_t)len, cnt, off) ? 1 : (parse_v2(d, (size_t)len, cnt, off) ? 2 : 0);
    if (!ver) { free(d); return -3; }
    void** dst[4] = {&out->spheres, &out->planes, &out->lights, &out->triangles};
    uint32_t* num[4] = {&out->S, &out->P, &out->L, &out->T};
    for (int k = 0; k < 4; k++) {
        size_t nb = (size_t)cnt[k] * kRec[k];
        /* 16-byte aligned like the reference's cl_float3-bearing structs */
        if (nb && posix_memalign(dst[k], 16, nb) != 0) { rtb_map_free(out); free(d); return -4; }
        if (nb) memcpy(*dst[k], d + off[k], nb);
        *num[k] = cnt[k];
    }
    out->version = ver;
    free(d);
    return 0;
}

int rtb_map_save(const char* path, const rtb_map* m, int version) {
    if (!path || !m) return -1;
    const void* src[4] = {m->spheres, m->planes, m->lights, m->triangles};
    const uint32_t cnt[4] = {m->S, m->P, m->L, m->T};
    if (version == 1 && (m->S > 255 || m->P > 255 || m->L > 255 || m->T)) return -3; /* not expressible (cpu_obj.c:62-69) */
    if (version != 1 && version != 2) return -1;
    FILE* fp = fopen(path, "wb");
    if (!fp) return -2;
    if (version == 2) {
        fwrite("RTM2", 1, 4, fp);
        fwrite(cnt, 4, 4, fp);
    }
    for (int k = 0; k < (version == 1 ? 3 : 4); k++) {
        if (version == 1) { unsigned char c = (unsigned char)cnt[k]; fwrite(&c, 1, 1, fp); }
        if (cnt[k]) fwrite(src[k], kRec[k], cnt[k], fp);
    }
    int bad = ferror(fp);
    return (fclose(fp) != 0 || bad) ? -2 : 0;
}

void rtb_map_free(rtb_map* m) {
    if (!m) return;
    free(m->spheres); free(m->planes); free(m->lights); free(m->triangles);
    memset(m, 0, sizeof *m);
}
