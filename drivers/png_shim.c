/* drivers/png_shim.c -- the eight libpng entry points of include/shim/png.h, over csrc/png_io.c.
 * Only what /root/reference/src/cpu_ray.c:108-165 (png_dump) needs: 8-bit RGB, non-interlaced,
 * rows handed over with png_set_rows and written by png_write_png.  Driver-side glue; not on the hot path. */
#include <png.h>
#include <stdlib.h>
#include <string.h>

#include "../example-gui-opencl-raytracer_b200/csrc/png_io.h"

struct rtb_png_struct {
    FILE* fp;
};
struct rtb_png_info {
    png_uint_32 w, h;
    int depth, ctype, interlace;
    png_bytepp rows;
};

png_structp png_create_write_struct(png_const_charp ver, png_voidp eptr, png_error_ptr efn, png_error_ptr wfn) {
    (void)ver; (void)eptr; (void)efn; (void)wfn;
    return (png_structp)calloc(1, sizeof(struct rtb_png_struct));
}
png_infop png_create_info_struct(png_structp p) {
    (void)p;
    return (png_infop)calloc(1, sizeof(struct rtb_png_info));
}
void png_set_IHDR(png_structp p, png_infop i, png_uint_32 w, png_uint_32 h, int depth, int ctype, int interlace,
                  int comp, int filt) {
    (void)p; (void)comp; (void)filt;
    i->w = w; i->h = h; i->depth = depth; i->ctype = ctype; i->interlace = interlace;
}
png_voidp png_malloc(png_structp p, png_alloc_size_t n) {
    (void)p;
    void* m = malloc(n ? n : 1);
    if (!m) { fprintf(stderr, "png shim: out of memory\n"); abort(); } /* libpng's png_malloc never returns NULL */
    return m;
}
void png_free(png_structp p, png_voidp m) { (void)p; free(m); }
void png_init_io(png_structp p, FILE* fp) { p->fp = fp; }
void png_set_rows(png_structp p, png_infop i, png_bytepp rows) { (void)p; i->rows = rows; }
void png_write_png(png_structp p, png_infop i, int transforms, png_voidp params) {
    (void)transforms; (void)params;
    if (!p->fp || !i->rows || i->depth != 8 || i->ctype != PNG_COLOR_TYPE_RGB || i->interlace != PNG_INTERLACE_NONE) {
        fprintf(stderr, "png shim: only 8-bit non-interlaced RGB is supported\n");
        return;
    }
    const size_t stride = (size_t)i->w * 3;
    uint8_t* flat = (uint8_t*)malloc((stride * i->h) != 0 ? stride * i->h : 1);
    if (!flat) return;
    for (png_uint_32 r = 0; r < i->h; r++) memcpy(flat + stride * r, i->rows[r], stride);
    int rc = rtpng_write_rgb8_fp(p->fp, flat, i->w, i->h);
    if (rc != RTPNG_OK) fprintf(stderr, "png shim: write failed (%d)\n", rc);
    free(flat);
}
