/* png_io.h -- minimal PNG codec over zlib for the image-array upload path.
 * Stands in for the libpng calls of the reference's cl_wrap_load_images
 * (/root/reference/src/opencl_wrap.c:229-326): libpng is not installed in this image, zlib is.
 * Reads PNGs of any colour type far enough to report (w, h, bit depth, colour type, interlace); pixel data is
 * produced only for 8-bit RGB, the one format the reference accepts (:299-307), non-interlaced or Adam7 (libpng's
 * png_read_image de-interlaces, so the reference takes both).  Chunk CRCs of IHDR/IDAT and the zlib Adler-32 trailer
 * are verified: a damaged or truncated file is RTPNG_ECORRUPT, never a half-decoded texture. */
#ifndef RT_PNG_IO_H
#define RT_PNG_IO_H
#include <stdint.h>
#include <stdio.h>
#ifdef __cplusplus
extern "C" {
#endif
enum { RTPNG_OK = 0, RTPNG_EOPEN = -1, RTPNG_ENOTPNG = -2, RTPNG_EFORMAT = -3, RTPNG_ECORRUPT = -4, RTPNG_ENOMEM = -5 };
/* IHDR fields.  colour_type uses the PNG numbering (2 = RGB). */
int rtpng_info(const char* path, uint32_t* w, uint32_t* h, int* bit_depth, int* colour_type, int* interlace);
/* Decode an 8-bit RGB file into w*h RGBA8 texels with A = 255 (png_set_filler(255, AFTER)), row 0 = top. */
int rtpng_read_rgb8_to_rgba(const char* path, uint8_t* dst, uint32_t w, uint32_t h);
/* Same texels through a decoded-texel cache: cache_dir NULL or "" = no cache.  *hit (may be
 * NULL) is set to 1 when the texels came from the cache.  The entry is keyed by the source's real path and
 * validated against its size, mtime and dimensions. */
int rtpng_read_rgb8_to_rgba_cached(const char* path, uint8_t* dst, uint32_t w, uint32_t h, const char* cache_dir, int* hit);
/* Write w*h RGB8 pixels (row 0 = top) as a non-interlaced 8-bit RGB PNG. */
int rtpng_write_rgb8(const char* path, const uint8_t* rgb, uint32_t w, uint32_t h);
/* Same, onto an open stream (left open). */
int rtpng_write_rgb8_fp(FILE* fp, const uint8_t* rgb, uint32_t w, uint32_t h);
#ifdef __cplusplus
}
#endif
#endif
