/* include/shim/png.h -- build shim, NOT libpng.
 *
 * The reference's src/cpu_ray.c:108-165 (png_dump) writes the frame with eight libpng calls; libpng and its
 * headers are not installed in this image.  This header declares exactly those names with
 * libpng's signatures; drivers/png_shim.c implements them over the library's own zlib-based encoder
 * (csrc/png_io.c), so cpu_ray.c compiles and links untouched.  On a machine with libpng, use libpng.
 */
#ifndef RTB_SHIM_PNG_H
#define RTB_SHIM_PNG_H
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PNG_LIBPNG_VER_STRING "1.6.0-rtb200-shim"
#define PNG_COLOR_TYPE_RGB 2
#define PNG_INTERLACE_NONE 0
#define PNG_COMPRESSION_TYPE_DEFAULT 0
#define PNG_FILTER_TYPE_DEFAULT 0
#define PNG_TRANSFORM_IDENTITY 0

typedef unsigned char png_byte;
typedef png_byte* png_bytep;
typedef png_byte** png_bytepp;
typedef uint32_t png_uint_32;
typedef size_t png_alloc_size_t;
typedef void* png_voidp;
typedef const char* png_const_charp;
typedef struct rtb_png_struct png_struct;
typedef png_struct* png_structp;
typedef struct rtb_png_info png_info;
typedef png_info* png_infop;
typedef void (*png_error_ptr)(png_structp, png_const_charp);

png_structp png_create_write_struct(png_const_charp user_png_ver, png_voidp error_ptr, png_error_ptr error_fn,
                                    png_error_ptr warn_fn);
png_infop png_create_info_struct(png_structp png_ptr);
void png_set_IHDR(png_structp png_ptr, png_infop info_ptr, png_uint_32 width, png_uint_32 height, int bit_depth,
                  int color_type, int interlace_method, int compression_method, int filter_method);
png_voidp png_malloc(png_structp png_ptr, png_alloc_size_t size);
void png_free(png_structp png_ptr, png_voidp ptr);
void png_init_io(png_structp png_ptr, FILE* fp);
void png_set_rows(png_structp png_ptr, png_infop info_ptr, png_bytepp row_pointers);
void png_write_png(png_structp png_ptr, png_infop info_ptr, int transforms, png_voidp params);

#ifdef __cplusplus
}
#endif
#endif
