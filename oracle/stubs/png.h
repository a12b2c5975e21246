// Stand-in for <png.h> (libpng is not installed in this image).  The upstream program's
// snapshot writer only runs when verbose != 0; with these no-ops it writes an empty file.
// It touches no arithmetic on the measured path.
#pragma once
#include <cstdio>
typedef unsigned char* png_bytep;
typedef struct png_stub_s { int unused; }* png_structp;
typedef struct png_stub_i { int unused; }* png_infop;
#define PNG_LIBPNG_VER_STRING "stub"
#define PNG_COLOR_TYPE_RGB 2
#define PNG_INTERLACE_NONE 0
#define PNG_COMPRESSION_TYPE_DEFAULT 0
#define PNG_FILTER_TYPE_DEFAULT 0
static inline png_structp png_create_write_struct(const char*, void*, void*, void*) { return nullptr; }
static inline png_infop png_create_info_struct(png_structp) { return nullptr; }
static inline void png_init_io(png_structp, FILE*) {}
static inline void png_set_IHDR(png_structp, png_infop, int, int, int, int, int, int, int) {}
static inline void png_write_info(png_structp, png_infop) {}
static inline void png_write_image(png_structp, png_bytep*) {}
static inline void png_write_end(png_structp, png_infop) {}
static inline void png_destroy_write_struct(png_structp*, png_infop*) {}
