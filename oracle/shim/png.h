/* TEST INFRASTRUCTURE ONLY.
 * Stub of the nine libpng entry points the reference's image/wpng.h:60-87 names, so the
 * reference's hot-path headers compile unmodified in a container without libpng.
 * The oracle harness never calls write_png, so these are declarations that abort if reached. */
#pragma once
#include <csetjmp>
#include <cstdio>
#include <cstdlib>
typedef struct png_struct_def { jmp_buf jb; } png_struct;
typedef png_struct* png_structp;
typedef struct png_info_def { int unused; } png_info;
typedef png_info* png_infop;
typedef unsigned char* png_bytep;
#define PNG_LIBPNG_VER_STRING "stub"
#define PNG_COLOR_TYPE_RGB 2
#define PNG_INTERLACE_NONE 0
#define PNG_COMPRESSION_TYPE_BASE 0
#define PNG_FILTER_TYPE_BASE 0
#define png_jmpbuf(p) ((p)->jb)
static inline png_structp png_create_write_struct(const char*, void*, void*, void*) { std::abort(); }
static inline png_infop png_create_info_struct(png_structp) { std::abort(); }
static inline void png_destroy_write_struct(png_structp*, png_infop*) { std::abort(); }
static inline void png_init_io(png_structp, FILE*) { std::abort(); }
static inline void png_set_IHDR(png_structp, png_infop, int, int, int, int, int, int, int) { std::abort(); }
static inline void png_write_info(png_structp, png_infop) { std::abort(); }
static inline void png_write_image(png_structp, png_bytep*) { std::abort(); }
static inline void png_write_end(png_structp, png_infop) { std::abort(); }
