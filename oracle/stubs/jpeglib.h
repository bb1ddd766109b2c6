/* Stub jpeglib.h (ours, not libjpeg's): just enough forward declarations for the reference's
 * helper *headers* to parse when oracle/Makefile compiles the reference hot path
 * (gainmapmath.cpp + ultrahdr.cpp:1-558). No libjpeg function is declared, called or linked. */
#ifndef UHDR_B200_STUB_JPEGLIB_H
#define UHDR_B200_STUB_JPEGLIB_H
#include <stddef.h>
typedef unsigned char JOCTET;
typedef int boolean;
struct jpeg_common_struct { int unused; };
struct jpeg_compress_struct { int unused; };
struct jpeg_decompress_struct { int unused; };
typedef struct jpeg_common_struct* j_common_ptr;
typedef struct jpeg_compress_struct* j_compress_ptr;
typedef struct jpeg_decompress_struct* j_decompress_ptr;
#endif
