/* Stub libheif/heif.h (ours): forward declarations only, so heifr.h parses. The reference hot
 * path (ultrahdr.cpp:1-558) uses nothing from libheif. */
#ifndef UHDR_B200_STUB_HEIF_H
#define UHDR_B200_STUB_HEIF_H
#include <stddef.h>
#include <stdint.h>
struct heif_image;
struct heif_context;
struct heif_image_handle;
struct heif_encoder;
struct heif_reader;
struct heif_writer;
enum heif_channel { heif_channel_Y = 0 };
enum heif_colorspace { heif_colorspace_undefined = 99 };
enum heif_chroma { heif_chroma_undefined = 99 };
enum heif_compression_format { heif_compression_undefined = 0 };
struct heif_error { int code; int subcode; const char* message; };
#endif
