/* png_io.h -- minimal PNG (8-bit RGBA, non-interlaced) writer/reader over zlib for the raytrace app.
 * Stands in for cocogfx SaveImage / LoadImage / CompareImages (reference
 * third_party/cocogfx/src/imageutil.cpp:30-186), whose libpng dependency is absent here. */
#ifndef SRT_PNG_IO_H
#define SRT_PNG_IO_H
#include <stdint.h>

/* Save an A8R8G8B8 image.  `bits` points at the first row to write and `pitch` (bytes, may be
 * negative) steps to the next one -- the SaveImage(bits + (H-1)*pitch, W, H, -pitch) idiom of
 * tests/regression/draw3d/main.cpp:383-386 therefore flips the image vertically.  0 on success. */
int png_save_argb(const char* path, const uint8_t* bits, uint32_t width, uint32_t height, int64_t pitch);

/* Load an 8-bit RGB/RGBA non-interlaced PNG as A8R8G8B8 (malloc'ed, top row first). 0 on success. */
int png_load_argb(const char* path, uint32_t** pixels, uint32_t* width, uint32_t* height);

/* Per-channel |a-b| > tolerance counts one error per pixel (imageutil.cpp:133-186).
 * Returns the error count, or -1 if a file cannot be read or the sizes differ. */
int png_compare(const char* file_a, const char* file_b, int tolerance);
#endif
