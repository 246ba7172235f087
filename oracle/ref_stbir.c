/* oracle/ref_stbir.c -- TEST INFRASTRUCTURE. Exposes the two stb_image_resize2 entry points the reference calls in
 * generateMipLevels (src/scene/Texture.cpp:193-208), compiled from the reference's vendored header where it lies
 * (oracle/Makefile adds -I/root/reference/ext/stb). Built into oracle/_ref/libstbir.so (git-ignored). */
#define STB_IMAGE_RESIZE_IMPLEMENTATION
#include "stb_image_resize2.h"

/* srgb != 0: stbir_resize_uint8_srgb, else stbir_resize_uint8_linear; layout = channel count as the reference passes it */
int ref_stbir_resize(const unsigned char *in, int iw, int ih, unsigned char *out, int ow, int oh, int channels, int srgb)
{
    const stbir_pixel_layout layout = (stbir_pixel_layout)channels;
    unsigned char *r = srgb ? stbir_resize_uint8_srgb(in, iw, ih, 0, out, ow, oh, 0, layout)
                            : stbir_resize_uint8_linear(in, iw, ih, 0, out, ow, oh, 0, layout);
    return r != 0;
}
