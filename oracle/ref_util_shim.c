/* ref_util_shim.c - TEST INFRASTRUCTURE.  Compiles the reference's own conv_to_rgbi (defined in
 * the header util/util.h:33-107 of /root/reference, included where it lies, nothing copied) into
 * oracle/_ref/libref_util.so so that tests can pin orc_conv_to_rgbi against it on the CPU. */
#include <util.h>

int ref_conv_to_rgbi(int* sizes_x, int* sizes_y, int num_components, const int* ssx, const int* ssy,
                     const uint8_t* p0, const uint8_t* p1, const uint8_t* p2, uint8_t* out)
{
    struct jpeggpu_subsampling ss;
    for (int c = 0; c < JPEGGPU_MAX_COMP; ++c) {
        ss.x[c] = ssx[c];
        ss.y[c] = ssy[c];
    }
    return conv_to_rgbi(sizes_x, sizes_y, num_components, ss, p0, p1, p2, out);
}
