/*
 * Symbol-level parity of the convolution shim (shim/src/convolution/convolution-run.c) against the
 * reference's own object code: the unmodified src/convolution/convolution-run.c is compiled from where
 * it lies with its public symbols renamed ref_* on the command line (shim/Makefile), and every symbol
 * src/convolution/convolution-run.h:45-46,58-60,70-71 declares is called on both sides with the same
 * inputs and compared bit for bit.  TEST INFRASTRUCTURE.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "convolution-run.h"

double ref_apply_kernel_to_padded_point(ImageWithPadding *padded_img, Image *kernel, int pointX, int pointY);
void ref_apply_kernel_to_padded_image(ImageWithPadding *padded_img, Image *kernel, Args *args, ImageWithPadding *buffer);
void ref_run_on_padded_image(ImageWithPadding **padded_img, Image *kernel, Args *args, ImageWithPadding **buffer);

static uint64_t rng_state = 0xD1B54A32D192ED03ull;
static double rnd(void) {
    rng_state ^= rng_state >> 12; rng_state ^= rng_state << 25; rng_state ^= rng_state >> 27;
    return (double) ((rng_state * 0x2545F4914F6CDD1Dull) >> 11) * (1.0 / 9007199254740992.0);
}

static int failures = 0, checks = 0;
static void same_padded(const char *what, const ImageWithPadding *a, const ImageWithPadding *b) {
    /* all real rows incl. their pad columns (pad rows are aliases of the edge rows) */
    for (int y = 0; y < a->inner_height; ++y) {
        ++checks;
        if (memcmp(a->image[y + a->padding], b->image[y + b->padding], sizeof(double) * (size_t) a->width) != 0)
            if (failures++ < 10) printf("MISMATCH %s row %d (w=%d h=%d p=%d)\n", what, y, a->inner_width, a->inner_height, a->padding);
    }
}

static ImageWithPadding *random_padded(int w, int h, int p) {
    Image *img = init_image(w, h, 0.0);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) img->image[y][x] = rnd() * 2 - 1;
    ImageWithPadding *pad = add_padding(img, p);
    free_image(img);
    return pad;
}

static ImageWithPadding *clone_padded(ImageWithPadding *src) {
    ImageWithPadding *dst = init_padded_image(src->inner_width, src->inner_height, src->padding);
    copy_padded_image(src, dst);
    return dst;
}

int main(void) {
    const int shapes[][5] = {   /* w, h, padding, kw, kh (kw, kh <= 2p+1 as the reference requires) */
        {37, 19, 1, 3, 3}, {64, 5, 2, 5, 5}, {300, 70, 3, 7, 7}, {33, 40, 2, 3, 5}, {50, 3, 3, 5, 2},
        {1, 1, 1, 3, 3}, {2, 7, 1, 1, 3}, {129, 33, 2, 4, 4}};
    for (unsigned s = 0; s < sizeof shapes / sizeof shapes[0]; ++s) {
        const int w = shapes[s][0], h = shapes[s][1], p = shapes[s][2], kw = shapes[s][3], kh = shapes[s][4];
        Image *kernel = init_image(kw, kh, 0.0);
        for (int y = 0; y < kh; ++y)
            for (int x = 0; x < kw; ++x) kernel->image[y][x] = rnd() * 2 - 1;
        Args args;
        memset(&args, 0, sizeof args);
        args.number_of_processes = 1; args.width = w; args.height = h;
        ImageWithPadding *img = random_padded(w, h, p);
        /* stale pad columns: the reference reads them as they are */
        img->image[p][0] = 7.5; img->image[p + h - 1][img->width - 1] = -3.25;

        /* apply_kernel_to_padded_point on every pixel of small images, a sample of large ones */
        const int stride = (w * h > 2000) ? 37 : 1;
        for (int i = 0; i < w * h; i += stride) {
            const int x = i % w, y = i / w;
            const double got = apply_kernel_to_padded_point(img, kernel, x, y);
            const double want = ref_apply_kernel_to_padded_point(img, kernel, x, y);
            ++checks;
            if (memcmp(&got, &want, sizeof got) != 0)
                if (failures++ < 10) printf("MISMATCH point (%d,%d) shape %u: %a vs %a\n", x, y, s, got, want);
        }

        /* apply_kernel_to_padded_image: inner region of the buffer written, its pad columns kept */
        ImageWithPadding *ba = init_padded_image(w, h, p), *bb = init_padded_image(w, h, p);
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < ba->width; ++x) { ba->image[y + p][x] = 42.0 + x; bb->image[y + p][x] = 42.0 + x; }
        apply_kernel_to_padded_image(img, kernel, &args, ba);
        ref_apply_kernel_to_padded_image(img, kernel, &args, bb);
        same_padded("apply_kernel_to_padded_image", ba, bb);

        /* run_on_padded_image: 0..4 iterations, caller's pointers swapped iff odd */
        for (int it = 0; it <= 4; ++it) {
            args.number_of_iterations = it;
            ImageWithPadding *ia = clone_padded(img), *ib = clone_padded(img);
            ImageWithPadding *ua = init_padded_image(w, h, p), *ub = init_padded_image(w, h, p);
            ImageWithPadding *ia0 = ia, *ib0 = ib;
            run_on_padded_image(&ia, kernel, &args, &ua);
            ref_run_on_padded_image(&ib, kernel, &args, &ub);
            ++checks;
            if ((ia == ia0) != (ib == ib0)) { if (failures++ < 10) printf("MISMATCH pointer swap, %d iterations\n", it); }
            same_padded("run_on_padded_image cur", ia, ib);
            if (it > 0) same_padded("run_on_padded_image prev", ua, ub);
        }
    }
    printf("%s: %d checks, %d mismatches\n", failures ? "FAILED" : "PASSED", checks, failures);
    return failures ? 1 : 0;
}
