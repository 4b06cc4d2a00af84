/*
 * Source-level drop-in for the reference's src/convolution/convolution-run.c.
 *
 * Overlay it on src/convolution/convolution-run.c of hg-c-benchmarks, add this
 * repository's include/ to the include path and link with -lhgb200.  The
 * reference's main (src/2d-convolution.c) and its gtest suite
 * (test/convolutionTests.cpp includes this file as C++) then run the stencil on
 * the B200 through the C ABI of include/hgb200.h.  Valid as C11 and C++11.
 * No CPU fallback: a library error ends the program through bail_out().
 *
 * Marshalling: an ImageWithPadding keeps inner_height separately malloc'ed rows
 * of `width` doubles; its top/bottom pad rows are aliases of the first/last real
 * row (convolution-util.c:394-400) and are therefore never copied.
 */
#include "convolution-run.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "hgb200.h"

static void hgb_conv_or_die(int rc) {
    if (rc != HGB_OK) {
        static char msg[1200];
        snprintf(msg, sizeof msg, "libhgb200 (%d): %s\n", rc, hgb_last_error());
        bail_out(msg);
    }
}

static double *hgb_rows_gather(const ImageWithPadding *img) {
    const size_t pitch = (size_t) img->width;
    double *flat = (double *) malloc(sizeof(double) * pitch * (size_t) img->inner_height);
    if (flat == NULL) bail_out((char *) "out of host memory while staging the image\n");
    for (int y = 0; y < img->inner_height; ++y)
        memcpy(flat + pitch * (size_t) y, img->image[y + img->padding], sizeof(double) * pitch);
    return flat;
}

static void hgb_rows_scatter(ImageWithPadding *img, const double *flat) {
    const size_t pitch = (size_t) img->width;
    for (int y = 0; y < img->inner_height; ++y)
        memcpy(img->image[y + img->padding], flat + pitch * (size_t) y, sizeof(double) * pitch);
}

static double *hgb_kernel_flat(const Image *kernel) {
    double *k = (double *) malloc(sizeof(double) * (size_t) kernel->width * (size_t) kernel->height);
    if (k == NULL) bail_out((char *) "out of host memory while staging the kernel\n");
    for (int y = 0; y < kernel->height; ++y)
        memcpy(k + (size_t) y * kernel->width, kernel->image[y], sizeof(double) * (size_t) kernel->width);
    return k;
}

void run_on_padded_image(ImageWithPadding **padded_img, Image *kernel, Args *args, ImageWithPadding **buffer) {
    const int n = args->number_of_iterations;
    if (n <= 0) return;
    ImageWithPadding *img = *padded_img, *buf = *buffer;
    double *cur = hgb_rows_gather(img);
    double *prev = (double *) malloc(sizeof(double) * (size_t) img->width * (size_t) img->inner_height);
    double *k = hgb_kernel_flat(kernel);
    if (prev == NULL) bail_out((char *) "out of host memory while staging the buffer\n");
    hgb_conv_or_die(hgb_conv_run(cur, prev, img->inner_width, img->inner_height, img->padding, k,
                                 kernel->width, kernel->height, n, args->number_of_processes));
    /* The reference ping-pongs between the two structs and swaps the caller's pointers
     * once per iteration: state n ends in `img` when n is even, in `buf` when n is odd,
     * and *padded_img always names the struct holding state n. */
    if (n % 2 == 0) {
        hgb_rows_scatter(img, cur);
        hgb_rows_scatter(buf, prev);
    } else {
        hgb_rows_scatter(buf, cur);
        hgb_rows_scatter(img, prev);
        *padded_img = buf;
        *buffer = img;
    }
    free(k);
    free(prev);
    free(cur);
}

void apply_kernel_to_padded_image(ImageWithPadding *padded_img, Image *kernel, Args *args, ImageWithPadding *buffer) {
    /* one pass; only the inner region of `buffer` is written, its pad columns stay as they are */
    double *cur = hgb_rows_gather(padded_img);
    double *prev = (double *) malloc(sizeof(double) * (size_t) padded_img->width * (size_t) padded_img->inner_height);
    double *k = hgb_kernel_flat(kernel);
    if (prev == NULL) bail_out((char *) "out of host memory while staging the buffer\n");
    hgb_conv_or_die(hgb_conv_run(cur, prev, padded_img->inner_width, padded_img->inner_height,
                                 padded_img->padding, k, kernel->width, kernel->height, 1,
                                 args->number_of_processes));
    for (int y = 0; y < padded_img->inner_height; ++y)
        memcpy(&ACCESS_IMAGE(buffer, 0, y), cur + (size_t) y * padded_img->width + padded_img->padding,
               sizeof(double) * (size_t) padded_img->inner_width);
    free(k);
    free(prev);
    free(cur);
}

double apply_kernel_to_padded_point(ImageWithPadding *padded_img, Image *kernel, int pointX, int pointY) {
    /* a 1 x (2p+1) image whose rows are the rows the point's window covers and whose pad
     * columns are the window's columns: its middle row sees exactly the reference's taps */
    const int p = padded_img->padding, span = 2 * p + 1;
    double *cur = (double *) malloc(sizeof(double) * (size_t) span * span);
    double *prev = (double *) malloc(sizeof(double) * (size_t) span * span);
    double *k = hgb_kernel_flat(kernel);
    if (cur == NULL || prev == NULL) bail_out((char *) "out of host memory\n");
    for (int r = 0; r < span; ++r)
        memcpy(cur + (size_t) r * span, padded_img->image[pointY + r] + pointX, sizeof(double) * (size_t) span);
    hgb_conv_or_die(hgb_conv_run(cur, prev, 1, span, p, k, kernel->width, kernel->height, 1, 1));
    const double val = cur[(size_t) p * span + p];
    free(k);
    free(prev);
    free(cur);
    return val;
}

time_t benchmark(ImageWithPadding **image, Image *kernel, Args *args, ImageWithPadding **buffer) {
    printf("Starting Kernel...\n");
    TIC(0);
    run_on_padded_image(image, kernel, args, buffer);
    const time_t elapsed = TOC(0);   /* host-inclusive, like the reference's own clock */
    printf("Kernel time: %zu.%06zus\n", (size_t) (elapsed / 1000000), (size_t) (elapsed % 1000000));
    return elapsed;
}

Image *create_kernel(Args *args) {
    if (!args->opt_kernel_from_file) return get_default_kernel();
    return read_image_from_fd(fopen(args->kernel_file_path, "r"), args);
}

Image *create_image(Args *args) {
    if (!args->opt_image_from_file) return init_image(args->width, args->height, 1.0);
    return read_image_from_fd(fopen(args->image_file_path, "r"), args);
}
