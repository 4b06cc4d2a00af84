/*
 * ref_bridge_conv.c -- flat-array bridge onto the UNMODIFIED reference
 * convolution sources.  TEST INFRASTRUCTURE ONLY (see oracle/hg_oracle.c).
 *
 * Compiled by oracle/Makefile together with /root/reference/src/util/util.c,
 * src/convolution/convolution-util.c and src/convolution/convolution-run.c
 * into oracle/_ref/libhgref_conv.so.  It builds the reference's own structs
 * with the reference's own allocators and calls the reference's
 * run_on_padded_image (src/convolution/convolution-run.c:41-49); nothing here
 * computes a pixel.
 */
#include "convolution/convolution-run.h"

#include <string.h>

static void rows_in(ImageWithPadding *img, const double *flat) {
    for (int y = 0; y < img->inner_height; ++y)
        memcpy(img->image[y + img->padding], flat + (size_t)y * img->width,
               sizeof(double) * (size_t)img->width);
}

static void rows_out(ImageWithPadding *img, double *flat) {
    for (int y = 0; y < img->inner_height; ++y)
        memcpy(flat + (size_t)y * img->width, img->image[y + img->padding],
               sizeof(double) * (size_t)img->width);
}

/* same contract as hgb_conv_run / oracle_conv_run: cur <- state n, prev <- state n-1 */
int ref_conv_run_flat(double *cur, double *prev, int w, int h, int p, const double *k, int kw,
                      int kh, int iterations, int procs) {
    ImageWithPadding *img = init_padded_image(w, h, p);
    ImageWithPadding *buf = init_padded_image(w, h, p);
    Image *kernel = init_image(kw, kh, 0.0);
    if (!img || !buf || !kernel) return -1;
    for (int y = 0; y < kh; ++y)
        for (int x = 0; x < kw; ++x) kernel->image[y][x] = k[y * kw + x];
    rows_in(img, cur);
    rows_in(buf, prev);
    Args args;
    memset(&args, 0, sizeof(args));
    args.number_of_iterations = iterations;
    args.number_of_processes = procs < 1 ? 1 : procs;
    args.width = w;
    args.height = h;
    run_on_padded_image(&img, kernel, &args, &buf);
    rows_out(img, cur);
    rows_out(buf, prev);
    /* free_padded_image under-frees (convolution-util.c:344-349); free rows here */
    for (int y = 0; y < h; ++y) { free(img->image[y + p]); free(buf->image[y + p]); }
    free(img->image); free(img); free(buf->image); free(buf);
    free_image(kernel);
    return 0;
}

/* reference add_padding + remove_padding + get_checksum on flat data */
int ref_add_padding_flat(const double *dense, double *flat, int w, int h, int p) {
    Image *im = init_image(w, h, 0.0);
    for (int y = 0; y < h; ++y) memcpy(im->image[y], dense + (size_t)y * w, sizeof(double) * (size_t)w);
    ImageWithPadding *pi = add_padding(im, p);
    /* also prove the aliasing: pad rows must BE the edge rows */
    int aliased = 1;
    for (int y = 0; y < p; ++y)
        if (pi->image[y] != pi->image[p] || pi->image[h + p + y] != pi->image[h + p - 1]) aliased = 0;
    rows_out(pi, flat);
    for (int y = 0; y < h; ++y) free(pi->image[y + p]);
    free(pi->image); free(pi);
    free_image(im);
    return aliased ? 0 : -3;
}

double ref_checksum_flat(const double *flat, int w, int h, int p) {
    ImageWithPadding *pi = init_padded_image(w, h, p);
    rows_in(pi, flat);
    Image *im = remove_padding(pi);
    double c = get_checksum(im);
    free_image(im);
    for (int y = 0; y < h; ++y) free(pi->image[y + p]);
    free(pi->image); free(pi);
    return c;
}

/* reference text reader (height-then-width quirk, convolution-util.c:9-45) */
int ref_read_image(const char *path, double *dense_out, int cap, int *w, int *h) {
    Args args;
    memset(&args, 0, sizeof(args));
    FILE *fd = fopen(path, "r");
    if (!fd) return -1;
    Image *im = read_image_from_fd(fd, &args);
    fclose(fd);
    if (!im) return -2;
    *w = im->width; *h = im->height;
    if ((long)im->width * im->height > cap) { free_image(im); return -4; }
    for (int y = 0; y < im->height; ++y)
        memcpy(dense_out + (size_t)y * im->width, im->image[y], sizeof(double) * (size_t)im->width);
    free_image(im);
    return 0;
}

int ref_default_kernel(double *k25) {
    Image *k = get_default_kernel();
    for (int y = 0; y < 5; ++y) for (int x = 0; x < 5; ++x) k25[y * 5 + x] = k->image[y][x];
    free_image(k);
    return 0;
}
