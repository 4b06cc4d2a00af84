// hgb_io.cpp -- image files for the convolution path.  Host code only.
//
// Text format = the reference's (src/convolution/convolution-util.c:9-45, files 2d-convolution.input,
// kernel.txt, laplace.txt): first line "<height> <width>" -- HEIGHT FIRST, as the reference code
// reads it (its header comment says the opposite) -- then one line per row, values separated by
// single spaces, parsed with strtol/strtod.  The reference reads lines through a 16384-byte buffer,
// so rows wider than ~16k characters (any 16384-pixel image) cannot come from a file there; here
// lines have no length limit.  Everything else is kept: tokens are split on ' ' only (a trailing
// "\n" token parses as 0.0, exactly like strtok+strtod do), surplus tokens and lines are ignored,
// missing ones are errors (returned, where the reference calls bail_out -> exit(2)).
//
// Binary format (new; 16k x 16k text would be ~5 GB): "HGBIMG1\n", int32 height, int32 width
// (same order as the text header), then height*width little-endian binary64 values, row-major.
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/hgb200.h"

namespace hgb {
void set_error(const char *fmt, ...);
}
using hgb::set_error;

namespace {
// one line of any length, newline kept (like fgets); false at end of file with nothing read
bool read_line(FILE *f, std::string &line) {
    line.clear();
    char chunk[65536];
    while (fgets(chunk, sizeof chunk, f)) {
        line += chunk;
        if (!line.empty() && line.back() == '\n') return true;
    }
    return !line.empty();
}

// strtok(buf, " ") semantics: skip runs of spaces, token ends at the next space or the terminator
const char *next_token(const char *&cursor) {
    while (*cursor == ' ') ++cursor;
    if (*cursor == '\0') return nullptr;
    const char *tok = cursor;
    while (*cursor != ' ' && *cursor != '\0') ++cursor;
    return tok;
}
}  // namespace

extern "C" {

void hgb_image_free(double *data) { free(data); }

int hgb_image_read_text(const char *path, int *width, int *height, double **data) {
    if (!path || !width || !height || !data) { set_error("hgb_image_read_text: NULL argument"); return HGB_ERR_ARG; }
    FILE *f = fopen(path, "r");
    if (!f) { set_error("cannot open %s: %s", path, strerror(errno)); return HGB_ERR_ARG; }
    std::string line;
    if (!read_line(f, line)) { fclose(f); set_error("%s is empty", path); return HGB_ERR_ARG; }
    const char *cur = line.c_str();
    const char *tok = next_token(cur);
    if (!tok) { fclose(f); set_error("height couldn't be read from file"); return HGB_ERR_ARG; }
    const long h = strtol(tok, nullptr, 10);
    tok = next_token(cur);
    if (!tok) { fclose(f); set_error("width couldn't be read from file"); return HGB_ERR_ARG; }
    const long w = strtol(tok, nullptr, 10);
    if (h <= 0 || w <= 0 || h > (1 << 20) || w > (1 << 20)) {
        fclose(f);
        set_error("image extents %ld x %ld out of range", h, w);
        return HGB_ERR_ARG;
    }
    double *img = static_cast<double *>(calloc((size_t)h * (size_t)w, sizeof(double)));
    if (!img) { fclose(f); set_error("out of host memory for a %ld x %ld image", h, w); return HGB_ERR_ARG; }
    for (long y = 0; y < h; ++y) {
        if (!read_line(f, line)) {
            fclose(f); free(img);
            set_error("the file did not have enough lines to read");
            return HGB_ERR_ARG;
        }
        cur = line.c_str();
        for (long x = 0; x < w; ++x) {
            tok = next_token(cur);
            if (!tok) {
                fclose(f); free(img);
                set_error("the file dimensions were not correct, less number of lines than given width");
                return HGB_ERR_ARG;
            }
            img[(size_t)y * w + x] = strtod(tok, nullptr);
        }
    }
    fclose(f);
    *width = (int)w;
    *height = (int)h;
    *data = img;
    return HGB_OK;
}

int hgb_image_write_text(const char *path, const double *data, int width, int height) {
    if (!path || !data || width <= 0 || height <= 0) { set_error("hgb_image_write_text: bad argument"); return HGB_ERR_ARG; }
    FILE *f = fopen(path, "w");
    if (!f) { set_error("cannot open %s: %s", path, strerror(errno)); return HGB_ERR_ARG; }
    fprintf(f, "%d %d\n", height, width);
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x)
            fprintf(f, x ? " %.17g" : "%.17g", data[(size_t)y * width + x]);   // round-trips every double
        fputc('\n', f);
    }
    const bool ok = fclose(f) == 0;
    if (!ok) { set_error("write to %s failed", path); return HGB_ERR_ARG; }
    return HGB_OK;
}

static const char kMagic[8] = {'H', 'G', 'B', 'I', 'M', 'G', '1', '\n'};

int hgb_image_write_binary(const char *path, const double *data, int width, int height) {
    if (!path || !data || width <= 0 || height <= 0) { set_error("hgb_image_write_binary: bad argument"); return HGB_ERR_ARG; }
    FILE *f = fopen(path, "wb");
    if (!f) { set_error("cannot open %s: %s", path, strerror(errno)); return HGB_ERR_ARG; }
    const int32_t hw[2] = {height, width};
    const size_t count = (size_t)width * (size_t)height;
    bool ok = fwrite(kMagic, 1, 8, f) == 8 && fwrite(hw, 4, 2, f) == 2 && fwrite(data, 8, count, f) == count;
    ok = (fclose(f) == 0) && ok;
    if (!ok) { set_error("write to %s failed", path); return HGB_ERR_ARG; }
    return HGB_OK;
}

int hgb_image_read_binary(const char *path, int *width, int *height, double **data) {
    if (!path || !width || !height || !data) { set_error("hgb_image_read_binary: NULL argument"); return HGB_ERR_ARG; }
    FILE *f = fopen(path, "rb");
    if (!f) { set_error("cannot open %s: %s", path, strerror(errno)); return HGB_ERR_ARG; }
    char magic[8];
    int32_t hw[2];
    if (fread(magic, 1, 8, f) != 8 || memcmp(magic, kMagic, 8) != 0 || fread(hw, 4, 2, f) != 2 || hw[0] <= 0 ||
        hw[1] <= 0) {
        fclose(f);
        set_error("%s is not an HGBIMG1 file", path);
        return HGB_ERR_ARG;
    }
    // same extent clamp as the text reader; and the payload must really be in the file before
    // anything is allocated for it (a crafted header must not size the buffer)
    if (hw[0] > (1 << 20) || hw[1] > (1 << 20)) {
        fclose(f);
        set_error("%s: image extent %d x %d out of range", path, hw[0], hw[1]);
        return HGB_ERR_ARG;
    }
    const size_t count = (size_t)hw[0] * (size_t)hw[1];
    {
        const long here = ftell(f);
        long end = -1;
        if (here >= 0 && fseek(f, 0, SEEK_END) == 0) end = ftell(f);
        if (here < 0 || end < 0 || fseek(f, here, SEEK_SET) != 0 ||
            (unsigned long long)(end - here) < (unsigned long long)count * sizeof(double)) {
            fclose(f);
            set_error("%s is truncated", path);
            return HGB_ERR_ARG;
        }
    }
    double *img = static_cast<double *>(malloc(count * sizeof(double)));
    if (!img) { fclose(f); set_error("out of host memory"); return HGB_ERR_ARG; }
    if (fread(img, 8, count, f) != count) {
        fclose(f); free(img);
        set_error("%s is truncated", path);
        return HGB_ERR_ARG;
    }
    fclose(f);
    *height = hw[0];
    *width = hw[1];
    *data = img;
    return HGB_OK;
}

}  // extern "C"
