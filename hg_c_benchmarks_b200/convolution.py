"""Host-side mirror of the reference's convolution interface
(src/convolution/convolution-run.h, src/convolution/convolution-util.h).

Same names, argument meaning and error behaviour as the reference so parity tests
read like the reference's own gtest cases.  Layout helpers (padding, borders,
checksum) are host code as in the reference; the stencil itself only ever runs on
the GPU through libhgb200.so.
"""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import check, dptr, lib


class Image:
    """struct Image {int width; int height; double **image;}  (convolution-util.h:33-37)"""

    def __init__(self, width, height, default_value=0.0):
        self.width, self.height = int(width), int(height)
        self.image = np.full((self.height, self.width), float(default_value), dtype=np.float64)


class ImageWithPadding:
    """struct ImageWithPadding (convolution-util.h:39-46).  `rows` are the allocated rows
    image[padding .. padding+inner_height-1]; the pad rows above/below ALIAS the first/last
    real row exactly as init_padded_image builds them (convolution-util.c:394-400)."""

    def __init__(self, inner_width, inner_height, padding):
        self.padding = int(padding)
        self.inner_width, self.inner_height = int(inner_width), int(inner_height)
        self.width = self.inner_width + 2 * self.padding
        self.height = self.inner_height + 2 * self.padding
        self.rows = np.zeros((self.inner_height, self.width), dtype=np.float64)

    def _row(self, y):   # image[y] with the aliasing
        return min(max(y - self.padding, 0), self.inner_height - 1)

    def field(self, x, y):                 # ACCESS_FIELD (convolution-util.h:17)
        return self.rows[self._row(y), x]

    def set_field(self, x, y, v):
        self.rows[self._row(y), x] = v

    def access(self, x, y):                # ACCESS_IMAGE (convolution-util.h:16)
        return self.field(x + self.padding, y + self.padding)

    def set_access(self, x, y, v):
        self.set_field(x + self.padding, y + self.padding, v)

    def as_matrix(self):
        """the full (height, width) view a reader of image[y][x] sees"""
        idx = [self._row(y) for y in range(self.height)]
        return self.rows[idx, :]


class Args:
    """struct arguments (convolution-util.h:19-31); only the fields the kernels read."""

    def __init__(self, number_of_iterations=4000, number_of_processes=1, width=1024, height=1024):
        self.number_of_iterations = int(number_of_iterations)
        self.number_of_processes = int(number_of_processes)
        self.width, self.height = int(width), int(height)
        self.debug = False


def init_image(width, height, default_value):
    return Image(width, height, default_value)


def init_padded_image(inner_width, inner_height, padding):
    return ImageWithPadding(inner_width, inner_height, padding)


def update_borders(padded_img):
    """left/right pad columns <- edge pixels (convolution-util.c:281-294)"""
    p, w = padded_img.padding, padded_img.inner_width
    if p:
        padded_img.rows[:, :p] = padded_img.rows[:, p:p + 1]
        padded_img.rows[:, p + w:] = padded_img.rows[:, p + w - 1:p + w]


def add_padding(img, padding):
    out = ImageWithPadding(img.width, img.height, padding)
    out.rows[:, padding:padding + img.width] = img.image
    update_borders(out)
    return out


def remove_padding(padded_img):
    p = padded_img.padding
    out = Image(padded_img.inner_width, padded_img.inner_height, 0.0)
    out.image[...] = padded_img.rows[:, p:p + padded_img.inner_width]
    return out


def copy_padded_image(padded_from, padded_to):
    """0 on success, -1 if an argument is None, -2 on extent mismatch (convolution-util.c:357-373)"""
    if padded_from is None or padded_to is None:
        return -1
    if padded_from.width != padded_to.width or padded_from.height != padded_to.height:
        return -2
    padded_to.rows[...] = padded_from.rows
    return 0


def get_checksum(img):
    """sequential row-major sum (convolution-util.c:63-71)"""
    val = 0.0
    for v in img.image.ravel().tolist():
        val += v
    return val


def _read(fn, path):
    w, h, data = C.c_int(), C.c_int(), C.POINTER(C.c_double)()
    check(fn(str(path).encode(), C.byref(w), C.byref(h), C.byref(data)))
    try:
        img = Image(w.value, h.value)
        img.image[...] = np.ctypeslib.as_array(data, shape=(h.value, w.value))
    finally:
        lib().hgb_image_free(data)
    return img


def read_image(path):
    """read_image_from_fd (convolution-util.c:9-45): "<height> <width>" then one line per row.
    Same token rules as the reference, no 16384-byte line limit; errors raise instead of exit(2)."""
    return _read(lib().hgb_image_read_text, path)


def write_image(path, img):
    check(lib().hgb_image_write_text(str(path).encode(), dptr(np.ascontiguousarray(img.image)), img.width,
                                     img.height))


def read_image_binary(path):
    """HGBIMG1: magic, int32 height, int32 width, raw little-endian doubles (for 16k x 16k images)"""
    return _read(lib().hgb_image_read_binary, path)


def write_image_binary(path, img):
    check(lib().hgb_image_write_binary(str(path).encode(), dptr(np.ascontiguousarray(img.image)), img.width,
                                       img.height))


def get_default_kernel():
    """convolution-util.c:135-151"""
    k = Image(5, 5, 0.0)
    k.image[...] = np.where((np.add.outer(np.arange(5), np.arange(5)) % 2) == 0, 2.0, -2.0)
    k.image[2, 2] = -1.0
    return k


def run_on_padded_image(padded_img, kernel, args, buffer):
    """void run_on_padded_image(ImageWithPadding **padded_img, Image *kernel, Args *args,
    ImageWithPadding **buffer)  -- reference src/convolution/convolution-run.c:41-49.

    `padded_img` and `buffer` are one-element lists standing in for the `**`: after the
    call padded_img[0] is the newest state and buffer[0] the one before it, as with the
    reference's pointer swap.  Runs on the GPU; args.number_of_processes = GPU count.
    """
    img, buf = padded_img[0], buffer[0]
    n = args.number_of_iterations
    if n <= 0:
        return
    kern = np.ascontiguousarray(kernel.image, dtype=np.float64)
    check(lib().hgb_conv_run(dptr(img.rows), dptr(buf.rows), img.inner_width, img.inner_height,
                             img.padding, dptr(kern), kernel.width, kernel.height, n,
                             args.number_of_processes))
    # hgb_conv_run leaves state n in img.rows and n-1 in buf.rows; the reference's swap leaves
    # the newest state reachable through *padded_img, which is what the caller reads next
    padded_img[0], buffer[0] = img, buf


def run_flat(cur, prev, inner_w, inner_h, padding, kernel, iterations, ngpus=1):
    """hgb_conv_run on flat host arrays: `cur` (inner_h rows of inner_w + 2*padding doubles) holds the
    input and receives state n; `prev` receives state n-1, or None when it is not wanted.  One GPU and a
    tall image: streamed in row chunks, PCIe copies under the iterations (set_streaming)."""
    kern = np.ascontiguousarray(kernel, dtype=np.float64)
    check(lib().hgb_conv_run(dptr(cur), dptr(prev) if prev is not None else None, int(inner_w), int(inner_h),
                             int(padding), dptr(kern), kern.shape[1], kern.shape[0], int(iterations), int(ngpus)))


def set_streaming(chunk_rows=1):
    """0: hgb_conv_run never streams; 1: automatic chunking (default); else valid rows per chunk"""
    check(lib().hgb_set_conv_streaming(int(chunk_rows)))


def unpin():
    """no longer needed: hgb_conv_run releases its page-lock registrations before it returns"""
    check(lib().hgb_conv_unpin())


def apply_kernel_to_padded_image(padded_img, kernel, args, buffer):
    """one stencil pass into the INNER region of `buffer`, pad columns untouched
    (reference src/convolution/convolution-run.c:52-62)."""
    tmp_cur = padded_img.rows.copy()
    tmp_prev = np.empty_like(tmp_cur)
    kern = np.ascontiguousarray(kernel.image, dtype=np.float64)
    check(lib().hgb_conv_run(dptr(tmp_cur), dptr(tmp_prev), padded_img.inner_width,
                             padded_img.inner_height, padded_img.padding, dptr(kern), kernel.width,
                             kernel.height, 1, args.number_of_processes))
    p, w = padded_img.padding, padded_img.inner_width
    buffer.rows[:, p:p + w] = tmp_cur[:, p:p + w]


class Conv2D:
    """Device-resident image.  ngpus=None -> one process per GPU (this rank owns one row band)."""

    def __init__(self, inner_w, inner_h, padding, ngpus=1):
        self.w, self.h, self.p = int(inner_w), int(inner_h), int(padding)
        self._h = C.c_void_p()
        self.rank_mode = ngpus is None
        if self.rank_mode:
            check(lib().hgb_conv_create_rank(C.byref(self._h), self.w, self.h, self.p))
        else:
            check(lib().hgb_conv_create(C.byref(self._h), self.w, self.h, self.p, int(ngpus)))

    def band(self):
        r0, rows = C.c_int(), C.c_int()
        check(lib().hgb_conv_band(self._h, C.byref(r0), C.byref(rows)))
        return r0.value, rows.value

    def set_kernel(self, kernel):
        kernel = np.ascontiguousarray(kernel, dtype=np.float64)
        kh, kw = kernel.shape
        check(lib().hgb_conv_set_kernel(self._h, dptr(kernel), kw, kh))
        return self

    def set_mode(self, mode):
        """CONV_EXACT (bit-identical, default) or CONV_FAST (fused multiply-add taps; floating-point images
        only -- an integer-valued image keeps the exact kernels); choose it before upload()"""
        check(lib().hgb_conv_set_mode(self._h, int(mode)))
        return self

    def mode_in_effect(self):
        return lib().hgb_conv_mode_in_effect(self._h)

    def tune(self, variant=0, band_rows=0):
        check(lib().hgb_conv_tune(self._h, int(variant), int(band_rows)))
        return self

    def set_depth(self, depth=0):
        """iterations per halo exchange (multi-GPU); returns the depth in effect"""
        rc = lib().hgb_conv_set_depth(self._h, int(depth))
        if rc < 0:
            check(rc)
        return rc

    def set_exchange(self, kind):
        check(lib().hgb_conv_set_exchange(self._h, int(kind)))
        return self

    def upload(self, rows):
        """rows: (band rows, w+2p) float64 incl. pad columns (whole image unless rank mode)"""
        assert rows.shape == (self.band()[1], self.w + 2 * self.p), rows.shape
        check(lib().hgb_conv_upload(self._h, dptr(rows)))
        return self

    def step(self, iterations=1):
        ms = C.c_float(0.0)
        check(lib().hgb_conv_step(self._h, int(iterations), C.byref(ms)))
        return ms.value

    def download(self, want_prev=False, out=None):
        shape = (self.band()[1], self.w + 2 * self.p)
        cur = np.empty(shape, dtype=np.float64) if out is None else out
        assert cur.shape == shape
        prev = np.empty(shape, dtype=np.float64) if want_prev else None
        check(lib().hgb_conv_download(self._h, dptr(cur), dptr(prev)))
        return (cur, prev) if want_prev else cur

    # ---- device-resident lifecycle (reference src/2d-convolution.c:41-74 without host round trips) ----
    def upload_dense(self, dense):
        """add_padding on the device: dense is the unpadded (band rows, w) image"""
        assert dense.shape == (self.band()[1], self.w), dense.shape
        check(lib().hgb_conv_upload_dense(self._h, dptr(dense)))
        return self

    def download_dense(self, out=None):
        """remove_padding of the newest state"""
        shape = (self.band()[1], self.w)
        dense = np.empty(shape, dtype=np.float64) if out is None else out
        assert dense.shape == shape
        check(lib().hgb_conv_download_dense(self._h, dptr(dense)))
        return dense

    def backup(self):
        """copy_padded_image(img, backup), device to device"""
        check(lib().hgb_conv_backup(self._h))
        return self

    def restore(self):
        """copy_padded_image(backup, img), device to device"""
        check(lib().hgb_conv_restore(self._h))
        return self

    def checksum(self):
        c = C.c_double(0.0)
        check(lib().hgb_conv_checksum(self._h, C.byref(c)))
        return c.value

    def close(self):
        if self._h:
            lib().hgb_conv_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
