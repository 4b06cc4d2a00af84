/*
 * hgb200.h -- C ABI of libhgb200.so: the B200 (sm_100a) implementation of the
 * hg-c-benchmarks hot path (all-pairs FP64 nbody step, 2D convolution stencil).
 *
 * Plain C, plain pointers and sizes.  Every function returns HGB_OK (0) or a
 * negative HGB_ERR_* code and never exits, throws or falls back to the CPU;
 * hgb_last_error() describes the last failure on the calling thread.  The
 * library owns device memory, streams, pinned staging and NCCL communicators
 * only; callers allocate and free every host buffer (as the reference's
 * callers do: src/nbody.c:18-19,45, src/2d-convolution.c:41-43,92).
 *
 * Not re-entrant: one host thread drives the library (the reference's run
 * functions are likewise called from a single thread and fan out inside).
 *
 * "reference" below = fendor/hg-c-benchmarks; file:line are relative to it.
 */
#ifndef HGB200_H
#define HGB200_H

#ifdef __cplusplus
extern "C" {
#endif

#define HGB_OK 0
#define HGB_ERR_ARG (-1)    /* bad argument (null pointer, negative size, kernel larger than padding) */
#define HGB_ERR_CUDA (-2)   /* a CUDA runtime call or kernel failed */
#define HGB_ERR_NCCL (-3)   /* NCCL missing or a collective failed */
#define HGB_ERR_STATE (-4)  /* call order (e.g. step before upload, comm not initialised) */
#define HGB_ERR_NOGPU (-5)  /* no CUDA device: there is NO CPU fallback */

/* nbody arithmetic modes */
#define HGB_NBODY_EXACT 0 /* bit-identical to the reference: unfused IEEE ops, j = 0..n-1 in order */
#define HGB_NBODY_FAST 1  /* FMA-contracted, rsqrt-Newton; ~1e-15 relative per step, not bit-identical */
/* position exchange between target shards (multi-GPU) */
#define HGB_EXCHANGE_NCCL 0 /* ncclAllGather after the step kernel */
#define HGB_EXCHANGE_P2P 1  /* step kernel stores each result into every peer's buffer over NVLink */

/* ---- library ---------------------------------------------------------- */
const char *hgb_version(void);
const char *hgb_last_error(void);
int hgb_device_count(void);                  /* >= 0, or HGB_ERR_NOGPU */
int hgb_init(void);                          /* optional; everything initialises lazily */
void hgb_finalize(void);                     /* frees streams, staging, communicators */
unsigned long long hgb_launch_count(void);   /* kernels launched by this library so far */
int hgb_flush_l2(void);                      /* overwrite a buffer larger than L2 on every local device */
/* measured roofline denominators (bench.py).  The FP64 probe is a dependent-free DFMA loop (8 chains
 * per thread) sustained for ~1 s after a warm-up, so that it reads the settled SM clock;
 * *sm_mhz_est = the clock that rate implies at 64 DFMA lanes per SM per cycle. */
int hgb_measure_fp64_peak(double *tflops, double *sm_mhz_est);
int hgb_measure_hbm_copy(double *gbs);                         /* device copy, read+write bytes */

/* ---- one process per GPU (torchrun): communicator over NCCL ------------ */
/* Rank 0 calls hgb_comm_unique_id, the launcher broadcasts the bytes, every rank
 * calls hgb_comm_init_rank after cudaSetDevice.  world == 1 needs no NCCL. */
#define HGB_UNIQUE_ID_BYTES 128
int hgb_comm_unique_id(void *id_out, int capacity);
int hgb_comm_init_rank(const void *id, int id_bytes, int rank, int world);
int hgb_comm_rank(void);
int hgb_comm_world(void);
int hgb_comm_barrier(void);
void hgb_comm_destroy(void);

/* ---- nbody: drop-in --------------------------------------------------- */
/*
 * Replaces  void run(Float3D *planets, Float3D *buffer, int number_of_planets,
 *                    int iterations, int number_of_processes)
 * (src/nbody/nbody-run.h:46, src/nbody/nbody-run.c:27-37).  planets_aos and
 * buffer_aos are the caller's Float3D arrays (AoS {x,y,z} doubles,
 * src/nbody/nbody-util.h:19-23).  Identical semantics, including the
 * reference's LOCAL pointer swap: on return state k is in planets when k is
 * even and in buffer when k is odd (k = iterations, iterations-1); with
 * iterations == 0 or n == 0 nothing is touched.  `ngpus` takes the place of
 * number_of_processes: GPUs to shard targets over (clamped to the device
 * count; env HGB_NGPUS overrides).  Uses the mode set by hgb_set_nbody_mode
 * (default HGB_NBODY_EXACT).
 */
int hgb_nbody_run(double *planets_aos, double *buffer_aos, int n, int iterations, int ngpus);
int hgb_set_nbody_mode(int mode);
int hgb_set_exchange(int kind);

/* ---- nbody: device-resident (kernel-only timing, many steps) ---------- */
typedef struct hgb_nbody hgb_nbody;
/* single process driving `ngpus` local devices (targets sharded evenly) */
int hgb_nbody_create(hgb_nbody **out, int n, int ngpus);
/* one process per GPU: this rank owns target shard hgb_comm_rank() of hgb_comm_world() */
int hgb_nbody_create_rank(hgb_nbody **out, int n);
int hgb_nbody_config(hgb_nbody *h, int mode, int exchange);
/* kernel variant; 0 = default (128-thread CTAs, 6 per SM, unroll 2): threads per CTA {64,128,256}, sources per
 * unrolled group/2 {1,2,4}, resident warps per SM the register allocation allows {24,28,32,40}
 * (28: 64/128-thread CTAs; 64-thread CTAs: (unroll, warps) in {(2,28), (2,24), (4,24), (1,28)}) */
int hgb_nbody_tune(hgb_nbody *h, int block, int unroll, int warps);
/* Pair-symmetric EXACT step (csrc/nbody_sym.cuh): every unordered pair is evaluated once and added to
 * both bodies, each target still summing its sources in the order of accel() (src/nbody/nbody-run.c:
 * 17-25), so the result is bit-identical to hgb_nbody_tune's kernel at 16 instead of 29 FP64
 * instructions per interaction.  variant: 0 = off, 1 = default, else R*100 + U*10 + CTAs per SM;
 * used on single-GPU handles in EXACT mode when n >= min_n (the wavefront needs a large n to fill
 * the GPU).  New handles pick the kernel by bodies per GPU (standard / 3002 / 2002 / 6242; env
 * HGB_NBODY_SYM = 0 | 1 | variant and HGB_NBODY_SYM_MIN_N override); variant 1 here asks for the first generation
 * from min_n up on one GPU; hgb_nbody_tune with a non-zero argument switches the pair-symmetric kernels off.
 * Variants 3000 + C (3002 default; 3010 + C with span counters): csrc/nbody_crow.cuh, a CTA of four warps per row.
 * Variants 4242 / 6242 (5242 / 7242 with span counters): csrc/nbody_seg.cuh, the first generation's warp on the segmented
 * schedule, plain / with software-pipelined groups (the one-GPU default from n = 655 360; on request also dealt over
 * several GPUs in blocks of 64 targets).  The thresholds live in auto_variant() (csrc/hgb_nbody.cu). */
int hgb_nbody_set_symmetric(hgb_nbody *h, int variant, int min_n);
/* Variants 2000 + C (C = 1..3 CTAs per SM; 2 = 2002; 2010 + C: with clock64 span counters) select the
 * second-generation pair-symmetric kernel (csrc/nbody_tri.cuh): forces evaluated per 32 x 8 minitile, the
 * ordered accumulation decoupled from them (row pass in registers, column pass through a shared-memory
 * transpose), rows cut into segments.  hgb_nbody_set_segment tunes its schedule: `seg` sources per item
 * (multiple of 32), `rows_per_tile` rows whose items are handed out together, `order_weight` of the segment
 * index in the ticket order.  0 = the measured default of the kernel in effect, from tri_params() in
 * csrc/hgb_nbody.cu: seg 16384 (8192 for cooperative rows on one GPU); rows per tile, one GPU / several: 8 / 16
 * (nbody_tri), 1 / 8 (nbody_crow), 4 / 4 (nbody_seg); weight 32 / 64, 128 / 32, 64 / 64.  It also runs on single-process multi-GPU
 * handles with HGB_EXCHANGE_P2P: tile-rows of targets are dealt over the GPUs, the chain words cross NVLink
 * as tagged stores, the last row broadcasts the finished sums (no position exchange). */
int hgb_nbody_set_segment(hgb_nbody *h, int seg, int rows_per_tile, int order_weight);
/* Rows of targets per turn of a GPU when the pair-symmetric wavefront is dealt over several GPUs (rounded to a
 * multiple of rows_per_tile; 0 = default: 16 for nbody_tri, 8 for nbody_crow / nbody_seg): rows I with I / rows_per_turn = t (mod GPUs) run on GPU t, so the
 * chain words (reference src/nbody/nbody-run.c:19-21: the running sums of accel()) cross NVLink once per turn. */
int hgb_nbody_set_deal(hgb_nbody *h, int rows_per_turn);
/* the variant hgb_nbody_step will use on this handle as configured now (0: the standard kernel) */
int hgb_nbody_symmetric(const hgb_nbody *h);
/* "<kernel> | <how targets/pairs are spread over the GPUs>" of the next hgb_nbody_step on this handle
 * as configured now (bench.py's config.kernel / config.parallelism) */
int hgb_nbody_describe(const hgb_nbody *h, char *buf, int capacity);
/* FP64-pipe instructions that kernel issues per interaction (SASS count, tools/sass_loop.py):
 * the numerator of the pipe-utilisation figure bench.py reports */
double hgb_nbody_fp64_instr(const hgb_nbody *h);
/* diagnostics: clock64 spans summed over all warps of the last step run with a profiling variant
 * (1226, 1242): tag wait, chunk loads, chunk release, predicate-free loop, predicated steps, whole
 * workers, chunks opened */
int hgb_nbody_sym_counters(hgb_nbody *h, unsigned long long *out7);
int hgb_nbody_upload(hgb_nbody *h, const double *pos_aos);       /* all n bodies */
/* `iterations` steps; *ms (may be NULL) = device time, CUDA events, max over local devices */
int hgb_nbody_step(hgb_nbody *h, int iterations, float *ms);
int hgb_nbody_download(hgb_nbody *h, double *cur_aos, double *prev_aos); /* either may be NULL */
/* accel() of `count` chosen targets against the current state (subsample parity at n = 2^20) */
int hgb_nbody_accel_subset(hgb_nbody *h, const int *targets, int count, double *out_aos);
/* extension, no reference counterpart: U = -G sum_{i<j} (|d|^2+EPS)^(-1/2) */
int hgb_nbody_energy(hgb_nbody *h, double *energy);
/* The same potential energy for callers that may not allocate device memory.  No counterpart in the
 * reference (it has no energy at all, and nothing of it runs on a GPU).
 *
 * Every other entry point needs a handle, and making one calls cudaMalloc.  That is not open to an
 * in-situ hook in a host code that gave all of the GPU's memory to its own pool at start-up (the
 * cudaMalloc would fail), to a monitoring call made while that code's step loop is in flight
 * (cudaMalloc / cudaFree synchronise the whole device and stall it), or to a plugin that may only
 * touch device pointers it was handed.  This call allocates and frees nothing, on the device or in
 * pinned host memory, and copies nothing:
 *   pos_aos_dev  n bodies as Float3D AoS (x, y, z doubles; src/nbody/nbody-util.h), in DEVICE memory
 *   energy_dev   one double in DEVICE memory, written with U
 *   stream       a cudaStream_t (NULL = the default stream) of the CURRENT device; both kernels are
 *                queued on it and the call returns without waiting -- read *energy_dev after that
 *                stream has been synchronised.  The pointers must be usable from the current device.
 * The per-block partial sums go to a fixed 64 KiB array built into the library (each GPU that loads
 * the library carries it until the process exits), which is why n is limited:
 * n > HGB_STATIC_ENERGY_MAX_BODIES is refused with HGB_ERR_ARG, nothing launched.
 * - The first call on each device looks that array's address up (cudaGetSymbolAddress, which may
 *   load the module); make that first call before starting any stream capture.
 * - All calls on a device share the array.  It is never cleared and need not be (each call writes
 *   every slot it later reads), but two calls that overlap on the GPU -- different streams, or
 *   different host threads -- overwrite each other's partials and both results are garbage.
 *   Order such calls on one stream or with events.
 * On one GPU the value is bit-identical to hgb_nbody_energy of the same positions. */
#define HGB_STATIC_ENERGY_MAX_BODIES 1048576
int hgb_nbody_energy_static(const double *pos_aos_dev, int n, double *energy_dev, void *stream);
/* ---- nbody: integrator extension --------------------
 * The reference moves positions only (buffer[i] = G*acc, src/nbody/nbody-run.c:22-24) and has no
 * velocity, mass, dt or energy.  These entry points add them as a separate, clearly non-reference
 * mode whose CPU checker is oracle_nbody_integrate (oracle/hg_oracle.c); hgb_nbody_step is unaffected.
 * One step, per target i, sources j = 0..n-1 in order, every operation one IEEE rounding (no FMA):
 *   a_i = G * sum_j d_ij * (m_j * f_ij)     d, s, f exactly as pair_wise_accel (:6-15)
 *   v_i = v_i + dt * a_i ;  x_i = x_i + dt * v_i        (symplectic Euler)
 * vel_aos: Float3D-like AoS [3n] (NULL = at rest); mass: [n] (NULL = all 1.0). */
int hgb_nbody_set_dynamics(hgb_nbody *h, const double *vel_aos, const double *mass, double dt);
int hgb_nbody_integrate(hgb_nbody *h, int steps, float *ms);
int hgb_nbody_download_velocities(hgb_nbody *h, double *vel_aos);
/* K = 1/2 sum m_i |v_i|^2,  U = -G sum_{i<j} m_i m_j (|d|^2+EPS)^(-1/2); warp-shuffle reductions */
int hgb_nbody_total_energy(hgb_nbody *h, double *kinetic, double *potential);
int hgb_nbody_destroy(hgb_nbody *h);
/* pair_wise_accel(p1, p2[j], &acc) for j = 0..count-1 in order (src/nbody/nbody-run.c:6-15):
 * acc_inout[3] += d * factor, no G scaling.  Symbol-level drop-in, one GPU thread. */
int hgb_nbody_pairwise(const double *p1, const double *p2_aos, int count, double *acc_inout);
/* diagnostic: the kernel's branch-free RN(1/RN(sqrt t)) against the IEEE intrinsics on `count`
 * random + hard-case operands; *mismatches must come back 0 */
int hgb_selftest_inv_sqrt(unsigned long long count, unsigned long long seed,
                          unsigned long long *mismatches, double *first_bad);

/* convolution arithmetic modes */
#define HGB_CONV_EXACT 0 /* bit-identical to the reference: one multiply and one add per tap, row-major tap order */
#define HGB_CONV_FAST 1  /* one fused multiply-add per tap (same order): floating-point images only, <= 1e-12
                          * relative to the reference after 10 iterations, not bit-identical; an image whose
                          * inner pixels are all integer-valued keeps the EXACT kernels (classified at upload; pad columns are not looked at) */

/* ---- convolution: drop-in --------------------------------------------- */
/*
 * Replaces  void run_on_padded_image(ImageWithPadding **padded_img, Image *kernel,
 *                                    Args *args, ImageWithPadding **buffer)
 * (src/convolution/convolution-run.h:70-71, src/convolution/convolution-run.c:41-49),
 * i.e. `iterations` x { apply_kernel_to_padded_image (:52-62); update_borders
 * (src/convolution/convolution-util.c:281-294); swap }.
 * cur / prev: inner_h rows of (inner_w + 2*padding) doubles -- the reference's rows
 * image[padding .. padding+inner_h-1] INCLUDING their pad columns (read as they
 * are on the first iteration); the aliased top/bottom pad rows are not passed
 * (convolution-util.c:394-400 makes them the edge rows).  kernel: kh rows of kw
 * doubles, correlation, needs kw <= 2*padding+1 and kh <= 2*padding+1.
 * On return cur = state `iterations`, prev = state iterations-1 (state 0 is the
 * input `cur`; the input contents of `prev` never matter) -- what the caller of
 * the reference sees through its swapped pointers.  iterations == 0 touches
 * nothing.  Bit-identical to the reference for finite data.
 */
int hgb_conv_run(double *cur, double *prev, int inner_w, int inner_h, int padding,
                 const double *kernel, int kw, int kh, int iterations, int ngpus);
/* arithmetic mode of hgb_conv_run (default HGB_CONV_EXACT; env HGB_CONV_FAST=1 selects FAST) */
int hgb_set_conv_mode(int mode);
/* hgb_conv_run on one GPU streams tall images in row chunks so that the PCIe copies run under the
 * iterations of other chunks (bit-identical: chunks overlap by iterations*padding rows).  `prev` may be
 * NULL when state n-1 is not wanted (halves the device-to-host traffic).  chunk_rows: 0 = never stream,
 * 1 = automatic (default; env HGB_CONV_STREAM_ROWS), else valid rows per chunk.  Asynchronous copies need
 * page-locked memory: buffers that are not page-locked yet are registered (cudaHostRegister) for the
 * duration of the call and released before it returns, so the caller may free them right away (as
 * run_on_padded_image's callers do).  Registering costs time on every call: a caller that reuses its
 * buffers should page-lock them itself (cudaHostAlloc / cudaHostRegister / torch pinned memory), which
 * the library detects and leaves alone. */
int hgb_set_conv_streaming(int chunk_rows);
/* kept for callers of the earlier ABI, where registrations outlived the call; nothing is left to release */
int hgb_conv_unpin(void);

/* ---- convolution: device-resident ------------------------------------- */
typedef struct hgb_conv hgb_conv;
int hgb_conv_create(hgb_conv **out, int inner_w, int inner_h, int padding, int ngpus);
int hgb_conv_create_rank(hgb_conv **out, int inner_w, int inner_h, int padding);
int hgb_conv_set_kernel(hgb_conv *h, const double *kernel, int kw, int kh);
/* HGB_CONV_EXACT (default) or HGB_CONV_FAST; FAST must be chosen before the upload.  hgb_conv_mode_in_effect
 * tells what the next step will really use (EXACT when the uploaded image was integer-valued). */
int hgb_conv_set_mode(hgb_conv *h, int mode);
int hgb_conv_mode_in_effect(const hgb_conv *h);
/* variant bit 0: generic kernel instead of the specialised one; bit 1: no boundary/interior overlap
 * in the multi-GPU step; bit 2: no two-iterations-per-pass kernel for 3x3; bit 3: that kernel with 4 columns per thread instead
 * of 2; band_rows: output rows
 * per CTA (0 = per-kernel default) */
int hgb_conv_tune(hgb_conv *h, int variant, int band_rows);
/* iterations per halo exchange (communication-avoiding deep halo, multi-GPU only): depth*p rows
 * travel once per `depth` iterations, the fringe is recomputed in between; 0 = automatic (8, or 16 for
 * padding 1 where the non-final iterations of a cycle run two per pass; capped by 32/p and the band height).  Returns the depth in effect (>= 1) or a negative error. */
int hgb_conv_set_depth(hgb_conv *h, int depth);
int hgb_conv_set_exchange(hgb_conv *h, int kind);           /* HGB_EXCHANGE_NCCL (send/recv) or _P2P (peer copies) */
/* rank mode: each rank passes ONLY its own band: rows [row0, row0+rows) of the image */
int hgb_conv_band(hgb_conv *h, int *row0, int *rows);
int hgb_conv_upload(hgb_conv *h, const double *cur);
int hgb_conv_step(hgb_conv *h, int iterations, float *ms);
int hgb_conv_download(hgb_conv *h, double *cur, double *prev /* either may be NULL */);
/* device-resident image lifecycle -- what src/2d-convolution.c:41-74 does with
 * add_padding / copy_padded_image / remove_padding around the benchmark, without the host round trips:
 * upload_dense = add_padding (convolution-util.c:258-269): `dense` is inner_h rows of inner_w doubles
 * (rank mode: this rank's band), pad columns are produced on the device;
 * download_dense = remove_padding (:271-279) of the newest state;
 * backup / restore = copy_padded_image (:357-373) to and from a device-side copy of the newest state. */
int hgb_conv_upload_dense(hgb_conv *h, const double *dense);
int hgb_conv_download_dense(hgb_conv *h, double *dense);
int hgb_conv_backup(hgb_conv *h);
int hgb_conv_restore(hgb_conv *h);
/* get_checksum(remove_padding(img)) (convolution-util.c:63-71): sequential row-major sum */
int hgb_conv_checksum(hgb_conv *h, double *checksum);
int hgb_conv_destroy(hgb_conv *h);

/* ---- convolution: image files (host code, no GPU needed) ----
 * Text = the reference's format (read_image_from_fd, src/convolution/convolution-util.c:9-45;
 * 2d-convolution.input, kernel.txt, laplace.txt): "<height> <width>" -- height FIRST, as the code
 * reads it -- then one space-separated line per row; same strtok/strtol/strtod token rules, but no
 * 16384-byte line limit, and errors are returned where the reference calls bail_out.
 * Binary = "HGBIMG1\n", int32 height, int32 width, height*width little-endian doubles.
 * *data is malloc'ed by the library: release it with hgb_image_free. */
int hgb_image_read_text(const char *path, int *width, int *height, double **data);
int hgb_image_write_text(const char *path, const double *data, int width, int height);
int hgb_image_read_binary(const char *path, int *width, int *height, double **data);
int hgb_image_write_binary(const char *path, const double *data, int width, int height);
void hgb_image_free(double *data);

#ifdef __cplusplus
}
#endif
#endif /* HGB200_H */
