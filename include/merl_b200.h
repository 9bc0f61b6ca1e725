/*
 * merl_b200.h -- C ABI of libmerl_b200: batched eval / pdf / sample / backward for the MERL
 * and customized_measurment measured-BRDF plugins on NVIDIA B200 (sm_100a).
 *
 * What each entry point replaces.  The reference (WenCao1101/mitsuba_customization) is
 * described only by /root/reference/README.md:1 ("Add Merl, customized_measurment brdf
 * pluggin for Mitsuba 0.6 Mitsuba 3.0"); its plugin sources are un-vendored submodules
 * (/root/reference/.SUBMODULES.json:8-23), so no file:line inside them can be cited.  The
 * interfaces named below are therefore the host-renderer virtuals those plugins must
 * override:
 *   Mitsuba 3   BSDF::eval / pdf / eval_pdf / sample (BSDFContext, SurfaceInteraction3f,
 *               wo in the local frame, Mask active)            -> merl_eval* / merl_pdf / merl_sample
 *   Mitsuba 0.6 BSDF::eval / pdf / sample (BSDFSamplingRecord)  -> the *_aos entry points
 *   plugin ctor reading `filename` (MERL .binary)              -> merl_table_create_from_file
 *   Dr.Jit AD of the table gather (scatter_reduce)             -> merl_backward (+ merl_grad_clear_counts)
 *   parameters_changed({"data"}) after an optimiser step       -> merl_table_refresh_sampler / _update_cells
 *
 * Conventions
 *   - Every query/result pointer is a DEVICE pointer owned by the caller (a Dr.Jit
 *     CUDAArray<float>::data(), a torch tensor's data_ptr(), a 0.6 wavefront buffer).
 *     The library never allocates, frees or retains them.  *_host entry points take HOST
 *     pointers instead and do their own staged copies.
 *   - SoA: a Vector3f is three float arrays x[n], y[n], z[n] (Dr.Jit layout).
 *     AoS: float4 per query (xyz + pad) for the 0.6 wavefront stage.
 *   - All launches are asynchronous on `stream` (a cudaStream_t; NULL = default stream);
 *     no hidden device synchronisation.  A table handle may be used from any thread; it
 *     changes only through merl_table_set_cdf / _refresh_sampler / _update_cells (or the
 *     caller writing d_cells), which are ordered on their stream like any launch and must
 *     not overlap queries on other streams.
 *   - Return 0 on success, a negative merl_status otherwise; merl_last_error() returns a
 *     thread-local message.  No exception crosses this boundary.  There is no CPU
 *     fallback: without a CUDA device every compute entry point fails with MERL_ERR_CUDA.
 */
#ifndef MERL_B200_H
#define MERL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MERL_B200_ABI_VERSION 2

typedef enum merl_status {
    MERL_OK = 0,
    MERL_ERR_ARG = -1,     /* NULL / misaligned / out-of-range argument */
    MERL_ERR_IO = -2,      /* file missing, short or not MERL layout    */
    MERL_ERR_CUDA = -3,    /* CUDA runtime error (message has details)  */
    MERL_ERR_NOMEM = -4,
    MERL_ERR_UNSUPPORTED = -5
} merl_status;

typedef struct merl_table merl_table_t;
typedef void* merl_stream_t; /* cudaStream_t */

/* Plugin parameters (the Properties a Mitsuba plugin would read). */
typedef struct merl_table_params {
    int32_t nh, nd, np;        /* table resolution; 0,0,0 = take from the file header      */
    int32_t theta_h_sqrt;      /* 1 = MERL nonlinear theta_half bins, 0 = linear            */
    int32_t ni;                /* sampler: incident-elevation slices, power of two <= 64    */
    int32_t clamp_negative;    /* 1 = negative (invalid) cells evaluate to 0                */
    int32_t cdf_nm, cdf_nq;    /* sampler CDF quadrature: sub-steps in theta_h, steps in phi_h */
    int32_t l2_persist;        /* 1 = launches carry a persisting-L2 access window on the table */
    int32_t reserved;
    float cdf_floor;           /* share of solid-angle-uniform mass mixed into every row    */
    float spec_weight;         /* probability of the tabulated half-vector lobe             */
    double scale[3];           /* RED/GREEN/BLUE_SCALE                                      */
} merl_table_params_t;

typedef struct merl_table_info {
    int32_t nh, nd, np, ni;
    int32_t theta_h_sqrt, device;
    int64_t ncell;
    int64_t table_bytes;       /* float4 table: ncell * 16                                   */
    const void* d_cells;       /* device float4[ncell]: the differentiable parameter         */
    const float* d_cdf;        /* device float[ni][nh+1]                                     */
    const float* d_dens;       /* device float[ni][nh]                                       */
    float spec_weight;
    int32_t smem_eval_bytes, smem_sample_bytes;
} merl_table_info_t;

/* ---- tables ---------------------------------------------------- */
void merl_table_params_default(merl_table_params_t* p);            /* MERL: 90/90/180, scales /1500 */
int merl_table_create_from_file(const char* path, const merl_table_params_t* p, int device,
                                merl_table_t** out);
/* planar_rgb: [3][nh*nd*np] doubles in .binary units (host memory). */
int merl_table_create_from_host(const double* planar_rgb, const merl_table_params_t* p, int device,
                                merl_table_t** out);
int merl_table_destroy(merl_table_t* t);
int merl_table_info(const merl_table_t* t, merl_table_info_t* out);
/* Replace the sampler CDF (host float[ni][nh+1]); the density table is re-derived. */
int merl_table_set_cdf(merl_table_t* t, const float* cdf_host);
/* After the caller has modified the float4 table in place on the device (d_cells is the
 * differentiable parameter): rebuild the sampler CDF rows and densities from it, asynchronously
 * on `stream`.  Mitsuba 3: parameters_changed({"data"}). */
int merl_table_refresh_sampler(merl_table_t* t, merl_stream_t stream);
/* The same when the new values live in ANOTHER device buffer (float4[ncell], same layout as
 * d_cells): Dr.Jit optimisers replace a parameter array instead of writing into it, so the
 * Mitsuba 3 host's parameters_changed({"data"}) hands the new array here.  Device-to-device copy
 * into d_cells (skipped when cells4_device == d_cells), then the sampler refresh; asynchronous
 * on `stream`. */
int merl_table_update_cells(merl_table_t* t, const void* cells4_device, merl_stream_t stream);
/* Copy tables back to the host; any pointer may be NULL.  Synchronous. */
int merl_table_download(const merl_table_t* t, float* cells4_host, float* cdf_host, float* dens_host);

/* ---- table files either side of the loader ---------------------------------
 * merl_file_report: host-only scan of a MERL-layout .binary file (header, size, per-channel range,
 * negative = below-horizon markers, non-finite values); fails with MERL_ERR_IO when the header is
 * implausible or the size does not match it.  Works without a GPU.
 * merl_table_save_packed / merl_table_create_from_packed: a sidecar holding the table as it lives
 * on the device (prescaled float4 cells, sampler CDF rows, the parameters, a checksum), so loading
 * skips the double -> float4 repack and the CDF quadrature; the resulting table is bit-identical. */
typedef struct merl_file_report {
    int32_t dims[3];
    int32_t is_merl_layout;    /* dims multiply to 90*90*180                                  */
    int64_t ncell, file_bytes, expected_bytes;
    int64_t negative[3], nonfinite[3];
    double min_valid[3], max_valid[3];   /* over finite, non-negative values, raw file units   */
} merl_file_report_t;
int merl_file_report(const char* path, merl_file_report_t* out);
int merl_table_save_packed(const merl_table_t* t, const char* path);
int merl_table_create_from_packed(const char* path, int device, merl_table_t** out);

/* ---- batched queries, SoA device pointers (rows A1-A6, A10) ----------------------------
 * `active` may be NULL (all lanes active) or a device array of n bytes (Dr.Jit Mask).
 * Lanes that are inactive or have wi.z <= 0 or wo.z <= 0 produce zeros. */
int merl_bins(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
              const float* wox, const float* woy, const float* woz, const uint8_t* active, int64_t n,
              int32_t* idx_out, merl_stream_t stream);
int merl_eval(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
              const float* wox, const float* woy, const float* woz, const uint8_t* active, int64_t n,
              float* r, float* g, float* b, merl_stream_t stream);
int merl_pdf(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
             const float* wox, const float* woy, const float* woz, const uint8_t* active, int64_t n,
             float* pdf, merl_stream_t stream);
int merl_eval_pdf(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                  const float* wox, const float* woy, const float* woz, const uint8_t* active,
                  int64_t n, float* r, float* g, float* b, float* pdf, merl_stream_t stream);
/* sample1 -> u1 (lobe choice), sample2 -> (u2x, u2y).  lobe: 0 = cosine, 1 = tabulated. */
int merl_sample(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                const float* u1, const float* u2x, const float* u2y, const uint8_t* active, int64_t n,
                float* wox, float* woy, float* woz, float* pdf, float* wr, float* wg, float* wb,
                uint32_t* lobe, merl_stream_t stream);
/* grad_cells4 (device float4[ncell], caller-zeroed) += d(eval)/d(cells) . (gr, gg, gb).
 * .w accumulates the number of contributing queries.  Channels of invalid (negative, clamped) cells
 * receive no gradient: they are kept at zero in grad_cells4 (batches with at least as many queries
 * as the table has cells scatter unmasked and zero those channels in one pass afterwards, so
 * grad_cells4 must not be shared with gradients from other sources).
 * flags bit0 MERL_BWD_WARP_AGGREGATE: fold lanes of a warp that hit the same cell with __match_any_sync
 *   before the atomic.  OFF by default in every host of this repository (0 = plain per-lane RED): on
 *   uniformly random batches collisions inside a warp are rare and the fold costs 4 %; set it for
 *   coherent batches (gradients of a rendered image: neighbouring pixels share cells).
 * flags bit1 MERL_BWD_DEFER_MASK: scatter unmasked whatever the batch size and leave the validity mask
 *   to the caller, who applies it ONCE after the last batch: merl_peerbuf_allreduce(.., table, ..)
 *   fuses it into the multi-GPU exchange, merl_grad_apply_mask does it on one GPU.  This is the
 *   intended form when gradients accumulate over several batches per exchange.
 * flags bit2 MERL_BWD_CTA_AGGREGATE: coherent batches (see merl_backward notes in DESIGN.md section 5):
 *   a per-CTA shared-memory accumulator folds the queries of a CTA that hit the same cell before one
 *   vector RED per distinct cell and flush. */
#define MERL_BWD_WARP_AGGREGATE 1u
#define MERL_BWD_DEFER_MASK 2u
#define MERL_BWD_CTA_AGGREGATE 4u
int merl_backward(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                  const float* wox, const float* woy, const float* woz, const float* gr,
                  const float* gg, const float* gb, const uint8_t* active, int64_t n,
                  float* grad_cells4, uint32_t flags, merl_stream_t stream);

/* Zero the hit-count lane (.w) of a gradient table, leaving the pure gradient of the float4 cells
 * (the .w lane of the table holds validity bits, which an optimiser must not move): what the
 * Mitsuba 3 host's dr::CustomOp hands back to Dr.Jit.  Asynchronous on `stream`. */
int merl_grad_clear_counts(const merl_table_t* t, float* grad_cells4, merl_stream_t stream);
/* Zero the channels of invalid (negative, clamped) cells in a gradient table that was accumulated
 * with MERL_BWD_DEFER_MASK.  Asynchronous on `stream`. */
int merl_grad_apply_mask(const merl_table_t* t, float* grad_cells4, merl_stream_t stream);

/* ---- AoS entry points for the Mitsuba 0.6 wavefront BSDF stage --------------------------
 * wi4 / wo4: float4 per query (xyz, w ignored).  out4 = (r, g, b, pdf). */
int merl_eval_pdf_aos(const merl_table_t* t, const float* wi4, const float* wo4, int64_t n, float* out4,
                      merl_stream_t stream);
/* u4 = (u1, u2x, u2y, unused); wo4_out = (wo.xyz, pdf); weight4_out = (w.rgb, lobe as float). */
int merl_sample_aos(const merl_table_t* t, const float* wi4, const float* u4, int64_t n, float* wo4_out,
                    float* weight4_out, merl_stream_t stream);

/* ---- host-buffer entry points (scalar-variant plugins, end-to-end timing) ---------------
 * Same semantics as merl_eval_pdf / merl_sample but every pointer is HOST memory (pinned
 * gives full PCIe rate).  Copies and kernels are pipelined in chunks over internal streams;
 * returns after the results are in the host buffers. */
int merl_eval_pdf_host(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                       const float* wox, const float* woy, const float* woz, int64_t n, float* r,
                       float* g, float* b, float* pdf);
int merl_sample_host(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                     const float* u1, const float* u2x, const float* u2y, int64_t n, float* wox,
                     float* woy, float* woz, float* pdf, float* wr, float* wg, float* wb, uint32_t* lobe);

/* ---- the same eval+pdf on host buffers, for callers that must not allocate --------------------
 * No counterpart in the reference: this entry point exists only because of how this library is
 * hosted.  merl_eval_pdf_host cudaMallocs its device staging buffers on first use and regrows them
 * for a larger batch.  That is wrong for three kinds of caller: a renderer that has already given
 * the whole device to its own pool (Dr.Jit's or torch's caching allocator, a scene of a hundred
 * tables), where a late cudaMalloc fails or forces the pool to shrink; a frame or optimiser loop
 * that cannot afford the device-wide synchronisation that cudaMalloc and cudaFree imply; and a
 * scalar-variant plugin queried from many worker threads, where the first query on each new table
 * would stall all of them while buffers are made.
 * merl_eval_pdf_host_static stages through a fixed array that is part of the library image instead
 * (40 bytes per query, MERL_STATIC_STAGE_MAX_QUERIES queries = 2.5 MiB, resident on every GPU the
 * library is loaded on until the process exits).  It calls no cudaMalloc / cudaFree / pool or
 * stream-creation routine: copy in on `stream`, one eval+pdf launch (two when n is not a multiple
 * of 4), copy out, cudaStreamSynchronize(stream).  Same arguments, results (bit for bit) and
 * errors as merl_eval_pdf_host, plus:
 *   - n > MERL_STATIC_STAGE_MAX_QUERIES is refused with MERL_ERR_ARG; split the batch.
 *   - The first call on each device looks up the array's device address and the kernel's occupancy
 *     figure (host-side bookkeeping only).  Make that first call outside any stream capture.
 *   - All calls on a device share the one array, which keeps the previous call's data; the entry
 *     point overwrites every element it reads, so nothing leaks between calls.  Calls from one
 *     process are serialised by a lock held until the copy-back has finished.  Work that would
 *     touch the array concurrently on the GPU by any other route would race. */
#define MERL_STATIC_STAGE_MAX_QUERIES 65536
int merl_eval_pdf_host_static(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                              const float* wox, const float* woy, const float* woz, int64_t n, float* r,
                              float* g, float* b, float* pdf, merl_stream_t stream);

/* ---- wavefront path tracer harness (BASELINE configs 2 and 5) ------------
 * The callers either side of the BSDF stage, headless: spheres under a constant environment map,
 * pinhole camera, path integrator with emitter sampling + BSDF sampling combined by the power
 * heuristic, Philox4x32-10 random streams keyed by (pixel, sample, bounce).  Each bounce is a
 * wavefront: trace -> (sort by material) -> BSDF stage (the eval_pdf / sample kernels above on
 * float4 AoS buffers) -> shade.  oracle/render_oracle.c states the same algorithm on the CPU.
 * film4: device float4[width*height], caller-zeroed; every finished path ADDS radiance/spp_total
 * to its pixel (.w counts samples/spp_total), so ranks that render disjoint sample ranges
 * [spp_begin, spp_end) of the same spp_total sum to the full image with one allreduce. */
typedef struct merl_sphere {
    float cx, cy, cz, radius;
    int32_t material;          /* index into tables[] */
    int32_t pad[3];
} merl_sphere_t;

typedef struct merl_render_params {
    int32_t width, height;
    int32_t spp_total;         /* normalisation                                         */
    int32_t spp_begin, spp_end;/* sample indices this call renders for every pixel       */
    int32_t max_depth;         /* surface interactions per path (>= 1)                   */
    int32_t wavefront;         /* paths per pass (264 B of workspace each); 0 = up to 1 << 26, at most 1/4 of the free device memory */
    int32_t reserved;
    uint64_t seed;
    float cam_pos[3], cam_target[3], cam_up[3];
    float fov_x_deg;
    float env[3];              /* constant environment radiance (RGB)                    */
    /* By-table sharding (BASELINE config 5 "sharded by table"), the allreduce-only
     * form: every rank traces every path of [spp_begin, spp_end) -- the SAME range on all ranks --, runs
     * the BSDF stage only for tables m with m % table_shard_world == table_shard_rank (tables[m] of the
     * other materials may be NULL: a rank holds only its share), and the stage outputs (48 B per path
     * slot) are summed over the ranks through `stage_exchange` once per bounce.  Every rank ends with the
     * full film; no film exchange.  table_shard_world <= 1: off (tables replicated, ranks split samples). */
    int32_t table_shard_world, table_shard_rank;
    struct merl_peerbuf* stage_exchange;   /* >= 12 * (paths per pass) floats; required when sharding */
} merl_render_params_t;

typedef struct merl_render_stats {
    int64_t paths, bsdf_queries, launches;
} merl_render_stats_t;

int merl_render(const merl_table_t* const* tables, int ntables, const merl_sphere_t* spheres_host, int nspheres,
                const merl_render_params_t* p, float* film4, merl_render_stats_t* stats_out /* may be NULL */,
                merl_stream_t stream);

/* merl_render replays one CUDA graph per wavefront pass (default).  0 = enqueue every kernel on the
 * caller's stream instead (same results; for A/B timing and for hosts that are themselves capturing). */
int merl_set_render_graphs(int on);

/* ---- multi-GPU exchange step: table gradients and film buffers --------------
 * The forward path shards by query batch with no collective.  The backward path keeps a private
 * gradient table per GPU and sums them with ONE allreduce; a renderer sums (or gathers) film
 * tiles the same way.  NCCL is bound at run time (dlopen of libnccl.so.2: the copy already loaded
 * by the host process, e.g. PyTorch's, else the system one), so the library loads without it.
 * One process per GPU: rank 0 calls merl_comm_unique_id, the host ships the 128 bytes to the
 * other ranks by whatever it has (MPI, torch.distributed, a file), every rank calls
 * merl_comm_create.  merl_allreduce_sum_f32 is in place, asynchronous on `stream`. */
typedef struct merl_comm merl_comm_t;
#define MERL_COMM_ID_BYTES 128
int merl_comm_unique_id(void* id_out /* MERL_COMM_ID_BYTES */);
int merl_comm_create(const void* id, int nranks, int rank, int device, merl_comm_t** out);
int merl_comm_destroy(merl_comm_t* c);
int merl_allreduce_sum_f32(merl_comm_t* c, float* buf, int64_t count, merl_stream_t stream);

/* ---- exchange buffers: ONE kernel over NVLink peer memory instead of a library allreduce -------
 * A gradient table (or a film) that the ranks of a communicator sum in place.  merl_peerbuf_create
 * is COLLECTIVE (every rank calls it with the same count, in the same order relative to other
 * collectives of `c`): it allocates `count` floats (rounded up to a multiple of 4, zeroed) on the
 * communicator's device and
 *   MERL_XCHG_PEER (default): maps every rank's buffer into every other rank's address space (CUDA
 *     IPC; one process per GPU, peers reachable over NVLink/NVSwitch or PCIe P2P).
 *     merl_peerbuf_allreduce is then one persistent kernel per rank: rank r pulls slab r of all R
 *     buffers, sums in rank order, optionally applies mask_table's validity mask, pushes the result
 *     into slab r of all R buffers; two cross-GPU flag barriers.  All ranks end with identical bits.
 *     No host-side state per call: capturable in a CUDA graph.
 *   MERL_XCHG_NCCL: ncclMemAlloc + registration with the communicator (symmetric window, else
 *     ncclCommRegister); merl_peerbuf_allreduce is ncclAllReduce in place (+ the mask pass).
 *   MERL_XCHG_NVLS: the buffer is an NCCL symmetric window with a multicast mapping (NCCL >= 2.28,
 *     NVSwitch); merl_peerbuf_allreduce is this library's own kernel: rank r reads slab r with
 *     multimem.ld_reduce (the switch adds the R copies), applies the mask, writes it back to all R
 *     copies with multimem.st; multicast flag barriers.  MERL_ERR_UNSUPPORTED when the build had no
 *     NCCL device headers, the loaded NCCL is not exactly the release of those headers, or there is no multicast over all ranks
 *     (every rank gets the same answer, so a host can fall back to another transport collectively).
 * merl_peerbuf_allreduce(count = floats to sum, multiple of 4; mask_table = NULL or the table whose
 * gradient this is, then count must be 4 * ncell) is asynchronous on `stream`; every rank must call it
 * with the same count, and exchanges of one communicator must be issued in the same order everywhere.
 * merl_peerbuf_info.status != 0: a cross-GPU barrier timed out (a rank died or skipped a call). */
typedef struct merl_peerbuf merl_peerbuf_t;
#define MERL_XCHG_PEER 0
#define MERL_XCHG_NCCL 1
#define MERL_XCHG_NVLS 2
typedef struct merl_peerbuf_info {
    float* data;               /* device float[count], owned by the library */
    int64_t count;
    int32_t transport, nranks, rank, status;
    char registration[64];     /* how the buffer was made visible to the other ranks */
} merl_peerbuf_info_t;
int merl_peerbuf_create(merl_comm_t* c, int64_t count, int transport, merl_peerbuf_t** out);
int merl_peerbuf_destroy(merl_peerbuf_t* b);   /* collective */
int merl_peerbuf_info(const merl_peerbuf_t* b, merl_peerbuf_info_t* out);
int merl_peerbuf_allreduce(merl_peerbuf_t* b, int64_t count, const merl_table_t* mask_table, merl_stream_t stream);

/* ---- diagnostics ------------------------------------------------------------------------ */
const char* merl_last_error(void);
int merl_abi_version(void);
/* Number of kernels this library has launched in this process (all threads). */
int64_t merl_launch_count(void);
/* Selects the CDF search of merl_sample: 0 = per-lane bisection, 1 = warp-cooperative
 * ballot search.  Both give identical results; see DESIGN.md §4 for the measurements. */
int merl_set_sample_search(int mode);

#ifdef __cplusplus
}
#endif
#endif /* MERL_B200_H */
