/* vortex_b200.h — C ABI of libvortex-b200.so, the B200 runtime backend for the Vortex ray-tracing path.
 *
 * Plain C: pointers, sizes and integers only.  Three groups of entry points:
 *
 *  (1) vx_dev_init()  — the ONE symbol the reference's stub dispatcher binds
 *      (reference: runtime/common/callbacks.h:74, looked up by runtime/stub/vortex.cpp:58-97 after
 *      dlopen("libvortex-$VORTEX_DRIVER.so")).  It fills the 16-slot callback table below, whose slot
 *      order and signatures are the reference's callbacks_t (runtime/common/callbacks.h:23-72).
 *      With VORTEX_DRIVER=b200 an unmodified Vortex host application runs on this backend.
 *
 *  (2) vx_* — the same 17 runtime calls + 4 upload helpers the reference declares in
 *      runtime/include/vortex.h:80-145, exported directly so a host can also link this library
 *      in place of the stub (no dlopen hop).  Names, argument meaning and error behaviour follow
 *      the reference glue (runtime/common/callbacks.inc:20-225, runtime/stub/utils.cpp:25-122):
 *      0 = success, non-zero (-1) = failure, null handles / zero sizes / out-of-range offsets rejected.
 *
 *  (3) vx_b200_* — extensions that have no reference counterpart (multi-GPU partition, device
 *      pointer export for NCCL, timing, per-ray hit-record queries, statistics).
 *
 * What vx_start() does here: it reads the kernel_arg_t bytes the host uploaded into the arguments
 * buffer, ignores the RISC-V kernel image, and launches hand-written sm_100a CUDA kernels that
 * consume the reference's scene buffers byte for byte:
 *   - 192-byte kernel_arg_t  -> layout A, the software ray-caster (tests/regression/raycast/common.h:120-147)
 *   - 216-byte kernel_arg_t  -> layout B, the simx RT-unit path   (tests/regression/raytracing/common.h:164-195)
 * There is no CPU fallback: every entry point fails (non-zero) when no CUDA device is usable.
 */
#ifndef VORTEX_B200_H
#define VORTEX_B200_H

#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void *vx_device_h;
typedef void *vx_buffer_h;

/* ---- capability ids (reference: runtime/include/vortex.h:30-44) ---- */
#define VX_CAPS_VERSION 0x0
#define VX_CAPS_NUM_THREADS 0x1
#define VX_CAPS_NUM_WARPS 0x2
#define VX_CAPS_NUM_CORES 0x3
#define VX_CAPS_CACHE_LINE_SIZE 0x4
#define VX_CAPS_GLOBAL_MEM_SIZE 0x5
#define VX_CAPS_LOCAL_MEM_SIZE 0x6
#define VX_CAPS_ISA_FLAGS 0x7
#define VX_CAPS_NUM_MEM_BANKS 0x8
#define VX_CAPS_MEM_BANK_SIZE 0x9
#define VX_CAPS_NUM_CLUSTERS 0xA
#define VX_CAPS_SOCKET_SIZE 0xB
#define VX_CAPS_ISSUE_WIDTH 0xC
#define VX_CAPS_CLOCK_RATE 0xD
#define VX_CAPS_PEAK_MEM_BW 0xE

/* ---- memory access flags (reference: vortex.h:74-78) ---- */
#define VX_MEM_READ 0x1
#define VX_MEM_WRITE 0x2
#define VX_MEM_READ_WRITE 0x3
#define VX_MEM_PIN_MEMORY 0x4

#define VX_MAX_TIMEOUT (24 * 60 * 60 * 1000)

/* ---- device configuration registers this backend understands (reference: hw/VX_types.toml:16-19,
 *      generated VX_types.h: VX_DCR_BASE_*).  0x006..0x009 carry the RT-unit scene pointers the
 *      raytracing host writes (tests/regression/raytracing/tracer.cpp:252-256). ---- */
#define VX_DCR_BASE_STARTUP_ADDR0 0x001
#define VX_DCR_BASE_STARTUP_ADDR1 0x002
#define VX_DCR_BASE_STARTUP_ARG0 0x003
#define VX_DCR_BASE_STARTUP_ARG1 0x004
#define VX_DCR_BASE_MPM_CLASS 0x005
#define VX_DCR_BASE_RTX_TLAS_PTR 0x006
#define VX_DCR_BASE_RTX_BLAS_PTR 0x007
#define VX_DCR_BASE_RTX_BVH_PTR 0x008
#define VX_DCR_BASE_RTX_TRI_PTR 0x009
/* b200-private DCRs (outside every range the reference assigns: base 0x001-0x00A, DXA 0x100-0x280) */
#define VX_DCR_B200_RT_MODE 0x0F0    /* bit0 shadow rays, bit1 diffuse bounces (SURVEY 8d extensions); default 0 */
#define VX_DCR_B200_RT_LAYOUT 0x0F1  /* 0 = infer from argument size (192 -> A, 216 -> B), 1 = A, 2 = B */
#define VX_DCR_B200_RT_ORDER 0x0F2   /* 0 = near-first BLAS order (fast), 1 = the reference's far-first order (strict) */
#define VX_DCR_B200_RT_CAMERA 0x0F3  /* layout B only: 0 = the hard-coded camera of raytracing/kernel.cpp:28-39, 1 = kernel_arg_t pinhole camera */
#define VX_DCR_B200_RT_KERNEL 0x0F4  /* render kernel: 0 = default (layout A: one pixel per lane over the 4-wide records, env VORTEX_B200_DIRECT=0 selects the warp-scheduled ray state machine; layout B: one ray per lane over the cursor-carrying records), 1 = round 1's one-pixel-per-lane kernels (BVH2 walk / records without cursors), 2 + n = layout-B shader program n */

#define VX_B200_MODE_SHADOW 0x1u
#define VX_B200_MODE_DIFFUSE 0x2u

/* ---- performance counter addresses answered by vx_mpm_query (reference: VX_types.h VX_CSR_*) ---- */
#define VX_CSR_MPM_BASE 0xB00
#define VX_CSR_MCYCLE 0xB00
#define VX_CSR_MINSTRET 0xB02

/* ---- (1) the dispatcher hook ------------------------------------------------------------------ */
typedef struct {
  int (*dev_open)(vx_device_h *hdevice);
  int (*dev_close)(vx_device_h hdevice);
  int (*dev_caps)(vx_device_h hdevice, uint32_t caps_id, uint64_t *value);
  int (*mem_alloc)(vx_device_h hdevice, uint64_t size, int flags, vx_buffer_h *hbuffer);
  int (*mem_reserve)(vx_device_h hdevice, uint64_t address, uint64_t size, int flags, vx_buffer_h *hbuffer);
  int (*mem_free)(vx_buffer_h hbuffer);
  int (*mem_access)(vx_buffer_h hbuffer, uint64_t offset, uint64_t size, int flags);
  int (*mem_address)(vx_buffer_h hbuffer, uint64_t *address);
  int (*mem_info)(vx_device_h hdevice, uint64_t *mem_free, uint64_t *mem_used);
  int (*copy_to_dev)(vx_buffer_h hbuffer, const void *host_ptr, uint64_t dst_offset, uint64_t size);
  int (*copy_from_dev)(void *host_ptr, vx_buffer_h hbuffer, uint64_t src_offset, uint64_t size);
  int (*start)(vx_device_h hdevice, vx_buffer_h hkernel, vx_buffer_h harguments);
  int (*ready_wait)(vx_device_h hdevice, uint64_t timeout);
  int (*dcr_read)(vx_device_h hdevice, uint32_t addr, uint32_t *value);
  int (*dcr_write)(vx_device_h hdevice, uint32_t addr, uint32_t value);
  int (*mpm_query)(vx_device_h hdevice, uint32_t addr, uint32_t core_id, uint64_t *value);
} callbacks_t; /* reference: runtime/common/callbacks.h:23-72 — same slot order */

int vx_dev_init(callbacks_t *callbacks); /* reference: runtime/common/callbacks.h:74 */

/* ---- (2) the runtime API, exported directly (reference: runtime/include/vortex.h:80-145) -------- */
int vx_dev_open(vx_device_h *hdevice);                                    /* vortex.h:81  */
int vx_dev_close(vx_device_h hdevice);                                    /* vortex.h:84  */
int vx_dev_caps(vx_device_h hdevice, uint32_t caps_id, uint64_t *value);  /* vortex.h:87  */
int vx_mem_alloc(vx_device_h hdevice, uint64_t size, int flags, vx_buffer_h *hbuffer);                    /* vortex.h:90  */
int vx_mem_reserve(vx_device_h hdevice, uint64_t address, uint64_t size, int flags, vx_buffer_h *hbuffer); /* vortex.h:93  */
int vx_mem_free(vx_buffer_h hbuffer);                                     /* vortex.h:96  */
int vx_mem_access(vx_buffer_h hbuffer, uint64_t offset, uint64_t size, int flags);                         /* vortex.h:99  */
int vx_mem_address(vx_buffer_h hbuffer, uint64_t *address);               /* vortex.h:102 */
int vx_mem_info(vx_device_h hdevice, uint64_t *mem_free, uint64_t *mem_used);                              /* vortex.h:105 */
int vx_copy_to_dev(vx_buffer_h hbuffer, const void *host_ptr, uint64_t dst_offset, uint64_t size);         /* vortex.h:108 */
int vx_copy_from_dev(void *host_ptr, vx_buffer_h hbuffer, uint64_t src_offset, uint64_t size);             /* vortex.h:111 */
int vx_start(vx_device_h hdevice, vx_buffer_h hkernel, vx_buffer_h harguments);                            /* vortex.h:114 */
int vx_ready_wait(vx_device_h hdevice, uint64_t timeout);                 /* vortex.h:117 */
int vx_dcr_read(vx_device_h hdevice, uint32_t addr, uint32_t *value);     /* vortex.h:120 */
int vx_dcr_write(vx_device_h hdevice, uint32_t addr, uint32_t value);     /* vortex.h:123 */
int vx_mpm_query(vx_device_h hdevice, uint32_t addr, uint32_t core_id, uint64_t *value);                   /* vortex.h:126 */
int vx_upload_kernel_bytes(vx_device_h hdevice, const void *content, uint64_t size, vx_buffer_h *hbuffer); /* vortex.h:131, stub/utils.cpp:25-61 */
int vx_upload_kernel_file(vx_device_h hdevice, const char *filename, vx_buffer_h *hbuffer);                /* vortex.h:134 */
int vx_upload_bytes(vx_device_h hdevice, const void *content, uint64_t size, vx_buffer_h *hbuffer);        /* vortex.h:137, stub/utils.cpp:88-106 */
int vx_upload_file(vx_device_h hdevice, const char *filename, vx_buffer_h *hbuffer);                       /* vortex.h:140 */
int vx_dump_perf(vx_device_h hdevice, FILE *stream);                                                       /* vortex.h:146 */

/* ---- (3) b200 extensions ----------------------------------------------------------------------- */

/* Which CUDA device the NEXT vx_dev_open() binds (default: env VORTEX_B200_DEVICE, else LOCAL_RANK, else 0). */
int vx_b200_select_device(int cuda_device);

/* How many GPUs of this box the NEXT vx_dev_open() spans (default: env VORTEX_B200_GPUS, else 1).  The vx_* API stays
 * single-device (the reference host opens ONE handle and reads ONE framebuffer: runtime/stub/vortex.cpp:53-97,
 * tests/regression/raycast/tracer.cpp:231-240) — with N > 1 the backend
 *   - keeps one arena per GPU with identical Vortex addresses;
 *   - vx_copy_to_dev: large buffers enter the box ONCE, slice i/N through GPU i's own PCIe link, and are replicated to the other
 *     GPUs over NVLink (peer copies that follow the slices; VORTEX_B200_COLLECTIVE=nccl: one ncclAllGather instead);
 *   - vx_start: 8-row bands of the image round-robin over the GPUs, one persistent kernel per GPU on its own stream;
 *   - vx_copy_from_dev(framebuffer): every GPU copies its own bands straight into the caller's buffer (one strided copy per
 *     GPU over its own PCIe link); any other read of the framebuffer first gathers the bands on GPU 0 over NVLink;
 *   - vx_b200_trace_rays / _host: rays sharded by index.
 * Devices used: first .. first+N-1 (first = vx_b200_select_device / VORTEX_B200_DEVICE, default 0 when N > 1). */
int vx_b200_select_gpus(int n_gpus);

/* Asynchronous readback (N >= 1): after this call every vx_start whose framebuffer is `hframebuffer` is followed, on the GPU
 * streams, by the copy of the finished image into `host_ptr` (page-locked memory of at least width*height*4 bytes; each GPU
 * copies its own bands), so that vx_ready_wait returns with the image already in host memory and the vx_copy_from_dev of the
 * whole framebuffer into that same pointer has nothing left to do.  host_ptr == NULL cancels. */
int vx_b200_set_readback(vx_device_h hdevice, vx_buffer_h hframebuffer, void *host_ptr, uint64_t bytes);

/* Write 512 MB of scratch on every GPU of the device (stream-ordered): flushes the 126 MB L2 between timed iterations. */
int vx_b200_flush_l2(vx_device_h hdevice);

/* Screen-space partition for multi-GPU runs: the image is cut into bands of `band_rows` rows and this
 * device renders band j iff j % world == rank.  Pixels of other bands are left untouched.
 * world == 1 (default) renders everything.  Env defaults: VORTEX_B200_RANK / VORTEX_B200_WORLD. */
int vx_b200_set_partition(vx_device_h hdevice, uint32_t rank, uint32_t world, uint32_t band_rows);

/* Multi-GPU framebuffer over NVLink peer memory (one process per GPU).  The owner of the final image exports its framebuffer
 * buffer (a CUDA IPC handle of the device arena + the buffer's offset in it); every other rank passes that to
 * vx_b200_set_remote_framebuffer, after which its render kernels store their bands' pixels DIRECTLY into the owner's
 * framebuffer over NVLink/NVSwitch while they trace (no separate gather step, no staging copy).  The image is complete on
 * the owner once every rank's vx_ready_wait has returned (synchronise the ranks, e.g. a barrier, before reading it).
 * handle == NULL detaches.  `bytes` must cover width*height*4 of every later vx_start. */
int vx_b200_export_buffer(vx_buffer_h hbuffer, void *ipc_handle_64, uint64_t *offset);
int vx_b200_set_remote_framebuffer(vx_device_h hdevice, const void *ipc_handle_64, uint64_t offset, uint64_t bytes);

/* Raw CUDA device pointer behind a Vortex buffer (for NCCL / torch interop on the same process). */
int vx_b200_device_ptr(vx_buffer_h hbuffer, uint64_t *cuda_ptr);

/* Tell the runtime that a buffer's device memory was written without vx_copy_to_dev (an NCCL receive into
 * vx_b200_device_ptr, a peer's store): the GPU-side records derived from it are rebuilt at the next vx_start. */
int vx_b200_mark_written(vx_buffer_h hbuffer);

/* Measurement aid (no reference counterpart): bandwidth of random gathers of `record_bytes` (32 or 128) over a scratch buffer
 * of `buffer_bytes` — L2-resident when it is well under the 126 MB L2, HBM-bound beyond.  This is the "L2 -> SM gather" roofline
 * denominator of the traversal kernels; bench.py reports it next to the HBM and FP32 ones. */
int vx_b200_probe_gather(vx_device_h hdevice, uint64_t buffer_bytes, uint32_t record_bytes, double *gbs);

/* The CUDA stream (cudaStream_t as integer) all of this device's kernels and copies are issued on. */
int vx_b200_stream(vx_device_h hdevice, uint64_t *cuda_stream);

/* Device time of the kernels of the last completed vx_start()/vx_b200_trace_* (CUDA events on the launch stream). */
int vx_b200_last_kernel_ms(vx_device_h hdevice, float *ms);

typedef struct {
  uint64_t rays;            /* TLASIntersect-equivalent queries issued by the last run (primary + bounce + shadow) */
  uint64_t box_tests;       /* child boxes slab-tested   (only when statistics are enabled) */
  uint64_t tri_tests;       /* triangles Moller-Trumbore tested */
  uint64_t blas_entered;    /* instance (BLAS) entries */
  uint64_t shaded_hits;     /* closest-hit shading invocations */
  uint64_t kernel_launches; /* CUDA kernels launched by the last run */
  uint64_t prep_launches;   /* of those: scene re-flatten kernels (0 when the cached GPU layout was reused) */
  float kernel_ms;          /* == vx_b200_last_kernel_ms */
  float prep_ms;            /* device time of the re-flatten pass inside the last run */
  /* RT-unit style counters (the shape of RTUnit::PerfStats, sim/simx/rt_unit.h:15-45), statistics mode only:
   * one step = one traversal iteration (node fetch + box tests, or one leaf); SIMT efficiency of the run =
   * lane_steps / (32 * warp_steps), the figure simx prints as rt_total_simt_efficiency / rt_total_warps. */
  uint64_t lane_steps;      /* traversal iterations summed over lanes   (~ rt_total_thread_latency) */
  uint64_t warp_steps;      /* traversal iterations issued by warps      (~ rt_total_warp_latency) */
  uint64_t warps;           /* warps that ran                            (~ rt_total_warps) */
  /* one vx_device_h over several GPUs (VORTEX_B200_GPUS / vx_b200_select_gpus): the counters above are sums over the GPUs,
   * kernel_ms / prep_ms the maxima; group_ms is the device time from the moment every GPU may start to the moment the last
   * one (and any readback copy) finished, taken by CUDA events on GPU 0's stream which joins the others */
  uint32_t gpus;
  float group_ms;
  float kernel_ms_min, kernel_ms_max;   /* fastest / slowest GPU of the run */
  /* RT-unit latency distribution (statistics mode; the shape of RTUnit::PerfStats::rt_latency_dist[warp_statuses][ray_statuses],
   * sim/simx/rt_unit.h:24, in lane-iterations of the ray state machine instead of simulated cycles).  Row = what the warp was
   * doing: 0 fetching work (GEN), 1 shading (POST), 2 traversing (TRAV).  Column = what a lane of that warp was doing meanwhile:
   * 0 awaiting_processing (traversal finished, waiting for POST), 1 awaiting_scheduling (empty slot, waiting for GEN),
   * 2 awaiting_mf (holds a node of the other kind than the one the warp is stepping), 3 executing_op (takes part in the step),
   * 4 trace_complete (retired: no work left for this lane). */
  uint64_t latency_dist[3][5];
  /* traversal steps per ray, log2 buckets: [0] = 0 steps, [k] = 2^(k-1) .. 2^k - 1 steps, [15] = 2^14 and more */
  uint64_t steps_hist[16];
} vx_b200_stats_t;

/* Statistics of the last completed run.  Work counters other than `rays` are gathered only after
 * vx_b200_enable_stats(dev, 1) (the counting kernels are slower; never enable it while timing). */
int vx_b200_enable_stats(vx_device_h hdevice, int enable);
int vx_b200_get_stats(vx_device_h hdevice, vx_b200_stats_t *stats);

/* Closest-hit records instead of colours.  `harguments` holds a kernel_arg_t exactly as for vx_start.
 *  - trace_primary: one ray per pixel (GenerateRay), hits[y*W+x]
 *  - trace_rays   : `count` caller-supplied rays (6 floats each: origin, direction) read from hrays
 * Each hit is the reference's 24-byte ray_hit_t {dist, bx, by, bz, blasIdx, triIdx}
 * (tests/regression/raycast/common.h:24-29); misses keep dist = 1e30.  Asynchronous like vx_start:
 * follow with vx_ready_wait. */
int vx_b200_trace_primary(vx_device_h hdevice, vx_buffer_h harguments, vx_buffer_h hhits);
int vx_b200_trace_rays(vx_device_h hdevice, vx_buffer_h harguments, vx_buffer_h hrays, uint64_t count, vx_buffer_h hhits);

/* The same records for rays and hits that live in HOST memory: equivalent to vx_copy_to_dev(rays) + vx_b200_trace_rays +
 * vx_ready_wait + vx_copy_from_dev(hits), but processed in chunks (VORTEX_B200_HOST_CHUNK rays, default 2^20) whose upload,
 * traversal and download legs overlap on separate streams, and without a device buffer for the whole batch.  Synchronous: the
 * hit records are in `host_hits` on return.  Page-locked host memory is needed for the legs to overlap (pageable memory
 * works, serialised by the driver's staging copies).  vx_b200_get_stats afterwards holds the sums over the chunks. */
int vx_b200_trace_rays_host(vx_device_h hdevice, vx_buffer_h harguments, const void *host_rays, uint64_t count, void *host_hits);

/* GPU BVH build: builds the layout-A BVH2 of `tri_count` triangles that already sit in `htri` (tri_t records
 * starting at triangle index tri_first) into `hbvh` (bvh_node_t records from node index node_first; 2*tri_count-1 are written)
 * and `htriIdx` (uint32 slots from slot_first; values are GLOBAL triangle indices tri_first+i in leaf order) — the byte layouts
 * of tests/regression/raycast/common.h:32-44 and scene.cpp:51, i.e. what BVH::build (raycast/bvh.cpp:23-66) leaves in those
 * buffers, so a host can point blas_node_t.bvh_offset at node_first and proceed exactly as after its own CPU build (or
 * download the buffers and hand them to any other Vortex backend).  Linear BVH (Morton sort + Karras radix tree), exact
 * vertex min/max boxes, one-triangle leaves.  Synchronous.  *nodes_used receives 2*tri_count-1. */
int vx_b200_build_bvh(vx_device_h hdevice, vx_buffer_h htri, uint64_t tri_first, uint64_t tri_count, vx_buffer_h hbvh,
                      uint64_t node_first, vx_buffer_h htriIdx, uint64_t slot_first, uint64_t *nodes_used);

/* vx_b200_build_bvh without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, same kernels and the same bytes in `hbvh` / `htriIdx` as vx_b200_build_bvh, but the builder's scratch arrays
 * (Morton keys, sort buffers, radix-tree links, boxes) live in a file-scope static __device__ array of 8 MiB that every GPU
 * which loads the library holds until exit, so tri_count is limited to 65 536 triangles: larger counts are rejected with an
 * error (vx_b200_last_error names the limit) and nothing is written.  The call itself performs no cudaMalloc / cudaFree /
 * event or stream creation and calls no allocating library routine.  All calls on one GPU share the array, which keeps the
 * previous call's contents (what must start at zero is cleared by the call): calls that overlap on the GPU — two host
 * threads building on the same GPU through different device handles — race; one handle serialises its calls.  Synchronous. */
int vx_b200_build_bvh_static(vx_device_h hdevice, vx_buffer_h htri, uint64_t tri_first, uint64_t tri_count, vx_buffer_h hbvh,
                             uint64_t node_first, vx_buffer_h htriIdx, uint64_t slot_first, uint64_t *nodes_used);

/* GPU BVH build for layout B (the RT unit's buffers): the same radix tree collapsed to a BVH4 and quantized into
 * bvh_quantized_node_t records (52 B, raytracing/common.h:58-67; what BVH::build + quantize leave in qBvh, raytracing/bvh.cpp:
 * 38-258) written to `hqbvh` from node index node_first; node indices inside the records are relative to node_first, which is
 * what the host stores in blas_node_t.bvh_offset.  Layout B has no triangle index array, so the tri_t records [tri_first,
 * tri_first+tri_count) of `htri` and the matching tri_ex_t records of `htriEx` (64 B each; may be NULL) are PERMUTED IN PLACE
 * into leaf order; a leaf's leftFirst is the global index tri_first + position.  One-triangle leaves; every child box is
 * verified to contain the child under the RT unit's own decode (origin + ldexp(q, e), rt_traversal.cpp:61-67).
 * Synchronous.  *nodes_used = nodes written (<= 2*tri_count; the call fails if `hqbvh` is smaller), root_box6 (may be NULL) =
 * the mesh's bounds {lo.xyz, hi.xyz} for the host's TLAS. */
int vx_b200_build_qbvh(vx_device_h hdevice, vx_buffer_h htri, uint64_t tri_first, uint64_t tri_count, vx_buffer_h htriEx,
                       vx_buffer_h hqbvh, uint64_t node_first, uint64_t *nodes_used, float *root_box6);

/* Human-readable description of the last error on this thread ("" if none). */
const char *vx_b200_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* VORTEX_B200_H */
