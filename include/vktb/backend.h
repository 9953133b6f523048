/* volkit-b200 -- the drop-in boundary.
 *
 * C-ABI of libvktb200.so: the B200 (sm_100a) backend for the voxel-parallel StructuredVolume core of
 * Kniggi/volkit.  Each entry point replaces one internal `*_cuda` backend function (or one Memory
 * primitive) of the reference; the reference-side shims that bind them are shown in INTEGRATION.md.
 * File:line citations are relative to the reference tree.
 *
 * Conventions
 *   - plain pointers and sizes only; every function returns 0 on success, non-zero on failure and
 *     leaves a per-thread message retrievable with vktb_last_error(); nothing throws.
 *   - `vktb_vol.data` is a DEVICE pointer obtained from vktb_alloc() (the algorithms never touch host
 *     memory and there is no CPU fallback: without a usable GPU every entry point fails loudly).
 *   - volumes are dense, x fastest: byte offset = ((z*dimX*dimY) + y*dimX + x) * bytesPerVoxel
 *     (src/vkt/StructuredVolume.cpp:311-317), little endian.
 *   - algorithms are asynchronous on the backend's per-device compute stream; results are visible to
 *     the host after the next device-to-host vktb_memcpy() (which orders itself after them) or
 *     vktb_sync() -- the same contract as the reference, where the D2H cudaMemcpy is the sync point
 *     (SURVEY.md 3.1).
 *   - ranges are half open [first,last); an empty range is a no-op.  As in the reference, dst bounds
 *     are the caller's responsibility (README.md:140).
 */
#ifndef VKTB_BACKEND_H
#define VKTB_BACKEND_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)   /* the library is built with -fvisibility=hidden */
#endif

#define VKTB_ABI_VERSION 1

/* numeric values of vkt::DataFormat / vktDataFormat (include/cpp/vkt/common.hpp:53-66) */
enum vktb_format {
    VKTB_UNSPECIFIED = 0, VKTB_INT8 = 1, VKTB_INT16 = 2, VKTB_INT32 = 3,
    VKTB_UINT8 = 4, VKTB_UINT16 = 5, VKTB_UINT32 = 6, VKTB_FLOAT32 = 7
};

/* vkt::CopyKind (include/cpp/vkt/Memory.hpp:15-21) */
enum vktb_copy_kind { VKTB_H2H = 0, VKTB_H2D = 1, VKTB_D2H = 2, VKTB_D2D = 3 };

/* the ten arithmetic ops in the order of include/cpp/vkt/Arithmetic.hpp */
enum vktb_arith_op {
    VKTB_SUM = 0, VKTB_DIFF, VKTB_PROD, VKTB_QUOT, VKTB_ABSDIFF,
    VKTB_SAFESUM, VKTB_SAFEDIFF, VKTB_SAFEPROD, VKTB_SAFEQUOT, VKTB_SAFEABSDIFF
};

/* vkt::FilterMode (include/cpp/vkt/Resample.hpp:14-18) plus the opt-in true trilinear filter */
enum vktb_filter {
    VKTB_NEAREST   = 0,
    VKTB_LINEAR    = 1,  /* reference-faithful: what StructuredVolumeView::sampleLinear computes */
    VKTB_TRILINEAR = 2   /* mathematically correct trilinear; NOT what the reference computes */
};

/* POD twin of StructuredVolumeView (src/vkt/StructuredVolumeView.hpp:19-33, 214-219) */
typedef struct vktb_vol {
    void*   data;      /* device pointer */
    int32_t dims[3];   /* x, y, z */
    int32_t format;    /* enum vktb_format */
    float   lo, hi;    /* voxel mapping */
} vktb_vol;

/* ---- library / device ------------------------------------------------------------------------- */
int         vktb_abi_version(void);
const char* vktb_last_error(void);                 /* per-thread, never NULL */
int         vktb_device_count(int* n);
int         vktb_set_device(int ordinal);          /* device used by the calling thread (default 0) */
int         vktb_get_device(int* ordinal);
/* Several GPUs of one box behind this same ABI, in ONE process.  vktb_set_devices names the
 * device group (all ids need peer access to each other: NVLink / NVSwitch); the calling thread's device becomes ids[0].
 * From then on vktb_alloc places an allocation of at least VKTB_SHARD_MIN_MB MiB (default 1024) as ONE address range
 * whose physical memory is striped over the group in equal contiguous parts (z is the slowest axis, so a part is a
 * z-slab), accessible from every device: the single data pointer of the reference's ManagedBuffer
 * (include/cpp/vkt/ManagedBuffer.hpp:168-198) spans the GPUs.  Then
 *   - vktb_fill_range, vktb_copy_range, vktb_arith_range, vktb_transform_range1/2 and vktb_resample(_window) whose
 *     operands are all striped run on every device at once, each on the dst planes that start in its stripe; a source
 *     plane or Resample tap in a neighbour's stripe is a peer load over NVLink inside the kernel (no halo copy);
 *   - vktb_memcpy to / from a striped range drives every device's copy engine on its own byte range concurrently --
 *     migrate() of a sharded volume;
 *   - vktb_histogram_range of a striped volume counts on every device (the planes that start in its stripe) and adds the
 *     partial bins on the calling device;
 *   - every other entry point runs on ids[0] and reaches the other stripes through peer access (correct, not scaled).
 * The caller sees one stream as before: the devices' work is forked from and joined to it by events.
 * vktb_set_devices(ids, 1) returns to one device.  Striped volumes must be freed before the group changes. */
int         vktb_set_devices(const int* ids, int n);
int         vktb_get_devices(int* ids, int capacity, int* n);
int         vktb_is_striped(const void* ptr, int* stripes);   /* 1 if ptr lies in a striped allocation */
/* number of kernels this library has launched so far (all threads); evidence for "gpu_launches" */
uint64_t    vktb_launch_count(void);
/* the compute stream (a cudaStream_t) of the calling thread's device; vktb_set_stream(NULL) restores
 * the backend-owned stream */
void*       vktb_get_stream(void);
int         vktb_set_stream(void* cuda_stream);
int         vktb_sync(void);                       /* waits for compute + copy streams of the device */

/* ---- memory: replaces Allocate_cuda / Free_cuda / cudaMemcpy / MemsetRange_cuda ----------------
 * src/vkt/Memory_cuda.hpp:16-32, src/vkt/Memory.cpp:40-74, src/vkt/Memory_cuda.cu:15-49 */
int vktb_alloc(void** ptr, size_t bytes);          /* pooled, stream-ordered device allocation */
int vktb_free(void* ptr);
/* what InputStream::read / OutputStream::write (src/vkt/InputStream.cpp:27-39,
 * OutputStream.cpp:25-36) bottom out in under the GPU policy, where the reference hands the DEVICE pointer of
 * volume.getData() to DataSource::read (fread into device memory).  The producer / consumer is called once per
 * 16 MiB chunk with a slot of the pinned staging ring, so a chunk is read from the file while the previous chunks'
 * DMAs run (stream_in), or written while the next chunks arrive (stream_out); no host copy of the volume exists.
 * *sum (may be NULL) receives the sum of the callback's return values: the caller compares it with `bytes` the way
 * the reference compares DataSource::read's return value.  Synchronous with respect to the callbacks; stream_in's
 * data is ordered before later work on the current stream. */
typedef size_t (*vktb_read_fn)(void* ctx, void* buf, size_t len);
typedef size_t (*vktb_write_fn)(void* ctx, const void* buf, size_t len);
int vktb_stream_in(void* dev_dst, size_t bytes, vktb_read_fn read, void* ctx, size_t* sum);
int vktb_stream_out(const void* dev_src, size_t bytes, vktb_write_fn write, void* ctx, size_t* sum);

/* Blocking contract of ManagedBuffer::migrate (include/cpp/vkt/ManagedBuffer.hpp:168-198): on return
 * the SOURCE may be freed or overwritten.  H2D from pageable memory is pipelined through pinned staging
 * (returns once the last chunk left the source; the DMA tail is ordered before later kernels);
 * H2D from pinned memory (vktb_host_alloc or foreign) is a single DMA the call waits for.  D2H is complete on
 * return. */
int vktb_memcpy(void* dst, const void* src, size_t bytes, int copy_kind);
int vktb_memset_range(void* dst, const void* pattern, size_t dst_bytes, size_t pattern_bytes);
/* pinned host memory from a reuse pool: what vkt::Allocate hands out under the CPU policy so that
 * migrate() becomes one full-speed DMA */
int vktb_host_alloc(void** ptr, size_t bytes);
int vktb_host_free(void* ptr);
int vktb_host_owns(const void* ptr);               /* 1 if ptr lies in a live vktb_host_alloc block, else 0 */
int vktb_mem_stats(size_t* device_bytes_in_use, size_t* device_bytes_reserved,
                   size_t* pinned_bytes_in_use, size_t* pinned_bytes_reserved);
int vktb_trim_pools(void);                         /* give cached device / pinned memory back */

/* ---- algorithms -------------------------------------------------------------------------------- */
/* FillRange_cuda(StructuredVolumeView, Vec3i, Vec3i, float)  src/vkt/Fill_cuda.hpp:13, Fill_cuda.cu:22-55;
 * semantics of FillRange_serial, src/vkt/Fill_serial.hpp:20-26 */
int vktb_fill_range(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], float value);

/* CopyRange_cuda  src/vkt/Copy_cuda.hpp:11-17; semantics of CopyRange_serial, src/vkt/Copy_serial.hpp:13-82:
 * dst[p - first + off] = src[clamp(p, 0, srcDims-1)], raw bytes if format and mapping agree, else by value */
int vktb_copy_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3],
                    const int32_t last[3], const int32_t dst_offset[3]);

/* BrickDecompose_cuda(Array3D<StructuredVolume>&, StructuredVolume&, Vec3i brickSize, Vec3i haloNeg, Vec3i haloPos)
 * (src/vkt/Decompose_cuda.hpp:16-22; semantics of BrickDecompose_serial, src/vkt/Decompose_serial.hpp:15-48: one
 * CopyRange(dest[index], source, first, last) per brick).  n CopyRange calls from ONE source in a single launch:
 * box i copies [firsts[3i..], lasts[3i..]) of src into dsts[i] at offs[3i..], src indices clamped as in
 * vktb_copy_range.  Boxes whose format or mapping differs from the source's are converted one launch each. */
int vktb_copy_range_batch(int n, const vktb_vol* dsts, const vktb_vol* src, const int32_t* firsts,
                          const int32_t* lasts, const int32_t* offs);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  vktb_copy_range_batch
 * with the same arguments, semantics and results, but nothing on its path allocates (no cudaMalloc / cudaFree /
 * stream-ordered allocation, no host heap): the box table, which vktb_copy_range_batch allocates per call, lives in
 * a file-scope device array of vktb_copy_range_batch_noalloc_max() = 4096 records (288 KiB, resident on every GPU
 * that loads the library until the process exits).  Limits: n <= 4096 (larger batches fail with a message; split
 * them or use vktb_copy_range_batch); every box must have the source's format and mapping (bytewise boxes only).
 * All calls share the array, which keeps the previous call's table: nothing in it needs clearing (a call overwrites
 * the records it uses, stream-ordered before its kernels), but calls whose kernels OVERLAP on the GPU -- different
 * streams or host threads -- race on it; issue them on one stream. */
int vktb_copy_range_batch_noalloc(int n, const vktb_vol* dsts, const vktb_vol* src, const int32_t* firsts,
                                  const int32_t* lasts, const int32_t* offs);
int vktb_copy_range_batch_noalloc_max(void);

/* {Sum..SafeAbsDiff}Range_cuda  src/vkt/Arithmetic_cuda.hpp:10-98; semantics of ArithmeticOp,
 * src/vkt/Arithmetic_serial.hpp:15-45: dst[p + off] = map(f(unmap(s1[p]), unmap(s2[p]))), p absolute */
int vktb_arith_range(int op, const vktb_vol* dst, const vktb_vol* s1, const vktb_vol* s2,
                     const int32_t first[3], const int32_t last[3], const int32_t dst_offset[3]);

/* TransformRange_cuda (unary / binary)  src/vkt/Transform_cuda.hpp:12-30 (empty stubs in the reference);
 * semantics of TransformRange_serial, src/vkt/Transform_serial.hpp:15-103.  A host function pointer cannot
 * run on the device: `op` is a device-op handle from the registry below. */
int vktb_transform_range1(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], int op);
int vktb_transform_range2(const vktb_vol* vol1, const vktb_vol* vol2, const int32_t first[3],
                          const int32_t last[3], int op);

/* Resample_cuda(StructuredVolume&, StructuredVolume&, FilterMode)  src/vkt/Resample_cuda.hpp:12-16;
 * semantics of Resample_serial, src/vkt/Resample_serial.hpp:26-71 (incl. the equal-dims conversion branch) */
int vktb_resample(const vktb_vol* dst, const vktb_vol* src, int filter);
/* z-slab form for sharded volumes: computes dst planes [dst_z0, dst_z0 + dst->dims[2]) of a volume of
 * dims dst_full from a src window holding planes [src_z0, src_z0 + src->dims[2]) of a volume of dims
 * src_full.  vktb_resample_window_src_range() tells which src planes a dst slab needs (incl. the +1 tap). */
int vktb_resample_window(const vktb_vol* dst, const vktb_vol* src, int filter,
                         int32_t dst_z0, const int32_t dst_full[3],
                         int32_t src_z0, const int32_t src_full[3]);
int vktb_resample_window_src_range(int filter, int32_t dst_z0, int32_t dst_z1, int32_t dst_dimz,
                                   int32_t src_dimz, int32_t* src_z0, int32_t* src_z1);
/* same with the full dims: drops the extra plane reference-faithful Linear requests for its unclamped +1 x tap
 * when the x/y maps show that tap can never leave its plane (e.g. exact 2:1 down-sampling) */
int vktb_resample_window_src_range3(int filter, int32_t dst_z0, int32_t dst_z1, const int32_t dst_full[3],
                                    const int32_t src_full[3], int32_t* src_z0, int32_t* src_z1);

/* ---- Flip and Histogram ----------------------------------------------------------
 * FlipRange_cuda(StructuredVolume& dest, StructuredVolume& source, Vec3i first, Vec3i last, Vec3i dstOffset, Axis)
 * (src/vkt/Flip_cuda.hpp:11-18); semantics of FlipRange_serial AS WRITTEN (src/vkt/Flip_serial.hpp:11-63): loops
 * over [0, last-first) halved on the axis, partner index last-1-i, z mirrored when axis == Y (so Y mirrors y and z,
 * Z mirrors nothing), raw bytes exchanged crosswise at + dst_offset.  axis 0/1/2 = vkt::Axis X/Y/Z.  dest may be
 * source (in place, dst_offset 0).  Both volumes must have the same bytes per voxel. */
int vktb_flip_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3], const int32_t last[3],
                    const int32_t dst_offset[3], int axis);

/* ScanRange_cuda(StructuredVolume& dst, StructuredVolume& src, Vec3i first, Vec3i last,
 * Vec3i dstOffset) (src/vkt/Scan_cuda.hpp:11-17); semantics of ScanRange_serial (src/vkt/Scan_serial.hpp:18-152): dst
 * receives CopyRange(dst, src, first, last, dst_offset) unless dst is src (same data pointer), then the 3-D
 * inclusive prefix sum of the raw voxel integers, modulo 2^(8 bytesPerVoxel), over [first, last) + dst_offset,
 * accumulating from `first` towards `last` on each axis (first > last = right to left; in place only). */
int vktb_scan_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3], const int32_t last[3],
                    const int32_t dst_offset[3]);
/* ComputeHistogramRange_cuda(StructuredVolume&, Histogram&, Vec3i first, Vec3i last)
 * (src/vkt/Histogram_cuda.hpp:12-17); semantics of ComputeHistogramRange_serial (src/vkt/Histogram_serial.hpp:18-51):
 * bins (DEVICE pointer, num_bins 64-bit counters = Histogram::getBinCounts() under the GPU policy) are zeroed, then
 * bin (size_t)((value - lo) * (num_bins / (hi - lo))) is incremented per voxel, binary32 arithmetic.  Voxels whose
 * index falls outside [0, num_bins) -- where the reference writes out of bounds -- are dropped; their number is
 * returned through dropped_host when it is not NULL (this makes the call synchronous). */
int vktb_histogram_range(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], uint64_t* bins,
                         size_t num_bins, uint64_t* dropped_host);

/* ScaleRange_cuda / RotateRange_cuda (src/vkt/Scale_cuda.hpp, Rotate_cuda.hpp: empty bodies in
 * the reference); semantics of ScaleRange_serial (src/vkt/Scale_serial.hpp:15-56) and RotateRange_serial
 * (src/vkt/Rotate_serial.hpp:41-94), which differ only in the matrix: every voxel p of the WHOLE dst volume is mapped
 * to q = M (p - center) + center in binary32; when first <= q < last on all axes the voxel receives
 * sampleLinear((int)q.x, (int)q.y, (int)q.z) of src (integer parameters: the reference's 8-tap expression with
 * frac 0, clamps as written), otherwise it keeps its value.  matrix = col0.xyz, col1.xyz, col2.xyz (vkt::Mat3f);
 * Scale passes diag(1/s), Rotate the inverse rotation built from axis and angle on the host.  dst may be src. */
int vktb_affine_range(const vktb_vol* dst, const vktb_vol* src, const int32_t first[3], const int32_t last[3],
                      const float matrix[9], const float center[3]);

/* ApplyFilterRange_cuda(StructuredVolume& dest, StructuredVolume& source, Vec3i first, Vec3i last, Filter, AddressMode)
 * (src/vkt/Filter_cuda.hpp:12-21, an empty body in the reference); semantics of ApplyFilterRange_serial
 * (src/vkt/Filter_serial.hpp:15-64) over [0, src dims) as vkt::ApplyFilter calls it (src/vkt/Filter.cpp:48-66): border
 * voxels (a coordinate equal to 0 or dims - 1) receive map(0), all others the binary32 sum of weight * voxel over
 * the filter window in z, y, x order (multiply and add rounded separately).  weights: filter_dims[0] *
 * filter_dims[1] * filter_dims[2] HOST floats, x fastest (vkt::Filter's data); each filter dim must be 1, 2 or 3
 * (the reference reads outside the volume with larger ones).  address_mode is accepted and ignored, as in the
 * reference.  dst may be src (the reference filters a copy of the source). */
int vktb_apply_filter(const vktb_vol* dst, const vktb_vol* src, const float* weights, const int32_t filter_dims[3],
                      int address_mode);

/* ComputeAggregatesRange_cuda(StructuredVolume&, Aggregates&, Vec3i first, Vec3i last)
 * (src/vkt/Aggregates_cuda.hpp:14-21, an empty body in the reference); semantics of ComputeAggregatesRange_serial
 * (src/vkt/Aggregates_serial.hpp:20-83).  vktb_aggregates has the layout of vkt::Aggregates / vktAggregates_t
 * (include/cpp/vkt/Aggregates.hpp:14-25).  Contract:
 *   min, max, argmin, argmax   exact: strict compares, the first voxel in x-fastest order wins, NaN never enters,
 *                              FLT_MAX / -FLT_MAX and argmin = argmax = (0,0,0) when no voxel qualifies;
 *   sum, mean, var, stddev, prod   the reference accumulates these sequentially in binary32, which no parallel
 *                              evaluation reproduces bit for bit.  Returned here: the real-number moments to about
 *                              1e-7 relative (binary64 accumulation), pushed through the reference's final steps
 *                              (mean = float(double(sum) / numElems) with numElems = the WHOLE volume's voxel
 *                              count, var = sum (v - mean)^2 / numElems, stddev = sqrtf(var)).  The reference's
 *                              own result lies within its a-priori bound n * 2^-24 * sum |terms| of that value;
 *                              prod additionally within FLT_MIN where the running binary32 product goes subnormal.
 * `out` is a HOST pointer; the call is synchronous.  An empty range gives the reference's initial values. */
typedef struct vktb_aggregates
{
    float   min, max, mean, stddev, var, sum, prod;
    int32_t argmin[3], argmax[3];
} vktb_aggregates;
int vktb_aggregates_range(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], vktb_aggregates* out);

/* The same pass as combinable moments, for volumes sharded into z-slabs (one partial per rank, exchanged as bytes):
 * sums of d = v - shift and d^2, prod = prod_mant * 2^prod_exp, extrema with their absolute voxel coordinates in
 * `vol` (has_min / has_max = 0 when no voxel qualified).  vktb_aggregates_finish folds n partials IN ORDER (the
 * earlier partial wins ties, so pass them in z order after moving argmin/argmax to whole-volume coordinates) and
 * applies the final steps above with num_elems = the whole volume's voxel count.  Pure host arithmetic. */
typedef struct vktb_agg_partial
{
    double   sum_d, sum_d2, prod_mant;
    int64_t  prod_exp;
    uint64_t count;
    float    min, max, shift;
    int32_t  has_min, has_max;
    int32_t  argmin[3], argmax[3];
    int32_t  pad;
} vktb_agg_partial;
int vktb_aggregates_partial(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], vktb_agg_partial* out);
int vktb_aggregates_finish(const vktb_agg_partial* parts, int n, uint64_t num_elems, vktb_aggregates* out);

/* ---- SURVEY 8(f) rank 4, last row: headless ray-marcher ----------------------------------------------------------
 * The arithmetic of the reference's RayMarchingKernel (src/vkt/Render_kernel.hpp:95-162: absorption + emission,
 * opacity correction 1 - (1 - a)^dt, front-to-back compositing, optional sRGB) as the reference's host path sets it up
 * (src/vkt/Render.cpp:434-489: Nearest-filtered, Clamp-addressed volume over normalised coordinates, RGBA look-up
 * table as a Nearest 1-D texture, bbox = [0, dims * dist]) -- written into a DEVICE buffer.  The reference's own
 * output path is a GLUT / GL window with progressive jittered accumulation (unavailable on this box); the headless
 * form renders ONE frame of pixel-centre rays of visionaray's pinhole camera (look_at(eye, center, up),
 * perspective(fovy), primary_ray(x, y)), so an image is reproducible.
 *   vol          UInt8 / UInt16 / Int16 / UInt32 (voxels are normalised by the texel type's numeric limits)
 *   dist         voxel spacing (NULL = 1, 1, 1)
 *   lut_dev      lut_n RGBA float quadruples in DEVICE memory, or NULL / 0: colour = (v, v, v, v)
 *   dt           ray step (RenderState::dtRayMarching, default 1); srgb = RenderState::sRGB (default 1)
 *   rgba8_dev    width * height * 4 bytes, row 0 = bottom row (as the reference's render target), visionaray's
 *                float -> unorm<8> rule (truncation); rgba32f_dev: the float image; either may be NULL
 * Asynchronous on the compute stream like every algorithm. */
typedef struct vktb_camera
{
    float eye[3], center[3], up[3];
    float fovy_degrees;
} vktb_camera;
int vktb_render_raymarch(const vktb_vol* vol, const float dist[3], const vktb_camera* cam, const float* lut_dev, int lut_n,
                         float dt, int srgb, int width, int height, uint8_t* rgba8_dev, float* rgba32f_dev);
/* the viewer's default camera: 45 degrees, pinhole_camera::view_all(bbox) (src/vkt/Render.cpp:856-865); host math */
int vktb_render_default_camera(const int32_t dims[3], const float dist[3], vktb_camera* cam);

/* ---- Host-resident volumes, streamed ---------------------------------------------------------------
 * Same semantics and results as the entry points above, but every vktb_vol.data is a HOST pointer (pinned:
 * vktb_host_alloc or cudaHostRegister; pageable memory works without overlap).  The range is cut into z-slabs and
 * upload, kernel and download of consecutive slabs overlap on two copy engines (PCIe is full duplex), instead of
 * the three serial whole-buffer phases of ManagedBuffer::migrate() (include/cpp/vkt/ManagedBuffer.hpp:168-198)
 * -> kernel -> migrate() back.  Volumes larger than device memory are fine; the volumes stay on the host and
 * the result is complete in host memory on return.  The range must lie inside all volumes; dst may be one of
 * the sources (in place, offset 0) but must not overlap them otherwise. */
int vktb_hosted_arith_range(int op, const vktb_vol* dst, const vktb_vol* src1, const vktb_vol* src2,
                            const int32_t first[3], const int32_t last[3], const int32_t dst_offset[3]);
int vktb_hosted_transform_range1(const vktb_vol* vol, const int32_t first[3], const int32_t last[3], int op);
int vktb_hosted_resample(const vktb_vol* dst, const vktb_vol* src, int filter);

/* ---- Transform device-op registry --------------------------------------------------------------
 * Built-in ops are addressed by name; a host TransformUnaryOp / TransformBinaryOp pointer is bound to a
 * device op once, after which the unchanged vkt::Transform(volume, hostFn) signature resolves it.
 * Unbound host pointers make Transform fail with InvalidValue under the GPU policy (no CPU fallback). */
int vktb_unary_op_by_name(const char* name, int* op);     /* "xor_coord", "invert", "checker8", "threshold" */
int vktb_binary_op_by_name(const char* name, int* op);    /* "or_assign", "swap" */
/* A user's OWN op (the reference's contract is an arbitrary TransformUnaryOp / TransformBinaryOp,
 * include/cpp/vkt/Transform.hpp:16-25): written once with VKTB_UNARY_OP / VKTB_BINARY_OP of vktb/transform_op.cuh, it
 * compiles for the host into the plain function and for sm_100a into the kernel vktb_unary_<name> /
 * vktb_binary_<name> that carries it inlined.  `image` is that cubin / PTX / fatbin (nvcc -cubin -gencode
 * arch=compute_100a,code=sm_100a --fmad=false); `_source` takes the CUDA C text and compiles it with NVRTC when
 * libnvrtc can be loaded (vktb_nvrtc_available()).  The returned handle is used like a built-in one: bind the host
 * function to it and the unchanged vkt::Transform(volume, hostFn) runs on the GPU. */
int vktb_register_unary_op_image(const void* image, size_t size, const char* name, int* op);
int vktb_register_binary_op_image(const void* image, size_t size, const char* name, int* op);
int vktb_register_unary_op_source(const char* cuda_source, const char* name, int* op);
int vktb_register_binary_op_source(const char* cuda_source, const char* name, int* op);
int vktb_nvrtc_available(void);
int vktb_bind_unary_op(const void* host_fn, int op);
int vktb_bind_binary_op(const void* host_fn, int op);
int vktb_lookup_unary_op(const void* host_fn, int* op);   /* non-zero if unbound */
int vktb_lookup_binary_op(const void* host_fn, int* op);

/* ---- scalar codec on the device (1 thread): used by tests to pin the float/convert emulation ----- */
/* whether map_dst's division by (hi - lo) runs in the reciprocal-and-two-FMA form for this mapping: 1 after the
 * exhaustive proof against __fdiv_rn (all 2^32 dividends, once per divisor and process) succeeded, else 0 */
int vktb_debug_div_proven(float lo, float hi, int* proven);
int vktb_debug_map_unmap(int format, float lo, float hi, const float* values_dev, uint32_t* mapped_dev,
                         float* unmapped_dev, size_t n);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* VKTB_BACKEND_H */
