/* TEST INFRASTRUCTURE ONLY -- see vkt_oracle.c.  Plain-C restatement of the reference's serial
 * CPU path for the StructuredVolume hot path.  Never linked into or imported by the product. */
#ifndef VKT_ORACLE_H
#define VKT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* numeric values of vkt::DataFormat (reference include/cpp/vkt/common.hpp:53-66) */
enum {
    VKO_UNSPECIFIED = 0, VKO_INT8 = 1, VKO_INT16 = 2, VKO_INT32 = 3,
    VKO_UINT8 = 4, VKO_UINT16 = 5, VKO_UINT32 = 6, VKO_FLOAT32 = 7
};

/* arithmetic op ids, in the order of reference include/cpp/vkt/Arithmetic.hpp */
enum {
    VKO_SUM = 0, VKO_DIFF, VKO_PROD, VKO_QUOT, VKO_ABSDIFF,
    VKO_SAFESUM, VKO_SAFEDIFF, VKO_SAFEPROD, VKO_SAFEQUOT, VKO_SAFEABSDIFF
};

/* same field order as vktb_vol (include/vktb/backend.h) and RefVol (oracle/ref_shim.cpp) */
typedef struct vko_vol {
    void*   data;
    int32_t dims[3];
    int32_t format;
    float   lo, hi;
} vko_vol;

int   vko_abi_version(void);
int   vko_bytes_per_voxel(int format);
void  vko_map_voxel(uint8_t* dst, float value, int format, float lo, float hi);
float vko_unmap_voxel(const uint8_t* src, int format, float lo, float hi);

int vko_fill_range(const vko_vol* vol, const int32_t first[3], const int32_t last[3], float value);
int vko_copy_range(const vko_vol* dst, const vko_vol* src, const int32_t first[3],
                   const int32_t last[3], const int32_t off[3]);
int vko_arith_range(int op, const vko_vol* dst, const vko_vol* s1, const vko_vol* s2,
                    const int32_t first[3], const int32_t last[3], const int32_t off[3]);
int vko_transform_range1(const vko_vol* vol, const int32_t first[3], const int32_t last[3], int op);
int vko_transform_range2(const vko_vol* v1, const vko_vol* v2, const int32_t first[3],
                         const int32_t last[3], int op);
/* filter 0 = Nearest, 1 = Linear (reference-faithful, "mode R"), 2 = true trilinear ("mode T",
 * a restatement of sampleLinear with float parameters and the six intended clamps) */
int vko_resample(const vko_vol* dst, const vko_vol* src, int filter);
/* windowed form used to check z-slab sharding and full-size volumes plane by plane: computes dst
 * planes [dz0, dz1) of a (virtual) dst volume of dims dst_dims_full from a src window that starts at
 * global src plane sz0 of a (virtual) src volume of dims src_dims_full; dst->data / src->data point
 * at the first plane of the window and dst->dims[2] / src->dims[2] hold the window heights. */
int vko_resample_window(const vko_vol* dst, const vko_vol* src, int filter,
                        int32_t dz0, const int32_t dst_dims_full[3],
                        int32_t sz0, const int32_t src_dims_full[3]);

/* z-window forms for plane-by-plane checks of full-size volumes (bench.py): the same loop bodies on windows of
 * planes, absolute coordinates preserved; each is tested against its whole-volume twin (tests/test_oracle_windows.py) */
int vko_transform_range1_window(const vko_vol* vol, const int32_t first[3], const int32_t last[3], int op,
                                int32_t z_origin);
int vko_affine_window(const vko_vol* dst, int32_t dz0, const vko_vol* src, int32_t sz0, const int32_t src_full[3],
                      const int32_t first[3], const int32_t last[3], const float m[9], const float c[3]);
int vko_apply_filter_window(const vko_vol* dst, int32_t dz0, const vko_vol* src, int32_t sz0, const int32_t full[3],
                            const float* weights, const int32_t fdims[3]);

/* Transform with a caller-supplied op (layout of vkt::VoxelView, include/cpp/vkt/Voxel.hpp:15-21) */
typedef struct vko_voxel_view { uint8_t* bytes; int32_t dataFormat; float mappingLo, mappingHi; } vko_voxel_view;
typedef void (*vko_unary_fn)(int32_t x, int32_t y, int32_t z, vko_voxel_view voxel);
int vko_transform_fn1(const vko_vol* vol, const int32_t first[3], const int32_t last[3], vko_unary_fn fn, int32_t z_origin);

/* axis 0/1/2 = vkt::Axis X/Y/Z; restated as written, see vkt_oracle.c */
int vko_flip_range(const vko_vol* dst, const vko_vol* src, const int32_t first[3], const int32_t last[3],
                   const int32_t off[3], int axis);
/* bins[num_bins] (zeroed first); voxels whose bin index falls outside [0, num_bins) are counted in *dropped */
int vko_histogram_range(const vko_vol* vol, const int32_t first[3], const int32_t last[3], uint64_t* bins,
                        size_t num_bins, uint64_t* dropped);

/* Scale / Rotate (range = [0, source dims)) and the shared body with an explicit matrix (col0, col1, col2) */
void vko_scale_matrix(const float factor[3], float m[9]);
void vko_rotate_matrix(const float axis[3], float angle, float m[9]);
int vko_affine_range(const vko_vol* dst, const vko_vol* src, const int32_t first[3], const int32_t last[3],
                     const float m[9], const float center[3]);
int vko_scale(const vko_vol* dst, const vko_vol* src, const float factor[3], const float center[3]);
int vko_rotate(const vko_vol* dst, const vko_vol* src, const float axis[3], float angle, const float center[3]);
/* ApplyFilter over [0, source dims); weights x fastest, filter dims 1..3 */
int vko_apply_filter(const vko_vol* dst, const vko_vol* src, const float* weights, const int32_t fdims[3]);

/* OutputStream::writeRange / InputStream::readRange as byte streams, RawFile's file-name parser */
size_t vko_rows_gather(const vko_vol* vol, const int32_t first[3], const int32_t last[3], uint8_t* bytes);
size_t vko_rows_scatter(const vko_vol* vol, const int32_t first[3], const int32_t last[3], const uint8_t* bytes);
void vko_rawfile_parse(const char* name, int32_t dims[3], int32_t* format);

/* vkt::Aggregates (reference include/cpp/vkt/Aggregates.hpp:14-25) */
typedef struct vko_aggregates
{
    float   min, max, mean, stddev, var, sum, prod;
    int32_t argmin[3], argmax[3];
} vko_aggregates;
int vko_aggregates_range(const vko_vol* vol, const int32_t first[3], const int32_t last[3], vko_aggregates* out);

/* 3-D prefix sum of the raw voxel integers mod 2^(8 bpv); dst->data == src->data means in place */
int vko_scan_range(const vko_vol* dst, const vko_vol* src, const int32_t first[3], const int32_t last[3],
                   const int32_t off[3]);

/* BrickDecomposeResize's layout (num_bricks[3]; brick_dims = 3 ints per brick in x-fastest brick order, may be
 * NULL) -> total voxels of all bricks; vko_brick_decompose writes the bricks back to back into out */
size_t vko_brick_layout(const int32_t src_dims[3], const int32_t brick[3], const int32_t halo_neg[3],
                        const int32_t halo_pos[3], int32_t num_bricks[3], int32_t* brick_dims);
int vko_brick_decompose(const vko_vol* src, const int32_t brick[3], const int32_t halo_neg[3],
                        const int32_t halo_pos[3], uint8_t* out, size_t out_bytes);

/* Headless ray-marcher (vkt_render_oracle.c): absorption + emission along pixel-centre rays of a pinhole camera,
 * Nearest-filtered volume and RGBA look-up table (lut: lut_n x 4 floats, or NULL = grey ramp), one frame.  Output
 * rgba32f (w * h * 4 floats, row 0 = bottom row as in the reference's render target) and optionally rgba8. */
int vko_render_supported_format(int format);
int vko_render_raymarch(const vko_vol* vol, const float* lut, int lut_n, const float eye[3], const float center[3],
                        const float up[3], float fovy_degrees, int w, int h, float dt, int srgb, const float dist[3],
                        float* rgba32f, uint8_t* rgba8);
void vko_render_default_camera(const int32_t dims[3], const float dist[3], float eye[3], float center[3], float up[3]);

/* synthetic inputs of SURVEY.md 8(d): element i of a volume with seed s gets the low bits of
 * splitmix64(i ^ s*0x9E3779B97F4A7C15); Float32 gets (h>>40)*2^-24 in [0,1). */
void vko_synth_fill(void* data, size_t first_voxel, size_t n_voxels, int format, uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif
