/* vortex_b200_host.h — C ABI of libvortex-b200-host.so: the host-side scene preparation of the ray-tracing path.
 *
 * A Vortex raycast host builds its acceleration structures on the CPU and uploads them through vx_copy_to_dev
 * (reference: tests/regression/raycast/scene.cpp:17-82 Scene::init, bvh.cpp:23-66 BVH build, bvh.cpp:181-212 TLAS build,
 * tracer.cpp:152-196 Tracer::setup).  This library is that host step for applications that switch to the b200 backend:
 * it produces the SAME byte layouts (tri_t, tri_ex_t, triIdx, bvh_node_t, blas_node_t, tlas_node_t, texels,
 * kernel_arg_t — raycast/common.h:19-147) from caller-supplied triangle arrays, deterministically, so that the buffers can
 * be consumed unchanged by the b200 kernels, by simx, or by the reference's CPU path.  Plain C: pointers and sizes only.
 * No CUDA dependency (the device side re-flattens triangles itself at vx_start).
 */
#ifndef VORTEX_B200_HOST_H
#define VORTEX_B200_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef void *vxh_scene_h;

typedef struct {
  const void *tri;    uint64_t tri_bytes;     /* tri_t[]      36 B each  (geometry.h:1401-1405) */
  const void *triEx;  uint64_t triEx_bytes;   /* tri_ex_t[]   60 B each  (raycast/common.h:19-22) */
  const void *triIdx; uint64_t triIdx_bytes;  /* uint32_t[]   leaf order (global triangle indices, scene.cpp:51) */
  const void *bvh;    uint64_t bvh_bytes;     /* bvh_node_t[] 2*N allocated as the reference does (scene.cpp:32) */
  const void *blas;   uint64_t blas_bytes;    /* blas_node_t[] 160 B each */
  const void *tlas;   uint64_t tlas_bytes;    /* tlas_node_t[] 2*M-1 */
  const void *tex;    uint64_t tex_bytes;     /* 0x00RRGGBB texels of all meshes, concatenated */
  uint32_t tlas_root;
  uint32_t bvh_nodes_used;
} vxh_scene_view_t;

/* Scene::init for caller-supplied meshes: concatenates the meshes, builds one binned-SAH BVH2 per mesh.
 * tris: 9 floats per triangle; triex: 15 floats per triangle (N0 N1 N2 uv0 uv1 uv2); texels: tex_w[i]*tex_h[i] uint32 per mesh;
 * transforms: optional 16 floats per mesh (row-major instance transform), NULL = identity.  threads <= 0: all cores. */
vxh_scene_h vxh_scene_create(uint32_t n_meshes, const uint32_t *tri_counts, const float *tris, const float *triex,
                             const uint32_t *tex_w, const uint32_t *tex_h, const uint32_t *texels, const float *reflectivity,
                             const float *transforms, int threads);
void vxh_scene_apply_transform(vxh_scene_h scene, const float *m16);   /* Scene::applyTransform (scene.cpp:216-220) */
void vxh_scene_build(vxh_scene_h scene);                               /* Scene::build -> TLAS::build (bvh.cpp:181-212) */
void vxh_scene_view(vxh_scene_h scene, vxh_scene_view_t *view);
void vxh_scene_free(vxh_scene_h scene);

/* Tracer::setup (tracer.cpp:152-196): optional scene rotation, TLAS build, framing camera, viewplane, lights.
 * Writes the 192-byte kernel_arg_t with every address field zero (the caller fills them from vx_mem_address). */
void vxh_setup(vxh_scene_h scene, void *kernel_arg_192, uint32_t width, uint32_t height, uint32_t spp, uint32_t max_depth,
               float camera_vfov_deg, float zoom, int rotate_scene, const float *light_pos, const float *light_color,
               const float *ambient, const float *background);


/* ---- layout B: the RT-unit ("raytracing") host path -------------------------------------------------------------------
 * Scene::init + BVH (BVH4 build + quantize) + TLAS (build + quantize) + Tracer::setup of the reference raytracing host
 * (tests/regression/raytracing/scene.cpp:17-119, bvh.cpp:12-260,262-623, tracer.cpp:195-235), emitting its byte layouts
 * (raytracing/common.h:20-99,164-195): triangles physically in leaf order, 52-byte quantized BVH4 nodes, quantized TLAS
 * (node 0 = root), blas_node_t, material_info_t, kernel_arg_t (216 B).  Deterministic (the reference builder is not),
 * conservative quantization verified against the RT unit's decode (sim/simx/rt_traversal.cpp:61-67). */
typedef void *vxh_rtx_scene_h;

typedef struct {
  const void *tri;    uint64_t tri_bytes;     /* tri_t[] 36 B, reordered into leaf order */
  const void *triEx;  uint64_t triEx_bytes;   /* tri_ex_t[] 64 B (texId = global material index), same order */
  const void *triIdx; uint64_t triIdx_bytes;  /* uint32_t[] identity (unused by the RT unit) */
  const void *mat;    uint64_t mat_bytes;     /* material_info_t[] 88 B */
  const void *tex;    uint64_t tex_bytes;
  const void *bvh;    uint64_t bvh_bytes;     /* unquantized helper tree (uploaded by the reference host, unused by the RT unit) */
  const void *qbvh;   uint64_t qbvh_bytes;    /* bvh_quantized_node_t[] 52 B, 2*N allocated */
  const void *blas;   uint64_t blas_bytes;    /* blas_node_t[] 160 B */
  const void *tlas;   uint64_t tlas_bytes;    /* QUANTIZED TLAS nodes, 52 B */
  uint32_t tlas_root, bvh_nodes_used, tlas_nodes_used, _pad;
} vxh_rtx_scene_view_t;

/* triex16: 16 words per triangle (15 floats N0 N1 N2 uv0 uv1 uv2 + uint32 texId local to its mesh); mats: n_mats[i]
 * material_info_t records per mesh (tex_offset = byte offset into `tex`); reflectivity/transforms: per mesh or NULL. */
vxh_rtx_scene_h vxh_rtx_scene_create(uint32_t n_meshes, const uint32_t *tri_counts, const float *tris, const uint32_t *triex16,
                                     const uint32_t *n_mats, const void *mats, const uint8_t *tex, uint64_t tex_bytes,
                                     const float *reflectivity, const float *transforms, int threads);
void vxh_rtx_scene_apply_transform(vxh_rtx_scene_h scene, const float *m16);
void vxh_rtx_scene_build(vxh_rtx_scene_h scene);
void vxh_rtx_scene_view(vxh_rtx_scene_h scene, vxh_rtx_scene_view_t *view);
void vxh_rtx_scene_free(vxh_rtx_scene_h scene);
void vxh_rtx_setup(vxh_rtx_scene_h scene, void *kernel_arg_216, uint32_t width, uint32_t height, uint32_t spp, uint32_t max_depth,
                   float camera_vfov_deg, float zoom, const float *light_pos, const float *light_color, const float *ambient,
                   const float *background);

/* ---- asset path: what the raytracing host reads and writes around the scene builder ----
 * vxh_obj_load: Wavefront OBJ (+ its mtllib and the PNG textures its map_Kd statements name, looked up next to the .obj)
 *   -> one mesh in exactly the arrays vxh_rtx_scene_create takes.  Mirrors Mesh::Mesh(objFile)
 *   (tests/regression/raytracing/mesh.cpp:130-293): faces in file order, un-indexed vertices, quads split along the shorter
 *   diagonal, larger polygons ear-clipped, texId = the face's material id (0 without one), missing normals / UVs = 0,
 *   one texture per distinct map_Kd name; material_info_t.diffuse_tex_id / tex_width / tex_height as the reference sets them,
 *   tex_offset = byte offset of the material's texture inside `tex`, reflectivity and illum-independent fields zero.
 * vxh_png_load: 8-bit RGB texels packed 0x00RRGGBB like Surface::loadImage (surface.cpp:40-52); non-interlaced PNG of any
 *   colour type / bit depth; alpha dropped.  *texels is malloc'ed: release with vxh_free.
 * vxh_ppm_write: format 3 = the reference's text writer byte for byte when flip_vertical = 1 (tracer.cpp:15-33); format 6 =
 *   binary PPM (the same picture, ~4x smaller and without the formatting cost that dominates the reference's wall time).
 * Errors: NULL / non-zero, text in vxh_asset_last_error(). */
typedef void *vxh_asset_mesh_h;
typedef struct {
  const float *tris;        /* 9 floats per triangle */
  const uint32_t *triex16;  /* 16 words per triangle: N0 N1 N2 uv0 uv1 uv2 (floats) + texId */
  uint32_t tri_count;
  uint32_t n_mats;
  const void *mats;         /* material_info_t[n_mats], 88 B each */
  const uint8_t *tex;       /* the mesh's textures, concatenated, 4 B per texel */
  uint64_t tex_bytes;
  uint32_t n_textures, _pad;
} vxh_asset_mesh_view_t;
vxh_asset_mesh_h vxh_obj_load(const char *obj_path);
void vxh_asset_mesh_view(vxh_asset_mesh_h mesh, vxh_asset_mesh_view_t *view);
const char *vxh_asset_mesh_texture_name(vxh_asset_mesh_h mesh, uint32_t index);
void vxh_asset_mesh_free(vxh_asset_mesh_h mesh);
int vxh_png_load(const char *path, uint32_t *width, uint32_t *height, uint32_t **texels);
void vxh_free(void *p);
int vxh_ppm_write(const char *path, const uint32_t *pixels, uint32_t width, uint32_t height, int format, int flip_vertical);
const char *vxh_asset_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
