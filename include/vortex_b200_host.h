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

#ifdef __cplusplus
}
#endif
#endif
