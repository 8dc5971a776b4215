// ref_assets.cpp — TEST INFRASTRUCTURE (oracle): calls the reference's own asset loaders so that the product's OBJ / PNG
// path (vortex-raytracing_b200/host/asset_io.cpp) can be compared with them array for array.
// Compiled by oracle/Makefile together with tests/regression/raytracing/mesh.cpp and surface.cpp taken where they lie
// under /root/reference (with their vendored tiny_obj_loader.h / stb_image.h); nothing is copied into this repository.
//   Mesh::Mesh(objFile)     tests/regression/raytracing/mesh.cpp:215-293
//   Surface::Surface(file)  tests/regression/raytracing/surface.cpp:18-55
#include <cstdint>
#include <cstring>
#include "mesh.h"

#define REFA_EXPORT extern "C" __attribute__((visibility("default")))

REFA_EXPORT void *refa_mesh_load(const char *obj_path) { return new Mesh(obj_path); }   // aborts on failure, like the app
REFA_EXPORT void refa_mesh_free(void *h) { delete (Mesh *)h; }
REFA_EXPORT uint32_t refa_mesh_tri_count(void *h) { return (uint32_t)((Mesh *)h)->tri().size(); }
REFA_EXPORT uint32_t refa_mesh_mat_count(void *h) { return (uint32_t)((Mesh *)h)->materials().size(); }
REFA_EXPORT uint32_t refa_mesh_tex_count(void *h) { return (uint32_t)((Mesh *)h)->textures().size(); }
REFA_EXPORT uint32_t refa_sizes(uint32_t which) { return which == 0 ? sizeof(tri_t) : which == 1 ? sizeof(tri_ex_t) : sizeof(material_info_t); }
REFA_EXPORT void refa_mesh_copy(void *h, void *tri, void *triex, void *mats) {
  Mesh *m = (Mesh *)h;
  memcpy(tri, m->tri().data(), m->tri().size() * sizeof(tri_t));
  memcpy(triex, m->triEx().data(), m->triEx().size() * sizeof(tri_ex_t));
  memcpy(mats, m->materials().data(), m->materials().size() * sizeof(material_info_t));
}
REFA_EXPORT void refa_mesh_tex_dims(void *h, uint32_t i, uint32_t *w, uint32_t *hgt) {
  Surface *s = ((Mesh *)h)->textures()[i];
  *w = s->width(), *hgt = s->height();
}
REFA_EXPORT void refa_mesh_tex_copy(void *h, uint32_t i, void *dst) {
  Surface *s = ((Mesh *)h)->textures()[i];
  memcpy(dst, s->pixels(), (size_t)s->width() * s->height() * 4);
}
REFA_EXPORT void *refa_png_load(const char *path, uint32_t *w, uint32_t *h) {
  Surface *s = new Surface(path);
  *w = s->width(), *h = s->height();
  return s;
}
REFA_EXPORT void refa_png_copy(void *surf, void *dst) {
  Surface *s = (Surface *)surf;
  memcpy(dst, s->pixels(), (size_t)s->width() * s->height() * 4);
}
REFA_EXPORT void refa_png_free(void *surf) { delete (Surface *)surf; }
