// TEMPORARY: placeholders until host_scene.cpp / host_field.cpp / host_exr.cpp land (same round).
#include <cstdio>
#include <cstdlib>
#include "../../include/vspg.h"
extern "C" {
static int nyi(char *err, size_t n) { if (err && n) snprintf(err, n, "not built yet"); return VSPG_ERR_UNSUPPORTED; }
int vspg_host_scene_parse(const char *, vspg_host_scene **, char *err, size_t n) { return nyi(err, n); }
int vspg_host_scene_parse_string(const char *, const char *, vspg_host_scene **, char *err, size_t n) { return nyi(err, n); }
void vspg_host_scene_free(vspg_host_scene *) {}
const vspg_scene_desc *vspg_host_scene_desc(const vspg_host_scene *) { return nullptr; }
const vspg_guiding_desc *vspg_host_scene_guiding(const vspg_host_scene *) { return nullptr; }
const vspg_integrator_params *vspg_host_scene_integrator(const vspg_host_scene *) { return nullptr; }
int vspg_host_scene_spp(const vspg_host_scene *) { return 0; }
int vspg_host_scene_seed(const vspg_host_scene *) { return 0; }
const char *vspg_host_scene_outfile(const vspg_host_scene *) { return ""; }
}
