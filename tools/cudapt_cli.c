/* cudapt_cli -- headless front door of the CUDA backend (the reference only has the SDL/ImGui
 * viewer, src/App.cpp, and a CLI sketch in architecture-redesign-gameplan.md:762-797).  Plain C over include/cudapt.h:
 * also the proof that the header is a C header.
 *
 *   cudapt_cli [--scene app | --obj FILE] [--size WxH] [--spp N] [--bounces N] [--philox] [--jitter]
 *              [--exposure X] [--auto-exposure] [--tonemap none|reinhard|aces] [--srgb] [--device D]
 *              [--look-at px,py,pz,tx,ty,tz,vfov] [--lens RADIUS] [--env FILE.hdr|.pfm [--env-size N]]
 *              --out FILE.png|.ppm|.pfm|.exr
 *
 * --scene app : the 38 spheres of App.cpp:101-122 with the reference integrator (default).
 * --obj FILE  : triangles of a Wavefront OBJ ("v" and "f" records; polygons are fanned), one grey Lambert material,
 *               mesh (wavefront) integrator; implies --philox.
 * --env FILE  : latitude-longitude HDR panorama instead of the procedural sky (cudapt_load_environment).
 * Build: cc -std=c99 -O2 -Iinclude -o build/cudapt_cli tools/cudapt_cli.c -Lsoftware-path-tracer_b200 -lcudapt \
 *           -Wl,-rpath,'$ORIGIN/../software-path-tracer_b200'
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "cudapt.h"

static void die(cudapt_ctx* ctx, const char* what) {
  fprintf(stderr, "cudapt_cli: %s: %s\n", what, cudapt_last_error(ctx));
  exit(1);
}
#define CHECK(call) \
  do {              \
    if ((call) != CUDAPT_OK) die(ctx, #call); \
  } while (0)

/* App.cpp:101-122 */
static uint32_t app_spheres(float* out /* 38 x 4 */) {
  uint32_t n = 0;
  int x, y;
  const float big[2][4] = {{0.f, -1.f, 5.f, 1.f}, {0.f, -102.f, 5.f, 100.f}};
  memcpy(out, big, sizeof(big));
  n = 2;
  for (x = -5; x <= 5; x += 2)
    for (y = -5; y <= 5; y += 2) {
      out[4 * n + 0] = (float)x, out[4 * n + 1] = (float)y, out[4 * n + 2] = 10.f, out[4 * n + 3] = 0.5f;
      ++n;
    }
  return n;
}

static int load_obj(const char* path, float** v_out, uint32_t* nv_out, uint32_t** t_out, uint32_t* nt_out) {
  FILE* f = fopen(path, "r");
  char line[1024];
  size_t nv = 0, nt = 0, cv = 1024, ct = 1024;
  float* v = (float*)malloc(cv * 3 * sizeof(float));
  uint32_t* t = (uint32_t*)malloc(ct * 3 * sizeof(uint32_t));
  if (!f || !v || !t) return 0;
  while (fgets(line, sizeof(line), f)) {
    if (line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) {
      if (nv == cv) v = (float*)realloc(v, (cv *= 2) * 3 * sizeof(float));
      if (sscanf(line + 2, "%f %f %f", &v[3 * nv], &v[3 * nv + 1], &v[3 * nv + 2]) == 3) ++nv;
    } else if (line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) {
      long idx[64];
      int k = 0;
      char* p = strtok(line + 2, " \t\r\n");
      while (p && k < 64) {
        long i = strtol(p, NULL, 10); /* "i", "i/j", "i//k", "i/j/k": the vertex index comes first */
        idx[k++] = i < 0 ? (long)nv + i : i - 1;
        p = strtok(NULL, " \t\r\n");
      }
      for (int j = 1; j + 1 < k; ++j) {
        if (nt == ct) t = (uint32_t*)realloc(t, (ct *= 2) * 3 * sizeof(uint32_t));
        t[3 * nt] = (uint32_t)idx[0], t[3 * nt + 1] = (uint32_t)idx[j], t[3 * nt + 2] = (uint32_t)idx[j + 1];
        ++nt;
      }
    }
  }
  fclose(f);
  for (size_t i = 0; i < 3 * nt; ++i)
    if (t[i] >= nv) return 0;
  *v_out = v, *nv_out = (uint32_t)nv, *t_out = t, *nt_out = (uint32_t)nt;
  return nv > 0 && nt > 0;
}

int main(int argc, char** argv) {
  cudapt_ctx* ctx = NULL;
  cudapt_config cfg;
  cudapt_postprocess post;
  cudapt_camera cam;
  int have_cam = 0;
  const char *out = NULL, *obj = NULL, *env = NULL;
  unsigned w = 640, h = 360, spp = 16, env_size = 256;
  cudapt_default_config(&cfg);
  cudapt_default_postprocess(&post);
  cudapt_default_camera(&cam);
  for (int i = 1; i < argc; ++i) {
    const char* a = argv[i];
    const char* nx = i + 1 < argc ? argv[i + 1] : "";
    if (!strcmp(a, "--out")) out = nx, ++i;
    else if (!strcmp(a, "--obj")) obj = nx, ++i;
    else if (!strcmp(a, "--scene")) ++i; /* only "app" */
    else if (!strcmp(a, "--env")) env = nx, ++i;
    else if (!strcmp(a, "--env-size")) env_size = (unsigned)atoi(nx), ++i;
    else if (!strcmp(a, "--size")) { if (sscanf(nx, "%ux%u", &w, &h) != 2) w = h = 0; ++i; }
    else if (!strcmp(a, "--spp")) spp = (unsigned)atoi(nx), ++i;
    else if (!strcmp(a, "--bounces")) cfg.max_bounces = atoi(nx), ++i;
    else if (!strcmp(a, "--device")) cfg.device = atoi(nx), ++i;
    else if (!strcmp(a, "--philox")) cfg.rng_mode = CUDAPT_RNG_PHILOX;
    else if (!strcmp(a, "--jitter")) cfg.jitter = 1, cfg.rng_mode = CUDAPT_RNG_PHILOX;
    else if (!strcmp(a, "--exposure")) post.exposure = (float)atof(nx), ++i;
    else if (!strcmp(a, "--auto-exposure")) post.auto_exposure = 1;
    else if (!strcmp(a, "--srgb")) post.srgb = 1;
    else if (!strcmp(a, "--look-at")) { /* the reference's placeholder Camera: position, target, vertical fov */
      float q[7];
      const float up[3] = {0.f, 1.f, 0.f};
      if (sscanf(nx, "%f,%f,%f,%f,%f,%f,%f", &q[0], &q[1], &q[2], &q[3], &q[4], &q[5], &q[6]) != 7 ||
          cudapt_camera_look_at(&cam, q, q + 3, up, q[6]) != CUDAPT_OK) {
        fprintf(stderr, "cudapt_cli: bad --look-at\n");
        return 2;
      }
      have_cam = 1, ++i;
    } else if (!strcmp(a, "--lens")) cam.lens_radius = (float)atof(nx), cfg.rng_mode = CUDAPT_RNG_PHILOX, have_cam = 1, ++i;
    else if (!strcmp(a, "--tonemap")) {
      post.tonemap = !strcmp(nx, "aces") ? CUDAPT_TONEMAP_ACES : !strcmp(nx, "reinhard") ? CUDAPT_TONEMAP_REINHARD : CUDAPT_TONEMAP_NONE;
      ++i;
    } else {
      fprintf(stderr, "cudapt_cli: unknown option %s\n", a);
      return 2;
    }
  }
  if (!out || !w || !h || !spp) {
    fprintf(stderr, "usage: cudapt_cli [--scene app | --obj FILE] [--size WxH] [--spp N] ... --out FILE.png|.ppm|.pfm|.exr\n");
    return 2;
  }
  if (obj) {
    cfg.rng_mode = CUDAPT_RNG_PHILOX;
    cfg.samples_per_pass = spp < 64 ? (int)spp : 64;
  }
  if (cudapt_create(&cfg, &ctx) != CUDAPT_OK) die(NULL, "cudapt_create");
  CHECK(cudapt_set_resolution(ctx, w, h));
  if (have_cam) CHECK(cudapt_set_camera(ctx, &cam));
  if (obj) {
    float* v;
    uint32_t *t, nv, nt;
    if (!load_obj(obj, &v, &nv, &t, &nt)) {
      fprintf(stderr, "cudapt_cli: cannot read triangles from %s\n", obj);
      return 1;
    }
    CHECK(cudapt_upload_mesh(ctx, v, nv, t, nt, NULL));
    CHECK(cudapt_build_accel(ctx));
    free(v), free(t);
  } else {
    float sph[38 * 4];
    const uint32_t n = app_spheres(sph);
    CHECK(cudapt_upload_spheres(ctx, sph, NULL, n));
  }
  if (env) CHECK(cudapt_load_environment(ctx, env, env_size, NULL));
  CHECK(cudapt_set_postprocess(ctx, &post));
  CHECK(cudapt_render(ctx, spp));
  CHECK(cudapt_write_image(ctx, out));
  {
    cudapt_stats st;
    CHECK(cudapt_get_stats(ctx, &st));
    printf("%s: %ux%u, %u spp, %llu paths, %llu rays, %.3f ms on the device (%.1f Msamples/s)\n", out, w, h, st.frame_count,
           (unsigned long long)st.paths, (unsigned long long)st.rays, st.render_ms, st.render_ms > 0 ? st.paths / st.render_ms / 1e3 : 0.0);
  }
  cudapt_destroy(ctx);
  return 0;
}
