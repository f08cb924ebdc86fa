// Environment-map files for cudapt_load_environment (reference anchor: the commented-out
// Scene::setEnvironmentMap of libs/render/src/Scene.cpp:124-169, which read an HDR file as float RGB through
// OpenImageIO and handed it to a SkyBox).  No image library here: a portable float map (.pfm, "PF"/"Pf") and a Radiance
// picture (.hdr / .pic, RGBE, flat or new-style run-length scan lines) are read by hand, taken as a LATITUDE-LONGITUDE
// panorama and resampled once, on the host and in double precision, into the octahedral map the kernels look up
// (cudapt_upload_environment) -- the file format is data on the way in, the device never sees it.
#include "env_load.h"

#include <cmath>
#include <cstdio>
#include <cstring>

namespace cudapt {
namespace {

bool ends_with(const std::string& s, const char* suffix) {
  const size_t n = std::strlen(suffix);
  if (s.size() < n) return false;
  for (size_t i = 0; i < n; ++i) {
    char a = s[s.size() - n + i], b = suffix[i];
    if (a >= 'A' && a <= 'Z') a = (char)(a - 'A' + 'a');
    if (a != b) return false;
  }
  return true;
}

// next whitespace-delimited token of a PFM header
bool pfm_token(FILE* f, char* out, size_t cap) {
  int c;
  do c = std::fgetc(f);
  while (c == ' ' || c == '\n' || c == '\r' || c == '\t');
  size_t n = 0;
  while (c != EOF && c != ' ' && c != '\n' && c != '\r' && c != '\t') {
    if (n + 1 < cap) out[n++] = (char)c;
    c = std::fgetc(f);
  }
  out[n] = 0;
  return n > 0;
}

bool load_pfm(FILE* f, LatLongImage* im, std::string* err) {
  char tok[64];
  if (!pfm_token(f, tok, sizeof(tok)) || (std::strcmp(tok, "PF") != 0 && std::strcmp(tok, "Pf") != 0)) {
    *err = "not a PFM file";
    return false;
  }
  const int channels = tok[1] == 'F' ? 3 : 1;
  char tw[64], th[64], ts[64];
  if (!pfm_token(f, tw, sizeof(tw)) || !pfm_token(f, th, sizeof(th)) || !pfm_token(f, ts, sizeof(ts))) {
    *err = "truncated PFM header";
    return false;
  }
  const long w = std::atol(tw), h = std::atol(th);
  const double scale = std::atof(ts);
  if (w <= 0 || h <= 0 || w > 32768 || h > 32768 || scale == 0.0) {
    *err = "bad PFM header";
    return false;
  }
  const bool little = scale < 0.0;  // exactly one whitespace byte after the scale has been consumed by pfm_token
  std::vector<unsigned char> raw((size_t)w * h * channels * 4);
  if (std::fread(raw.data(), 1, raw.size(), f) != raw.size()) {
    *err = "truncated PFM data";
    return false;
  }
  im->width = (uint32_t)w;
  im->height = (uint32_t)h;
  im->rgb.resize((size_t)w * h * 3);
  for (long y = 0; y < h; ++y)  // the file stores the BOTTOM row first
    for (long x = 0; x < w; ++x)
      for (int c = 0; c < 3; ++c) {
        const unsigned char* p = &raw[(((size_t)(h - 1 - y) * w + x) * channels + (channels == 3 ? c : 0)) * 4];
        uint32_t u = little ? (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24
                            : (uint32_t)p[3] | (uint32_t)p[2] << 8 | (uint32_t)p[1] << 16 | (uint32_t)p[0] << 24;
        float v;
        std::memcpy(&v, &u, 4);
        im->rgb[((size_t)y * w + x) * 3 + c] = v;
      }
  return true;
}

// Radiance RGBE: value = mantissa * 2^(e - 136), e = 0 -> black (Ward, "Real Pixels", Graphics Gems II)
void rgbe_to_float(const unsigned char* p, float* out) {
  if (p[3] == 0) {
    out[0] = out[1] = out[2] = 0.0f;
    return;
  }
  const float s = std::ldexp(1.0f, (int)p[3] - 136);
  out[0] = (float)p[0] * s, out[1] = (float)p[1] * s, out[2] = (float)p[2] * s;
}

bool load_hdr(FILE* f, LatLongImage* im, std::string* err) {
  char line[512];
  if (!std::fgets(line, sizeof(line), f) || (std::strncmp(line, "#?RADIANCE", 10) != 0 && std::strncmp(line, "#?RGBE", 6) != 0)) {
    *err = "not a Radiance picture";
    return false;
  }
  bool format_ok = false;
  for (;;) {  // header lines up to the empty one
    if (!std::fgets(line, sizeof(line), f)) {
      *err = "truncated Radiance header";
      return false;
    }
    if (line[0] == '\n' || (line[0] == '\r' && line[1] == '\n')) break;
    if (std::strncmp(line, "FORMAT=32-bit_rle_rgbe", 22) == 0) format_ok = true;
  }
  if (!format_ok) {
    *err = "Radiance picture is not 32-bit_rle_rgbe";
    return false;
  }
  long w = 0, h = 0;
  if (!std::fgets(line, sizeof(line), f) || std::sscanf(line, "-Y %ld +X %ld", &h, &w) != 2 || w <= 0 || h <= 0 || w > 32768 ||
      h > 32768) {
    *err = "unsupported Radiance resolution line (only -Y H +X W)";
    return false;
  }
  im->width = (uint32_t)w;
  im->height = (uint32_t)h;
  im->rgb.resize((size_t)w * h * 3);
  std::vector<unsigned char> scan((size_t)w * 4);
  for (long y = 0; y < h; ++y) {
    unsigned char head[4];
    if (std::fread(head, 1, 4, f) != 4) {
      *err = "truncated Radiance data";
      return false;
    }
    const bool rle = w >= 8 && w < 32768 && head[0] == 2 && head[1] == 2 && (head[2] & 0x80) == 0;
    if (rle) {
      if ((((long)head[2]) << 8 | head[3]) != w) {
        *err = "Radiance scan line length mismatch";
        return false;
      }
      for (int c = 0; c < 4; ++c) {  // each component run-length encoded on its own
        long x = 0;
        while (x < w) {
          int n = std::fgetc(f);
          if (n == EOF) {
            *err = "truncated Radiance data";
            return false;
          }
          if (n > 128) {  // run
            n -= 128;
            const int v = std::fgetc(f);
            if (v == EOF || x + n > w) {
              *err = "bad Radiance run";
              return false;
            }
            for (int k = 0; k < n; ++k) scan[(size_t)(x++) * 4 + c] = (unsigned char)v;
          } else {  // literal
            if (n == 0 || x + n > w) {
              *err = "bad Radiance literal";
              return false;
            }
            for (int k = 0; k < n; ++k) {
              const int v = std::fgetc(f);
              if (v == EOF) {
                *err = "truncated Radiance data";
                return false;
              }
              scan[(size_t)(x++) * 4 + c] = (unsigned char)v;
            }
          }
        }
      }
    } else {  // flat scan line: the four bytes read were the first pixel
      std::memcpy(scan.data(), head, 4);
      if (std::fread(scan.data() + 4, 1, (size_t)(w - 1) * 4, f) != (size_t)(w - 1) * 4) {
        *err = "truncated Radiance data";
        return false;
      }
    }
    for (long x = 0; x < w; ++x) rgbe_to_float(&scan[(size_t)x * 4], &im->rgb[((size_t)y * w + x) * 3]);
  }
  return true;
}

}  // namespace

bool load_latlong_image(const char* path, LatLongImage* im, std::string* err) {
  const std::string p(path);
  const bool pfm = ends_with(p, ".pfm"), hdr = ends_with(p, ".hdr") || ends_with(p, ".pic");
  if (!pfm && !hdr) {
    *err = "'" + p + "' is neither .pfm nor .hdr";
    return false;
  }
  FILE* f = std::fopen(path, "rb");
  if (!f) {
    *err = "cannot open '" + p + "'";
    return false;
  }
  const bool ok = pfm ? load_pfm(f, im, err) : load_hdr(f, im, err);
  std::fclose(f);
  return ok;
}

// Octahedral texel (i, j) -> direction (the inverse of sky_radiance()'s mapping, common.cuh) -> panorama coordinates
// u = 1/2 + atan2(d.x, d.z) / 2pi (the image centre looks down +Z, the default camera's view direction; +X is to its
// right), v = acos(d.y) / pi (top row = +Y) -> bilinear sample, wrapping in u and clamping in v.
void latlong_to_octahedral(const LatLongImage& im, uint32_t size, float* out_rgb) {
  const double kPi = 3.14159265358979323846;
  const long W = (long)im.width, H = (long)im.height;
  for (uint32_t j = 0; j < size; ++j)
    for (uint32_t i = 0; i < size; ++i) {
      const double px = ((double)i + 0.5) / (double)size * 2.0 - 1.0, pz = ((double)j + 0.5) / (double)size * 2.0 - 1.0;
      const double y = 1.0 - std::fabs(px) - std::fabs(pz);
      double x = px, z = pz;
      if (y < 0.0) {
        x = (px < 0.0 ? -1.0 : 1.0) * (1.0 - std::fabs(pz));
        z = (pz < 0.0 ? -1.0 : 1.0) * (1.0 - std::fabs(px));
      }
      const double len = std::sqrt(x * x + y * y + z * z);
      const double dx = x / len, dy = y / len, dz = z / len;
      const double u = 0.5 + std::atan2(dx, dz) / (2.0 * kPi);
      const double v = std::acos(dy < -1.0 ? -1.0 : (dy > 1.0 ? 1.0 : dy)) / kPi;
      const double fx = u * (double)W - 0.5, fy = v * (double)H - 0.5;
      const double x0f = std::floor(fx), y0f = std::floor(fy);
      const double ax = fx - x0f, ay = fy - y0f;
      long x0 = (long)x0f, x1 = x0 + 1, y0 = (long)y0f, y1 = y0 + 1;
      x0 = ((x0 % W) + W) % W, x1 = ((x1 % W) + W) % W;
      y0 = y0 < 0 ? 0 : (y0 > H - 1 ? H - 1 : y0), y1 = y1 < 0 ? 0 : (y1 > H - 1 ? H - 1 : y1);
      for (int c = 0; c < 3; ++c) {
        const double c00 = im.rgb[((size_t)y0 * W + x0) * 3 + c], c10 = im.rgb[((size_t)y0 * W + x1) * 3 + c];
        const double c01 = im.rgb[((size_t)y1 * W + x0) * 3 + c], c11 = im.rgb[((size_t)y1 * W + x1) * 3 + c];
        out_rgb[((size_t)j * size + i) * 3 + c] = (float)((c00 * (1.0 - ax) + c10 * ax) * (1.0 - ay) + (c01 * (1.0 - ax) + c11 * ax) * ay);
      }
    }
}

}  // namespace cudapt
