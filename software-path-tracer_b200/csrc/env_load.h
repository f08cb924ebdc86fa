// Environment-map files (env_load.cpp): .pfm / .hdr latitude-longitude panoramas -> octahedral texels.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <string>
#include <vector>

namespace cudapt {

struct LatLongImage {
  uint32_t width = 0, height = 0;
  std::vector<float> rgb;  // row-major, top row first, 3 floats per pixel
};

bool load_latlong_image(const char* path, LatLongImage* im, std::string* err);
void latlong_to_octahedral(const LatLongImage& im, uint32_t size, float* out_rgb /* size * size * 3 */);

}  // namespace cudapt
