// image_io.cpp -- floor texture decode and frame output.  JPEG/PNG codecs are the third-party stb headers the
// reference vendors (include/stb_image.h v2.30, include/stb_image_write.h v1.16); they are compiled from where
// they lie (-I$(STB_DIR), see ../Makefile) and never copied here.  Without them (RTB_HAVE_STB undefined) texture
// loading and PNG output report failure; binary / text output do not need them.
#include <cstdio>
#include <cstring>
#include <fstream>

#include "scene_host.h"

#ifdef RTB_HAVE_STB
#define STB_IMAGE_IMPLEMENTATION
#define STB_IMAGE_STATIC
#include "stb_image.h"
#define STB_IMAGE_WRITE_IMPLEMENTATION
#define STB_IMAGE_WRITE_STATIC
#include "stb_image_write.h"
#endif

namespace rtb {

// stbi_loadf(path, &w, &h, &channels, 4) as load_texture_to_gpu / load_texture_cpu call it (src/main.cu:18, :54):
// RGBA float, sRGB bytes linearised by stb with gamma 2.2
bool load_texels(const std::string& path, std::vector<float>& rgba, int& w, int& h) {
#ifdef RTB_HAVE_STB
    int ch = 0;
    float* d = stbi_loadf(path.c_str(), &w, &h, &ch, 4);
    if (!d) return false;
    rgba.assign(d, d + (size_t)w * h * 4);
    stbi_image_free(d);
    return true;
#else
    (void)path; (void)rgba; (void)w; (void)h;
    return false;
#endif
}

// BinarySaver (src/camera.cu:128-153): int width, int height, then RGB bytes row-major -- written in ONE call
// instead of three one-byte writes per pixel.
bool save_binary(const std::string& path, int W, int H, const unsigned char* rgb) {
    std::ofstream f(path, std::ios::binary);
    if (!f) return false;
    f.write(reinterpret_cast<const char*>(&W), sizeof(int));
    f.write(reinterpret_cast<const char*>(&H), sizeof(int));
    f.write(reinterpret_cast<const char*>(rgb), (std::streamsize)W * H * 3);
    return (bool)f;
}

// FileSaver (src/camera.cu:58-74): "P3\nW H\n255\n" then "r g b\n" per pixel
bool save_ppm_text(const std::string& path, int W, int H, const unsigned char* rgb) {
    FILE* f = std::fopen(path.c_str(), "w");
    if (!f) return false;
    std::fprintf(f, "P3\n%d %d\n255\n", W, H);
    for (size_t i = 0; i < (size_t)W * H; ++i) std::fprintf(f, "%d %d %d\n", rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
    return std::fclose(f) == 0;
}

// PNGSaver (src/camera.cu:94-126): stbi_write_png(path, W, H, 3, data, W*3)
bool save_png(const std::string& path, int W, int H, const unsigned char* rgb) {
#ifdef RTB_HAVE_STB
    return stbi_write_png(path.c_str(), W, H, 3, rgb, W * 3) != 0;
#else
    (void)path; (void)W; (void)H; (void)rgb;
    return false;
#endif
}

}  // namespace rtb
