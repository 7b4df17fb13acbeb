// scenes.h -- the four scene builders of the reference's scenes.h, restated on rtow::api.
//
//   randomScene()      scenes.h:29-120     Y-up "one weekend" scene
//   randomScene_NED()  scenes.h:122-218    the one main() renders when OUTPUT_RANDOM_SCENE (main.cpp:132)
//   cornellBox()       scenes.h:220-353
//   cornellBox_NED()   scenes.h:355-483    the default world (main.cpp:136)
//
// Each consumes the global engine (rtow::api::unif01) in the same order as the reference, so
// after rtow::api::seed(S) the flattened scene equals the flattened reference scene built
// after randomNumberGenerator.seed(S) -- tests/test_oracle_vs_reference.py checks that array by array.
// Unused locals of the reference that allocate but never reach the scene are not reproduced
// unless they draw random numbers (none do).
//
// Image textures: the reference decodes ./input_images/*.jpg with its vendored stb_image
// (scenes.h:47,140).  load_image() decodes the same files with rtow/jpeg_decode.h, which reproduces
// stb_image's integer arithmetic byte for byte (tests/test_jpeg_decode.py), so the flattened scene
// equals the reference's without any override.  The reference's paths are relative to the CWD; when
// the file is not there, load_image() also looks in image_dir() (rtow_host_set_image_dir, or the
// environment variable RTOW_INPUT_IMAGES: a directory that holds the .jpg files themselves).
// Binary PPM (P6) is read as well.  When nothing can be loaded the builders substitute the glass
// sphere, as the reference does when stbi_load returns NULL (scenes.h:104-109,202-207).
#pragma once

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "jpeg_decode.h"
#include "scene_api.h"

namespace rtow {
namespace api {

inline unsigned char *load_ppm(const char *path, int *nx, int *ny) {
    FILE *f = std::fopen(path, "rb");
    if (!f) return nullptr;
    int w = 0, h = 0, maxv = 0;
    char magic[3] = {0, 0, 0};
    if (std::fscanf(f, "%2s", magic) != 1 || std::strcmp(magic, "P6") != 0) { std::fclose(f); return nullptr; }
    int got = 0, vals[3];
    while (got < 3) {
        int c = std::fgetc(f);
        if (c == EOF) { std::fclose(f); return nullptr; }
        if (c == '#') { while (c != '\n' && c != EOF) c = std::fgetc(f); continue; }
        if (c == ' ' || c == '\t' || c == '\n' || c == '\r') continue;
        std::ungetc(c, f);
        if (std::fscanf(f, "%d", &vals[got]) != 1) { std::fclose(f); return nullptr; }
        ++got;
    }
    w = vals[0]; h = vals[1]; maxv = vals[2];
    std::fgetc(f);   // the single whitespace byte after maxval
    if (w <= 0 || h <= 0 || maxv != 255) { std::fclose(f); return nullptr; }
    size_t n = (size_t)w * h * 3;
    unsigned char *px = new unsigned char[n];
    if (std::fread(px, 1, n, f) != n) { delete[] px; std::fclose(f); return nullptr; }
    std::fclose(f);
    *nx = w; *ny = h;
    return px;
}

// Optional override: tests/bench hand the builders pixels decoded elsewhere.
struct ImageOverride { unsigned char *data = nullptr; int nx = 0, ny = 0; };
inline ImageOverride &image_override() { static ImageOverride o; return o; }

// Where else to look for "./input_images/<file>" (empty: the CWD-relative path only, like the reference).
inline std::string &image_dir() {
    static std::string d = [] { const char *e = std::getenv("RTOW_INPUT_IMAGES"); return std::string(e ? e : ""); }();
    return d;
}

inline unsigned char *load_image_file(const std::string &path, int *nx, int *ny, int *nn) {
    if (unsigned char *p = load_ppm(path.c_str(), nx, ny)) { *nn = 3; return p; }
    return jpeg::load_file(path.c_str(), nx, ny, nn);
}

inline unsigned char *load_image(const char *path, int *nx, int *ny, int *nn) {
    *nn = 3;
    ImageOverride &o = image_override();
    if (o.data) { *nx = o.nx; *ny = o.ny; return o.data; }
    std::string s(path);
    std::vector<std::string> tries{s};
    const size_t slash = s.find_last_of('/');
    const std::string base = slash == std::string::npos ? s : s.substr(slash + 1);
    if (!image_dir().empty()) tries.push_back(image_dir() + "/" + base);
    for (const std::string &t : tries) {
        int c = 0;
        if (unsigned char *p = load_image_file(t, nx, ny, &c)) {
            if (c == 3) { *nn = 3; return p; }
            delete[] p;          // ImageTexture indexes RGB triples (texture.h:72-76): a grey file cannot stand in
        }
        const size_t dot = t.rfind('.');
        if (dot != std::string::npos)
            if (unsigned char *p = load_ppm((t.substr(0, dot) + ".ppm").c_str(), nx, ny)) return p;
    }
    return nullptr;
}

namespace detail {
// The ~97 small spheres both random scenes scatter: one draw picks the material, two place the
// centre, then 1 (moving diffuse) or 4 (metal) more; rejected centres consume only the first 3.
template <class CenterFn>
inline int scatter_small_spheres(Hitable **list, int i, int limit, int xMod, int yMod, int startX, int startY,
                                 Texture *checker, float glassRadius, CenterFn centre) {
    int cx = startX, cy = startY;
    while (i < limit) {
        float pick = unif01();
        vec3 c = centre(cx, cy);
        if ((c - vec3(4, 0.2, 0)).length() > 0.9) {
            if (pick < 0.8) {
                list[i++] = new MovingSphere(c, c + vec3(0, 0.5 * unif01(), 0), 0.0, 1.0, 0.2, new Lambertian(checker));
            } else if (pick < 0.95) {
                // four draws, evaluated left to right inside one full-expression in the reference;
                // g++ evaluates these constructor arguments right-to-left (verified against the
                // reference build by tests/test_scene_api.py), so draw into named temporaries in
                // that order here
                float fuzz = 0.5 * (1 + unif01());
                float b = 0.5 * (1 + unif01());
                float g = 0.5 * (1 + unif01());
                float r = 0.5 * (1 + unif01());
                list[i++] = new Sphere(c, 0.2, new Metal(vec3(r, g, b), fuzz));
            } else {
                list[i++] = new Sphere(c, glassRadius, new Dielectric(1.5));
            }
        }
        if (cx < xMod) cx++;
        else { cx = -xMod; cy++; }
        if (cy > yMod) cy = -yMod;
    }
    return i;
}
}  // namespace detail

inline Hitable *randomScene() {
    const int n = 100;
    Hitable **list = new Hitable *[n + 1];
    Texture *checker = new CheckerTexture(new ConstantTexture(vec3(0.2, 0.2, 0.8)), new ConstantTexture(vec3(0.9, 0.9, 0.9)));
    Texture *marble = new NoiseTexture(true, 8.0f);
    Material *sun = new DiffuseLight(new ConstantTexture(vec3(20, 0, 0)));
    int nx = 0, ny = 0, nn = 0;
    unsigned char *texels = load_image("./input_images/earth1300x1300.jpg", &nx, &ny, &nn);

    int i = 0;
    list[i++] = new Sphere(vec3(0, 1000, 0), 1000, new Lambertian(marble));
    i = detail::scatter_small_spheres(list, i, n - 3, 5, 5, -5, -5, checker, 0.2f, [](int cx, int cy) {
        // vec3 constructor arguments: see the evaluation-order note above
        float z = cy + 0.9 * unif01();
        float x = cx + 0.9 * unif01();
        return vec3(x, -0.2, z);
    });
    if (texels) list[i++] = new Sphere(vec3(0, -2.0, 20), 2.0, new Lambertian(new ImageTexture(texels, nx, ny)));
    else list[i++] = new Sphere(vec3(5, -1.0, 0), 1.0, new Dielectric(1.5));
    list[i++] = new Sphere(vec3(-4, -1.0, 0), 1.0, new Lambertian(marble));
    list[i++] = new Sphere(vec3(4, -1.0, 0), 1.0, new Metal(vec3(0.7, 0.6, 0.5), 0.0));
    list[i++] = new Sphere(vec3(0, -400, 10), 100.0, sun);
    return new BvhNode(list, i, 0.0, 1.0);
}

inline Hitable *randomScene_NED() {
    const int n = 100;
    Hitable **list = new Hitable *[n + 1];
    Texture *checker = new CheckerTexture(new ConstantTexture(vec3(0.2, 0.2, 0.8)), new ConstantTexture(vec3(0.9, 0.9, 0.9)));
    Texture *marble = new NoiseTexture(true, 8.0f);
    Material *sun = new DiffuseLight(new ConstantTexture(vec3(20, 0, 0)));
    int nx = 0, ny = 0, nn = 0;
    unsigned char *texels = load_image("./input_images/1_earth_8k.jpg", &nx, &ny, &nn);

    const float R = 1000.0, zOff = -1 * R;
    int i = 0;
    list[i++] = new Sphere(vec3(0, 0, 0), R, new Lambertian(marble));
    i = detail::scatter_small_spheres(list, i, n - 3, 10, 10, 10, 10, checker, 0.5f, [zOff](int cx, int cy) {
        float y = cy * unif01();
        float x = cx * unif01();
        return vec3(x, y, zOff - 1);
    });
    if (texels) list[i++] = new Sphere(vec3(20.0, 0.0, zOff - 25.0), 10.0, new Lambertian(new ImageTexture(texels, nx, ny)));
    else list[i++] = new Sphere(vec3(20.0, 0.0, zOff - 25.0), 10.0, new Dielectric(1.5));
    list[i++] = new Sphere(vec3(100, -70, zOff), 50.0, new Lambertian(marble));
    list[i++] = new Sphere(vec3(100, 70, zOff), 50.0, new Metal(vec3(0.7, 0.6, 0.5), 0.0));
    list[i++] = new Sphere(vec3(0, 0, zOff - 500), 100.0, sun);
    return new BvhNode(list, i, 0.0, 1.0);
}

// Y-up Cornell box, 500-unit room (scenes.h:220-353).  The reference computes the panel
// coordinates from integer expressions with xCoord = yCoord = 0; the values below are those
// expressions evaluated (integer division included).
inline Hitable *cornellBox() {
    Hitable **list = new Hitable *[100];
    int i = 0;
    Material *white = new Lambertian(new ConstantTexture(vec3(0.73, 0.73, 0.73)));
    Material *red = new Lambertian(new ConstantTexture(vec3(0.65, 0.05, 0.05)));
    Material *blue = new Lambertian(new ConstantTexture(vec3(0.12, 0.12, 0.45)));
    Material *light = new DiffuseLight(new ConstantTexture(vec3(3, 3, 3)));

    list[i++] = new XZRectangle(-175, 175, 75, 425, -249, light);          // light panel
    list[i++] = new XZRectangle(-250, 250, 0, 500, -250, white);           // top
    list[i++] = new YZRectangle(-250, 250, 0, 500, -250, red);             // left
    list[i++] = new Translate(new FlipNormals(new XYRectangle(-250, 250, -250, 250, 500, white)), vec3(0, 0, 0));  // back
    list[i++] = new FlipNormals(new YZRectangle(-250, 250, 0, 500, 250, blue));   // right
    list[i++] = new FlipNormals(new XZRectangle(-250, 250, 0, 500, 250, white));  // bottom

    Material *core = new Lambertian(new ConstantTexture(vec3(0.9, 0.05, 0.10)));
    Hitable *shell = new Sphere(vec3(-150, 150, 200), 100.0, new Dielectric(1.5));
    Hitable *inner = new Sphere(vec3(-150, 150, 200), 99.0, core);
    list[i++] = new ConstantMedium(inner, 0.001, new ConstantTexture(vec3(0.8, 0.2, 0.1)));
    list[i++] = inner;
    list[i++] = shell;
    return new HitableList(list, i);
}

// NED Cornell box, 1000-unit room built from k=0 rectangles moved into place by Translate
// (scenes.h:355-483).
inline Hitable *cornellBox_NED() {
    Hitable **list = new Hitable *[100];
    int i = 0;
    Material *white = new Lambertian(new ConstantTexture(vec3(0.73, 0.73, 0.73)));
    Material *red = new Lambertian(new ConstantTexture(vec3(0.65, 0.05, 0.05)));
    Material *blue = new Lambertian(new ConstantTexture(vec3(0.12, 0.12, 0.45)));
    Material *light = new DiffuseLight(new ConstantTexture(vec3(3, 3, 3)));
    const int W = 1000, H = 1000;

    list[i++] = new Translate(new XYRectangle(0, W / 2, 0, H / 2, 0, light), vec3(-W / 4, -W / 4, (-H / 2) + 1));
    list[i++] = new Translate(new XYRectangle(0, W, 0, H, 0, white), vec3(-W / 2, -W / 2, -H / 2));                    // top
    list[i++] = new Translate(new FlipNormals(new XYRectangle(0, W, 0, H, 0, white)), vec3(-W / 2, -W / 2, H / 2));    // bottom
    list[i++] = new Translate(new XZRectangle(0, W, 0, H, 0, red), vec3(-W / 2, -W / 2, -H / 2));                      // left
    list[i++] = new Translate(new FlipNormals(new YZRectangle(0, W, 0, H, 0, white)), vec3(W / 2, -W / 2, -H / 2));    // back
    list[i++] = new Translate(new FlipNormals(new XZRectangle(0, W, 0, H, 0, blue)), vec3(-W / 2, W / 2, -H / 2));     // right

    list[i++] = new Translate(new Box(vec3(0, 0, 0), vec3(200, 200, 300), blue), vec3(0, -300, 0));
    list[i++] = new Translate(new Box(vec3(0, 0, 0), vec3(200, 200, 300), red), vec3(0, 150, 0));
    list[i++] = new Translate(new RotateY(new Box(vec3(0, 0, 0), vec3(160, 160, 160), blue), 18.0), vec3(0, 0, -200));

    Material *core = new Lambertian(new ConstantTexture(vec3(0.1, 0.7, 0.20)));
    Hitable *shell = new Sphere(vec3(0, 0, 150), 100.0, new Dielectric(1.5));
    Hitable *inner = new Sphere(vec3(0, 0, 150), 99.0, core);
    list[i++] = new ConstantMedium(inner, 0.001, new ConstantTexture(vec3(0.1, 0.7, 0.2)));
    list[i++] = inner;
    list[i++] = shell;
    return new HitableList(list, i);
}

}  // namespace api
}  // namespace rtow
