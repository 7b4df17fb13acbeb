// extra_scenes.h -- benchmark scenes of BASELINE.json that the reference ships only as a
// comment or not at all, written once as templates over the scene API so the very same
// builder runs on rtow::api (product) and on the reference's own classes (oracle).
//
//   cornellSmoke_NED<API>()  config 4: cornellBox_NED's room (scenes.h:372-442) with REACHABLE
//                            ConstantMedium volumes, after the commented "smoke box" block
//                            scenes.h:461-468 (two box volumes, light and dark, as in "The Next
//                            Week") plus a fog sphere and the shipped glass/medium sphere.
//   bvhStress<API>(n)        config 5: n random spheres, one Material object each (so the
//                            reference's HitRecord::materialPointer identifies the primitive),
//                            under the reference's BvhNode build.
//
// API must provide the class names of scenes.h plus `static double unif01()`.
#pragma once

#include <cstdint>

namespace rtow {

template <class API>
typename API::Hitable *cornellSmoke_NED() {
    using vec3 = typename API::vec3;
    using Hitable = typename API::Hitable;
    using Material = typename API::Material;
    typedef typename API::ConstantTexture CT;
    Hitable **list = new Hitable *[32];
    int i = 0;
    Material *white = new typename API::Lambertian(new CT(vec3(0.73, 0.73, 0.73)));
    Material *red = new typename API::Lambertian(new CT(vec3(0.65, 0.05, 0.05)));
    Material *blue = new typename API::Lambertian(new CT(vec3(0.12, 0.12, 0.45)));
    Material *light = new typename API::DiffuseLight(new CT(vec3(7, 7, 7)));
    const int W = 1000, H = 1000;

    list[i++] = new typename API::Translate(new typename API::XYRectangle(0, W / 2, 0, H / 2, 0, light), vec3(-W / 4, -W / 4, (-H / 2) + 1));
    list[i++] = new typename API::Translate(new typename API::XYRectangle(0, W, 0, H, 0, white), vec3(-W / 2, -W / 2, -H / 2));
    list[i++] = new typename API::Translate(new typename API::FlipNormals(new typename API::XYRectangle(0, W, 0, H, 0, white)), vec3(-W / 2, -W / 2, H / 2));
    list[i++] = new typename API::Translate(new typename API::XZRectangle(0, W, 0, H, 0, red), vec3(-W / 2, -W / 2, -H / 2));
    list[i++] = new typename API::Translate(new typename API::FlipNormals(new typename API::YZRectangle(0, W, 0, H, 0, white)), vec3(W / 2, -W / 2, -H / 2));
    list[i++] = new typename API::Translate(new typename API::FlipNormals(new typename API::XZRectangle(0, W, 0, H, 0, blue)), vec3(-W / 2, W / 2, -H / 2));

    // two smoke volumes bounded by rotated, translated boxes
    Hitable *b1 = new typename API::Translate(
        new typename API::RotateY(new typename API::Box(vec3(0, 0, 0), vec3(200, 200, 300), white), 15.0), vec3(0, -300, 150));
    Hitable *b2 = new typename API::Translate(
        new typename API::RotateY(new typename API::Box(vec3(0, 0, 0), vec3(200, 200, 300), white), -18.0), vec3(0, 150, 150));
    list[i++] = new typename API::ConstantMedium(b1, 0.01, new CT(vec3(1.0, 1.0, 1.0)));
    list[i++] = new typename API::ConstantMedium(b2, 0.01, new CT(vec3(0.0, 0.0, 0.0)));
    // a solid rotated box, as shipped (scenes.h:457-460)
    list[i++] = new typename API::Translate(
        new typename API::RotateY(new typename API::Box(vec3(0, 0, 0), vec3(160, 160, 160), blue), 18.0), vec3(0, 0, -200));
    // fog ball in the open
    list[i++] = new typename API::ConstantMedium(new typename API::Sphere(vec3(-150, 0, -300), 120.0, white), 0.004,
                                                 new CT(vec3(0.8, 0.8, 0.95)));
    // the shipped glass shell with an opaque core and an (unreachable) medium (scenes.h:472-479)
    Material *core = new typename API::Lambertian(new CT(vec3(0.1, 0.7, 0.20)));
    Hitable *shell = new typename API::Sphere(vec3(0, 0, 150), 100.0, new typename API::Dielectric(1.5));
    Hitable *inner = new typename API::Sphere(vec3(0, 0, 150), 99.0, core);
    list[i++] = new typename API::ConstantMedium(inner, 0.001, new CT(vec3(0.1, 0.7, 0.2)));
    list[i++] = inner;
    list[i++] = shell;
    return new typename API::HitableList(list, i);
}

// n spheres, centres ~U(-100,100)^3, radii ~U(0.05,0.30); draw order per sphere: x, y, z, r.
template <class API>
typename API::Hitable *bvhStress(int n) {
    using vec3 = typename API::vec3;
    using Hitable = typename API::Hitable;
    Hitable **list = new Hitable *[n > 0 ? n : 1];
    typename API::Texture *grey = new typename API::ConstantTexture(vec3(0.5, 0.5, 0.5));
    for (int i = 0; i < n; ++i) {
        float x = -100.0 + 200.0 * API::unif01();
        float y = -100.0 + 200.0 * API::unif01();
        float z = -100.0 + 200.0 * API::unif01();
        float r = 0.05 + 0.25 * API::unif01();
        list[i] = new typename API::Sphere(vec3(x, y, z), r, new typename API::Lambertian(grey));
    }
    return new typename API::BvhNode(list, n, 0.0, 1.0);
}

}  // namespace rtow
