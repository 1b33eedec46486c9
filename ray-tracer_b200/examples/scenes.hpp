// Scene builders of the reference's three example drivers plus the synthetic stress scene, written
// once against an "API" type bundle so the same construction code drives
//   * the product host mirror (ray-tracer_b200/host/*.hpp -> flatten -> C ABI -> CUDA), and
//   * the f64 oracle (oracle/oracle.hpp),
// i.e. both sides see the identical scene graph (BASELINE.json configs 1-5).
//
// Follows examples/book-one.rs:103-205 (randomScene), examples/cornell-box.rs:31-135,
// examples/main.rs:156-330 (finalScene).  The reference seeds scene randomness from the OS
// (thread_rng); here every draw comes from a Philox stream with a caller-supplied seed.
// earthmap.jpg (main.rs:266, a network download) is replaced by a procedural 1024x512 RGB8 texture.
#pragma once
#include <cmath>
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "../csrc/philox.h"

namespace rtscenes {

// Sequential uniform stream standing in for `thread_rng()` at scene-build time.
struct SceneRng {
    uint64_t seed;
    uint32_t n = 0;
    explicit SceneRng(uint64_t s) : seed(s) {}
    double next() {  // 53-bit uniform in [0,1)
        const rtb::Philox4 r = rtb::philox4x32_10(n++, 0x5CE9Eu, (uint32_t)seed, (uint32_t)(seed >> 32), 0xB00C0DEu, 1u);
        const uint64_t bits = ((uint64_t)r.x << 21) ^ (uint64_t)(r.y >> 11);
        return (double)(bits & ((1ull << 53) - 1)) * (1.0 / 9007199254740992.0);
    }
    double gen_range(double lo, double hi) { return lo + (hi - lo) * next(); }
};

struct CameraSpec {  // arguments of PerspectiveCamera::new (camera.rs:25-33) as each example passes them
    double eye[3], center[3], up[3], fov_degrees, focusDistance, lensRadius;
};

inline double to_radians(double deg) { return deg * 3.14159265358979323846 / 180.0; }

// Procedural stand-in for earthmap.jpg: lat/long checker plus gradients, 8-bit RGB, row 0 = top.
inline std::shared_ptr<std::vector<uint8_t>> syntheticEarth(uint32_t w = 1024, uint32_t h = 512) {
    auto img = std::make_shared<std::vector<uint8_t>>((size_t)w * h * 3);
    for (uint32_t j = 0; j < h; ++j)
        for (uint32_t i = 0; i < w; ++i) {
            const uint32_t checker = ((i >> 6) ^ (j >> 6)) & 1u;
            uint8_t* p = img->data() + ((size_t)j * w + i) * 3;
            p[0] = (uint8_t)(checker ? 230u : 30u + (i * 64u) / w);
            p[1] = (uint8_t)(40u + (j * 200u) / h);
            p[2] = (uint8_t)(checker ? 60u : 200u);
        }
    return img;
}

// main.rs:266 loads ./earthmap.jpg; callers that have a decoded image (rth_set_earth_image / orc_set_earth_image) install it
// here and book-two maps it exactly as main.rs:267-280 does; otherwise the synthetic texture stands in.
struct EarthImage {
    uint32_t width = 0, height = 0;
    std::shared_ptr<std::vector<uint8_t>> rgb;  // top-down 8-bit RGB
};
inline EarthImage& earthOverride() {
    static EarthImage e;
    return e;
}

template <class A>
struct Scenes {
    using Vec3 = typename A::Vec3;
    using Mat4 = typename A::Mat4;
    using BoundPtr = typename A::BoundPtr;
    using List = std::vector<BoundPtr>;

    template <class G, class M>
    static BoundPtr sprite(std::shared_ptr<G> g, std::shared_ptr<M> m, const Mat4& t) {
        return A::Sprite::builder().geometry(std::move(g)).material(std::move(m)).transform(t).build();
    }
    template <class G, class M>
    static BoundPtr sprite(std::shared_ptr<G> g, std::shared_ptr<M> m) {
        return A::Sprite::builder().geometry(std::move(g)).material(std::move(m)).build();
    }
    static std::shared_ptr<typename A::Sphere> sphere(double r) { return std::make_shared<typename A::Sphere>(r); }
    static std::shared_ptr<typename A::Rectangle> rect(double w, double h) { return std::make_shared<typename A::Rectangle>(w, h); }

    // ---- examples/book-one.rs:103-205 ---------------------------------------------------------
    static List bookOne(SceneRng& generator) {
        List scene;
        scene.push_back(sprite(sphere(1000.0), std::make_shared<typename A::Lambertian>(Vec3(0.5, 0.5, 0.5)),
                               Mat4::translation(Vec3(0.0, -1000.0, 0.0))));                                    // ground
        scene.push_back(sprite(sphere(2000.0), std::make_shared<typename A::DiffuseLight>(Vec3(0.5, 0.7, 1.0))));  // sky
        for (int a = -11; a < 11; ++a)
            for (int b = -11; b < 11; ++b) {
                const double whichMaterial = generator.gen_range(0.0, 1.0);
                const double cx = (double)a + 0.9 * generator.gen_range(0.0, 1.0);
                const double cz = (double)b + 0.9 * generator.gen_range(0.0, 1.0);
                const Vec3 center(cx, 0.2, cz);
                const double dx = cx - 4.0, dy = 0.2 - 0.2, dz = cz - 0.0;
                if (std::sqrt(dx * dx + dy * dy + dz * dz) > 0.9) {
                    if (whichMaterial < 0.3) {
                        const double r = generator.gen_range(0.0, 1.0), g = generator.gen_range(0.0, 1.0),
                                     bl = generator.gen_range(0.0, 1.0);
                        scene.push_back(sprite(sphere(0.2), std::make_shared<typename A::Lambertian>(Vec3(r * r, g * g, bl * bl)),
                                               Mat4::translation(center)));
                    } else if (whichMaterial < 0.6) {
                        const double r = generator.gen_range(0.5, 1.0), g = generator.gen_range(0.5, 1.0),
                                     bl = generator.gen_range(0.5, 1.0);
                        const double fuzziness = generator.gen_range(0.0, 0.5);
                        scene.push_back(sprite(sphere(0.2), std::make_shared<typename A::Metal>(Vec3(r, g, bl), fuzziness),
                                               Mat4::translation(center)));
                    } else {
                        scene.push_back(sprite(sphere(0.2), std::make_shared<typename A::Dielectric>(1.5), Mat4::translation(center)));
                    }
                }
            }
        scene.push_back(sprite(sphere(1.0), std::make_shared<typename A::Lambertian>(Vec3(0.4, 0.2, 0.1)), Mat4::translation(Vec3(-4.0, 1.0, 0.0))));
        scene.push_back(sprite(sphere(1.0), std::make_shared<typename A::Metal>(Vec3(0.7, 0.6, 0.5), 0.0), Mat4::translation(Vec3(4.0, 1.0, 0.0))));
        scene.push_back(sprite(sphere(1.0), std::make_shared<typename A::Dielectric>(1.5), Mat4::translation(Vec3(0.0, 1.0, 0.0))));
        return scene;
    }
    static CameraSpec bookOneCamera() { return {{13.0, 2.0, 3.0}, {0.0, 0.0, 0.0}, {0.0, 1.0, 0.0}, 20.0, 10.0, 0.05}; }  // book-one.rs:35-47

    // ---- examples/cornell-box.rs:31-135 -------------------------------------------------------
    static BoundPtr cubeSprite(double w, double h, double d, std::shared_ptr<typename A::Lambertian> m, const Mat4& t) {
        auto faces = A::Cube::make(w, h, d);
        auto bvh = A::BoundingVolumeHierarchyNode::make(List(faces.begin(), faces.end()));
        return sprite(bvh, std::move(m), t);
    }
    static List cornellBox() {
        auto red = std::make_shared<typename A::Lambertian>(Vec3(0.65, 0.05, 0.05));
        auto white = std::make_shared<typename A::Lambertian>(Vec3(0.73, 0.73, 0.73));
        auto green = std::make_shared<typename A::Lambertian>(Vec3(0.12, 0.45, 0.15));
        auto light = std::make_shared<typename A::DiffuseLight>(Vec3(15.0, 15.0, 15.0));
        const Vec3 ex(1.0, 0.0, 0.0), ey(0.0, 1.0, 0.0);
        List world;
        world.push_back(sprite(rect(555.0, 555.0), green,
                               Mat4::translation(Vec3(555.0, 555.0 / 2.0, 555.0 / 2.0)).multiplied(Mat4::rotation(to_radians(-90.0), ey))));
        world.push_back(sprite(rect(555.0, 555.0), red,
                               Mat4::translation(Vec3(0.0, 555.0 / 2.0, 555.0 / 2.0)).multiplied(Mat4::rotation(to_radians(90.0), ey))));
        world.push_back(sprite(rect(130.0, 105.0), light,
                               Mat4::translation(Vec3(555.0 / 2.0, 554.0, 555.0 / 2.0)).multiplied(Mat4::rotation(to_radians(90.0), ex))));
        world.push_back(sprite(rect(555.0, 555.0), white,
                               Mat4::translation(Vec3(555.0 / 2.0, 0.0, 555.0 / 2.0)).multiplied(Mat4::rotation(to_radians(-90.0), ex))));
        world.push_back(sprite(rect(555.0, 555.0), white,
                               Mat4::translation(Vec3(555.0 / 2.0, 555.0, 555.0 / 2.0)).multiplied(Mat4::rotation(to_radians(90.0), ex))));
        world.push_back(sprite(rect(555.0, 556.0), white,  // 555x556 as in cornell-box.rs:77
                               Mat4::translation(Vec3(555.0 / 2.0, 555.0 / 2.0, 555.0)).multiplied(Mat4::rotation(to_radians(180.0), ey))));
        world.push_back(cubeSprite(165.0, 165.0, 165.0, white,
                                   Mat4::translation(Vec3(212.5, 82.5, 147.5)).multiplied(Mat4::rotation(to_radians(-18.0), ey))));
        world.push_back(cubeSprite(165.0, 330.0, 165.0, white,
                                   Mat4::translation(Vec3(347.5, 165.0, 377.5)).multiplied(Mat4::rotation(to_radians(15.0), ey))));
        return world;
    }
    static CameraSpec cornellBoxCamera() {  // cornell-box.rs:139-150
        return {{555.0 / 2.0, 555.0 / 2.0, -800.0}, {555.0 / 2.0, 555.0 / 2.0, 0.0}, {0.0, 1.0, 0.0}, 40.0, 10.0, 0.0};
    }

    // ---- examples/main.rs:156-330 (book-two cover) ---------------------------------------------
    static List bookTwo(SceneRng& generator) {
        auto ground = std::make_shared<typename A::Lambertian>(Vec3(0.48, 0.83, 0.53));
        const int boxCountPerSide = 20;
        List cubes;
        for (int i = 0; i < boxCountPerSide; ++i)
            for (int j = 0; j < boxCountPerSide; ++j) {
                const double w = 100.0;
                const double x0 = -1000.0 + (double)i * w, y0 = 0.0, z0 = -1000.0 + (double)j * w;
                const double x1 = x0 + w, y1 = generator.gen_range(1.0, 101.0), z1 = z0 + w;
                cubes.push_back(cubeSprite(x1 - x0, y1 - y0, z1 - z0, ground,
                                           Mat4::translation(Vec3((x0 + x1) / 2.0, (y0 + y1) / 2.0, (z0 + z1) / 2.0))));
            }
        BoundPtr cubesNode = A::BoundingVolumeHierarchyNode::make(cubes);
        const Vec3 ex(1.0, 0.0, 0.0);
        BoundPtr light = sprite(rect(300.0, 265.0), std::make_shared<typename A::DiffuseLight>(Vec3(7.0, 7.0, 7.0)),
                                Mat4::translation(Vec3(273.0, 554.0, 279.5)).multiplied(Mat4::rotation(to_radians(90.0), ex)));
        BoundPtr movingSphere = sprite(sphere(50.0), std::make_shared<typename A::Lambertian>(Vec3(0.7, 0.3, 0.1)),
                                       Mat4::translation(Vec3(400.0, 400.0, 200.0)));
        BoundPtr glassSphere = sprite(sphere(50.0), std::make_shared<typename A::Dielectric>(1.5), Mat4::translation(Vec3(260.0, 150.0, 45.0)));
        BoundPtr metalSphere = sprite(sphere(50.0), std::make_shared<typename A::Metal>(Vec3(0.8, 0.8, 0.9), 1.0),
                                      Mat4::translation(Vec3(0.0, 150.0, 145.0)));
        BoundPtr blueSurface = sprite(sphere(70.0), std::make_shared<typename A::Dielectric>(1.5), Mat4::translation(Vec3(360.0, 150.0, 145.0)));
        BoundPtr blueMedium = sprite(std::make_shared<typename A::ConstantMedium>(sphere(70.0 - 1e-6), 0.03),
                                     std::make_shared<typename A::Isotropic>(Vec3(0.2, 0.4, 0.9)), Mat4::translation(Vec3(360.0, 150.0, 145.0)));
        BoundPtr fog = sprite(std::make_shared<typename A::ConstantMedium>(sphere(5000.0), 0.0001),
                              std::make_shared<typename A::Isotropic>(Vec3(1.0, 1.0, 1.0)));
        // earth: nearest-texel closure as in main.rs:267-280 (clamped instead of panicking at u == 1)
        const EarthImage& override_image = earthOverride();
        auto image = override_image.rgb ? override_image.rgb : syntheticEarth();
        const uint32_t iw = override_image.rgb ? override_image.width : 1024u, ih = override_image.rgb ? override_image.height : 512u;
        std::function<Vec3(double, double)> mapping = [image, iw, ih](double u, double v) {
            double fx = u * (double)iw, fy = (1.0 - v) * (double)ih;
            uint32_t px = fx > 0.0 ? (uint32_t)fx : 0u, py = fy > 0.0 ? (uint32_t)fy : 0u;  // `as u32` saturates
            if (px > iw - 1) px = iw - 1;
            if (py > ih - 1) py = ih - 1;
            const uint8_t* p = image->data() + ((size_t)py * iw + px) * 3;
            return Vec3((double)p[0] / 255.0, (double)p[1] / 255.0, (double)p[2] / 255.0);
        };
        auto earthTexture = std::make_shared<typename A::ImageTexture>(mapping);
        A::setBakeSize(*earthTexture, iw, ih);  // the device bakes the closure at the image's own resolution: nearest lookup stays exact
        BoundPtr earth = sprite(sphere(100.0), std::make_shared<typename A::Lambertian>(earthTexture), Mat4::translation(Vec3(400.0, 200.0, 400.0)));
        auto white = std::make_shared<typename A::Lambertian>(Vec3(0.73, 0.73, 0.73));
        List spheres;
        for (int k = 0; k < 1000; ++k) {
            const double x = generator.gen_range(0.0, 165.0), y = generator.gen_range(0.0, 165.0), z = generator.gen_range(0.0, 165.0);
            spheres.push_back(sprite(sphere(10.0), white, Mat4::translation(Vec3(x - 100.0, y + 270.0, z + 395.0))));
        }
        BoundPtr spheresNode = A::BoundingVolumeHierarchyNode::make(spheres);
        return {cubesNode, light, movingSphere, glassSphere, metalSphere, blueSurface, blueMedium, earth, fog, spheresNode};
    }
    static CameraSpec bookTwoCamera() {  // main.rs:49-61
        return {{555.0 / 2.0 + 200.0, 550.0 / 2.0, -600.0}, {555.0 / 2.0, 555.0 / 2.0, 0.0}, {0.0, 1.0, 0.0}, 40.0, 10.0, 0.0};
    }

    // ---- synthetic stress scene (BASELINE.json configs[3]; defined in SURVEY §8(d) C4) ----------
    static List stress(SceneRng& generator, uint32_t count) {
        List scene;
        scene.push_back(sprite(sphere(2000.0), std::make_shared<typename A::DiffuseLight>(Vec3(0.5, 0.7, 1.0))));
        for (uint32_t k = 0; k < count; ++k) {
            const double whichMaterial = generator.gen_range(0.0, 1.0);
            const Vec3 center(generator.gen_range(-50.0, 50.0), generator.gen_range(-50.0, 50.0), generator.gen_range(-50.0, 50.0));
            if (whichMaterial < 0.3) {
                const double r = generator.gen_range(0.0, 1.0), g = generator.gen_range(0.0, 1.0), bl = generator.gen_range(0.0, 1.0);
                scene.push_back(sprite(sphere(0.25), std::make_shared<typename A::Lambertian>(Vec3(r * r, g * g, bl * bl)), Mat4::translation(center)));
            } else if (whichMaterial < 0.6) {
                const double r = generator.gen_range(0.5, 1.0), g = generator.gen_range(0.5, 1.0), bl = generator.gen_range(0.5, 1.0);
                const double fuzziness = generator.gen_range(0.0, 0.5);
                scene.push_back(sprite(sphere(0.25), std::make_shared<typename A::Metal>(Vec3(r, g, bl), fuzziness), Mat4::translation(center)));
            } else {
                scene.push_back(sprite(sphere(0.25), std::make_shared<typename A::Dielectric>(1.5), Mat4::translation(center)));
            }
        }
        return scene;
    }
    static CameraSpec stressCamera() { return {{0.0, 0.0, 180.0}, {0.0, 0.0, 0.0}, {0.0, 1.0, 0.0}, 40.0, 10.0, 0.0}; }

    // ---- parity scene for the node / texture kinds none of the examples uses (not a BASELINE config) ----------------
    // Vec<Arc<dyn Bound>> as the WORLD ROOT, as an inner node and as a Sprite's geometry (src/geometry.rs:76-116,
    // src/optimize.rs:256-334); CheckerTexture, nested once, on a rectangle, on a sphere, under a Metal and as an emitter
    // (src/material.rs:217-245); a ConstantMedium bounded by a Cube instead of a Sphere (src/volume.rs:40-43 takes any Hit; the
    // book's smoke boxes); a thin lens.  The returned list is the world itself — no BoundingVolumeHierarchyNode::new around it.
    static List checkerLists(SceneRng& generator) {
        const Vec3 ex(1.0, 0.0, 0.0), ey(0.0, 1.0, 0.0), ez(0.0, 0.0, 1.0);
        auto solidTex = [](double r, double g, double b) { return std::shared_ptr<const typename A::Texture>(std::make_shared<typename A::SolidColor>(Vec3(r, g, b))); };
        auto checker = [](std::shared_ptr<const typename A::Texture> black, std::shared_ptr<const typename A::Texture> white) {
            return std::shared_ptr<const typename A::Texture>(std::make_shared<typename A::CheckerTexture>(std::move(black), std::move(white)));
        };
        auto inner = checker(solidTex(0.9, 0.9, 0.9), solidTex(0.8, 0.2, 0.2));
        auto nested = checker(solidTex(0.2, 0.3, 0.1), inner);  // CheckerTexture whose "white" child is a CheckerTexture
        List world;
        world.push_back(sprite(rect(60.0, 60.0), std::make_shared<typename A::Lambertian>(nested),
                               Mat4::rotation(to_radians(-90.0), ex)));  // floor, normal +y
        auto group = std::make_shared<typename A::HitList>();         // a Vec as an inner node
        group->items.push_back(sprite(sphere(3.0), std::make_shared<typename A::Lambertian>(inner), Mat4::translation(Vec3(-6.0, 3.0, 0.0))));
        group->items.push_back(sprite(sphere(2.0), std::make_shared<typename A::Metal>(nested, 0.1),
                                      Mat4::translation(Vec3(0.0, 2.0, 4.0)).multiplied(Mat4::rotation(to_radians(40.0), ez))));  // rotated: uv turns with it
        group->items.push_back(sprite(sphere(2.5), std::make_shared<typename A::Dielectric>(1.5), Mat4::translation(Vec3(5.0, 2.5, -2.0))));
        world.push_back(group);
        List small;
        for (int k = 0; k < 12; ++k) {
            const double x = generator.gen_range(-12.0, 12.0), z = generator.gen_range(-9.0, 9.0), which = generator.gen_range(0.0, 1.0);
            const double r = generator.gen_range(0.2, 1.0), gcol = generator.gen_range(0.2, 1.0), b = generator.gen_range(0.2, 1.0);
            if (which < 0.5) small.push_back(sprite(sphere(0.7), std::make_shared<typename A::Lambertian>(Vec3(r, gcol, b)), Mat4::translation(Vec3(x, 0.7, z))));
            else if (which < 0.8) small.push_back(sprite(sphere(0.7), std::make_shared<typename A::Metal>(Vec3(r, gcol, b), 0.3), Mat4::translation(Vec3(x, 0.7, z))));
            else small.push_back(sprite(sphere(0.7), std::make_shared<typename A::Dielectric>(1.5), Mat4::translation(Vec3(x, 0.7, z))));
        }
        world.push_back(A::BoundingVolumeHierarchyNode::make(small));  // a BVH inside the list
        auto tent = std::make_shared<typename A::HitList>();           // a Vec as the geometry of a Sprite
        tent->items.push_back(std::make_shared<typename A::TransformedGeometry>(
            rect(4.0, 5.0), Mat4::translation(Vec3(-1.2, 2.0, 0.0)).multiplied(Mat4::rotation(to_radians(-60.0), ey))));
        tent->items.push_back(std::make_shared<typename A::TransformedGeometry>(
            rect(4.0, 5.0), Mat4::translation(Vec3(1.2, 2.0, 0.0)).multiplied(Mat4::rotation(to_radians(60.0), ey))));
        world.push_back(sprite(tent, std::make_shared<typename A::Lambertian>(Vec3(0.3, 0.4, 0.8)),
                               Mat4::translation(Vec3(9.0, 0.5, 6.0)).multiplied(Mat4::rotation(to_radians(25.0), ey))));
        world.push_back(sprite(rect(20.0, 20.0), std::make_shared<typename A::DiffuseLight>(checker(solidTex(6.0, 6.0, 6.0), solidTex(1.0, 2.0, 6.0))),
                               Mat4::translation(Vec3(0.0, 20.0, 0.0)).multiplied(Mat4::rotation(to_radians(90.0), ex))));  // light, normal -y
        world.push_back(sprite(sphere(500.0), std::make_shared<typename A::DiffuseLight>(Vec3(0.15, 0.2, 0.3))));        // sky
        auto faces = A::Cube::make(5.0, 4.0, 5.0);
        world.push_back(sprite(std::make_shared<typename A::ConstantMedium>(A::BoundingVolumeHierarchyNode::make(List(faces.begin(), faces.end())), 0.35),
                               std::make_shared<typename A::Isotropic>(Vec3(0.9, 0.9, 0.95)),
                               Mat4::translation(Vec3(-3.0, 2.01, -7.0)).multiplied(Mat4::rotation(to_radians(30.0), ey))));  // smoke box
        return world;
    }
    static CameraSpec checkerListsCamera() { return {{0.0, 9.0, -26.0}, {0.0, 3.0, 0.0}, {0.0, 1.0, 0.0}, 35.0, 10.0, 0.02}; }

    // Scene by name: "book-one", "cornell-box", "book-two", "stress:<count>", "checker-lists".  Returns false on an unknown name.
    // *rootIsList: the world is the returned Vec itself (impl Hit / Bound for Vec), not a BoundingVolumeHierarchyNode over it.
    static bool byName(const std::string& name, uint64_t seed, List* out, CameraSpec* cam, bool* rootIsList = nullptr) {
        SceneRng g(seed);
        if (rootIsList) *rootIsList = false;
        if (name == "checker-lists") { *out = checkerLists(g); *cam = checkerListsCamera(); if (rootIsList) *rootIsList = true; return true; }
        if (name == "book-one") { *out = bookOne(g); *cam = bookOneCamera(); return true; }
        if (name == "cornell-box") { *out = cornellBox(); *cam = cornellBoxCamera(); return true; }
        if (name == "book-two") { *out = bookTwo(g); *cam = bookTwoCamera(); return true; }
        if (name.rfind("stress:", 0) == 0) { *out = stress(g, (uint32_t)std::stoul(name.substr(7))); *cam = stressCamera(); return true; }
        return false;
    }
};

}  // namespace rtscenes
