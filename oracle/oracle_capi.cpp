// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.hpp).  extern "C" surface loaded with ctypes by
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include <atomic>
#include <chrono>
#include <cstring>
#include <string>
#include <thread>

#include "oracle.hpp"
#include "../ray-tracer_b200/examples/scenes.hpp"

namespace {

struct OracleApi {
    using Vec3 = orc::Vec3;
    using Mat4 = orc::Mat4;
    using BoundPtr = orc::BoundPtr;
    using Sphere = orc::Sphere;
    using Rectangle = orc::Rectangle;
    using Cube = orc::Cube;
    using Sprite = orc::Sprite;
    using ConstantMedium = orc::ConstantMedium;
    using BoundingVolumeHierarchyNode = orc::BoundingVolumeHierarchyNode;
    using Lambertian = orc::Lambertian;
    using Metal = orc::Metal;
    using Dielectric = orc::Dielectric;
    using DiffuseLight = orc::DiffuseLight;
    using Isotropic = orc::Isotropic;
    using ImageTexture = orc::ImageTexture;
    static void setBakeSize(ImageTexture&, uint32_t, uint32_t) {}  // the oracle evaluates the closure itself
    using CheckerTexture = orc::CheckerTexture;
    using Texture = orc::Texture;
    using SolidColor = orc::SolidColor;
    using HitList = orc::HitList;
    using TransformedGeometry = orc::TransformedGeometry;
};

}  // namespace

extern "C" {

struct OrcPrimInfo {
    int32_t kind;                // 0 sphere, 1 rectangle, 2 constant medium (sphere boundary), 3 medium (other boundary)
    int32_t material_kind;       // -1 none, 0 lambertian, 1 metal, 2 dielectric, 3 diffuse light, 4 isotropic
    double world_from_local[16]; // composed, column-major like Mat4
    double param[4];             // sphere: r; rectangle: w,h; medium: boundary r, density
};

struct OrcCameraArgs {  // PerspectiveCamera::new arguments (camera.rs:25-33); fov in radians
    double eye[3], center[3], up[3], fov, aspect, focusDistance, lensRadius;
};

struct OrcRenderParams {
    uint32_t width, height, spp, max_depth;
    uint64_t seed;
    uint32_t row_begin, row_stride;
    int32_t sampler_mode;  // 0 faithful (rejection sampling, sequential stream), 1 paired (same variates as CUDA)
    int32_t threads;       // 0 = hardware concurrency
};

struct OrcStats {
    uint64_t paths, segments, box_tests, prim_tests, medium_draws;
    double seconds;
    int32_t threads;
};

struct OrcHit {
    int32_t hit;     // 0/1
    int32_t prim;    // leaf id (DFS order of the scene graph), -1 on miss
    int32_t stable;  // 1 if hit/miss and prim are unchanged under small perturbations of the ray
    int32_t material_kind;
    double t, p[3], n[3], u, v;
    double t_second;  // smallest t among the OTHER leaf primitives the ray hits (+inf if none): near-tie margin
    double t_sens, p_sens, n_sens;  // largest change of t / point / normal under the perturbations (conditioning of the query)
};

struct OrcScene {
    std::vector<orc::BoundPtr> list;
    orc::BoundPtr world;
    rtscenes::CameraSpec cam;
    std::vector<OrcPrimInfo> prims;
};

}  // extern "C"

namespace {

int materialKind(const orc::Material* m) {
    if (!m) return -1;
    if (dynamic_cast<const orc::Lambertian*>(m)) return 0;
    if (dynamic_cast<const orc::Metal*>(m)) return 1;
    if (dynamic_cast<const orc::Dielectric*>(m)) return 2;
    if (dynamic_cast<const orc::DiffuseLight*>(m)) return 3;
    if (dynamic_cast<const orc::Isotropic*>(m)) return 4;
    return -2;
}

// DFS enumeration of leaf primitives (left before right, list order), assigning ids = DFS rank.
void enumerate(const orc::Bound* node, const orc::Mat4& worldFromLocal, const orc::Material* material, bool inSprite,
               std::vector<OrcPrimInfo>& out) {
    if (!node) return;
    if (auto n = dynamic_cast<const orc::BoundingVolumeHierarchyNode*>(node)) {
        enumerate(n->left.get(), worldFromLocal, material, inSprite, out);
        enumerate(n->right.get(), worldFromLocal, material, inSprite, out);
    } else if (auto l = dynamic_cast<const orc::HitList*>(node)) {
        for (auto& v : l->items) enumerate(v.get(), worldFromLocal, material, inSprite, out);
    } else if (auto s = dynamic_cast<const orc::Sprite*>(node)) {
        // The OUTERMOST sprite's material wins, None included: Sprite::hit overwrites the inner
        // record's material unconditionally (sprite.rs:124).
        enumerate(s->geometry_.get(), worldFromLocal.multiplied(s->transform_.origin), inSprite ? material : s->material_.get(), true, out);
    } else if (auto t = dynamic_cast<const orc::TransformedGeometry*>(node)) {
        enumerate(t->geometry.get(), worldFromLocal.multiplied(t->transform.origin), material, inSprite, out);
    } else {
        OrcPrimInfo info{};
        std::memcpy(info.world_from_local, worldFromLocal.a, sizeof(info.world_from_local));
        info.material_kind = materialKind(material);
        if (auto sp = dynamic_cast<const orc::Sphere*>(node)) {
            info.kind = 0; info.param[0] = sp->radius; sp->id = (int)out.size();
        } else if (auto r = dynamic_cast<const orc::Rectangle*>(node)) {
            info.kind = 1; info.param[0] = r->width; info.param[1] = r->height; r->id = (int)out.size();
        } else if (auto m = dynamic_cast<const orc::ConstantMedium*>(node)) {
            if (auto bs = dynamic_cast<const orc::Sphere*>(m->boundary.get())) { info.kind = 2; info.param[0] = bs->radius; }
            else info.kind = 3;
            info.param[1] = m->density; m->id = (int)out.size();
        } else {
            info.kind = -1;
        }
        out.push_back(info);
    }
}

orc::PerspectiveCamera makeCamera(const OrcCameraArgs* c) {
    return orc::PerspectiveCamera({c->eye[0], c->eye[1], c->eye[2]}, {c->center[0], c->center[1], c->center[2]},
                                  {c->up[0], c->up[1], c->up[2]}, c->fov, c->aspect, c->focusDistance, c->lensRadius);
}

void fillHit(const std::optional<orc::HitRecord>& rec, OrcHit* o) {
    std::memset(o, 0, sizeof(*o));
    o->prim = -1;
    o->material_kind = -1;
    if (!rec) return;
    o->hit = 1;
    o->prim = rec->prim;
    o->material_kind = materialKind(rec->material);
    o->t = rec->t;
    o->p[0] = rec->intersection.x; o->p[1] = rec->intersection.y; o->p[2] = rec->intersection.z;
    o->n[0] = rec->normal.x; o->n[1] = rec->normal.y; o->n[2] = rec->normal.z;
    o->u = rec->u; o->v = rec->v;
}

}  // namespace

extern "C" {

OrcScene* orc_scene_create(const char* name, uint64_t seed) {
    auto scene = std::make_unique<OrcScene>();
    orc::BuildRng build(seed ^ 0xA11CE5EEDull);
    orc::tls_build_rng() = &build;
    bool rootIsList = false;
    const bool ok = rtscenes::Scenes<OracleApi>::byName(name, seed, &scene->list, &scene->cam, &rootIsList);
    if (ok && rootIsList) {  // the world is the Vec itself (geometry.rs:76-116)
        auto list = std::make_shared<orc::HitList>();
        list->items = scene->list;
        scene->world = list;
    } else if (ok) scene->world = orc::BoundingVolumeHierarchyNode::make(scene->list);  // book-one.rs:32
    orc::tls_build_rng() = nullptr;
    if (!ok || !scene->world) return nullptr;
    enumerate(scene->world.get(), orc::Mat4::identity(), nullptr, false, scene->prims);
    return scene.release();
}

void orc_scene_destroy(OrcScene* s) { delete s; }

// the image book-two maps onto its earth sphere (examples/main.rs:266-281); null restores the synthetic stand-in
void orc_set_earth_image(const uint8_t* rgb_top_down, uint32_t width, uint32_t height) {
    rtscenes::EarthImage& e = rtscenes::earthOverride();
    if (!rgb_top_down || !width || !height) { e = rtscenes::EarthImage(); return; }
    e.width = width; e.height = height;
    e.rgb = std::make_shared<std::vector<uint8_t>>(rgb_top_down, rgb_top_down + (size_t)width * height * 3);
}

void orc_scene_camera(const OrcScene* s, double aspect, OrcCameraArgs* out) {
    const auto& c = s->cam;
    for (int i = 0; i < 3; ++i) { out->eye[i] = c.eye[i]; out->center[i] = c.center[i]; out->up[i] = c.up[i]; }
    out->fov = rtscenes::to_radians(c.fov_degrees);
    out->aspect = aspect;
    out->focusDistance = c.focusDistance;
    out->lensRadius = c.lensRadius;
}

// Derived camera constants (lowerLeft, horizontal, vertical) — 9 doubles — for camera KATs.
void orc_camera_frame(const OrcCameraArgs* c, double* out9) {
    const orc::PerspectiveCamera cam = makeCamera(c);
    const orc::Vec3 v[3] = {cam.lowerLeft, cam.horizontal, cam.vertical};
    for (int i = 0; i < 3; ++i) { out9[3 * i] = v[i].x; out9[3 * i + 1] = v[i].y; out9[3 * i + 2] = v[i].z; }
}

uint32_t orc_scene_prim_count(const OrcScene* s) { return (uint32_t)s->prims.size(); }
void orc_scene_prims(const OrcScene* s, OrcPrimInfo* out) { std::memcpy(out, s->prims.data(), s->prims.size() * sizeof(OrcPrimInfo)); }

// Thread-per-core, rows interleaved y % threads == i, exactly the reference's driver loop
// (examples/book-one.rs:56-81).  sum / sumsq are [local_rows][width][3] f64 radiance sums over spp.
int orc_render(const OrcScene* s, const OrcCameraArgs* camArgs, const OrcRenderParams* p, double* sum, double* sumsq, OrcStats* stats) {
    if (!s || !camArgs || !p || !sum || p->row_stride == 0) return -1;
    const orc::PerspectiveCamera camera = makeCamera(camArgs);
    int threads = p->threads > 0 ? p->threads : (int)std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    std::vector<uint32_t> rows;
    for (uint32_t y = p->row_begin; y < p->height; y += p->row_stride) rows.push_back(y);
    std::vector<orc::Counters> counters((size_t)threads);
    const auto t0 = std::chrono::steady_clock::now();
    auto worker = [&](int i) {
        orc::PathRng g;
        g.mode = (orc::SamplerMode)p->sampler_mode;
        g.seed_lo = (uint32_t)p->seed; g.seed_hi = (uint32_t)(p->seed >> 32);
        for (size_t k = rows.size(); k-- > 0;) {  // (0..height).rev()
            const uint32_t y = rows[k];
            if ((int)(y % (uint32_t)threads) != i) continue;
            for (uint32_t x = 0; x < p->width; ++x) {
                orc::Vec3 acc, acc2;
                for (uint32_t sIdx = 0; sIdx < p->spp; ++sIdx) {
                    g.key0 = y * p->width + x; g.key1 = sIdx;
                    const orc::Vec3 c = orc::samplePixel(camera, *s->world, x, y, p->width, p->height, p->max_depth, g);
                    acc += c; acc2 += c * c;
                }
                double* o = sum + (k * p->width + x) * 3;
                o[0] = acc.x; o[1] = acc.y; o[2] = acc.z;
                if (sumsq) { double* q = sumsq + (k * p->width + x) * 3; q[0] = acc2.x; q[1] = acc2.y; q[2] = acc2.z; }
            }
        }
        counters[(size_t)i] = g.counters;
    };
    std::vector<std::thread> pool;
    for (int i = 0; i < threads; ++i) pool.emplace_back(worker, i);
    for (auto& t : pool) t.join();
    const auto t1 = std::chrono::steady_clock::now();
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->paths = (uint64_t)rows.size() * p->width * p->spp;
        for (auto& c : counters) {
            stats->segments += c.segments; stats->box_tests += c.box_tests; stats->prim_tests += c.prim_tests; stats->medium_draws += c.medium_draws;
        }
        stats->seconds = std::chrono::duration<double>(t1 - t0).count();
        stats->threads = threads;
    }
    return 0;
}

// Radiance of individual paths (pixel x,y + sample index), PAIRED or FAITHFUL.  out_rgb: n*3, out_segments: n (nullable).
int orc_paths(const OrcScene* s, const OrcCameraArgs* camArgs, const OrcRenderParams* p, const uint32_t* xy, const uint32_t* sample,
              uint64_t n, double* out_rgb, uint32_t* out_segments) {
    if (!s || !camArgs || !p) return -1;
    const orc::PerspectiveCamera camera = makeCamera(camArgs);
    orc::PathRng g;
    g.mode = (orc::SamplerMode)p->sampler_mode;
    g.seed_lo = (uint32_t)p->seed; g.seed_hi = (uint32_t)(p->seed >> 32);
    for (uint64_t i = 0; i < n; ++i) {
        const uint32_t x = xy[2 * i], y = xy[2 * i + 1];
        g.key0 = y * p->width + x; g.key1 = sample[i];
        const uint64_t before = g.counters.segments;
        const orc::Vec3 c = orc::samplePixel(camera, *s->world, x, y, p->width, p->height, p->max_depth, g);
        out_rgb[3 * i] = c.x; out_rgb[3 * i + 1] = c.y; out_rgb[3 * i + 2] = c.z;
        if (out_segments) out_segments[i] = (uint32_t)(g.counters.segments - before);
    }
    return 0;
}

// Closest-hit query on caller-supplied rays (6 doubles each: origin, direction), full record out.
// Medium free-flight draws are keyed by (ray index, primitive id) so they do not depend on traversal
// order.  Margin policy of the fixed-ray-set tests, all decided here in f64:
//   * `t_second`: the nearest hit among the other leaf primitives — a runner-up within `tie_rel` of the
//     winner is a near-tie (coplanar cube faces, a sphere resting on the ground) and is not in the bit-exact set;
//   * the ray is re-traced under 8 perturbations of relative size `rel_perturb` (a few f32 ulps of the
//     inputs); hit/miss or primitive changes make it unstable, and the largest change of t / p / n is
//     reported as the query's conditioning (`*_sens`).
// render::color (src/render.rs:5-29) on caller rays: ray i evaluated spp times, random streams keyed (i, sample) like a pixel of
// an n x 1 image; out_sum: n*3 sums, out_segments (nullable): n segment counts summed over the spp evaluations.
int orc_color_rays(const OrcScene* s, const double* rays, uint64_t n, uint32_t spp, uint32_t max_depth, uint64_t seed, int sampler_mode,
                   double* out_sum, uint64_t* out_segments) {
    if (!s || !rays || !out_sum) return -1;
    orc::PathRng g;
    g.mode = (orc::SamplerMode)sampler_mode;
    g.seed_lo = (uint32_t)seed; g.seed_hi = (uint32_t)(seed >> 32);
    for (uint64_t i = 0; i < n; ++i) {
        const double* r = rays + 6 * i;
        const orc::Ray ray{{r[0], r[1], r[2]}, {r[3], r[4], r[5]}};
        orc::Vec3 acc;
        const uint64_t before = g.counters.segments;
        for (uint32_t k = 0; k < spp; ++k) {
            g.key0 = (uint32_t)i; g.key1 = k;
            g.bounce = 0; g.seq = 0;
            orc::tls_rng() = &g;
            acc += orc::color(ray, *s->world, max_depth);
            orc::tls_rng() = nullptr;
        }
        out_sum[3 * i] = acc.x; out_sum[3 * i + 1] = acc.y; out_sum[3 * i + 2] = acc.z;
        if (out_segments) out_segments[i] = g.counters.segments - before;
    }
    return 0;
}

int orc_trace(const OrcScene* s, const double* rays, uint64_t n, uint64_t seed, double rel_perturb, double tie_rel, OrcHit* out) {
    if (!s || !rays || !out) return -1;
    orc::PathRng g;
    g.mode = orc::PAIRED;
    g.seed_lo = (uint32_t)seed; g.seed_hi = (uint32_t)(seed >> 32);
    std::vector<std::pair<double, int>> log;
    orc::tls_rng() = &g;
    for (uint64_t i = 0; i < n; ++i) {
        const double* r = rays + 6 * i;
        const orc::Ray ray{{r[0], r[1], r[2]}, {r[3], r[4], r[5]}};
        g.key0 = (uint32_t)i; g.key1 = (uint32_t)(i >> 32); g.bounce = 1;
        log.clear();
        g.leaf_log = &log;
        const auto rec = s->world->hit(ray);
        g.leaf_log = nullptr;
        fillHit(rec, &out[i]);
        int stable = 1;
        double second = orc::INF;
        if (rec)
            for (const auto& e : log)
                if (e.second != rec->prim && e.first < second) second = e.first;
        out[i].t_second = second;
        if (rec && second - rec->t <= tie_rel * std::max(std::fabs(rec->t), 1e-3)) stable = 0;
        double ts = 0.0, ps = 0.0, ns = 0.0;
        if (rel_perturb > 0.0) {
            const double so = rel_perturb * std::max(1.0, ray.origin.length());
            const double sd = rel_perturb * ray.direction.length();
            for (uint32_t k = 0; k < 8 && stable; ++k) {
                const orc::Philox4 q = orc::philox4x32_10((uint32_t)i, k, 0x57AB1Eu, 0, (uint32_t)(i >> 32), 7u);
                const orc::Philox4 q2 = orc::philox4x32_10((uint32_t)i, k, 0x57AB1Fu, 0, (uint32_t)(i >> 32), 7u);
                auto sgn = [](uint32_t w) { return 2.0 * orc::u24(w) - 1.0; };
                const orc::Ray pr{ray.origin + orc::Vec3(sgn(q.v[0]), sgn(q.v[1]), sgn(q.v[2])) * so,
                                  ray.direction + orc::Vec3(sgn(q2.v[0]), sgn(q2.v[1]), sgn(q2.v[2])) * sd};
                const auto prec = s->world->hit(pr);
                if ((bool)prec != (bool)rec) stable = 0;
                else if (rec && prec->prim != rec->prim) stable = 0;
                else if (rec) {
                    ts = std::max(ts, std::fabs(prec->t - rec->t));
                    ps = std::max(ps, (prec->intersection - rec->intersection).length());
                    ns = std::max(ns, (prec->normal - rec->normal).length());
                }
            }
        }
        out[i].t_sens = ts; out[i].p_sens = ps; out[i].n_sens = ns;
        out[i].stable = stable;
    }
    orc::tls_rng() = nullptr;
    return 0;
}

// Resolve exactly as the example drivers print pixels (examples/book-one.rs:90-100):
// mean -> sqrt -> *255 -> min(255) -> truncation; rows top to bottom.  NaN -> 255 (f64::min semantics, A.12).
void orc_resolve_u8(const double* sum, uint32_t width, uint32_t height, uint32_t spp, uint8_t* out_rgb_top_down) {
    for (uint32_t row = 0; row < height; ++row) {
        const uint32_t y = height - 1 - row;
        for (uint32_t x = 0; x < width; ++x)
            for (int c = 0; c < 3; ++c) {
                const double mean = sum[((size_t)y * width + x) * 3 + c] / (double)spp;
                double v = std::sqrt(mean) * 255.0;
                v = (v != v) ? 255.0 : std::min(v, 255.0);
                if (v < 0.0) v = 0.0;  // `as usize` saturates
                out_rgb_top_down[((size_t)row * width + x) * 3 + c] = (uint8_t)v;
            }
    }
}

// ---- small entry points for the math / sampling known-answer tests ------------------------------
void orc_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    const orc::Philox4 r = orc::philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    std::memcpy(out, r.v, sizeof(r.v));
}
void orc_mat4_translation(const double v[3], double out[16]) { std::memcpy(out, orc::Mat4::translation({v[0], v[1], v[2]}).a, 128); }
void orc_mat4_rotation(double radians, const double axis[3], double out[16]) {
    std::memcpy(out, orc::Mat4::rotation(radians, {axis[0], axis[1], axis[2]}).a, 128);
}
void orc_mat4_multiplied(const double a[16], const double b[16], double out[16]) {
    orc::Mat4 ma, mb; std::memcpy(ma.a, a, 128); std::memcpy(mb.a, b, 128);
    std::memcpy(out, ma.multiplied(mb).a, 128);
}
int orc_mat4_inversed(const double a[16], double out[16], double* det) {
    orc::Mat4 m; std::memcpy(m.a, a, 128);
    *det = m.determinant();
    auto inv = m.inversed();
    if (!inv) return 0;
    std::memcpy(out, inv->a, 128);
    return 1;
}
void orc_vec4_transformed(const double v[4], const double m[16], double out[4]) {
    orc::Mat4 mm; std::memcpy(mm.a, m, 128);
    const orc::Vec4 r = orc::Vec4{v[0], v[1], v[2], v[3]}.transformed(mm);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
// Cube::new face normals / centres in world space (KAT: outward axes).
void orc_cube_faces(double w, double h, double d, double* out_normals18, double* out_centres18) {
    auto faces = orc::Cube::make(w, h, d);
    for (size_t i = 0; i < 6; ++i) {
        auto tg = std::dynamic_pointer_cast<const orc::TransformedGeometry>(faces[i]);
        const orc::Vec3 n = orc::xyz0(orc::Vec3(0, 0, 1)).transformed(tg->transform.origin).xyz();
        const orc::Vec3 c = orc::xyz1(orc::Vec3(0, 0, 0)).transformed(tg->transform.origin).xyz();
        out_normals18[3 * i] = n.x; out_normals18[3 * i + 1] = n.y; out_normals18[3 * i + 2] = n.z;
        out_centres18[3 * i] = c.x; out_centres18[3 * i + 1] = c.y; out_centres18[3 * i + 2] = c.z;
    }
}
// Single-primitive hit KATs.  kind: 0 sphere(r), 1 rectangle(w,h), 2 aabb(min,max in param[0..5]).
int orc_prim_hit(int kind, const double* param, const double ray6[6], OrcHit* out) {
    const orc::Ray ray{{ray6[0], ray6[1], ray6[2]}, {ray6[3], ray6[4], ray6[5]}};
    if (kind == 0) { fillHit(orc::Sphere(param[0]).hit(ray), out); return out->hit; }
    if (kind == 1) { fillHit(orc::Rectangle(param[0], param[1]).hit(ray), out); return out->hit; }
    if (kind == 2) {
        const orc::AABB b{{param[0], param[1], param[2]}, {param[3], param[4], param[5]}};
        std::memset(out, 0, sizeof(*out));
        out->hit = b.hit(ray) ? 1 : 0;
        return out->hit;
    }
    return -1;
}
// ConstantMedium<Sphere>(r, density)::hit for one ray with the free-flight draw of path key `key` (A.9 KATs).
int orc_medium_hit(double r, double density, const double ray6[6], int sampler_mode, uint64_t seed, uint32_t key, OrcHit* out) {
    orc::PathRng g;
    g.mode = (orc::SamplerMode)sampler_mode;
    g.seed_lo = (uint32_t)seed; g.seed_hi = (uint32_t)(seed >> 32);
    g.key0 = key; g.key1 = 0; g.bounce = 1;
    orc::tls_rng() = &g;
    orc::ConstantMedium m(std::make_shared<orc::Sphere>(r), density);
    m.id = 0;
    const orc::Ray ray{{ray6[0], ray6[1], ray6[2]}, {ray6[3], ray6[4], ray6[5]}};
    fillHit(m.hit(ray), out);
    orc::tls_rng() = nullptr;
    return out->hit;
}
// Scatter one ray off a material (distribution tests).  kind: 0 lambertian, 1 metal(fuzz), 2 dielectric(ior), 4 isotropic.
// Returns 1 if scattered; out6 = scattered origin+direction, att3 = attenuation.
int orc_scatter(int kind, double param, const double albedo[3], const double ray6[6], const double p[3], const double nrm[3],
                int sampler_mode, uint64_t seed, uint32_t key, double out6[6], double att3[3]) {
    orc::PathRng g;
    g.mode = (orc::SamplerMode)sampler_mode;
    g.seed_lo = (uint32_t)seed; g.seed_hi = (uint32_t)(seed >> 32);
    g.key0 = key; g.key1 = 0; g.bounce = 1;
    orc::tls_rng() = &g;
    const orc::Vec3 a(albedo[0], albedo[1], albedo[2]);
    std::unique_ptr<orc::Material> m;
    if (kind == 0) m = std::make_unique<orc::Lambertian>(a);
    else if (kind == 1) m = std::make_unique<orc::Metal>(a, param);
    else if (kind == 2) m = std::make_unique<orc::Dielectric>(param);
    else if (kind == 4) m = std::make_unique<orc::Isotropic>(a);
    else { orc::tls_rng() = nullptr; return -1; }
    orc::HitRecord rec;
    rec.t = 1.0; rec.intersection = {p[0], p[1], p[2]}; rec.normal = {nrm[0], nrm[1], nrm[2]};
    const orc::Ray in{{ray6[0], ray6[1], ray6[2]}, {ray6[3], ray6[4], ray6[5]}};
    auto s = m->scatter(in, rec);
    orc::tls_rng() = nullptr;
    if (!s) return 0;
    out6[0] = s->first.origin.x; out6[1] = s->first.origin.y; out6[2] = s->first.origin.z;
    out6[3] = s->first.direction.x; out6[4] = s->first.direction.y; out6[5] = s->first.direction.z;
    att3[0] = s->second.x; att3[1] = s->second.y; att3[2] = s->second.z;
    return 1;
}
// Camera ray for given screen (u, v) with the lens draw of path key (pixel, sample) — camera KATs.
void orc_camera_ray(const OrcCameraArgs* c, double u, double v, int sampler_mode, uint64_t seed, uint32_t key0, uint32_t key1, double out6[6]) {
    const orc::PerspectiveCamera cam = makeCamera(c);
    orc::PathRng g;
    g.mode = (orc::SamplerMode)sampler_mode;
    g.seed_lo = (uint32_t)seed; g.seed_hi = (uint32_t)(seed >> 32);
    g.key0 = key0; g.key1 = key1;
    orc::tls_rng() = &g;
    const orc::Ray r = cam.ray(u, v);
    orc::tls_rng() = nullptr;
    out6[0] = r.origin.x; out6[1] = r.origin.y; out6[2] = r.origin.z;
    out6[3] = r.direction.x; out6[4] = r.direction.y; out6[5] = r.direction.z;
}

}  // extern "C"
