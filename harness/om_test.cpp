// om_test — C++ tests of the host layer (include/ottomarcher.hpp), written the way the reference would test its own
// types (it ships no tests; the expected values are derived by hand from its formulas, file:line beside each).
//   om_test          host-only checks (no GPU needed)
//   om_test --gpu    + the checks that trace rays / render on the device
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

#include "ottomarcher.hpp"

using namespace ottomarcher;

static int g_failed = 0, g_run = 0;
#define CHECK(cond) do { g_run++; if (!(cond)) { g_failed++; std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); } } while (0)
template <class F> static om_status status_of(F f) { try { f(); } catch (const Error& e) { return e.status; } return OM_OK; }
static bool near(float a, float b, float tol = 1e-5f) { return std::fabs(a - b) <= tol * std::fmax(1.f, std::fabs(b)); }

static void host_checks() {
    // materials.rs:26-35
    const Material lam = Material::new_lambertian(Color::new_(0.1f, 0.2f, 0.3f)), met = Material::new_metal_fuzz(Color::new_(1, 1, 1), 2.0f), die = Material::new_dielectric(1.5f);
    CHECK(lam.mat_type == OM_MAT_LAMBERTIAN && lam.albedo[1] == 0.2f);
    CHECK(met.mat_type == OM_MAT_METAL && met.fuzz == 2.0f && Material::new_metal(Color::new_(1, 1, 1)).fuzz == 0.f);  // no clamp (materials.rs:29-34)
    CHECK(die.mat_type == OM_MAT_DIELECTRIC && die.ior == 1.5f);
    // traced.rs:26: new_with_radius = translate ^ scale, inverse stored beside it
    const Sphere s = Sphere::new_with_radius(Point3::new_(1, 2, 3), 2.f, lam);
    CHECK(s.m_local_to_world[0] == 2.f && s.m_local_to_world[3] == 1.f && s.m_local_to_world[7] == 2.f && s.m_local_to_world[11] == 3.f);
    CHECK(s.m_world_to_local[0] == 0.5f && s.m_world_to_local[3] == -0.5f);
    const Mat4x4 m = Mat4x4::translate(1, 2, 3) ^ Mat4x4::scale(2, 2, 2);
    CHECK(std::memcmp(m.m, s.m_local_to_world, sizeof m.m) == 0);
    const Mat4x4 mi = m.fast_homogenous_inverse();
    CHECK(std::memcmp(mi.m, s.m_world_to_local, sizeof mi.m) == 0);
    // traced.rs:148-170: u, v unit edge vectors and their lengths
    const Triangle t = Triangle::new3points(Point3::new_(0, 0, 0), Point3::new_(2, 0, 0), Point3::new_(0, 3, 0), lam);
    CHECK(t.u_length == 2.f && t.v_length == 3.f && t.u[0] == 1.f && t.v[1] == 1.f && near(std::fabs(t.uxv[2]), 1.f));
    // camera.rs:38-59: vfov 90, aspect 2, looking down -z from the origin: viewport 4 x 2 at focus distance 1
    const Camera c = Camera::new_(Point3::new_(0, 0, 0), Point3::new_(0, 0, -1), Vec3::new_(0, 1, 0), 90.f, 2.f, 0.f, 1.f);
    CHECK(near(c.viewport_height, 2.f) && near(c.viewport_width, 4.f) && c.lens_radius == 0.f);
    CHECK(near(c.horizontal[0], 4.f) && near(c.vertical[1], 2.f) && near(c.lower_left_corner[0], -2.f) && near(c.lower_left_corner[1], -1.f) && near(c.lower_left_corner[2], -1.f));
    // hits.rs:91-110: += routes every type to its own list
    HittableList w = HittableList::new_();
    w += s; w += Cube::new_with_length(Point3::new_(0, 0, 0), 1.f, met); w += t;
    w += Parallelogram::new3points(Point3::new_(0, 0, 0), Point3::new_(1, 0, 0), Point3::new_(0, 1, 0), lam);
    w += InfinitePlane::new_(Point3::new_(0, -1, 0), Vec3::new_(0, 1, 0), lam);
    w += MarchedSphere::new_(Point3::new_(0, 0, 5), 1.f, die); w += MarchedBox::new_(Point3::new_(3, 0, 5), Vec3::new_(1, 1, 1), lam);
    w += MarchedTorus::new_(Mat4x4::translate(0, 3, 0), Vec3::new_(0.5f, 0.1f, 0.f), die);
    CHECK(w.len() == 8 && w.spheres.size() == 1 && w.cubes.size() == 1 && w.triangles.size() == 1 && w.parallelograms.size() == 1 && w.infinite_planes.size() == 1 &&
          w.marched_spheres.size() == 1 && w.marched_boxes.size() == 1 && w.marched_tori.size() == 1);
    // a user's Marched type as data (hits.rs:96-100): local_sdf = |p| - 1 as a program; malformed programs are refused
    const SdfProgram unit_ball = SdfProgram().x().y().z().len3().c(1.f).sub();
    w += DynMarched::new_(Mat4x4::translate(0, 0, -10), Mat4x4::translate(0, 0, 10), 1.f, unit_ball, lam);
    CHECK(w.len() == 9 && w.marched_objects.size() == 1 && w.marched_objects[0].local_sdf.words.size() == 7);
    {
        HittableList bad;
        bad += DynMarched::new_(Mat4x4::identity(), Mat4x4::identity(), 1.f, SdfProgram().x().add(), lam);  // stack underflow
        CHECK(status_of([&] { bad.freeze(&c); }) == OM_ERR_INVALID);
        bad.clear();
        bad += DynMarched::new_(Mat4x4::identity(), Mat4x4::identity(), 1.f, SdfProgram().load(2), lam);  // register never written
        CHECK(status_of([&] { bad.freeze(&c); }) == OM_ERR_INVALID);
    }
    // native trait objects cannot cross to a device (hits.rs:91-95)
    int dummy = 0;
    CHECK(status_of([&] { w += static_cast<const void*>(&dummy); }) == OM_ERR_UNSUPPORTED);
    // freeze needs no device; the frozen scene knows what it holds
    const FrozenHittableList fz = w.freeze(&c);
    uint64_t n[8];
    check(om_scene_counts(fz.handle(), n));
    CHECK(n[0] == 1 && n[3] == 1 && n[7] == 1);
    uint64_t n_dyn = 0;
    check(om_scene_dyn_marched_count(fz.handle(), &n_dyn));
    CHECK(n_dyn == 1);
    w.clear();
    CHECK(w.len() == 0);
    // the generated random_scene arrives through the same lists (ground + small spheres + 3 big ones; torus, parallelogram, triangle, cube)
    const HittableList rs = HittableList::generate(0, 1);
    CHECK(rs.spheres.size() > 400 && rs.marched_tori.size() == 1 && rs.parallelograms.size() == 1 && rs.triangles.size() == 1 && rs.cubes.size() == 1);
    // argument errors surface as Error(OM_ERR_INVALID), not as a crash
    std::vector<render_thread::Pixel> px(10);
    std::atomic<uint64_t> ctr{0};
    CHECK(status_of([&] { render_thread::render(c, fz, 50, 0.001f, 100.f, 4, 8, 8, px, 0, 1, ctr); }) == OM_ERR_INVALID);
    px.resize(64);
    CHECK(status_of([&] { render_thread::render(c, fz, 50, 0.001f, 100.f, 4, 8, 8, px, 3, 2, ctr); }) == OM_ERR_INVALID);
}

static void gpu_checks() {
    const Material lam = Material::new_lambertian(Color::new_(0.5f, 0.5f, 0.5f));
    HittableList w;
    w += Sphere::new_with_radius(Point3::new_(0, 0, 0), 1.f, lam);
    w += InfinitePlane::new_(Point3::new_(0, -2, 0), Vec3::new_(0, 1, 0), lam);
    const FrozenHittableList fz = w.freeze(nullptr);
    HitRecord rec;
    // traced.rs:38-61: unit sphere from (0,0,-3) along +z: t = 2, point (0,0,-1), outward normal (0,0,-1)
    CHECK(fz.hit(Ray::new_(Point3::new_(0, 0, -3), Vec3::new_(0, 0, 1)), 0.001f, 100.f, &rec));
    CHECK(rec.t == 2.f && rec.point.z() == -1.f && rec.normal.z() == -1.f && rec.obj_id == OM_OBJ_ID(OM_TYPE_SPHERE, 0));
    // origin inside: the far root (traced.rs:52-55)
    CHECK(fz.hit(Ray::new_(Point3::new_(0, 0, 0), Vec3::new_(0, 0, 1)), 0.001f, 100.f, &rec) && rec.t == 1.f);
    // t_max rejection (traced.rs:51): nothing closer than 1.5 along that ray
    CHECK(!fz.hit(Ray::new_(Point3::new_(0, 0, -3), Vec3::new_(0, 0, 1)), 0.001f, 1.5f, &rec));
    // plane below (traced.rs:91-113): from (5,0,0) straight down, t = 2, normal faces the ray
    CHECK(fz.hit(Ray::new_(Point3::new_(5, 0, 0), Vec3::new_(0, -1, 0)), 0.001f, 100.f, &rec));
    CHECK(rec.t == 2.f && rec.normal.y() == 1.f && rec.obj_id == OM_OBJ_ID(OM_TYPE_INFINITE_PLANE, 0));
    // a ray parallel to the plane misses it (|n.d| < 1e-6, traced.rs:93)
    CHECK(!fz.hit(Ray::new_(Point3::new_(5, 0, 0), Vec3::new_(1, 0, 0)), 0.001f, 100.f, &rec));

    // a dyn marched unit ball at (0,0,10) (marched.rs:14-44 defaults over a program): from (0,0,5) along +z the march lands on
    // the near pole, numeric normal with the sign fix points back at the ray
    {
        HittableList wd;
        wd += DynMarched::new_(Mat4x4::translate(0, 0, -10), Mat4x4::translate(0, 0, 10), 1.f, SdfProgram().x().y().z().len3().c(1.f).sub(), lam);
        const FrozenHittableList fd = wd.freeze(nullptr);
        CHECK(fd.hit(Ray::new_(Point3::new_(0, 0, 5), Vec3::new_(0, 0, 1)), 0.001f, 100.f, &rec));
        CHECK(near(rec.t, 4.f) && near(rec.point.z(), 9.f) && rec.normal.z() == -1.f && rec.obj_id == OM_OBJ_ID(OM_TYPE_DYN_MARCHED, 0));
        CHECK(!fd.hit(Ray::new_(Point3::new_(0, 3, 5), Vec3::new_(0, 0, 1)), 0.001f, 100.f, &rec));
    }

    // render: the union of N threads' chunks equals one thread's image, bit for bit, and the counter ends at width*height*spp
    const Camera cam = Camera::new_(Point3::new_(13, 2, 3), Point3::new_(0, 0, 0), Vec3::new_(0, 1, 0), 20.f, 1.5f, 0.1f, 10.f);
    const HittableList world = HittableList::generate(0, 1);
    const FrozenHittableList frozen = world.freeze(&cam);
    const uint32_t W = 150, H = 100, spp = 24;
    for (uint32_t flags : {0u, (uint32_t)OM_RENDER_ADAPTIVE}) {
        render_thread::Options opt;
        opt.flags = flags; opt.pass_spp = 8; opt.seed = 5;
        std::vector<render_thread::Pixel> one(W * H), many(W * H);
        std::atomic<uint64_t> c1{0}, c3{0};
        render_thread::render(cam, frozen, 50, 0.001f, 100.f, spp, W, H, one, 0, 1, c1, opt);
        std::vector<std::thread> th;
        for (uint32_t i = 0; i < 3; i++) th.emplace_back([&, i] { render_thread::render(cam, frozen, 50, 0.001f, 100.f, spp, W, H, many, i, 3, c3, opt); });
        for (auto& t : th) t.join();
        CHECK(c1.load() == uint64_t(W) * H * spp && c3.load() == c1.load());
        CHECK(std::memcmp(one.data(), many.data(), one.size() * sizeof(one[0])) == 0);
        uint64_t traced = 0;
        for (const auto& p : one) traced += p.n;
        CHECK(flags ? traced < uint64_t(W) * H * spp : traced == uint64_t(W) * H * spp);  // adaptive retires pixels early (render_thread.rs:31-38)
    }
}

int main(int argc, char** argv) {
    const bool gpu = argc > 1 && std::strcmp(argv[1], "--gpu") == 0;
    try {
        host_checks();
        if (gpu) gpu_checks();
        else {
            // no device: render must fail loudly, never fall back to a CPU path
            if (om_device_count() <= 0) {
                const Camera c = Camera::new_(Point3::new_(0, 0, 0), Point3::new_(0, 0, -1), Vec3::new_(0, 1, 0), 90.f, 1.f, 0.f, 1.f);
                HittableList w; w += Sphere::new_with_radius(Point3::new_(0, 0, -2), 1.f, Material::new_dielectric(1.5f));
                const FrozenHittableList fz = w.freeze(&c);
                std::vector<render_thread::Pixel> px(16);
                std::atomic<uint64_t> ctr{0};
                CHECK(status_of([&] { render_thread::render(c, fz, 5, 0.001f, 100.f, 1, 4, 4, px, 0, 1, ctr); }) == OM_ERR_NO_DEVICE);
            }
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "om_test: unexpected exception: %s\n", e.what());
        return 2;
    }
    std::printf("om_test%s: %d checks, %d failed\n", gpu ? " --gpu" : "", g_run, g_failed);
    return g_failed ? 1 : 0;
}
