// om_main — the role of the reference's src/main.rs:122-214 played by a C++ host over ottomarcher.hpp:
// image and camera setup, scene construction with `world += ...`, the log thread that polls the shared
// sample counter twice a second, one render "thread" per GPU with the round-robin chunk assignment, and — in place
// of the SDL window (no display on this box) — the frame written to a BMP file (main.rs:473-476).
//
//   om_main [--width 1000] [--aspect 1.5] [--spp 200] [--depth 50] [--scene random_scene|basic_scene|<kind 0-5>]
//           [--n N] [--seed 1] [--threads G] [--pass 16] [--fixed] [--flags F] [--out frame.bmp] [--pixels raw.bin] [--quiet]
//
// Build: g++ -std=c++17 -O2 -Iinclude harness/om_main.cpp -Lraytracingoneweekend_b200 -lottomarcher_cuda -pthread
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "ottomarcher.hpp"

using namespace ottomarcher;

// main.rs:102-110
static HittableList basic_scene() {
    HittableList world = HittableList::new_();
    const Material mat_ground = Material::new_lambertian(Color::new_(0.5f, 0.5f, 0.5f));
    world += Sphere::new_with_radius(Point3::new_(0.f, 0.f, -2.f), 1.f, mat_ground);
    world += Sphere::new_with_radius(Point3::new_(-2.f, 0.f, -2.f), 1.f, mat_ground);
    world += Sphere::new_with_radius(Point3::new_(2.f, 0.f, -2.f), 1.f, mat_ground);
    return world;
}
// main.rs:37-100 draws the scene from thread_rng; the library's generator restates it with the seeded Philox stream
static HittableList random_scene(uint64_t seed) { return HittableList::generate(0, seed); }

// main.rs:112-117
static void print_progress(double progress) {
    const double progress100 = std::round(100.0 * progress * 100.0) / 100.0;
    const double frac = std::fmod(progress100, 1.0);
    const double whole = progress100 - frac;
    std::fprintf(stderr, "%3llu.%02llu%%\r", (unsigned long long)whole, (unsigned long long)(frac * 100.0));
}

int main(int argc, char** argv) {
    // IMAGE (main.rs:123-129)
    float aspect_ratio = 3.0f / 2.0f;
    uint32_t image_width = 1000, samples_per_pixel = 200, max_depth = 50, num_threads = 0, pass = 16, flags = OM_RENDER_ADAPTIVE;
    uint64_t seed = 1, n_param = 0;
    std::string scene = "random_scene", out = "", raw = "";
    bool quiet = false;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--width") image_width = (uint32_t)std::atoi(val());
        else if (a == "--aspect") aspect_ratio = (float)std::atof(val());
        else if (a == "--spp") samples_per_pixel = (uint32_t)std::atoi(val());
        else if (a == "--depth") max_depth = (uint32_t)std::atoi(val());
        else if (a == "--scene") scene = val();
        else if (a == "--n") n_param = std::strtoull(val(), nullptr, 10);
        else if (a == "--seed") seed = std::strtoull(val(), nullptr, 10);
        else if (a == "--threads") num_threads = (uint32_t)std::atoi(val());
        else if (a == "--pass") pass = (uint32_t)std::atoi(val());
        else if (a == "--fixed") flags &= ~uint32_t(OM_RENDER_ADAPTIVE);
        else if (a == "--flags") flags |= (uint32_t)std::strtoul(val(), nullptr, 0);
        else if (a == "--out") out = val();
        else if (a == "--pixels") raw = val();
        else if (a == "--quiet") quiet = true;
        else if (a == "--version") { std::printf("%s\n", om_version()); return 0; }
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    const float image_width_f = (float)image_width;
    const float image_height_f = image_width_f / aspect_ratio;
    const uint32_t image_height = (uint32_t)image_height_f;
    const uint32_t image_size = image_width * image_height;
    try {
        // camera (main.rs:131-143)
        const Camera camera = Camera::new_(Point3::new_(13.f, 2.f, 3.f), Point3::new_(0.f, 0.f, 0.f), Vec3::new_(0.f, 1.f, 0.f), 20.f, aspect_ratio, 0.1f, 10.f);
        HittableList world;
        if (scene == "random_scene") world = random_scene(seed);
        else if (scene == "basic_scene") world = basic_scene();
        else world = HittableList::generate((uint32_t)std::atoi(scene.c_str()), seed, n_param);
        std::atomic<uint64_t> samples_atomic{0};
        std::atomic<bool> failed{false};

        // Log thread (main.rs:150-168)
        const uint64_t total_samples = uint64_t(image_size) * samples_per_pixel;
        const auto start = std::chrono::steady_clock::now();
        std::thread log_thread([&]() {
            for (;;) {
                const uint64_t progress = samples_atomic.load(std::memory_order_relaxed);
                if (!quiet) print_progress(double(progress) / double(total_samples));
                if (total_samples == progress || failed.load()) { if (!quiet) print_progress(double(progress) / double(total_samples)); break; }
                std::this_thread::sleep_for(std::chrono::milliseconds(500));
            }
            if (!quiet) std::fprintf(stderr, "\n%lld seconds\n", (long long)std::chrono::duration_cast<std::chrono::seconds>(std::chrono::steady_clock::now() - start).count());
        });

        // one render thread per GPU (main.rs:170: num_cpus - 1 CPU threads)
        if (num_threads == 0) num_threads = (uint32_t)std::max(1, om_device_count());
        std::vector<render_thread::Pixel> pixels(image_size);  // main.rs:192
        const FrozenHittableList frozen = world.freeze(&camera);  // main.rs:199
        if (!quiet) std::fprintf(stderr, "Running %u threads\n", num_threads);
        std::vector<std::thread> handlers;
        std::string first_error;
        const auto t0 = std::chrono::steady_clock::now();
        for (uint32_t i = 0; i < num_threads; i++) {
            handlers.emplace_back([&, i]() {
                const float tmin = 0.001f, tmax = 100.0f;  // main.rs:205-206
                render_thread::Options opt;
                opt.seed = seed; opt.flags = flags; opt.pass_spp = pass;
                try {
                    render_thread::render(camera, frozen, max_depth, tmin, tmax, samples_per_pixel, image_width, image_height, pixels, i, num_threads, samples_atomic, opt);
                } catch (const std::exception& e) {
                    if (!failed.exchange(true)) first_error = e.what();
                }
            });
        }
        for (auto& h : handlers) h.join();  // (the reference never joins: its main thread runs the SDL loop instead)
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (failed.load()) { log_thread.join(); std::fprintf(stderr, "render failed: %s\n", first_error.c_str()); return 1; }
        // fixed-spp renders credit exactly total_samples; adaptive ones too (leftovers are credited at retirement)
        log_thread.join();
        uint64_t traced = 0;
        for (const auto& p : pixels) traced += p.n;
        if (!out.empty()) save_bmp(out, pixels, image_width, image_height);
        if (!raw.empty()) {
            FILE* f = std::fopen(raw.c_str(), "wb");
            if (!f || std::fwrite(pixels.data(), sizeof(render_thread::Pixel), pixels.size(), f) != pixels.size()) { std::fprintf(stderr, "cannot write %s\n", raw.c_str()); return 1; }
            std::fclose(f);
        }
        std::printf("{\"width\": %u, \"height\": %u, \"spp\": %u, \"threads\": %u, \"objects\": %zu, \"credited\": %llu, \"traced\": %llu, \"seconds\": %.4f, \"msamples_per_s\": %.2f}\n",
                    image_width, image_height, samples_per_pixel, num_threads, world.len(), (unsigned long long)samples_atomic.load(), (unsigned long long)traced, secs,
                    double(traced) / secs / 1e6);
    } catch (const Error& e) {
        std::fprintf(stderr, "om_main: %s (status %d)\n", e.what(), (int)e.status);
        return 1;
    }
    return 0;
}
