// ref_driver.cpp -- TEST INFRASTRUCTURE.  A command-line front end for the UNMODIFIED integrators of the
// reference, linked against the patched scalar build that oracle/build_ref_scalar.sh produces
// (oracle/_ref/scalar/libmitsuba.so + plugins).  It exists to pin the oracle at integrator level: it dumps
// what the reference's own code returns for given scenes and seeds; tests/golden/make_scalar_golden.py turns
// those dumps into committed fixtures (tests/golden/scalar_ref_*.npz) that tests/test_oracle_vs_scalar_ref.py
// replays against the oracle, here and on the GPU box.
//
//   ref_driver <variant> <scene.xml> samples <pixels.txt> <spp> <out.bin>
//       pixels.txt: one "x y seed" per line.  For every line: sampler->seed(seed), then spp times
//       SamplingIntegrator::render_sample (integrator.cpp:1139-1240) + sampler->advance() -- the loop body of
//       render_block (integrator.cpp:1110-1123).  out.bin: float32 [line][sample][4] = the aovs render_sample
//       fills: R,G,B,alpha|1 (rgb), I,-,-,- (mono), I,Q,U,V in the sensor frame (mono_polarized).
//   ref_driver <variant> <scene.xml> render1 <seed> <spp> <out.bin> [threads]
//       SamplingIntegrator::render_1 (integrator.cpp:399-743), thread_count = threads (default 1): float32 [16] (Mueller
//       matrix, row-major; Stokes vector = column 0) or [1] (mono).  Prints "RENDER1_SECONDS <s> THREADS <n>" to stderr: the
//       wall time of the render_1 call alone (bench.py --impl reference times the reference's own CPU path with it).
//   ref_driver <variant> <scene.xml> render <seed> <spp> <out.bin>
//       Integrator::render with one thread, developed film: float32 [H][W][C].
//
// render_sample is a protected member of SamplingIntegrator; the driver reaches it by compiling the reference's
// header with `protected` spelled `public` (access control only: no layout or code change).
#define protected public
#include <mitsuba/render/integrator.h>
#undef protected
#include <mitsuba/core/argparser.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/core/filesystem.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/jit.h>
#include <mitsuba/core/logger.h>
#include <mitsuba/core/profiler.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/core/thread.h>
#include <mitsuba/core/util.h>
#include <mitsuba/core/xml.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/sampler.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/sensor.h>
#include <chrono>
#include <cstdio>
#include <fstream>
#include <vector>

using namespace mitsuba;

static void write_floats(const char *path, const std::vector<float> &v) {
    FILE *f = fopen(path, "wb");
    if (!f || fwrite(v.data(), sizeof(float), v.size(), f) != v.size()) { perror(path); exit(3); }
    fclose(f);
}

template <typename Float, typename Spectrum>
void scene_static_accel_initialization() { Scene<Float, Spectrum>::static_accel_initialization(); }
template <typename Float, typename Spectrum>
void scene_static_accel_shutdown() { Scene<Float, Spectrum>::static_accel_shutdown(); }

template <typename Float, typename Spectrum>
void run(Object *scene_, int argc, char **argv) {
    MI_IMPORT_TYPES(Scene, Sensor, Sampler, SamplingIntegrator, ImageBlock, Film)
    auto *scene = dynamic_cast<Scene *>(scene_);
    if (!scene) Throw("Root element of the input file must be a <scene> tag!");
    Sensor *sensor = scene->sensors()[0];
    auto *integrator = dynamic_cast<SamplingIntegrator *>(scene->integrator());
    if (!integrator) Throw("the scene's integrator is not a SamplingIntegrator");
    Film *film = sensor->film();
    const std::string mode = argv[3];

    if (mode == "samples") {
        const uint32_t spp = (uint32_t) atoi(argv[5]);
        sensor->sampler()->set_sample_count(spp);
        size_t n_channels = film->prepare(integrator->aov_names());
        ref<Sampler> sampler = sensor->sampler()->fork();
        ref<ImageBlock> block = film->create_block(ScalarVector2u(0u), false, true);
        block->set_offset(film->crop_offset());
        std::vector<Float> aovs(std::max<size_t>(n_channels, 16), 0.f);
        const ScalarFloat diff_scale_factor = dr::rsqrt((ScalarFloat) spp);
        std::ifstream in(argv[4]);
        std::vector<float> out;
        long x, y;
        unsigned long seed;
        while (in >> x >> y >> seed) {
            sampler->seed((uint32_t) seed);
            Point2f pos_f((Float) (x + film->crop_offset().x()), (Float) (y + film->crop_offset().y()));
            for (uint32_t j = 0; j < spp; ++j) {
                std::fill(aovs.begin(), aovs.end(), 0.f);
                integrator->render_sample(scene, sensor, sampler, block, aovs.data(), pos_f, diff_scale_factor);
                sampler->advance();
                for (int c = 0; c < 4; ++c) out.push_back((float) aovs[c]);
            }
        }
        write_floats(argv[6], out);
    } else if (mode == "render1") {
        if constexpr (is_rgb_v<Spectrum>) {
            Throw("render_1 does not support rgb variants");
        } else {
            const uint32_t seed = (uint32_t) strtoul(argv[4], nullptr, 10), spp = (uint32_t) strtoul(argv[5], nullptr, 10);
            // optional 7th argument: host threads (default 1: the fixtures need one stream); with threads the reference splits the
            // samples into one range per thread (integrator.cpp:510-551).  The wall time of the call goes to stderr for bench.py.
            const size_t threads = argc > 7 ? (size_t) strtoul(argv[7], nullptr, 10) : 1;
            if (threads > 1) Thread::set_thread_count(threads);
            auto t0 = std::chrono::steady_clock::now();
            Spectrum s = integrator->render_1(scene, sensor, seed, spp, false, true, threads);
            fprintf(stderr, "RENDER1_SECONDS %.6f THREADS %zu\n",
                    std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(), threads);
            std::vector<float> out;
            if constexpr (is_polarized_v<Spectrum>) {
                for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out.push_back((float) s(r, c)[0]);
            } else {
                for (size_t k = 0; k < dr::array_size_v<Spectrum>; ++k) out.push_back((float) s[k]);
            }
            write_floats(argv[6], out);
        }
    } else if (mode == "render") {
        const uint32_t seed = (uint32_t) strtoul(argv[4], nullptr, 10), spp = (uint32_t) strtoul(argv[5], nullptr, 10);
        auto img = integrator->render(scene, sensor, seed, spp, true, true);
        std::vector<float> out(img.array().size());
        for (size_t i = 0; i < out.size(); ++i) out[i] = (float) img.array()[i];
        write_floats(argv[6], out);
    } else Throw("unknown mode %s", mode);
}

int main(int argc, char **argv) {
    if (argc < 7) {
        fprintf(stderr, "usage: ref_driver <variant> <scene.xml> samples <pixels.txt> <spp> <out.bin>\n"
                        "       ref_driver <variant> <scene.xml> render1|render <seed> <spp> <out.bin>\n");
        return 2;
    }
    Jit::static_initialization();
    Class::static_initialization();
    Thread::static_initialization();
    Logger::static_initialization();
    Bitmap::static_initialization();
    librender_nop();
    int rc = 0;
    const std::string variant = argv[1];
    try {
        Thread::thread()->logger()->set_log_level(Error);   // (negative Stokes components make ImageBlock::put warn per sample)
        Thread::set_thread_count(1);
        Profiler::static_initialization();
        color_management_static_initialization(false, false);
        MI_INVOKE_VARIANT(variant, scene_static_accel_initialization);
        ref<FileResolver> fr = Thread::thread()->file_resolver();
        fs::path base_path = util::library_path().parent_path();
        if (!fr->contains(base_path)) fr->append(base_path);
        fs::path filename(argv[2]);
        fs::path scene_dir = filename.parent_path();
        if (!scene_dir.empty() && !fr->contains(scene_dir)) fr->append(scene_dir);
        {
            std::vector<ref<Object>> parsed = xml::load_file(filename, variant, xml::ParameterList(), false, false);
            if (parsed.size() != 1) Throw("Root element of the input file is expanded into multiple objects, only a single object is expected");
            MI_INVOKE_VARIANT(variant, run, parsed[0].get(), argc, argv);
        }
        MI_INVOKE_VARIANT(variant, scene_static_accel_shutdown);
    } catch (const std::exception &e) {
        fprintf(stderr, "ref_driver: %s\n", e.what());
        rc = 1;
    }
    color_management_static_shutdown();
    Profiler::static_shutdown();
    Bitmap::static_shutdown();
    Logger::static_shutdown();
    Thread::static_shutdown();
    Class::static_shutdown();
    Jit::static_shutdown();
    return rc;
}
