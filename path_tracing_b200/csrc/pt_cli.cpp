// pt_cli.cpp -- headless driver with the reference CLI's surface (src/main_cli.cpp:42-256), on top of the C ABI.
//
//   reference flags (main_cli.cpp:54-73):  --spp (8)  --spl (8)  --mode pt|bdpt|ppm  --device gpu|cpu  --output  --input
//   additions (defaults reproduce the reference): --width --height (default: the scene's R line, main_cli.cpp:103)
//                --eye-depth --light-depth (default 4/4 = EYE_DEPTH / LIGHT_DEPTH, main_cli.cpp:19-20)  --seed  --gpus
//                --passes N --dump-every K: the viewer's progressive accumulation (src/main.cpp:399-531) headless: N renders
//                of --spp samples each (seed + pass) averaged; every K passes the running average is written next to
//                --output as <stem>_passNNNN<ext>, and the viewer's frame-to-frame RMS curve goes to <stem>_rms.txt
// --mode pt / ppm run the library's versions of run_cuda_pt / run_cuda_ppm (B2RenderParams.integrator; for pt the depth
// is --eye-depth, for ppm --spp is ignored as in the reference and --gpus must be 1).  --device cpu exits with a message
// (the CPU integrator is the reference's own cpu_bdpt.cpp; this library has no CPU execution path).  The default mode is
// "pt", the reference's own default (main_cli.cpp:47).
// The camera is Camera{eye, look_at, view_up, fov} as run_cpu_bdpt uses it (cpu_bdpt.cpp:190-200), not the CLI's
// hard-coded F = 50 (main_cli.cpp:158-160).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "b2pt.h"

static void usage() {
    printf("Usage: pt_cli [options]\n"
           "Options:\n"
           "  --spp <int>          Samples per pixel (default: 8)\n"
           "  --spl <int>          Samples per light (default: 8)\n"
           "  --mode <string>      Render mode: pt, bdpt, ppm (default: pt)\n"
           "  --device <string>    Compute device: gpu (cpu: not built) (default: gpu)\n"
           "  --output <string>    Output image path (.png, .ppm, .pfm)\n"
           "  --input <string>     Input scene file\n"
           "  --width/--height <int>            Film size (default: the scene's R line)\n"
           "  --eye-depth/--light-depth <int>   Subpath depths (default: 4)\n"
           "  --seed <int>         Philox key (default: 1)\n"
           "  --gpus <int>         GPUs of this node to shard the samples over (default: 1)\n"
           "  --strategies <string> reference (cpu_bdpt.cpp's set, default) | full (next-event + t=1 splats) | light | nee |\n"
           "                        mis (one light subpath per sample, every s/t strategy, balance heuristic)\n"
           "  --passes <int>       Progressive passes of --spp samples each, averaged (default: 1)\n"
           "  --dump-every <int>   Write the running average every K passes (default: 0 = only the final image)\n");
}

struct DumpCtx { std::string stem, ext; int every; };
static void dump_pass(int32_t pass, const float *film, int32_t w, int32_t h, void *user) {
    const DumpCtx *d = (const DumpCtx *) user;
    if (d->every <= 0 || (pass + 1) % d->every != 0) return;
    char name[32]; snprintf(name, sizeof(name), "_pass%04d", pass + 1);
    const std::string path = d->stem + name + d->ext;
    if (b2_write_image(path.c_str(), film, w, h) == B2_OK) printf("[Save] pass %d -> %s\n", pass + 1, path.c_str());
}

int main(int argc, char **argv) {
    int passes = 1, dump_every = 0, strategies = B2_STRATEGIES_REFERENCE;
    int spp = 8, spl = 8, width = 0, height = 0, eye_depth = 4, light_depth = 4, gpus = 1;
    unsigned long long seed = 1;
    std::string mode = "pt", output_file = "output.png", input_file = "../../input.txt", device = "gpu";
    for (int i = 1; i < argc; ++i) {
        const std::string arg = argv[i];
        auto next = [&]() -> const char * { return i + 1 < argc ? argv[++i] : nullptr; };
        if (arg == "--spp") { if (const char *v = next()) spp = atoi(v); }
        else if (arg == "--spl") { if (const char *v = next()) spl = atoi(v); }
        else if (arg == "--mode") { if (const char *v = next()) mode = v; }
        else if (arg == "--device") { if (const char *v = next()) device = v; }
        else if (arg == "--output") { if (const char *v = next()) output_file = v; }
        else if (arg == "--input") { if (const char *v = next()) input_file = v; }
        else if (arg == "--width") { if (const char *v = next()) width = atoi(v); }
        else if (arg == "--height") { if (const char *v = next()) height = atoi(v); }
        else if (arg == "--eye-depth") { if (const char *v = next()) eye_depth = atoi(v); }
        else if (arg == "--light-depth") { if (const char *v = next()) light_depth = atoi(v); }
        else if (arg == "--seed") { if (const char *v = next()) seed = strtoull(v, nullptr, 10); }
        else if (arg == "--gpus") { if (const char *v = next()) gpus = atoi(v); }
        else if (arg == "--strategies") {
            if (const char *v = next()) {
                const std::string sv = v;
                if (sv == "reference") strategies = B2_STRATEGIES_REFERENCE; else if (sv == "full") strategies = B2_STRATEGIES_FULL;
                else if (sv == "light") strategies = B2_STRATEGIES_LIGHT_TRACING; else if (sv == "nee") strategies = B2_STRATEGIES_NEE;
                else if (sv == "mis") strategies = B2_STRATEGIES_MIS;
                else { fprintf(stderr, "[Error] unknown --strategies '%s'\n", v); return 2; }
            }
        }
        else if (arg == "--passes") { if (const char *v = next()) passes = atoi(v); }
        else if (arg == "--dump-every") { if (const char *v = next()) dump_every = atoi(v); }
        else if (arg == "--help" || arg == "-h") { usage(); return 0; }
    }
    printf("====================================\n");
    printf(" Device : %s\n Mode   : %s\n SPP    : %d\n SPL    : %d (used in BDPT/PPM)\n Input  : %s\n Output : %s\n", device.c_str(), mode.c_str(), spp, spl,
           input_file.c_str(), output_file.c_str());
    printf("====================================\n");
    int integrator = B2_INTEGRATOR_BDPT;
    if (mode == "pt") integrator = B2_INTEGRATOR_PT; else if (mode == "ppm") integrator = B2_INTEGRATOR_PPM;
    else if (mode != "bdpt") { fprintf(stderr, "[Error] unknown mode '%s' (pt, bdpt, ppm)\n", mode.c_str()); return 2; }
    if (device != "gpu") { fprintf(stderr, "[Error] device '%s': this library has no CPU path (the CPU integrator is the reference's cpu_bdpt.cpp)\n", device.c_str()); return 2; }

    B2ParsedScene *ps = nullptr;
    if (b2_parse_scene_file_ex(input_file.c_str(), B2_PARSE_OBJ_DIRECTIVE, &ps) != B2_OK) { fprintf(stderr, "[Error] Cannot open input file: %s\n", input_file.c_str()); return -1; }
    if (width <= 0) width = ps->width;
    if (height <= 0) height = ps->height;
    printf("Eye Position: %g %g %g\nResolution: %d x %d\nTriangle: %zu\nBall: %zu\nLight: %zu\n", ps->camera.eye[0], ps->camera.eye[1], ps->camera.eye[2],
           width, height, ps->n_tris, ps->n_spheres, ps->n_lights);
    if (width <= 0 || height <= 0) { fprintf(stderr, "[Error] no resolution: the scene has no R line and no --width/--height was given\n"); b2_parsed_scene_free(ps); return -1; }

    int ndev = 0;
    b2_device_count(&ndev);
    if (ndev < 1) { fprintf(stderr, "[Error] no CUDA device\n"); b2_parsed_scene_free(ps); return 3; }
    if (gpus < 1) gpus = 1;
    if (gpus > ndev) { fprintf(stderr, "[Warn] --gpus %d but %d device(s) present; using %d\n", gpus, ndev, ndev); gpus = ndev; }

    printf("[Init] Transferring Data to CUDA (%d GPU%s) + BVH build...\n", gpus, gpus > 1 ? "s" : "");
    std::vector<B2Scene *> scenes(gpus, nullptr);
    int rc = B2_OK;
    for (int g = 0; g < gpus && rc == B2_OK; ++g) {
        rc = b2_scene_create(ps->tris, ps->n_tris, ps->spheres, ps->n_spheres, ps->materials, ps->n_materials, ps->lights, ps->n_lights, g, &scenes[g]);
        if (rc == B2_OK) rc = b2_scene_build_bvh(scenes[g], nullptr);
    }
    if (rc == B2_OK && gpus > 1) {                       // NCCL communicators before the clock of the render starts
        std::vector<int> devs(gpus);
        for (int g = 0; g < gpus; ++g) devs[g] = g;
        rc = b2_multi_init(devs.data(), gpus);
    }
    std::vector<float> film((size_t) 3 * width * height, 0.0f);
    B2Stats st; memset(&st, 0, sizeof(st));
    if (rc == B2_OK) {
        printf("[Render] Starting Render...\n");
        const auto t0 = std::chrono::steady_clock::now();
        B2RenderParams p; memset(&p, 0, sizeof(p));
        p.width = width; p.height = height; p.spp = spp; p.spl = spl; p.eye_depth = eye_depth; p.light_depth = light_depth;
        p.seed = seed; p.sample_offset = 0; p.sample_stride = 1; p.film_mode = B2_FILM_MEAN; p.strategies = strategies; p.integrator = integrator;
        if (passes <= 1) rc = b2_render_bdpt_multi(scenes.data(), gpus, &ps->camera, &p, film.data(), &st);
        else {
            const size_t dot = output_file.find_last_of('.');
            DumpCtx dc{ dot == std::string::npos ? output_file : output_file.substr(0, dot), dot == std::string::npos ? std::string(".png") : output_file.substr(dot), dump_every };
            std::vector<float> rms(passes, 0.0f);
            if (gpus == 1) rc = b2_render_bdpt_progressive(scenes[0], &ps->camera, &p, passes, film.data(), rms.data(), dump_pass, &dc, &st);
            else {                                           // several GPUs: one sharded render per pass, accumulated here as the viewer does
                std::vector<float> sum(film.size(), 0.0f), one(film.size());
                std::vector<uint8_t> cur8(film.size()), last8(film.size());
                for (int k = 0; k < passes && rc == B2_OK; ++k) {
                    B2RenderParams pk = p; pk.seed = seed + (unsigned long long) k;
                    B2Stats sk;
                    rc = b2_render_bdpt_multi(scenes.data(), gpus, &ps->camera, &pk, one.data(), &sk);
                    if (rc != B2_OK) break;
                    st.eye_samples += sk.eye_samples; st.closest_rays += sk.closest_rays; st.shadow_rays += sk.shadow_rays; st.path_samples += sk.path_samples; st.ms_total += sk.ms_total;
                    for (size_t i = 0; i < sum.size(); ++i) { sum[i] += one[i]; film[i] = sum[i] / (float) (k + 1); }
                    b2_film_to_rgb8(film.data(), width, height, cur8.data());
                    if (k > 0) b2_frame_rms(cur8.data(), last8.data(), cur8.size(), &rms[k]);
                    cur8.swap(last8);
                    dump_pass(k, film.data(), width, height, &dc);
                }
            }
            if (rc == B2_OK) {
                const std::string rp = dc.stem + "_rms.txt";
                if (FILE *f = fopen(rp.c_str(), "w")) { for (int k = 0; k < passes; ++k) fprintf(f, "%d %.9g\n", k + 1, rms[k]); fclose(f); printf("[Save] RMS history -> %s\n", rp.c_str()); }
                for (int k = 0; k < passes; ++k) printf("Render Iteration: %d  frame RMS %.6f\n", k + 1, rms[k]);
            }
        }
        const auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
        if (rc == B2_OK) {
            printf("\n[Render] Finished in %lld ms (device %.2f ms): %.2f M eye samples/s, %.2f Mrays/s (%llu closest + %llu shadow), %llu connections\n",
                   (long long) ms, st.ms_total, st.eye_samples / (st.ms_total * 1e3), (st.closest_rays + st.shadow_rays) / (st.ms_total * 1e3),
                   (unsigned long long) st.closest_rays, (unsigned long long) st.shadow_rays, (unsigned long long) st.path_samples);
        }
    }
    if (rc != B2_OK) fprintf(stderr, "[Error] %s\n", b2_last_error());
    else {
        printf("[Save] Writing to %s...\n", output_file.c_str());
        if (b2_write_image(output_file.c_str(), film.data(), width, height) == B2_OK) printf("[Success] Image saved!\n");
        else { fprintf(stderr, "[Error] Failed to save image: %s\n", b2_last_error()); rc = B2_ERR_IO; }
    }
    for (B2Scene *s : scenes) b2_scene_destroy(s);
    b2_parsed_scene_free(ps);
    return rc == B2_OK ? 0 : 1;
}
