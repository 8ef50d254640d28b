// vspg-pbrt: command-line front end over the C ABI, the counterpart of the reference's `pbrt scene.pbrt`
// (src/pbrt/cmd/pbrt.cpp:30-250) for the guidedvolpath render pass: parse -> stage to HBM -> render -> EXR.
// Options keep the reference's spelling where they are meaningful here: --spp, --seed, --outfile, --gpu-device,
// --quiet (cmd/pbrt.cpp:36-90).  Everything else about the run comes from the scene file.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vspg.h"

static void usage() {
    fprintf(stderr,
            "usage: vspg-pbrt [--spp N] [--seed N] [--outfile FILE] [--gpu-device N] [--fp32] [--quiet] scene.pbrt\n"
            "  renders Integrator \"guidedvolpath\" / \"guidedvolpathvspg\" scenes on one B200 and writes the film as EXR\n");
}

int main(int argc, char **argv) {
    std::string scene, outfile;
    int spp = -1, seed = -1, dev = 0;
    bool quiet = false, fp32 = false;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto val = [&](const char *name) -> const char * {
            size_t n = strlen(name);
            if (a.compare(0, n, name) == 0 && a.size() > n && a[n] == '=') return argv[i] + n + 1;
            if (a == name && i + 1 < argc) return argv[++i];
            return nullptr;
        };
        if (const char *v = val("--spp")) spp = atoi(v);
        else if (const char *v = val("--seed")) seed = atoi(v);
        else if (const char *v = val("--outfile")) outfile = v;
        else if (const char *v = val("--gpu-device")) dev = atoi(v);
        else if (a == "--quiet") quiet = true;
        else if (a == "--fp32") fp32 = true;
        else if (a == "--help" || a == "-h") { usage(); return 0; }
        else if (!a.empty() && a[0] == '-') { fprintf(stderr, "vspg-pbrt: unknown option %s\n", a.c_str()); usage(); return 1; }
        else scene = a;
    }
    if (scene.empty()) { usage(); return 1; }
    char err[1024];
    vspg_host_scene *hs = nullptr;
    if (vspg_host_scene_parse(scene.c_str(), &hs, err, sizeof err) != VSPG_OK) { fprintf(stderr, "vspg-pbrt: %s\n", err); return 1; }
    if (spp <= 0) spp = vspg_host_scene_spp(hs);
    if (seed < 0) seed = vspg_host_scene_seed(hs);
    if (outfile.empty()) outfile = vspg_host_scene_outfile(hs);
    vspg_ctx *ctx = nullptr;
    if (vspg_create(&ctx, dev) != VSPG_OK) { fprintf(stderr, "vspg-pbrt: %s\n", vspg_last_error(nullptr)); return 1; }
    auto ck = [&](int rc, const char *what) {
        if (rc != VSPG_OK) { fprintf(stderr, "vspg-pbrt: %s: %s\n", what, vspg_last_error(ctx)); exit(1); }
    };
    const vspg_scene_desc *sd = vspg_host_scene_desc(hs);
    ck(vspg_set_scene(ctx, sd), "set_scene");
    ck(vspg_set_guiding(ctx, vspg_host_scene_guiding(hs)), "set_guiding");
    ck(vspg_set_integrator(ctx, vspg_host_scene_integrator(hs)), "set_integrator");
    int W = sd->camera.width, H = sd->camera.height;
    const std::string trStore = vspg_host_scene_store_tr_file(hs);   // "storeTrBuffer" (vspg.cpp:190-193, 219-221)
    if (!trStore.empty()) ck(vspg_tr_store_begin(ctx), "tr_store_begin");
    auto t0 = std::chrono::steady_clock::now();
    const bool trainIsg = vspg_host_scene_train_isg(hs) != 0;      // no loaded ISG buffer: trained on the fly (vspg.cpp:161-178)
    const std::string isgStore = vspg_host_scene_store_isg_file(hs);   // "storeISGBuffer"
    if (trainIsg) {   // the reference's wave loop: waves of one sample index, Update() after waves 1, 2, 4, 8, ... (vspg.cpp:230-258)
        ck(vspg_isg_store_begin(ctx), "isg_store_begin");
        for (int wave = 1, nextUpdate = 1; wave <= spp; wave++) {
            ck(vspg_render(ctx, wave - 1, wave, 1, 0, 0, W, H, seed, nullptr), "render");
            if (wave == nextUpdate) { ck(vspg_isg_store_update(ctx), "isg_store_update"); nextUpdate *= 2; }
        }
    } else ck(vspg_render(ctx, 0, spp, 1, 0, 0, W, H, seed, nullptr), "render");
    std::vector<float> rgb((size_t)W * H * 3);
    ck(vspg_film_resolve(ctx, rgb.data()), "film_resolve");
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (vspg_exr_write_rgb(outfile.c_str(), rgb.data(), W, H, fp32 ? 0 : 1) != VSPG_OK) { fprintf(stderr, "vspg-pbrt: cannot write %s\n", outfile.c_str()); return 1; }
    if (!trStore.empty()) {   // TrBuffer::Store (cpu/trbuffer.h:52-73): float EXR, channels Transmittance.R/G/B
        std::vector<float> tr((size_t)W * H * 3), planes[3];
        ck(vspg_tr_store_download(ctx, tr.data(), nullptr), "tr_store_download");
        for (int k = 0; k < 3; k++) { planes[k].resize((size_t)W * H); for (size_t i = 0; i < (size_t)W * H; i++) planes[k][i] = tr[3 * i + k]; }
        const char *names[3] = {"Transmittance.R", "Transmittance.G", "Transmittance.B"};
        const float *pp[3] = {planes[0].data(), planes[1].data(), planes[2].data()};
        if (vspg_exr_write_channels(trStore.c_str(), W, H, 3, names, pp) != VSPG_OK) { fprintf(stderr, "vspg-pbrt: cannot write %s\n", trStore.c_str()); return 1; }
        if (!quiet) printf("vspg-pbrt: transmittance buffer -> %s\n", trStore.c_str());
    }
    if (trainIsg && !isgStore.empty()) {   // ImageSpaceGuidingBuffer::Store: the layers this path's own loader and the reference's read
        const size_t n = (size_t)W * H;
        std::vector<float> vsp(n), rec(8 * n), pl[8];
        ck(vspg_isg_store_download(ctx, vsp.data(), rec.data()), "isg_store_download");
        for (int k = 0; k < 8; k++) { pl[k].resize(n); for (size_t i = 0; i < n; i++) pl[k][i] = rec[8 * i + k]; }
        const bool var = vspg_host_scene_integrator(hs)->vsp_criterion_variance != 0;
        const std::string sc_ = var ? "surfContrib2nd" : "surfContrib", vc_ = var ? "volContrib2nd" : "volContrib";
        std::vector<std::string> nm = {"VSP.Y", "surfSpp.Y", "volSpp.Y", sc_ + ".R", sc_ + ".G", sc_ + ".B", vc_ + ".R", vc_ + ".G", vc_ + ".B"};
        const char *names[9]; const float *pp[9] = {vsp.data(), pl[0].data(), pl[1].data(), pl[2].data(), pl[3].data(), pl[4].data(), pl[5].data(), pl[6].data(), pl[7].data()};
        for (int k = 0; k < 9; k++) names[k] = nm[k].c_str();
        if (vspg_exr_write_channels(isgStore.c_str(), W, H, 9, names, pp) != VSPG_OK) { fprintf(stderr, "vspg-pbrt: cannot write %s\n", isgStore.c_str()); return 1; }
        if (!quiet) printf("vspg-pbrt: image-space guiding buffer (un-denoised) -> %s\n", isgStore.c_str());
    }
    if (!quiet) {
        vspg_counters c;
        vspg_counters_get(ctx, &c);
        printf("vspg-pbrt: %dx%d, %d spp, %llu paths in %.3f s (%.2f Mpaths/s), %llu kernel launches -> %s\n", W, H, spp, c.paths, sec,
               c.paths / sec * 1e-6, c.kernel_launches, outfile.c_str());
    }
    vspg_destroy(ctx);
    vspg_host_scene_free(hs);
    return 0;
}
