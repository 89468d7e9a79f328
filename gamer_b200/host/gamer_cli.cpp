// Headless command-line renderer, argument-compatible with the reference's `Gamer galaxy …`
// (ConsoleRenderer::RenderSingleGalaxy, source/consolerenderer.cpp:16-101; singleGalaxy.sh):
//   gamer_b200_cli galaxy <method> <cam xyz> <target xyz> <up xyz> <fov> <exposure> <gamma> <saturation>
//                         <ray step> <galaxy.gax> <size> <outfile>          (19 words, writes <outfile>.png)
// <method> is accepted as omp / thread / b200 — all run the CUDA path; there is no CPU renderer here.
// Extra commands for the Qt-free file layer: `gaxdump <file>` and `gaxcopy <in> <out>`.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <thread>

#include "gax.hpp"
#include "fits_write.hpp"
#include "png_write.hpp"
#include "rasterizer_b200.hpp"
#include "renderparams.hpp"

using namespace gamer;

static void PrintUsage() {
    std::cout << "Usage: gamer_b200_cli [ command ] [ parameters ]\n"
                 "Commands are :\n"
                 "   - galaxy : renders a single galaxy with (example) parameters:\n"
                 "       [ method = omp/thread/b200 ] [ camera = 1 0 0  ] [ target = 0 0 0  ] [ up = 0 1 0 ] [ fov = 90 ]"
                 "  [ exposure = 1 ] [ gamma = 1 ] [ saturation = 1 ] [ ray step = 0.025 ] [ galaxy gax file ]"
                 " [ size = 256 ] [ filename = test{.png} ]\n"
                 "   - skybox [ method = omp/thread/b200 ] [ Renderingparams ] [ galaxy file ] [ png size ]\n"
                 "   - gaxdump [ galaxy gax file ]\n"
                 "   - gaxcopy [ galaxy gax file ] [ out file ]\n"
                 "   environment: GAMER_B200_FITS=1 makes `galaxy` also write <filename>_red/_green/_blue.fits (linear planes)\n";
}

// QString::toFloat(): 0 on garbage
static float toFloat(const char* s) { return std::strtof(s, nullptr); }

static std::vector<int32_t> devices_from_env() {
    std::vector<int32_t> ids;
    if (const char* e = std::getenv("GAMER_B200_DEVICES")) {
        std::stringstream ss(e);
        std::string tok;
        while (std::getline(ss, tok, ','))
            if (!tok.empty()) ids.push_back(std::atoi(tok.c_str()));
    }
    if (ids.empty()) ids.push_back(0);
    return ids;
}

static int RenderSingleGalaxy(int n, char** param) {
    if (n != 19) {
        std::cout << n << "\nIncorrect usage/parameters for galaxy. Usage: \n";
        PrintUsage();
        return 1;
    }
    std::string method = param[1];
    for (char& c : method) c = char(std::tolower((unsigned char)c));
    if (method != "omp" && method != "thread" && method != "b200") {
        std::cout << "ERROR! Cannot recognize " << param[1] << "\nMust be 'omp', 'thread' or 'b200'\n";
        return 1;
    }
    RenderingParams rp;
    auto v3 = [&](int i) { return Vec3{toFloat(param[i]), toFloat(param[i + 1]), toFloat(param[i + 2])}; };
    rp.camera().setCamera(v3(2));
    rp.camera().setTarget(v3(5));
    rp.camera().setUp(v3(8));
    rp.camera().setPerspective(toFloat(param[11]));
    rp.setExposure(toFloat(param[12]));
    rp.setGamma(toFloat(param[13]));
    rp.setSaturation(toFloat(param[14]));
    rp.setRayStep(toFloat(param[15]));
    const std::string galaxyFile = param[16];
    rp.setSize(int(toFloat(param[17])));
    const std::string outFile = param[18];
    rp.spectra().PopulateSpectra();

    Galaxy galaxy;
    try {
        if (!galaxy.Load(galaxyFile)) {
            std::cout << "ERROR! Cannot open " << galaxyFile << "\n";
            return 1;
        }
    } catch (const std::exception& e) {
        std::cout << "ERROR! " << galaxyFile << ": " << e.what() << "\n";
        return 1;
    }
    try {
        RasterizerB200 rasterizer(devices_from_env());
        rasterizer.setRenderingParams(&rp);
        GalaxyInstance gi(&galaxy, galaxy.galaxyParams.name, Vec3{0, 0, 0}, Vec3{0, 1, 0}, 1, 0);
        rasterizer.AddGalaxy(&gi);
        std::cout << "Starting rendering on B200 (" << devices_from_env().size() << " device(s))." << std::endl;
        const auto t0 = std::chrono::steady_clock::now();
        rasterizer.Render();
        int prevP = 0;
        while (rasterizer.getState() != RasterizerB200::done) {
            const int curP = int(rasterizer.getPercentDone() * 1000.0);
            if (curP != prevP) {
                std::cout << "\r[ " << curP * 0.1 << "% ]" << std::flush;
                prevP = curP;
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(1));
        }
        std::cout << std::endl;
        if (rasterizer.lastStatus() != GAMER_OK) {
            std::cout << "ERROR! render failed (" << rasterizer.lastStatus() << "): " << rasterizer.lastError() << "\n";
            return 2;
        }
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::cout << "Rendering : " << ms << " ms (kernel " << rasterizer.lastKernelMs() << " ms)" << std::endl;
        if (rasterizer.AssembleImage() != GAMER_OK) {
            std::cout << "ERROR! post-process failed: " << rasterizer.lastError() << "\n";
            return 2;
        }
        if (!save_png_argb32(outFile + ".png", rasterizer.getImageShadowBuffer().data(), rasterizer.width(),
                             rasterizer.height())) {
            std::cout << "ERROR! cannot write " << outFile << ".png\n";
            return 1;
        }
        std::cout << "Image saved to file " << outFile << ".png" << std::endl;
        // the GUI's "Save FITS" (mainwindow.cpp:890-892) for the headless tool: GAMER_B200_FITS=1 also writes the linear
        // planes <outfile>_red.fits / _green.fits / _blue.fits
        const char* fits = std::getenv("GAMER_B200_FITS");
        if (fits && *fits && std::string(fits) != "0") {
            if (!save_fits_rgb(outFile, rasterizer.getRenderBuffer().data(), rasterizer.width(), rasterizer.height())) {
                std::cout << "ERROR! cannot write " << outFile << "_red/_green/_blue.fits\n";
                return 1;
            }
            std::cout << "Linear planes saved to " << outFile << "_{red,green,blue}.fits" << std::endl;
        }
    } catch (const std::exception& e) {
        std::cout << "ERROR! " << e.what() << "\n";
        return 2;
    }
    return 0;
}

// ConsoleRenderer::RenderSkybox (consolerenderer.cpp:103-172) + RenderQueue::RenderSkybox
// (galaxy/renderqueue.cpp:129-173): six fov-90 views from the camera position of the RenderingParams file
// along -z +z +y -y +x -x, saved as SkyboxZ-.png ... SkyboxX+.png (the reference's own file names).
static int RenderSkybox(int n, char** param) {
    if (n != 5) {
        std::cout << n << "\nIncorrect usage/parameters for galaxy. Usage: \n";
        PrintUsage();
        return 1;
    }
    std::string method = param[1];
    for (char& c : method) c = char(std::tolower((unsigned char)c));
    if (method != "omp" && method != "thread" && method != "b200") {
        std::cout << "ERROR! Cannot recognize " << param[1] << "\nMust be 'omp', 'thread' or 'b200'\n";
        return 1;
    }
    RenderingParams rp;
    try {
        if (!LoadRenderingParams(param[2], rp)) {
            std::cout << "ERROR! Cannot open " << param[2] << "\n";
            return 1;
        }
    } catch (const std::exception& e) {
        std::cout << "ERROR! " << param[2] << ": " << e.what() << "\n";
        return 1;
    }
    rp.setSize(int(toFloat(param[4])));
    Galaxy galaxy;
    try {
        if (!galaxy.Load(param[3])) {
            std::cout << "ERROR! Cannot open " << param[3] << "\n";
            return 1;
        }
    } catch (const std::exception& e) {
        std::cout << "ERROR! " << param[3] << ": " << e.what() << "\n";
        return 1;
    }
    const Vec3 planes[6] = {{0, 0, -1}, {0, 0, 1}, {0, 1, 0}, {0, -1, 0}, {1, 0, 0}, {-1, 0, 0}};
    const Vec3 ups[6] = {{0, 1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}, {0, 1, 0}, {0, 1, 0}};
    const char* names[6] = {"Z-", "Z+", "Y-", "Y+", "X-", "X+"};
    try {
        RasterizerB200 rasterizer(devices_from_env());
        GalaxyInstance gi(&galaxy, galaxy.galaxyParams.name, Vec3{0, 0, 0}, Vec3{0, 1, 0}, 1, 0);
        rasterizer.AddGalaxy(&gi);
        const Vec3 cam = rp.camera().m_camera;
        for (int i = 0; i < 6; ++i) {
            RenderingParams face = rp;
            face.camera().setCamera(cam);
            face.camera().setTarget(Vec3{cam.x + planes[i].x, cam.y + planes[i].y, cam.z + planes[i].z});
            face.camera().setUp(ups[i]);
            face.camera().setPerspective(90);
            rasterizer.setRenderingParams(&face);
            rasterizer.RenderStars(uint64_t(i));          // PrepareBuffer renders the star layer (rasterizer.cpp:174)
            rasterizer.Render();
            while (rasterizer.getState() != RasterizerB200::done) std::this_thread::sleep_for(std::chrono::milliseconds(1));
            if (rasterizer.lastStatus() != GAMER_OK || rasterizer.AssembleImage() != GAMER_OK) {
                std::cout << "ERROR! render failed: " << rasterizer.lastError() << "\n";
                return 2;
            }
            const std::string file = std::string("Skybox") + names[i] + ".png";
            if (!save_png_argb32(file, rasterizer.getImageShadowBuffer().data(), rasterizer.width(), rasterizer.height())) {
                std::cout << "ERROR! cannot write " << file << "\n";
                return 1;
            }
            std::cout << "Image saved to file Skybox" << names[i] << ".png" << std::endl;
        }
    } catch (const std::exception& e) {
        std::cout << "ERROR! " << e.what() << "\n";
        return 2;
    }
    return 0;
}

static int GaxDump(const char* path) {
    Galaxy g;
    if (!g.Load(path)) {
        std::cout << "ERROR! Cannot open " << path << "\n";
        return 1;
    }
    const GalaxyParams& p = g.galaxyParams;
    std::printf("displayName=%s\nname=%s\naxis=%.17g %.17g %.17g\nbulgeDust=%.17g\nbulgeAxis=%.17g %.17g %.17g\n",
                g.displayName.c_str(), p.name.c_str(), p.axis[0], p.axis[1], p.axis[2], p.bulgeDust, p.bulgeAxis[0],
                p.bulgeAxis[1], p.bulgeAxis[2]);
    std::printf("winding=%.17g %.17g noArms=%.17g arms=%.17g %.17g %.17g %.17g twirl=%.17g warp=%.17g %.17g\n", p.windingB,
                p.windingN, p.noArms, p.arm1, p.arm2, p.arm3, p.arm4, p.innerTwirl, p.warpAmplitude, p.warpScale);
    std::printf("components=%zu\n", g.componentParams.size());
    for (const ComponentParams& c : g.componentParams)
        std::printf("%s|%.17g|%s|%.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g|%s\n", c.className.c_str(),
                    c.strength, c.spectrum.c_str(), c.arm, c.z0, c.r0, c.active, c.delta, c.winding, c.scale, c.noiseOffset,
                    c.noiseTilt, c.ks, c.inner, c.name.c_str());
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) {
        PrintUsage();
        return 1;
    }
    const std::string cmd = argv[1];
    try {
        if (cmd == "galaxy") return RenderSingleGalaxy(argc - 1, argv + 1);
        if (cmd == "skybox") return RenderSkybox(argc - 1, argv + 1);
        if (cmd == "rpdump" && argc == 3) {
            RenderingParams rp;
            if (!LoadRenderingParams(argv[2], rp)) return 1;
            std::printf("camera=%.9g %.9g %.9g target=%.9g %.9g %.9g up=%.9g %.9g %.9g fov=%.9g\n", rp.m_camera.m_camera.x,
                        rp.m_camera.m_camera.y, rp.m_camera.m_camera.z, rp.m_camera.m_target.x, rp.m_camera.m_target.y,
                        rp.m_camera.m_target.z, rp.m_camera.m_up.x, rp.m_camera.m_up.y, rp.m_camera.m_up.z, rp.m_camera.m_perspective);
            std::printf("size=%d preview=%d exposure=%.9g gamma=%.9g saturation=%.9g noStars=%d starSize=%.17g spread=%.17g "
                        "strength=%.17g rayStep=%.9g spectra=%zu galaxy=%s\n", rp.m_size, rp.m_previewSize, rp.m_exposure, rp.m_gamma,
                        rp.m_saturation, rp.m_noStars, rp.m_starSize, rp.m_starSizeSpread, rp.m_starStrength, rp.m_rayStep,
                        rp.m_spectra.m_spectra.size(), rp.m_currentGalaxy.c_str());
            return 0;
        }
        if (cmd == "gaxdump" && argc == 3) return GaxDump(argv[2]);
        if (cmd == "pngtest" && argc == 5) {  // deterministic pattern through the PNG encoder (CPU test of the writer)
            const int w = std::atoi(argv[3]), h = std::atoi(argv[4]);
            std::vector<uint32_t> img(size_t(w) * h);
            for (int y = 0; y < h; ++y)
                for (int x = 0; x < w; ++x)
                    img[size_t(y) * w + x] = 0xFF000000u | (uint32_t(x * 7 + y) & 0xFF) << 16 | (uint32_t(y * 5) & 0xFF) << 8 |
                                             (uint32_t(x ^ y) & 0xFF);
            return save_png_argb32(argv[2], img.data(), w, h) ? 0 : 1;
        }
        if (cmd == "fitstest" && argc == 5) {  // deterministic pattern through the FITS writer (CPU test of the writer)
            const int w = std::atoi(argv[3]), h = std::atoi(argv[4]);
            std::vector<float> rgb(size_t(w) * h * 3);
            for (int y = 0; y < h; ++y)
                for (int x = 0; x < w; ++x)
                    for (int c = 0; c < 3; ++c)
                        rgb[3 * (size_t(y) * w + x) + c] = float(x) * 0.37f - float(y) * 1.25f + float(c) * 1000.0f + 1e-3f * float(x * y);
            return save_fits_rgb(argv[2], rgb.data(), w, h) ? 0 : 1;
        }
        if (cmd == "gaxcopy" && argc == 4) {
            Galaxy g;
            if (!g.Load(argv[2])) return 1;
            return g.Save(argv[3]) ? 0 : 1;
        }
    } catch (const std::exception& e) {
        std::cout << "ERROR! " << e.what() << "\n";
        return 1;
    }
    PrintUsage();
    return 1;
}
