// vxrt_main — headless stand-in for src/bin/main.rs: builds a Game, creates the renderer through
// render::create_instance, then runs the main loop body (main.rs:41-54: tick, draw_frame) for a fixed number of
// frames, printing the reference's "avg ms / max ms" line and writing the last frame as a PPM.
//
//   vxrt_main --world words.u32 --dims NX NY NZ --noise noise.rgba [--size W H] [--frames N] [--mode 0|1|2]
//             [--camera x y z heading pitch sun] [--hold forward] [--dt 0.016] [--out frame.ppm] [--block block.bin]
//             [--world-dims WX WY WZ]  the file holds a world larger than the window (chunks around it)
//             [--stream]               the window follows the camera (TerrainUploadManager, pipeline.rs:174-189)
//             [--dump-world file]      materials (u32) then occupancy field (u8) of the device window after the last frame
#include <chrono>
#include <cstdlib>
#include <fstream>
#include <iostream>

#include "vxrt_host.hpp"

using namespace raytrace;

static std::vector<char> slurp(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    return std::vector<char>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

int main(int argc, char** argv) {
    try {
        RenderOptions o;
        std::string world_path, noise_path, out_path, block_path, hold, dump_path;
        int frames = 1, wd[3] = {0, 0, 0};
        float dt = 0.016f;
        std::unique_ptr<Game> game(new Game());
        for (int i = 1; i < argc; i++) {
            const std::string a = argv[i];
            auto need = [&](int n) { if (i + n >= argc) throw std::runtime_error("missing value after " + a); };
            if (a == "--world") { need(1); world_path = argv[++i]; }
            else if (a == "--noise") { need(1); noise_path = argv[++i]; }
            else if (a == "--out") { need(1); out_path = argv[++i]; }
            else if (a == "--block") { need(1); block_path = argv[++i]; }
            else if (a == "--hold") { need(1); hold = argv[++i]; }
            else if (a == "--frames") { need(1); frames = std::atoi(argv[++i]); }
            else if (a == "--dt") { need(1); dt = (float)std::atof(argv[++i]); }
            else if (a == "--mode") { need(1); o.mode = std::atoi(argv[++i]); }
            else if (a == "--dims") { need(3); o.nx = std::atoi(argv[++i]); o.ny = std::atoi(argv[++i]); o.nz = std::atoi(argv[++i]); }
            else if (a == "--world-dims") { need(3); wd[0] = std::atoi(argv[++i]); wd[1] = std::atoi(argv[++i]); wd[2] = std::atoi(argv[++i]); }
            else if (a == "--stream") o.streaming = true;
            else if (a == "--dump-world") { need(1); dump_path = argv[++i]; }
            else if (a == "--size") { need(2); o.width = std::atoi(argv[++i]); o.height = std::atoi(argv[++i]); }
            else if (a == "--camera") {
                need(6);
                float v[6]; for (float& x : v) x = (float)std::atof(argv[++i]);
                game.reset(new Game(v[0], v[1], v[2], v[3], v[4], v[5]));
            } else throw std::runtime_error("unknown argument " + a);
        }
        if (world_path.empty() || noise_path.empty()) throw std::runtime_error("--world and --noise are required");
        const std::vector<char> words = slurp(world_path), noise = slurp(noise_path);
        if (!wd[0]) { wd[0] = o.nx; wd[1] = o.ny; wd[2] = o.nz; }
        if (words.size() != (size_t)wd[0] * wd[1] * wd[2] * 4) throw std::runtime_error("world file size does not match --dims / --world-dims");
        if (noise.size() != (size_t)constants::BLUE_NOISE_SIZE) throw std::runtime_error("noise file must be 512*512*4 bytes");
        game->borrow_world_mut().store_window(reinterpret_cast<const uint32_t*>(words.data()), wd[0], wd[1], wd[2]);
        if (!hold.empty()) game->hold(hold, true);

        std::printf("Creating renderer (and world.)\n");
        const auto t0 = std::chrono::steady_clock::now();
        auto pipeline = create_instance(*game, o, reinterpret_cast<const uint8_t*>(noise.data()));
        std::printf("Created in %fs.\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());

        double sum_ms = 0, max_ms = 0;
        for (int f = 0; f < frames; f++) {                               // main.rs:41-54
            const auto tf = std::chrono::steady_clock::now();
            if (f) game->tick(dt);
            pipeline->draw_frame(*game);
            const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tf).count();
            sum_ms += ms; if (ms > max_ms) max_ms = ms;
        }
        const VxrtFrameStats s = pipeline->stats();
        std::printf("%.3fms / %.3fms   (last frame on device: trace %.3f, denoise %.3f, finalize %.3f ms; %llu rays, %llu steps)\n",
                    sum_ms / frames, max_ms, s.trace_ms, s.denoise_ms, s.finalize_ms, (unsigned long long)s.rays, (unsigned long long)s.steps);
        if (!out_path.empty()) {
            std::ofstream f(out_path, std::ios::binary);
            f << "P6\n" << o.width << " " << o.height << "\n255\n";
            const std::vector<uint8_t>& px = pipeline->frame();
            for (size_t p = 0; p < px.size(); p += 4) f.write(reinterpret_cast<const char*>(&px[p]), 3);
        }
        if (!dump_path.empty()) {
            std::vector<uint32_t> m((size_t)o.nx * o.ny * o.nz); std::vector<uint8_t> f(m.size());
            check(pipeline->handle(), vxrt_read_world(pipeline->handle(), m.data(), f.data()));
            std::ofstream df(dump_path, std::ios::binary);
            df.write(reinterpret_cast<const char*>(m.data()), (std::streamsize)(m.size() * 4));
            df.write(reinterpret_cast<const char*>(f.data()), (std::streamsize)f.size());
        }
        if (!block_path.empty()) {
            std::ofstream f(block_path, std::ios::binary);
            f.write(reinterpret_cast<const char*>(&pipeline->uniform_data()), sizeof(VxrtFrameParams));
        }
        return 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "vxrt_main: %s\n", e.what());
        return 1;
    }
}
